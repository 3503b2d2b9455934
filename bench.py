#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path on synthetic scenes (datasets are not available offline).

metric : LM iterations/s of the full-batch continuous-time bundle adjustment (BASELINE.json `metric`).  Default workload = BASELINE
         configs[2], the north-star problem (optimize_gv_continuous_time: 1 h, 10k frames, 200k+ landmarks, 5 M observations); it fits one
         GPU.  N > 1: the SAME problem sharded by time segment over N ranks (strong scaling); `--scaling weak` makes the trajectory N x
         longer instead.  One "step" = one trust-region LM iteration over the whole batch (residual + Jacobian evaluation of every term,
         landmark Schur complement, bordered band factorisation, candidate evaluation, accept / reject).
value  : device-resident rate (state and measurements already in HBM), CUDA events on the solver's stream.  The timed region starts from
         the scene's initial (perturbed) state; accepted / rejected step counts are printed; if more than 20 % of the timed steps were
         rejected (the LM has converged inside the region) the region is timed again with every valid candidate accepted, so that each
         timed iteration does the full work of an accepted step (`timed_mode`).
e2e    : the same metric through the C-ABI with HOST buffers: every step re-uploads state + measurements (ctba_invalidate_device), runs
         ctba_solve(1 iteration) and reads the state back.
also   : N = 1 only — the device-resident rate of configs[0] (G+V+I driver shape) and configs[1] (V+I, EuRoC-shaped).
roofline      : HBM roofline of k_reproj (the per-observation residual + Jacobian kernel north_star's 60 % target names).
roofline_band : FP64 tensor-pipe roofline of the band factorisation (the phase that dominates the iteration): algorithmic flops of the
                block band Cholesky, nbc (nbw+1)(nbw+2)/2 block products of 2*32^3 flop, over the measured solve-phase time and the measured
                DMMA peak (profiles/fp64_peaks.json).
--impl reference : the reference's CPU path restated (oracle): Jet evaluation with Ceres' DynamicAutoDiff schedule + sparse exact LM step
         (landmark Schur complement + band Cholesky) on all host threads, on the WHOLE configured scene, one LM iteration per step.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "rpg_vision-based_slam_b200"
WORKLOAD = {1: "configs[0] optimize_continuous_time indoor_forward_3-shaped (G+V+I)", 2: "configs[1] optimize_vi_continuous_time EuRoC MH_01-shaped",
            3: "configs[2] optimize_gv_continuous_time 1h/10k frames/5M obs",
            4: "configs[3] optimize_discrete_time (per-frame SE3 + IMU pre-integration + GP) on the configs[2] scene with a 100 Hz IMU",
            5: "configs[4] knot-spacing / spline-order sweep at 2M observations (GV shape, 1440 s)"}
BASE = {1: (50.0, 20000), 2: (180.0, 40000), 3: (3600.0, 270000), 4: (3600.0, 270000), 5: (1440.0, 108000)}   # duration, landmarks of the configured scene


def clocks_sampler(stop, out, dev):
    q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    while not stop.is_set():
        try:
            o = subprocess.run(["nvidia-smi", "-i", str(dev), f"--query-gpu={q}", "--format=csv,noheader,nounits"], capture_output=True, text=True,
                               timeout=5).stdout
            f = [x.strip() for x in o.strip().split(",")]
            if len(f) >= 6:
                out.append(f)
        except Exception:
            pass
        stop.wait(0.2)


def clocks_summary(samples):
    if not samples:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    sm = [float(s[0]) for s in samples]
    reasons = []
    for i, name in enumerate(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]):
        if any(s[2 + i].lower().startswith("active") for s in samples):
            reasons.append(name)
    return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(samples[0][1]), "reasons": reasons}


SCENE_OVER = {}   # config 5: {"dt_s": ..., "order": ...}
DT_PERTURB = (0.005, 0.1)   # config 4: initial pose error (m, deg) of the discrete-time states


def cpu_reference_leg(scene_mod, cfg, steps, warmup, threads, budget_s):
    """The reference's CPU path restated (oracle/): ONE LM iteration per step on the WHOLE configured scene — Jacobian evaluation with the
    Ceres DynamicAutoDiff schedule (stride-4 Jets, one residual block per frame with all its landmarks: what the reference executes),
    landmark Schur complement + bordered band Cholesky (oracle/sparse_lm.h, an exact sparse step like SPARSE_NORMAL_CHOLESKY, with an
    elimination order that is favourable to the CPU), candidate evaluation, accept / reject.  Steps beyond the time budget are not run:
    the line reports the steps that were."""
    from oracle.binding import OracleSpline, OracleDiscrete
    pkg = importlib.import_module(PKG)
    sc = scene_mod.config_scene(cfg, **SCENE_OVER)
    if cfg == 4:
        o = OracleDiscrete()
        pkg.load_dt_scene(o, scene_mod.dt_scene_from(sc, perturb=DT_PERTURB))
        c = o.counts(); c["accel"] = c["imu"]
    else:
        o = OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
        pkg.load_scene(o, sc)
        c = o.counts()
    t_begin = time.perf_counter()
    times, phases, done_w = [], [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        s = o.optimizeSparse(1, threads, True)
        dt = time.perf_counter() - t0
        tm = s["timings"]
        # a Solve of one iteration evaluates the Jacobian at the start AND after an accepted step; one LM iteration owns one Jacobian
        # evaluation, one build + factorisation and one candidate evaluation
        t_iter = tm["eval_jac_s"] / max(tm["jac_evals"], 1) + tm["eval_res_s"] + tm["build_s"] + tm["factor_s"]
        if i < warmup:
            done_w += 1
        else:
            times.append(t_iter); phases.append(tm)
        elapsed = time.perf_counter() - t_begin
        if i + 1 >= warmup + 1 and elapsed + dt > budget_s:   # at least one timed step
            break
    t_it = float(np.mean(times))
    ph = {k: float(np.mean([p[k] / (max(p["jac_evals"], 1) if k == "eval_jac_s" else 1.0) for p in phases])) for k in ("eval_jac_s", "eval_res_s", "build_s", "factor_s")}
    return dict(sample=f"whole scene ({c['frames']} frames, {c['obs']} observations, {c['accel']} IMU {'factors' if cfg == 4 else 'samples'}, {c['gp']} GP samples): {len(times)} timed LM "
                       f"iteration(s) after {done_w} warm-up; Jacobians by Jet autodiff with Ceres' schedule, sparse exact step (Schur + band Cholesky)",
                t_iter=t_it, lm_per_s=1.0 / t_it, mobs_per_s=c["obs"] / ph["eval_jac_s"] / 1e6, threads=threads, steps_done=len(times), warmup_done=done_w,
                phases_s=ph, half_bandwidth=int(phases[-1]["half_bandwidth"]), n_band=int(phases[-1]["n_band"]))


def reset_state(g, sc):
    """The scene's initial (perturbed) state back into the handle."""
    g.setKnots(sc["knot_q"], sc["knot_p"])
    g.set3DPoints(sc["landmarks"])
    g.setGlobals(sc["scale"], sc["p_bp"], sc["g"], sc["t_s_gp"])
    c = np.zeros(10); c[4] = 1.0
    g.setCorrections(c)
    if sc["imu_t_ns"].size:
        nb = g.numBiasKnots()
        g.setBiases(np.tile(sc["init_acc_bias"], (nb, 1)), np.tile(sc["init_gyro_bias"], (nb, 1)))


def timed_lm(torch, g, sc, stream, steps, warmup, barrier):
    """K LM iterations from the initial state, device resident.  Returns (ms_total, summary, mode)."""
    def run(mode):
        g.setOption("bench_accept_all", 1 if mode == "accept_all" else 0)
        reset_state(g, sc)
        g.optimize(0)                      # uploads the state, evaluates; no iteration
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        s = g.optimize(steps)
        e1.record(stream)
        barrier()
        return e0.elapsed_time(e1), s
    g.optimize(max(warmup, 3))             # warm-up: structure analysis, graph capture, clocks
    ms, s = run("lm")
    mode = "lm"
    if s["num_unsuccessful_steps"] > 0.2 * steps:
        ms, s = run("accept_all")
        mode = "accept_all"
    g.setOption("bench_accept_all", 0)
    return ms, s, mode


def small_config_rate(torch, pkg, scene_mod, cfg, dev, steps=10):
    sc = scene_mod.config_scene(cfg)
    g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=dev)
    pkg.load_scene(g, sc)
    stream = torch.cuda.Stream()
    g.setStream(stream.cuda_stream)
    g.setOption("bench_no_termination", 1)
    ms, s, mode = timed_lm(torch, g, sc, stream, steps, 3, torch.cuda.synchronize)
    n = g.counts()
    out = {"workload": WORKLOAD[cfg], "ms_per_step": ms / steps, "lm_it_per_s": 1e3 * steps / ms, "obs": n["obs"], "timed_mode": mode,
           "accepted": s["num_successful_steps"], "rejected": s["num_unsuccessful_steps"],
           "lm_breakdown_ms": {"eval": 1e3 * s["time_eval_s"] / steps, "assemble": 1e3 * s["time_build_s"] / steps, "solve": 1e3 * s["time_solve_s"] / steps}}
    g.close()
    return out


def bench_dt(a, scene_mod, threads, rank, world, local_rank):
    """configs[3]: the discrete-time estimator (include/ctba_dt.h).  The path does not shard (no multi-GPU decomposition of the discrete-time
    problem is built): N > 1 runs N independent replicas, no collective, value = N x the slowest replica's rate ("replicas only")."""
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module(PKG)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    torch.cuda.set_device(local_rank)
    sc = scene_mod.config_scene(4)
    d = scene_mod.dt_scene_from(sc, perturb=DT_PERTURB)
    g = pkg.DiscreteTimeEstimator(local_rank)
    pkg.load_dt_scene(g, d)
    stream = torch.cuda.Stream()
    g.setStream(stream.cuda_stream)
    g.setOption("bench_no_termination", 1)
    F, L = len(d["cam_ts"]), len(d["landmarks"])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def reset():
        g.initializeStatesKeep(d["poses"], np.zeros((F, 9))); g.initializeLinearVelocity(); g.set3DPoints(d["landmarks"])
        c = np.zeros(9); c[4] = 1.0
        g.setCorrections(c, 0.0)

    def run(mode):
        g.setOption("bench_accept_all", 1 if mode == "accept_all" else 0)
        reset()
        g.optimize(0)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        s = g.optimize(a.steps)
        e1.record(stream)
        barrier()
        return e0.elapsed_time(e1), s

    stop = threading.Event(); samples = []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples, local_rank)); th.start()
    g.optimize(max(a.warmup, 3))
    ms_total, s = run("lm"); mode = "lm"
    if s["num_unsuccessful_steps"] > 0.2 * a.steps:
        ms_total, s = run("accept_all"); mode = "accept_all"
    g.setOption("bench_accept_all", 0)
    stop.set(); th.join()
    assert s["iterations"].shape[0] - 1 == a.steps
    n = g.counts()
    # end to end through the C ABI with host buffers: state and measurements uploaded again, one LM iteration, state read back
    n_imu_samples = sum(len(t) for (_, t, _, _) in d["imu"])
    h2d = 8 * (16 * F + 3 * L + 13) + n["obs"] * (32 + 4) + 8 * n["frames"] + n_imu_samples * 56 + n["gp"] * 36 + 4 * (n["frames"] + n["imu"] + n["gp"])
    d2h = 8 * (16 * F + 3 * L + 13)
    e2e_steps = max(3, min(a.steps, 10))
    reset(); g.invalidateDevice(); g.optimize(1); g.getStates()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        g.invalidateDevice()
        g.optimize(1)
        g.getStates(); g.get3DPoints(); g.getGlobals()
    torch.cuda.synchronize()
    t_e2e = (time.perf_counter() - t0) / e2e_steps
    dev = f"cuda:{local_rank}"
    tms = torch.tensor([ms_total, t_e2e * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms = [float(x) for x in tms.cpu()]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ms_step = ms_total / a.steps
    fp64 = {}
    try:
        fp64 = json.load(open(os.path.join(ROOT, "profiles", "fp64_peaks.json")))
    except Exception:
        pass
    nbc, nbw = int(g.getStat("band_block_columns")), int(g.getStat("band_half_width_blocks"))
    band_flops = float(nbc) * (nbw + 1) * (nbw + 2) / 2.0 * 2.0 * 32 ** 3
    dmma_peak = float(fp64.get("dmma_tflops", 37.1))
    solve_ms = 1e3 * s["time_solve_s"] / a.steps
    line = {"metric": "LM iterations/s", "value": world * 1e3 / ms_step, "unit": "it/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD[4] + (f" x {world} independent replicas (the discrete-time path does not shard)" if world > 1 else ""),
                       "value_is": "LM iterations/s of the configured problem" + (" x N replicas" if world > 1 else ""),
                       "l2": "the working set (band 2 x 0.5 GB, reprojection records 1.4 GB, pair lists 0.4 GB) is larger than L2; nothing is flushed inside the step",
                       **n, "states": F, "landmarks": L, "band_block_columns": nbc, "band_half_width_blocks": nbw, "initial_pose_error_m_deg": list(DT_PERTURB)},
            "timed_mode": mode, "accepted": int(s["num_successful_steps"]), "rejected": int(s["num_unsuccessful_steps"]), "gpu_launches": int(s["kernel_launches"]) * world,
            "clocks": clocks_summary(samples),
            "e2e": {"value": world * 1e3 / e2e_ms, "unit": "it/s", "h2d_bytes_per_step": int(h2d) * world, "d2h_bytes_per_step": int(d2h) * world},
            "lm_breakdown_ms": {"eval": 1e3 * s["time_eval_s"] / a.steps, "assemble": 1e3 * s["time_build_s"] / a.steps, "solve": solve_ms},
            "roofline": {"kernel": "band factorisation + substitutions (ctba_band.cu) on the 15-scalar state nodes: the largest phase of the iteration", "bound": "tensor",
                         "achieved": band_flops / (solve_ms * 1e-3) / 1e12, "peak": dmma_peak, "unit": "TFLOP/s", "frac": band_flops / (solve_ms * 1e-3) / 1e12 / dmma_peak,
                         "traffic": None, "algorithmic_flops": band_flops, "solve_ms": solve_ms,
                         "note": "FP64 DMMA peak from profiles/fp64_peaks.json; at 14 block diagonals the factorisation is a chain of ~4700 dependent 32-column steps and is launch / latency bound"}}
    g.close()
    if world == 1 and not (a.cpu_budget_s is not None and a.cpu_budget_s <= 0):
        r = cpu_reference_leg(scene_mod, 4, 1, 0, threads, a.cpu_budget_s or 100.0)
        line["cpu_baseline"] = {"value": r["lm_per_s"], "unit": "it/s", "cores": threads, "kind": "port", "sample": r["sample"], "mobs_per_s": r["mobs_per_s"], "phases_s": r["phases_s"]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", type=int, default=3)
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="N > 1: strong = the configured problem itself, sharded over N ranks; weak = ONE trajectory of N x the configured duration")
    ap.add_argument("--cpu-budget-s", type=float, default=None, help="time budget of the CPU leg (reference arm: 240 s; cpu_baseline: 100 s)")
    ap.add_argument("--no-also", action="store_true", help="skip the configs[0] / configs[1] rates")
    ap.add_argument("--dt", type=float, default=None, help="config 5: knot spacing in seconds (0.02 .. 0.2)")
    ap.add_argument("--order", type=int, default=None, help="config 5: spline order (4 or 6)")
    ap.add_argument("--opt", action="append", default=[], metavar="NAME=VALUE", help="ctba_set_option on the benchmarked handle (tuning runs, e.g. band_partitions=16)")
    ap.add_argument("--linear-solver", default="cholesky", choices=["cholesky", "pcg"], help="bordered band Cholesky (exact step) or block-Jacobi PCG on the same reduced system")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    scene_mod = importlib.import_module(PKG + ".scene")
    threads = os.cpu_count() or 1
    if a.dt is not None:
        SCENE_OVER["dt_s"] = a.dt
    if a.order is not None:
        SCENE_OVER["order"] = a.order
    if SCENE_OVER:
        WORKLOAD[a.config] += ", " + ", ".join(f"{k} = {v}" for k, v in SCENE_OVER.items())

    if a.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference_leg(scene_mod, a.config, a.steps, a.warmup, threads, a.cpu_budget_s or 240.0)
        line = {"impl": "reference", "metric": "LM iterations/s", "value": r["lm_per_s"], "unit": "it/s", "n_gpus": a.gpus, "steps": r["steps_done"],
                "warmup": r["warmup_done"], "steps_requested": a.steps, "warmup_requested": a.warmup, "ms_per_step": 1e3 * r["t_iter"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": {"workload": WORKLOAD[a.config]},
                "cpu_baseline": {"value": r["lm_per_s"], "unit": "it/s", "cores": threads, "kind": "port", "sample": r["sample"], "phases_s": r["phases_s"],
                                 "half_bandwidth": r["half_bandwidth"], "band_unknowns": r["n_band"]},
                "e2e": {"value": r["lm_per_s"], "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "mobs_per_s": r["mobs_per_s"]}
        print(json.dumps(line))
        return

    if a.config == 4:
        return bench_dt(a, scene_mod, threads, rank, world, local_rank)
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module(PKG)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    torch.cuda.set_device(local_rank)
    # Every rank is handed the complete problem through the same C-ABI calls as a single-GPU caller; the library keeps the rank's units
    # (time-segment sharding, NCCL exchange steps per LM iteration: DESIGN.md §7).
    weak = world > 1 and a.scaling == "weak"
    mult = world if weak else 1
    sc = (scene_mod.config_scene(a.config, duration_s=BASE[a.config][0] * world, n_landmarks=BASE[a.config][1] * world, **SCENE_OVER) if weak
          else scene_mod.config_scene(a.config, **SCENE_OVER))
    g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=local_rank)
    pkg.load_scene(g, sc)
    n_total = g.counts()
    plan1 = g.shardPlan(1, 0)          # structure of the reduced system: block columns, band width
    stream = torch.cuda.Stream()
    g.setStream(stream.cuda_stream)
    g.setOption("bench_no_termination", 1)
    if a.linear_solver == "pcg":
        g.setOption("linear_solver", 1)
    for kv in a.opt:
        k, v = kv.split("=")
        g.setOption(k, float(v))
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device=f"cuda:{local_rank}")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(pkg.BSplineSO3R3.commUniqueId()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        g.initCommNccl(world, rank, bytes(uid.cpu().numpy().tobytes()))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident LM iterations ----
    stop = threading.Event(); samples = []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples, local_rank)); th.start()
    ms_total, s, mode = timed_lm(torch, g, sc, stream, a.steps, a.warmup, barrier)
    n = g.counts()     # sharded: this rank's units
    steps_done = s["iterations"].shape[0] - 1
    assert steps_done == a.steps, (steps_done, a.steps)
    launches = s["kernel_launches"]
    pcg_stats = (g.getStat("pcg_iterations"), g.getStat("pcg_rel_residual")) if a.linear_solver == "pcg" else None

    # ---- residual+Jacobian kernels alone ----
    for _ in range(3):
        g.benchEvaluate(1)
    torch.cuda.synchronize()
    reps = 20
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(stream); g.benchEvaluate(reps); k1.record(stream); torch.cuda.synchronize()
    ms_eval = k0.elapsed_time(k1) / reps
    # the dominant (HBM-bound) evaluation kernel alone, L2 flushed between launches
    g.benchReproj(3, True)
    ms_reproj = g.benchReproj(20, True)
    ms_reproj_warm = g.benchReproj(20, False)
    stop.set(); th.join()

    # ---- end to end through the C-ABI with host buffers ----
    h2d, d2h = g.transferBytes()
    e2e_steps = max(3, min(a.steps, 10))
    reset_state(g, sc)
    g.invalidateDevice(); g.optimize(1); g.getKnots()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        g.invalidateDevice()          # state + (this rank's) measurements are uploaded again from host memory
        g.optimize(1)                 # one LM iteration
        g.getKnots(); g.get3DPoints()  # result back in host memory
    torch.cuda.synchronize()
    t_e2e = (time.perf_counter() - t0) / e2e_steps

    # ---- N > 1: the sharded (NCCL) solve against the SAME problem solved unsharded on rank 0, outside every timed region ----
    parity = None
    if world > 1:
        reset_state(g, sc)
        g.setOption("bench_accept_all", 0)
        sp = g.optimize(3)
        if rank == 0:
            u = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=local_rank)
            pkg.load_scene(u, sc)
            u.setOption("bench_no_termination", 1)
            su = u.optimize(3)
            a_, b_ = sp["iterations"], su["iterations"]
            parity = {"iterations": 3, "cost_rel": float(np.max(np.abs(a_[:, 0] - b_[:, 0]) / np.abs(b_[:, 0]))),
                      "accept_seq_equal": bool((a_[:, 6] == b_[:, 6]).all())}
            (qs, ps), (qu, pu) = g.getKnots(), u.getKnots()
            parity["knot_pos_max_abs_m"] = float(np.abs(ps - pu).max())
            u.close()
        else:
            g.getKnots()

    dev = f"cuda:{local_rank}"
    tms = torch.tensor([ms_total, t_e2e * 1e3, ms_eval, ms_reproj, ms_reproj_warm], dtype=torch.float64, device=dev)
    sums = torch.tensor([n["obs"], h2d, d2h, launches], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    ms_total, e2e_ms, ms_eval, ms_reproj, ms_reproj_warm = [float(x) for x in tms.cpu()]
    obs_all, h2d_all, d2h_all, launches_all = [int(x) for x in sums.cpu()]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ms_step = ms_total / a.steps
    peaks, fp64 = {}, {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    try:
        fp64 = json.load(open(os.path.join(ROOT, "profiles", "fp64_peaks.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # dominant evaluation kernel k_reproj: algorithmic bytes per observation = 48 B in (uv 16, landmark index 4, frame index 4,
    # landmark 24) + 64 B out (r 16, Jl = dr/dp_w 48): the pose / global Jacobian blocks are linear maps of Jl and are never
    # materialised.  Timed alone with CUDA events on its stream, L2 flushed between launches (rank 0's shard when N > 1).
    alg_bytes = 112.0 * n["obs"]
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "k_reproj_traffic.json"))).get(f"config{a.config}")
    except Exception:
        pass
    wl = WORKLOAD[a.config]
    if weak:
        wl += f" x {world}: ONE trajectory of {world} x the duration, sharded by time segment over {world} ranks (NCCL)"
    elif world > 1:
        wl += f", the one problem sharded by time segment over {world} ranks (NCCL)"
    solve_ms = 1e3 * s["time_solve_s"] / a.steps
    nbc, nbw = plan1["nbc"], plan1["nbw"]
    band_flops = float(nbc) * (nbw + 1) * (nbw + 2) / 2.0 * 2.0 * 32 ** 3
    dmma_peak = float(fp64.get("dmma_tflops", 37.1))
    bound = "hbm" if (traffic is None or traffic >= 0.5 * alg_bytes) else "l2"   # measured DRAM traffic far below the algorithmic bytes: the data lives in L2
    line = {"metric": "LM iterations/s", "value": mult * 1e3 / ms_step, "unit": "it/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl, "value_is": "LM iterations/s of the configured problem" + (" x N (weak: N x longer trajectory)" if weak else ""),
                       "l2": "LM step: the working set (band 2 x 1.2 GB + Jacobians 0.3 GB at configs[2]) is larger than L2, nothing is flushed inside the step; "
                             "roofline kernel: L2 flushed with a 256 MiB memset between launches", **n_total, "obs_all_ranks": obs_all,
                       "band_block_columns": nbc, "band_half_width_blocks": nbw},
            "timed_mode": mode, "accepted": int(s["num_successful_steps"]), "rejected": int(s["num_unsuccessful_steps"]),
            "gpu_launches": int(launches_all), "clocks": clocks_summary(samples),
            "e2e": {"value": mult * 1e3 / e2e_ms, "unit": "it/s", "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all},
            "mobs_per_s": obs_all / ms_eval / 1e3, "eval_ms": ms_eval,
            "lm_breakdown_ms": {"eval": 1e3 * s["time_eval_s"] / a.steps, "assemble": 1e3 * s["time_build_s"] / a.steps, "solve": solve_ms},
            "roofline": {"kernel": "k_reproj (per-observation residual + Jacobian)", "bound": bound, "achieved": alg_bytes / (ms_reproj * 1e-3) / 1e9,
                         "peak": peak, "unit": "GB/s", "frac": alg_bytes / (ms_reproj * 1e-3) / 1e9 / peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": ms_reproj, "launch_ms_l2_warm": ms_reproj_warm,
                         "evaluation_pass_gbs": alg_bytes / (ms_eval * 1e-3) / 1e9,
                         "peak_source": "MEASURED_PEAKS.json (burst copy)" if peaks else "fallback"},
            "roofline_band": {"kernel": "band factorisation + substitutions (k_band_panel / k_band_gather / k_band_border / k_gram_* / k_band_backsolve)",
                              "bound": "tensor", "achieved": band_flops / (solve_ms * 1e-3) / 1e12, "peak": dmma_peak, "unit": "TFLOP/s",
                              "frac": band_flops / (solve_ms * 1e-3) / 1e12 / dmma_peak, "traffic": None, "algorithmic_flops": band_flops,
                              "solve_ms": solve_ms, "note": "FP64 DMMA peak measured by profiles/tools/fp64_peaks.cu (profiles/fp64_peaks.json); algorithmic flops = "
                              "sequential block band Cholesky, the time partitions' redundant border work is not counted" + (" (N > 1: rank 0's share of the time, whole-problem flops)" if world > 1 else "")}}
    line["config"]["linear_solver"] = a.linear_solver
    if a.linear_solver == "pcg":
        line["pcg"] = {"iterations_last_solve": pcg_stats[0], "rel_residual_last_solve": pcg_stats[1]}
    if parity is not None:
        line["parity_vs_unsharded"] = parity
        if parity["cost_rel"] > 1e-8 or not parity["accept_seq_equal"]:
            print(json.dumps(line))
            raise SystemExit(f"sharded solve deviates from the unsharded solve: {parity}")
    if world == 1:
        if not a.no_also and a.config == 3:
            del g
            line["also"] = {f"cfg{c}": small_config_rate(torch, pkg, scene_mod, c, local_rank) for c in (1, 2) if c != a.config}
        if a.cpu_budget_s is not None and a.cpu_budget_s <= 0:
            print(json.dumps(line))
            return
        r = cpu_reference_leg(scene_mod, a.config, 1, 0, threads, a.cpu_budget_s or 100.0)
        line["cpu_baseline"] = {"value": r["lm_per_s"], "unit": "it/s", "cores": threads, "kind": "port", "sample": r["sample"], "mobs_per_s": r["mobs_per_s"],
                                "phases_s": r["phases_s"]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
