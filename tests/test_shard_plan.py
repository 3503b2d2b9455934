"""Host logic of the multi-GPU path: the time-segment shard planner, on planning-only handles (device -1, no GPU).

Properties checked:
  * every unit (landmark, IMU sample, GP sample, bias block) has exactly one owner, identically computed on every rank;
  * a rank's units touch only block columns between its left separator and the end of its right separator, so partial
    normal equations of neighbouring ranks overlap only inside a separator (what the interface all-reduce relies on);
  * the partition layout tiles the band; ownership ranges tile the columns;
  * with torch.distributed (gloo, world_size 2): ranks agree on the plan and their unit sets partition the problem.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT


def _planner(pkg, scene, **kw):
    sc = scene.make_scene(**kw)
    h = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=-1)
    pkg.load_scene(h, sc)
    return sc, h


def _node_maps(sc, n_bias):
    """knots and bias knots merged by time, ties: knot first (ctba_assemble.cu analyse_structure)."""
    K = sc["knot_p"].shape[0]
    dt, dtb = sc["dt_ns"], int(sc["dt_s_bias"] / 1e-9)
    nk, nb = np.zeros(K, int), np.zeros(max(n_bias, 1), int)
    ik = ib = node = 0
    while ik < K or ib < n_bias:
        if ib >= n_bias or (ik < K and ik * dt <= ib * dtb):
            nk[ik] = node; ik += 1
        else:
            nb[ib] = node; ib += 1
        node += 1
    return nk, nb


@pytest.mark.parametrize("world", [2, 3, 4])
def test_plan_partitions_units_and_columns(pkg, scene, world):
    sc, h = _planner(pkg, scene, order=6, duration_s=25.0 * world, n_landmarks=500 * world, track_len=12, imu_hz=100, gp_hz=10, dt_s_bias=0.5)
    plans = [h.shardPlan(world, r) for r in range(world)]
    p0 = plans[0]
    P, nbc, nbw = p0["P"], p0["nbc"], p0["nbw"]
    assert P % world == 0 and P >= world
    # layout tiles the band: interior, separator, interior, ...
    c = 0
    for q in range(P):
        assert p0["c0"][q] == c and p0["nI"][q] >= max(nbw, 2)
        c += p0["nI"][q] + (nbw if q < P - 1 else 0)
    assert c == nbc
    for r, p in enumerate(plans):
        for k in ("P", "nbc", "nbw", "c0", "nI", "F", "nodes", "hbn"):
            assert p[k] == p0[k]
        for k in ("lm_owner", "imu_owner", "gp_owner", "bias_owner"):
            assert (p[k] == p0[k]).all()                      # every rank computes the same owners
        assert p["q_lo"] == r * (P // world) and p["q_hi"] == (r + 1) * (P // world)
    # ownership ranges tile [0, n_pad)
    assert plans[0]["own_lo"] == 0 and plans[-1]["own_hi"] == 32 * nbc
    for a, b in zip(plans[:-1], plans[1:]):
        assert a["own_hi"] == b["own_lo"]
        assert b["bc_lo"] * 32 == b["own_lo"] and a["bc_hi"] == b["bc_lo"] + nbw      # overlap = exactly one separator
    # a rank's units touch only its own column range
    n_bias = h.numBiasKnots()
    nk, nb = _node_maps(sc, n_bias)
    N, dt = sc["order"], sc["dt_ns"]
    fs = (sc["frame_t_ns"] - sc["start_t_ns"]) // dt
    lm_lo = np.full(sc["landmarks"].shape[0], 1 << 30); lm_hi = np.full(sc["landmarks"].shape[0], -1)
    fi = np.repeat(np.arange(fs.size), np.diff(sc["obs_begin"]))
    np.minimum.at(lm_lo, sc["lm_index"], fs[fi]); np.maximum.at(lm_hi, sc["lm_index"], fs[fi])
    seen = lm_hi >= 0
    for r, p in enumerate(plans):
        mine = seen & (p0["lm_owner"] == r)
        assert mine.any()
        lo = 6 * nk[lm_lo[mine]] // 32
        hi = (6 * nk[lm_hi[mine] + N - 1] + 5) // 32
        assert lo.min() >= p["bc_lo"] and hi.max() < p["bc_hi"]
        assert ((lo >= p["own_lo"] // 32) & (lo < max(p["own_hi"] // 32, 1))).all()
    # IMU samples: knot window + 4 bias knots
    s_imu = (sc["imu_t_ns"] - int(sc["t_offset_cam_imu"] * 1e9) - sc["start_t_ns"]) // dt
    n_imu = p0["imu_owner"].size
    assert n_imu > 0 and set(np.unique(p0["imu_owner"])) == set(range(world))
    assert set(np.unique(p0["gp_owner"])) == set(range(world))
    assert (np.diff(p0["imu_owner"]) >= 0).all() and (np.diff(p0["gp_owner"]) >= 0).all()   # time-ordered units -> monotone owners


def test_plan_refuses_too_many_ranks(pkg, scene):
    sc, h = _planner(pkg, scene, order=6, duration_s=3.0, n_landmarks=60, track_len=8, imu_hz=0, gp_hz=10)
    with pytest.raises(pkg.CtbaError):
        h.shardPlan(8, 0)
    with pytest.raises(pkg.CtbaError):      # planning-only handles have no compute entry points
        h.cost()


_WORKER = r'''
import importlib, os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
pkg = importlib.import_module("rpg_vision-based_slam_b200"); scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
sc = scene.make_scene(order=6, duration_s=30.0, n_landmarks=900, track_len=10, imu_hz=50, gp_hz=10, dt_s_bias=0.5, seed=9)
h = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=-1)
pkg.load_scene(h, sc)
p = h.shardPlan(world, rank)
c = h.counts()
# each rank counts the units it owns; the sums must reproduce the whole problem exactly once
obs_owner = p["lm_owner"][sc["lm_index"]]
mine = torch.tensor([int((obs_owner == rank).sum()), int((p["imu_owner"] == rank).sum()), int((p["gp_owner"] == rank).sum()),
                     int((p["bias_owner"] == rank).sum())], dtype=torch.int64)
dist.all_reduce(mine)
assert mine.tolist() == [c["obs"], c["accel"], c["gp"], c["bias"]], (mine.tolist(), c)
# ranks agree on the layout and the owners
sig = torch.tensor([p["P"], p["nbc"], p["nbw"], int(p["lm_owner"].sum()), int(p["imu_owner"].sum())], dtype=torch.int64)
lo, hi = sig.clone(), sig.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
assert (lo == hi).all()
# the column ranges of neighbouring ranks overlap in exactly one separator
rng = torch.tensor([p["bc_lo"], p["bc_hi"], p["own_lo"], p["own_hi"]], dtype=torch.int64)
allr = [torch.zeros_like(rng) for _ in range(world)]
dist.all_gather(allr, rng)
for a, b in zip(allr[:-1], allr[1:]):
    assert int(a[1]) == int(b[0]) + p["nbw"] and int(a[3]) == int(b[2])
dist.destroy_process_group()
print("OK", rank)
'''


def test_plan_agrees_across_processes_gloo(tmp_path):
    """world_size-2 gloo run of the planner (the N>1 host path, CPU only)."""
    w = tmp_path / "worker.py"
    w.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29571", str(w), ROOT], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert out.stdout.count("OK") == 2
