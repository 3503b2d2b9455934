#!/usr/bin/env python
"""Discrete-time estimator at BASELINE configs[3] size (the configs[2] scene + 100 Hz IMU): phase times of the LM iteration from the
handle's own CUDA-event timers.  Usage: python profiles/tools/dt_timing.py [steps] [scale]   (scale < 1 shortens the trajectory)."""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("rpg_vision-based_slam_b200")
scene = importlib.import_module("rpg_vision-based_slam_b200.scene")

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
t0 = time.time()
sc = scene.config_scene(4, duration_s=3600.0 * scale, n_landmarks=int(270000 * scale))
d = scene.dt_scene_from(sc, perturb=(0.005, 0.1))
print(f"scene in {time.time() - t0:.1f} s: {len(d['cam_ts'])} states, {d['uv'].shape[0]} reprojections, {len(d['imu'])} IMU factors, {d['gp_t'].size} GP", flush=True)
g = pkg.DiscreteTimeEstimator(0)
pkg.load_dt_scene(g, d)
g.setOption("bench_no_termination", 1)
t0 = time.time()
s = g.optimize(3)
print(f"first solve (structure + 3 iterations) {time.time() - t0:.2f} s; costs {s['iterations'][:, 0]}", flush=True)
for mode in (0, 1):
    g.setOption("bench_accept_all", mode)
    sb0 = np.zeros((len(d["cam_ts"]), 9))
    g.initializeStatesKeep(d["poses"], sb0); g.initializeLinearVelocity(); g.set3DPoints(d["landmarks"])
    c = np.zeros(9); c[4] = 1.0
    g.setCorrections(c, 0.0)
    g.optimize(0)
    s = g.optimize(steps)
    n = s["iterations"].shape[0] - 1
    print(json.dumps({"accept_all": mode, "steps": n, "accepted": s["num_successful_steps"], "rejected": s["num_unsuccessful_steps"], "ms_per_step": 1e3 * s["time_total_s"] / n,
                      "eval_ms": 1e3 * s["time_eval_s"] / n, "assemble_ms": 1e3 * s["time_build_s"] / n, "solve_ms": 1e3 * s["time_solve_s"] / n,
                      "launches_per_step": s["kernel_launches"] / n, "nbc": g.getStat("band_block_columns"), "nbw": g.getStat("band_half_width_blocks"),
                      "cost": [float(s["iterations"][0, 0]), float(s["iterations"][-1, 0])]}), flush=True)
g.close()
