#!/usr/bin/env python
"""bench.py — headline benchmark (BASELINE.json metric) on synthetic scenes.  See DESIGN.md §Measurement."""
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


def clocks_sampler(stop, out):
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    while not stop.is_set():
        try:
            o = subprocess.run(["nvidia-smi", "-i", "0", f"--query-gpu={q}", "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
            out.append(o.strip().split(", "))
        except Exception:
            pass
        stop.wait(0.2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the workload (debug only)")
    a = ap.parse_args()
    import torch
    pkg = importlib.import_module(PKG)
    scene = importlib.import_module(PKG + ".scene")
    over = {}
    if a.scale != 1.0:
        base = dict(duration_s=180.0, n_landmarks=40000)
        over = dict(duration_s=base["duration_s"] * a.scale, n_landmarks=int(base["n_landmarks"] * a.scale))
    sc = scene.config_scene(a.config, **over)
    g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    pkg.load_scene(g, sc)
    n = g.counts()
    stream = torch.cuda.Stream()
    g.setStream(stream.cuda_stream)
    g.cost()
    for _ in range(a.warmup):
        g.benchEvaluate(1)
    torch.cuda.synchronize()
    stop = threading.Event(); samples = []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples)); th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        launches = g.benchEvaluate(a.steps)
        e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    stop.set(); th.join()
    sm = [float(s[0]) for s in samples if len(s) >= 2]
    line = {"metric": "residual+Jacobian Mobs/s", "value": n["obs"] / ms / 1e3, "unit": "Mobs/s", "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"configs[{a.config - 1}]", **n}, "gpu_launches": launches,
            "clocks": {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(samples[0][1]) if samples else None}}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
