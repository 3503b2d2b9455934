"""Solve-phase time of one LM iteration against the number of time partitions of the band factorisation (option band_partitions).
usage: python profiles/tools/sweep_partitions.py <cfg> P1 P2 ...   (0 = automatic)"""
import importlib, sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("rpg_vision-based_slam_b200"); scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
cfg = int(sys.argv[1])
parts = [int(a) for a in sys.argv[2:]] or [0]
sc = scene.config_scene(cfg)
g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
pkg.load_scene(g, sc)
g.optimize(3)
g.setOption("bench_no_termination", 1)
for P in parts:
    g.setOption("band_partitions", P)
    g.optimize(2)
    s = g.optimize(6)
    print(f"cfg {cfg} band_partitions {P:3d}: ms/it {1e3 * s['time_total_s'] / 6:8.3f}  eval {1e3 * s['time_eval_s'] / 6:7.3f} build {1e3 * s['time_build_s'] / 6:7.3f} "
          f"solve {1e3 * s['time_solve_s'] / 6:7.3f}  launches/it {s['kernel_launches'] / 6:.0f}  accepted {s['num_successful_steps']}", flush=True)
