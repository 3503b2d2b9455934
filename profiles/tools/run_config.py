import importlib, sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("rpg_vision-based_slam_b200"); scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 3
opts = dict(a.split('=') for a in sys.argv[2:])
t = time.time(); sc = scene.config_scene(cfg); print("scene", time.time() - t, flush=True)
g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
t = time.time(); pkg.load_scene(g, sc); print("load", time.time() - t, g.counts(), flush=True)
for k, v in opts.items(): g.setOption(k, float(v))
t = time.time(); c = g.cost(); print("cost", time.time() - t, c, flush=True)
t = time.time(); s = g.optimize(3); print("first solve (3 it)", time.time() - t, flush=True)
print(s["iterations"][:, [0, 1, 4, 5, 6]])
g.setOption("bench_no_termination", 1)
t = time.time(); s = g.optimize(5); dt = time.time() - t
print("5 it", dt, "ms/it", 1e3 * dt / 5, {k: 1e3 * s[k] / 5 for k in ("time_eval_s", "time_build_s", "time_solve_s")}, "launches", s["kernel_launches"], flush=True)
print(s["iterations"][:, [0, 1, 4, 5, 6]])
import torch
print("mem GB", torch.cuda.mem_get_info()[0] / 1e9, torch.cuda.mem_get_info()[1] / 1e9)
q, p = g.getKnots()
print("pos err vs gt (median, max) m:", np.median(np.abs(p - sc["gt_knot_p"])), np.abs(p - sc["gt_knot_p"]).max())
