import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("rpg_vision-based_slam_b200"); scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 2
sc = scene.config_scene(cfg)
g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
pkg.load_scene(g, sc)
for a in sys.argv[2:]:
    k, v = a.split("="); g.setOption(k, float(v))
g.setOption("bench_no_termination", 1)
s = g.optimize(2)
print(s["iterations"][:, [0, 1, 4, 5, 6]])
