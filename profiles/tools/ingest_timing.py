"""SURVEY §8(f) rank 3 decided by a number: wall-clock phases of drivers/optimize_gv_continuous_time on a configs[2]-sized dataset
(COLMAP binary export written here) — ingest, computeValidReprojections (host), upload, solve.
usage: python profiles/tools/ingest_timing.py [duration_s n_landmarks]      (default: the whole of configs[2])"""
import importlib, os, re, subprocess, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from scipy.spatial.transform import Rotation as Rot
import test_drivers as T
scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
kw = {}
if len(sys.argv) > 2:
    kw = dict(duration_s=float(sys.argv[1]), n_landmarks=int(sys.argv[2]))
t0 = time.time(); sc = scene.config_scene(3, **kw); print(f"scene generated in {time.time() - t0:.1f} s: {len(sc['frame_t_ns'])} frames, {len(sc['uv'])} observations", flush=True)
# the feature grid of computeValidReprojections asserts that every feature lies inside the image (camera.h:157-168): pixel noise pushes a few
# synthetic features over the border, they are dropped here like a feature detector would never report them
inside = (sc["uv"][:, 0] >= 0) & (sc["uv"][:, 0] < sc["width"]) & (sc["uv"][:, 1] >= 0) & (sc["uv"][:, 1] < sc["height"])
cnt = np.add.reduceat(inside.astype(np.int64), sc["obs_begin"][:-1]) if len(sc["uv"]) else np.zeros(0, np.int64)
sc["obs_begin"] = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
sc["uv"] = np.ascontiguousarray(sc["uv"][inside]); sc["lm_index"] = np.ascontiguousarray(sc["lm_index"][inside])
print(f"{int((~inside).sum())} features outside the image dropped")
d = tempfile.mkdtemp(prefix="ctba_ingest_")
t0 = time.time()
yml, res, _ = T._write_gv_dataset(d, sc, 6, Rot.from_quat([0, 0, 0, 1.0]).as_quat(), np.zeros(3), (False, True, False, True, True))
y = open(yml).read().replace("grid_cell_size: 30", "grid_cell_size: 20").replace("n_max_outer_iters: 3", "n_max_outer_iters: 1").replace("n_max_inner_iters: 4", "n_max_inner_iters: 50")
open(yml, "w").write(y)
sz = sum(os.path.getsize(os.path.join(d, "colmap", f)) for f in os.listdir(os.path.join(d, "colmap")))
print(f"dataset written in {time.time() - t0:.1f} s ({sz / 1e6:.0f} MB of COLMAP files)", flush=True)
exe = os.path.join(ROOT, "drivers", "bin", "optimize_gv_continuous_time")
if not os.path.exists(exe):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "drivers")])
t0 = time.time()
out = subprocess.run([exe, yml, "0", "6"], capture_output=True, text=True)
print(f"driver wall time {time.time() - t0:.2f} s, exit {out.returncode}")
for l in out.stdout.splitlines():
    if l.startswith("[timing]") or l.startswith("solve ") or "valid reprojections" in l or "Loaded:" in l:
        print(l)
if out.returncode:
    print(out.stderr[-2000:])
