"""Timeline of one band factorisation inside its CUDA graph (debug build of the library with -DCTBA_BAND_TRACE).
build:  python profiles/tools/band_trace.py build      (here: writes scratch/libctba_trace.so)
run  :  CTBA_LIB_PATH=scratch/libctba_trace.so python profiles/tools/band_trace.py <cfg> [option=value ...]   (on the GPU box)"""
import ctypes as C, glob, importlib, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
if sys.argv[1] == "build":
    csrc = os.path.join(ROOT, "rpg_vision-based_slam_b200", "csrc")
    os.makedirs(os.path.join(ROOT, "scratch"), exist_ok=True)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-DCTBA_BAND_TRACE",
                           "-shared", "-o", os.path.join(ROOT, "scratch", "libctba_trace.so"), *sorted(glob.glob(os.path.join(csrc, "*.cu")))])
    sys.exit(0)
pkg = importlib.import_module("rpg_vision-based_slam_b200"); scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
cfg = int(sys.argv[1])
sc = scene.config_scene(cfg)
g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
pkg.load_scene(g, sc)
for a in sys.argv[2:]:
    k, v = a.split("="); g.setOption(k, float(v))
g.setOption("bench_no_termination", 1)
g.optimize(3)
L = pkg.lib()
L.ctba_debug_band_trace(None, 1)
s = g.optimize(1)
buf = np.zeros((3, 8192, 2), dtype=np.uint64)
L.ctba_debug_band_trace(buf.ctypes.data_as(C.c_void_p), 0)
print("solve ms", 1e3 * s["time_solve_s"], "launches", s["kernel_launches"])
names = ["panel", "gather", "border"]
valid = [(buf[k, :, 1] > 0) & (buf[k, :, 0] < np.uint64(2**63)) for k in range(3)]
t0 = min(int(buf[k, valid[k], 0].min()) for k in range(3) if valid[k].any())
for k in range(3):
    v = np.nonzero(valid[k])[0]
    if v.size == 0:
        continue
    st = (buf[k, v, 0].astype(np.int64) - t0) * 1e-3; en = (buf[k, v, 1].astype(np.int64) - t0) * 1e-3
    print(f"{names[k]:7s}: {v.size} launches, duration us median {np.median(en - st):.2f} mean {np.mean(en - st):.2f} max {np.max(en - st):.2f}; first start {st.min():.1f} last end {en.max():.1f}")
    if k == 0:
        print("   panel start-to-start us: median", np.median(np.diff(st)), "mean", np.mean(np.diff(st)))
        P = (st, en, v)
# detail of steps 200..207 (level 0 and the interface level share step indices: the later launch wins; early steps belong to level 0 only when far from 0)
st, en, v = P
for t in range(200, 208):
    row = [f"t={t}"]
    for k in range(2):
        if valid[k][t]:
            row.append(f"{names[k]} [{(int(buf[k, t, 0]) - t0) * 1e-3:9.2f}, {(int(buf[k, t, 1]) - t0) * 1e-3:9.2f}]")
    print("  ".join(row))
for t in range(192, 216, 8):
    if valid[2][t]:
        print(f"border chunk t0={t}: [{(int(buf[2, t, 0]) - t0) * 1e-3:9.2f}, {(int(buf[2, t, 1]) - t0) * 1e-3:9.2f}]")
# ---- whole-factorisation view: where the solve time goes (level-0 steps are those with t above the interface level's step count) ----
st, en, v = P
lvl0 = v >= 150
d = np.diff(st[lvl0])
print("level-0 panel steps t>=150: n", lvl0.sum(), "span ms", (en[lvl0].max() - st[lvl0].min()) * 1e-3, "start-to-start us: p10/p50/p90/max",
      np.percentile(d, 10), np.percentile(d, 50), np.percentile(d, 90), d.max())
for a in range(150, int(v.max()), 150):
    m = (v >= a) & (v < a + 150)
    if m.sum() > 2:
        print(f"   t in [{a},{a + 150}): start-to-start median {np.median(np.diff(st[m])):.2f} us, panel duration median {np.median(en[m] - st[m]):.2f}")
vu = np.nonzero(valid[1])[0]; vu = vu[vu >= 150]
if vu.size:
    su = (buf[1, vu, 0].astype(np.int64) - t0) * 1e-3; eu = (buf[1, vu, 1].astype(np.int64) - t0) * 1e-3
    print("update kernel duration us (t>=150): p10/p50/p90", np.percentile(eu - su, 10), np.percentile(eu - su, 50), np.percentile(eu - su, 90))
vb = np.nonzero(valid[2])[0]
if vb.size:
    sb = (buf[2, vb, 0].astype(np.int64) - t0) * 1e-3; eb = (buf[2, vb, 1].astype(np.int64) - t0) * 1e-3
    print("border chunks:", vb.size, "duration us p10/p50/p90", np.percentile(eb - sb, 10), np.percentile(eb - sb, 50), np.percentile(eb - sb, 90),
          "busy fraction of the level-0 span", (eb - sb).sum() / (en[lvl0].max() - st[lvl0].min()), "last border end ms", eb.max() * 1e-3)
    # lag of the border pipeline behind the panel chain: chunk starting at step t can start after panel step t+7
    lag = [sb[i] - en[np.nonzero(v == min(vb[i] + 7, v.max()))[0][0]] for i in range(vb.size) if (v == min(vb[i] + 7, v.max())).any()]
    print("border chunk start minus end of its last panel step, us: p10/p50/p90", np.percentile(lag, 10), np.percentile(lag, 50), np.percentile(lag, 90))
low = v < 150
if low.any():
    print("interface-level steps (t<150, overwrite level 0): n", low.sum(), "span ms", (en[low].max() - st[low].min()) * 1e-3, "starts at ms", st[low].min() * 1e-3,
          "start-to-start median us", np.median(np.diff(st[low])))
print("first event 0, last panel/update/border end ms:", en.max() * 1e-3)
