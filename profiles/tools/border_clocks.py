"""Phase clocks of one k_band_border CTA inside the running graph (debug build -DCTBA_BAND_TIMING).
build: python profiles/tools/border_clocks.py build ; run (GPU box): CTBA_LIB_PATH=scratch/libctba_timing.so python profiles/tools/border_clocks.py 3"""
import ctypes as C, glob, importlib, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
if sys.argv[1] == "build":
    csrc = os.path.join(ROOT, "rpg_vision-based_slam_b200", "csrc")
    os.makedirs(os.path.join(ROOT, "scratch"), exist_ok=True)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-DCTBA_BAND_TIMING",
                           "-shared", "-o", os.path.join(ROOT, "scratch", "libctba_timing.so"), *sorted(glob.glob(os.path.join(csrc, "*.cu")))])
    sys.exit(0)
pkg = importlib.import_module("rpg_vision-based_slam_b200"); scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
sc = scene.config_scene(int(sys.argv[1]))
g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
pkg.load_scene(g, sc)
for a in sys.argv[2:]:
    k, v = a.split("="); g.setOption(k, float(v))
g.setOption("bench_no_termination", 1)
g.optimize(4)
buf = np.zeros((8, 8), dtype=np.int64)
pkg.lib().ctba_debug_border_clocks(buf.ctypes.data_as(C.c_void_p))
names = ["start", "Linv staged", "k loop done", "partials in smem + barrier", "combine + barrier", "triangular product + store", "barrier"]
print("one k_band_border CTA (grid index 30), 8 consecutive columns; cycles since the column started (1965 MHz: 1000 cycles = 0.51 us)")
for c in range(8):
    r = buf[c]
    print(f"column {c}: " + "  ".join(f"{names[k]} {int(r[k] - r[0])}" for k in range(1, 7)) + (f"  | next column starts {int(buf[c + 1][0] - r[0])}" if c < 7 else ""))
