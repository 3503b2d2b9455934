"""The C++ host mirror of BSplineSO3R3 (include/ctba_bspline_so3r3.hpp) compiled with g++ against libctba.so: host logic on a
planning-only handle (CPU), a fit_spline_to_colmap-shaped solve on the GPU."""
import os
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "rpg_vision-based_slam_b200", "csrc")


def _build(name, tmp_path, pkg):
    pkg.lib()                                    # fails loudly if libctba.so was not built
    exe = str(tmp_path / name)
    cmd = ["g++", "-std=c++14", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", name + ".cpp"),
           "-L", LIBDIR, "-lctba", "-Wl,-rpath," + LIBDIR, "-o", exe]
    subprocess.check_call(cmd)
    return exe


def test_cpp_mirror_host_logic(pkg, tmp_path):
    exe = _build("mirror_host", tmp_path, pkg)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and out.stdout.strip().endswith("OK"), out.stdout + out.stderr


@pytest.mark.gpu
def test_cpp_mirror_fit_on_gpu(pkg, tmp_path):
    exe = _build("mirror_fit", tmp_path, pkg)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.strip().endswith("OK"), out.stdout + out.stderr
