#!/usr/bin/env python
"""Generates tests/golden/*.npz: frozen outputs of the CPU oracle on small seeded scenes.

The reference holds no golden vectors for this path and cannot be built here, so these
files do not come from the reference itself: they freeze the oracle (checked in the same test module
against an independent numpy/scipy re-derivation of the functors and complex-step-free Richardson
differences) so that (a) a later edit of oracle/ cannot silently move the parity target and (b) the
GPU tests have committed vectors to compare with.  Re-run only on purpose:
    python tests/golden/make_golden.py
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CASES = {
    # name: (make_scene kwargs, LM iterations)
    "gvi_n6_euroc": (dict(order=6, duration_s=2.0, n_landmarks=40, track_len=8, imu_hz=50, gp_hz=10, dt_s_bias=0.2, cam="euroc", seed=101,
                          flags=(True, True, True, True, True), gravity_norm=True, outlier_frac=0.05), 6),
    "gv_n4_fpv": (dict(order=4, duration_s=2.0, n_landmarks=40, track_len=8, imu_hz=0, gp_hz=10, cam="fpv", seed=102,
                       flags=(True, True, False, True, True)), 6),
    "vi_n6_fov": (dict(order=6, duration_s=2.0, n_landmarks=40, track_len=8, imu_hz=50, gp_hz=0, dt_s_bias=0.2, cam="fov", seed=103,
                       flags=(False, True, True, True, True), gravity_norm=True), 6),
}


def scene_of(name):
    scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
    kw, iters = CASES[name]
    return scene.make_scene(**kw), iters


def main():
    from oracle.binding import OracleSpline
    pkg = importlib.import_module("rpg_vision-based_slam_b200")
    for name in CASES:
        sc, iters = scene_of(name)
        o = OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
        pkg.load_scene(o, sc)
        out = {}
        for kind in range(6):
            r, J, dropped = o.evaluate(kind)
            assert dropped == 0.0
            out[f"r{kind}"] = r
            out[f"J{kind}"] = J
        out["cost"] = o.cost()
        s = o.optimize(iters, 4)
        out["lm_iterations"] = s["iterations"]
        out["lm_termination"] = np.array([s["termination_type"]])
        q, p = o.getKnots()
        out["knot_q"], out["knot_p"] = q, p
        out["landmarks"] = o.get3DPoints()
        out["globals"] = o.getGlobals()
        a, g = o.getBiases()
        out["bias_a"], out["bias_g"] = a, g
        # scene fingerprint: guards against a silent change of the generator
        out["scene_fingerprint"] = np.array([sc["uv"].sum(), sc["knot_q"].sum(), sc["landmarks"].sum(), float(sc["uv"].shape[0])])
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, {k: v.shape for k, v in out.items() if k.startswith("J")}, "final cost", s["final_cost"])


if __name__ == "__main__":
    main()
