"""Generator of tests/golden/align_n{4,6}.npz: the frozen-spline world alignment problem (src/align_spline_to_global_frame.cpp) evaluated
and solved by the CPU oracle on the seeded case of tests/align_case.py.  Run from the repo root:  python tests/golden/make_golden_align.py
Stored: a fingerprint of the inputs, the residuals and the 3n x 11 Jacobian at a NON-default linearisation point, the LM trace from the
default point and the optimised blocks."""
import importlib
import os
import sys

import numpy as np
from scipy.spatial.transform import Rotation as Rot

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.dirname(HERE)):
    if p not in sys.path:
        sys.path.insert(0, p)

CASES = {"align_n6": 6, "align_n4": 4}
CORR10 = np.array([0.07, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 0.004])      # scale_corr, (R_bc, p_bc identity), t_cam_imu 0, t_s_gp_corr
WG_CORR = np.r_[Rot.from_rotvec([0.05, -0.1, 0.15]).as_quat(), [0.3, -0.2, 0.1]]
P_BP = np.array([0.08, -0.03, 0.15])


def case_of(name):
    from oracle import binding
    import align_case
    scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
    binding.build()
    N = CASES[name]
    return binding, align_case, align_case.make_alignment_case(scene, binding, N, seed=500 + N, duration_s=12.0)


def evaluate_at_point(binding, align_case, case):
    """residuals / Jacobians at the non-default point (scale_corr, t_s_gp_corr, R_w_g_corr, p_w_g_corr, p_bp all non-trivial)"""
    o = binding.OracleSpline(case["N"], case["dt_ns"], case["start_t_ns"])
    align_case.load_alignment(o, case, True, True)
    o.setGlobals(case["scale_init"], P_BP, np.array([0.0, 0.0, -9.81]), case["t_off_init"])
    o.setCorrections(CORR10)
    o.setWorldColmapTransformation(case["T_wg_init"], WG_CORR)
    return o.evaluate(7)


def main():
    for name in CASES:
        binding, align_case, case = case_of(name)
        r, J, _ = evaluate_at_point(binding, align_case, case)
        o = binding.OracleSpline(case["N"], case["dt_ns"], case["start_t_ns"])
        align_case.load_alignment(o, case, True, True)
        s = o.optimize(30)
        T, corr = o.getWorldColmapTransformation()
        fp = np.array([case["z"].sum(), case["q"].sum(), case["p"].sum(), float(len(case["t_ns"]))])
        np.savez_compressed(os.path.join(HERE, name + ".npz"), fingerprint=fp, r=r, J=J[:, 6 * case["N"]:], lm_iterations=s["iterations"],
                            T_wg=T, wg_corr=corr, globals=o.getGlobals())
        print(name, r.shape, len(s["iterations"]), s["final_cost"])


if __name__ == "__main__":
    main()
