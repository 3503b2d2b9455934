"""End-to-end LM parity: ctba_solve (device LM: Schur complement + bordered banded Cholesky) against the oracle LM
(dense normal equations of the FULL problem incl. landmarks, Jet autodiff).  north_star tolerances: per-iteration
cost 1e-8 relative, final cost 1e-6 relative, trajectory 1 mm / 0.01 deg."""
import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot

pytestmark = pytest.mark.gpu


def _run(pkg, oracle, sc, iters, **kw):
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(gpu, sc, **kw); pkg.load_scene(cpu, sc, **kw)
    sg = gpu.optimize(iters)
    so = cpu.optimize(iters, 8)
    return gpu, cpu, sg, so


def _check(gpu, cpu, sg, so, cost_tol=1e-8):
    ig, io = sg["iterations"], so["iterations"]
    assert ig.shape == io.shape, (ig.shape, io.shape, sg["termination_type"], so["termination_type"])
    assert sg["termination_type"] == so["termination_type"]
    assert (ig[:, 6] == io[:, 6]).all()                      # same accept/reject sequence
    np.testing.assert_allclose(ig[:, 0], io[:, 0], rtol=cost_tol)   # per-iteration cost
    np.testing.assert_allclose(ig[:, 5], io[:, 5], rtol=1e-6)       # trust-region radius trajectory
    ok = io[:, 6] > 0
    np.testing.assert_allclose(ig[1:, 3], io[1:, 3], rtol=1e-5, atol=1e-12)  # step norms
    assert abs(sg["final_cost"] - so["final_cost"]) <= 1e-6 * so["final_cost"]
    qg, pg = gpu.getKnots(); qo, po = cpu.getKnots()
    assert np.abs(pg - po).max() < 1e-3                      # 1 mm
    ang = (Rot.from_quat(qg).inv() * Rot.from_quat(qo)).magnitude()
    assert np.rad2deg(ang.max()) < 0.01                      # 0.01 deg
    np.testing.assert_allclose(gpu.get3DPoints(), cpu.get3DPoints(), atol=1e-3)
    np.testing.assert_allclose(gpu.getGlobals(), cpu.getGlobals(), atol=1e-5)
    ag, gg = gpu.getBiases(); ao, go = cpu.getBiases()
    if ao.size:
        np.testing.assert_allclose(ag, ao, atol=1e-5); np.testing.assert_allclose(gg, go, atol=1e-6)


def test_lm_full_gvi(pkg, scene, oracle):
    """optimize_continuous_time-shaped: vision + IMU + bias spline + GP + gravity norm, every global free."""
    sc = scene.make_scene(order=6, duration_s=4.0, n_landmarks=120, track_len=12, imu_hz=100, gp_hz=10, dt_s_bias=0.25,
                          flags=(True, True, True, True, True), gravity_norm=True, outlier_frac=0.02)
    gpu, cpu, sg, so = _run(pkg, oracle, sc, 12)
    assert so["num_successful_steps"] >= 5
    _check(gpu, cpu, sg, so)


def test_lm_vi(pkg, scene, oracle):
    """optimize_vi_continuous_time-shaped: vision + IMU, no GP (gauge left to LM damping)."""
    sc = scene.make_scene(order=6, duration_s=3.0, n_landmarks=100, track_len=10, imu_hz=100, gp_hz=0, dt_s_bias=0.2, seed=77,
                          flags=(False, True, True, True, True), gravity_norm=True)
    gpu, cpu, sg, so = _run(pkg, oracle, sc, 8)
    _check(gpu, cpu, sg, so)


def test_lm_gv_cubic(pkg, scene, oracle):
    """optimize_gv_continuous_time-shaped with a cubic spline: vision + GP, extrinsics free (ctor default), fisheye camera."""
    sc = scene.make_scene(order=4, duration_s=3.0, n_landmarks=100, track_len=10, imu_hz=0, gp_hz=10, cam="fpv", seed=5,
                          flags=(True, True, False, True, True))
    gpu, cpu, sg, so = _run(pkg, oracle, sc, 8)
    _check(gpu, cpu, sg, so)


def test_lm_gi_no_frames(pkg, scene, oracle):
    """optimize_gi_continuous_time-shaped (src/optimize_gi_continuous_time.cpp:303-431): global positions + IMU + bias cost, NO vision;
    p_bp, gravity and the spline / global-sensor time offset free; one iteration, relinearise, then the rest (outer / inner loop)."""
    sc = scene.make_scene(order=6, duration_s=4.0, n_landmarks=20, track_len=4, imu_hz=100, gp_hz=10, dt_s_bias=0.25, seed=41,
                          flags=(False, True, True, True, True), gravity_norm=True)
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(gpu, sc, with_frames=False); pkg.load_scene(cpu, sc, with_frames=False)
    assert gpu.counts()["frames"] == 0 and gpu.counts()["gp"] > 0 and gpu.counts()["accel"] > 0
    for h in (gpu, cpu):
        h.optimize(1); h.relinearize()
    sg = gpu.optimize(8); so = cpu.optimize(8, 8)
    _check(gpu, cpu, sg, so)


def test_lm_nested_interface_levels(pkg, scene, oracle):
    """The same LM parity with the band solver's nested interface reduction (odd-even over the separators, 3 levels here) and with a
    forced flat 5-partition plan: the linear solve is exact either way."""
    sc = scene.make_scene(order=4, duration_s=20.0, n_landmarks=420, track_len=6, imu_hz=0, gp_hz=10, seed=29, flags=(True, True, False, True, True))
    for opts in ({"band_nested": 1}, {"band_nested": 0, "band_partitions": 5}, {"band_nested": 1, "band_partitions": 6}):
        gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
        cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
        pkg.load_scene(gpu, sc); pkg.load_scene(cpu, sc)
        for k, v in opts.items():
            gpu.setOption(k, v)
        sg, so = gpu.optimize(5), cpu.optimize(5, 8)
        _check(gpu, cpu, sg, so)


def test_lm_sparse_frame_rate(pkg, scene, oracle):
    """3 Hz camera on 0.1 s knots (the config-3 regime: fewer than two frames per knot -> thread-per-task Schur expansion)."""
    sc = scene.make_scene(order=6, duration_s=8.0, cam_hz=3.0, n_landmarks=120, track_len=6, imu_hz=0, gp_hz=10, seed=13,
                          flags=(False, True, False, True, True))
    gpu, cpu, sg, so = _run(pkg, oracle, sc, 6)
    _check(gpu, cpu, sg, so)


def test_lm_fixed_globals_and_first_call(pkg, scene, oracle):
    """optimize(1) as the drivers' first call (always NO_CONVERGENCE), then a second Solve re-derives the Jacobi scaling;
    globals constant."""
    sc = scene.make_scene(order=6, duration_s=2.5, n_landmarks=80, track_len=8, imu_hz=50, gp_hz=10, dt_s_bias=0.2, seed=3,
                          flags=(False, False, False, False, False))
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(gpu, sc); pkg.load_scene(cpu, sc)
    s1g, s1o = gpu.optimize(1), cpu.optimize(1, 4)
    assert s1g["termination_type"] == 1 and s1o["termination_type"] == 1
    gpu.relinearize(); cpu.relinearize()
    s2g, s2o = gpu.optimize(6), cpu.optimize(6, 4)
    _check(gpu, cpu, s2g, s2o)
    g = gpu.getGlobals()
    assert g[0] == 0.0 and (g[5:10] == 0.0).all()     # constant parameter blocks untouched


def test_solve_idempotent_at_convergence(pkg, scene):
    """Size-independent property: a second Solve from the converged state does not move it (cost decrease < function tol)."""
    sc = scene.make_scene(order=6, duration_s=6.0, n_landmarks=400, track_len=12, imu_hz=100, gp_hz=10, dt_s_bias=0.25, seed=21)
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    pkg.load_scene(gpu, sc)
    s1 = gpu.optimize(40)
    assert s1["termination_type"] == 0
    assert s1["final_cost"] < 0.05 * s1["initial_cost"]
    s2 = gpu.optimize(5)
    assert abs(s2["final_cost"] - s1["final_cost"]) <= 1e-5 * s1["final_cost"]
    q, p = gpu.getKnots()
    # recovered trajectory close to the ground truth where it is observed (synthetic noise level)
    inner = slice(3, -4)
    assert np.abs(p[inner] - sc["gt_knot_p"][inner]).max() < 0.05


@pytest.mark.gpu
@pytest.mark.parametrize("shape", ["gvi", "gv"])
def test_pcg_step_equals_cholesky_step(pkg, scene, shape):
    """option linear_solver = 1 (block-Jacobi PCG on the reduced system, ctba_pcg.cu) against the bordered band Cholesky on the same handle
    inputs: same decisions and cost (1e-8) over the first iterations, the same cost afterwards (1e-5 / 1e-4), final knots within 1 mm."""
    kw = dict(order=6, duration_s=20.0, n_landmarks=900, track_len=12, seed=77, outlier_frac=0.02)
    if shape == "gvi":
        kw.update(imu_hz=100, gp_hz=10, dt_s_bias=0.5, flags=(False, True, True, True, True), gravity_norm=True)
    else:
        kw.update(imu_hz=0, gp_hz=10, flags=(False, True, False, True, True))
    sc = scene.make_scene(**kw)
    out = []
    for solver in (0, 1):
        g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
        pkg.load_scene(g, sc)
        g.setOption("linear_solver", solver)
        g.setOption("pcg_tol", 1e-12 if shape == "gvi" else 1e-11)    # the 1e6-weighted gravity-norm row of the G+V+I shape needs the tighter residual
        g.setOption("pcg_max_iters", 40000)
        s = g.optimize(5)
        out.append((s["iterations"], g.getKnots()[1], g.getStat("pcg_iterations") if solver else 0))
        g.close()
    (ia, pa, _), (ib, pb, nit) = out
    # The first iterations: same decisions, cost at the 1e-8 gate.  Later ones: an iterative solve stops at a residual, not at the exact step,
    # and near convergence (large trust region, condition ~1e10) the remaining step error shows in the cost at the 1e-6 level; a decision whose
    # ratio sits at the acceptance threshold may then fall the other way (seen on a B200: 2 runs of 6), so the later iterations are held to
    # the cost they reach, not to the path they take.
    k = 3
    assert len(ia) > k and len(ib) > k and (ia[:k, 6] == ib[:k, 6]).all()
    np.testing.assert_allclose(ib[:k, 0], ia[:k, 0], rtol=1e-8)
    n = min(len(ia), len(ib))
    same = np.cumprod(ia[:n, 6] == ib[:n, 6]).astype(bool)         # the common prefix of identical decisions
    np.testing.assert_allclose(ib[:n, 0][same], ia[:n, 0][same], rtol=1e-5)
    assert abs(ib[-1, 0] - ia[-1, 0]) <= 1e-4 * ia[-1, 0]
    # the 1 mm trajectory gate on the knots the measurements constrain.  The first / last knots of the spline support only the ends of the
    # first / last windows: they are (nearly) in the null space of the problem, where an iterative solve stopped at a residual and the exact
    # step legitimately differ by decimetres (seen on a B200: 0.19 m on the last supported knot of the G+V shape).
    inner = slice(3, -4)
    assert np.abs(pa[inner] - pb[inner]).max() < 1e-3, np.abs(pa[inner] - pb[inner]).max()
    assert 0 < nit <= 40000, nit


def _gappy_scene_with_duplicates(scene, seed=31):
    """Tracks with gaps (40 % of the observations dropped at random) and a few landmarks observed TWICE in one frame (a second, slightly
    shifted detection): the cases the landmark Schur gather treats apart — track bitmask / list search / same-frame pairs."""
    sc = scene.make_scene(order=6, duration_s=6.0, n_landmarks=260, track_len=14, imu_hz=0, gp_hz=10, seed=seed, flags=(False, True, False, True, True))
    rng = np.random.default_rng(seed)
    F = len(sc["frame_t_ns"])
    ob, uv, li = sc["obs_begin"], sc["uv"], sc["lm_index"]
    fi = np.repeat(np.arange(F), np.diff(ob))
    keep = rng.random(len(li)) > 0.4
    fi, uv, li = fi[keep], uv[keep], li[keep]
    dup = rng.choice(len(li), 25, replace=False)                       # the same landmark once more in the same frame
    fi = np.r_[fi, fi[dup]]; li = np.r_[li, li[dup]]; uv = np.vstack([uv, uv[dup] + rng.normal(size=(len(dup), 2)) * 0.7])
    o = np.lexsort((li, fi))
    fi, li, uv = fi[o], li[o], uv[o]
    ob2 = np.zeros(F + 1, dtype=np.int32); np.add.at(ob2, fi + 1, 1)
    sc = dict(sc, obs_begin=np.cumsum(ob2).astype(np.int32), uv=np.ascontiguousarray(uv), lm_index=li.astype(np.int32))
    return sc


@pytest.mark.parametrize("opts", [{}, {"schur_threads": 256}, {"schur_threads": 128}, {"schur_mask": 0}, {"schur_mask": 0, "schur_list_a": 0},
                                  {"schur_compact": 0}])
def test_lm_gappy_tracks_and_duplicate_observations(pkg, scene, oracle, opts):
    """Every variant of the landmark Schur gather (track-bitmask kernel at its three CTA sizes, list-order / observation-order partner
    blocks, the Y/Z path) on tracks with gaps and same-frame duplicates: LM trace against the oracle."""
    sc = _gappy_scene_with_duplicates(scene)
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(gpu, sc); pkg.load_scene(cpu, sc)
    for k, v in opts.items():
        gpu.setOption(k, v)
    sg, so = gpu.optimize(6), cpu.optimize(6, 8)
    assert so["num_successful_steps"] >= 2
    _check(gpu, cpu, sg, so)


def test_lm_tracks_longer_than_the_bitmask(pkg, scene, oracle):
    """Landmarks tracked over 70 frames: beyond the 64-bit track mask of k_frame_schur_m, so the gather falls back to the list search of
    k_frame_schur_c (with the list-ordered partner blocks); same LM trace as the oracle."""
    sc = scene.make_scene(order=4, duration_s=4.0, cam_hz=25.0, n_landmarks=60, track_len=70, imu_hz=0, gp_hz=10, seed=17, flags=(False, True, False, True, True))
    F = len(sc["frame_t_ns"])
    fi = np.repeat(np.arange(F), np.diff(sc["obs_begin"]))
    span = max(fi[sc["lm_index"] == l].max() - fi[sc["lm_index"] == l].min() for l in np.unique(sc["lm_index"]))
    assert span >= 64, span          # the scene really has tracks that the mask cannot hold
    gpu, cpu, sg, so = _run(pkg, oracle, sc, 5)
    _check(gpu, cpu, sg, so)
