"""computeValidReprojections (bspline_so3r3.h:406-521): the product's flat-array host implementation (ctba_compute_valid_reprojections,
on a planning-only handle — no GPU needed) against the oracle's restatement in the reference's own shape, plus properties the
procedure guarantees.  Bit-exact: flags and counts are integers."""
import numpy as np
import pytest


def _candidates(sc, rng, extra_frac=0.0, behind=0):
    """All 2-D points of the scene as candidates (frame-major CSR), optionally shuffled inside each frame."""
    pb = sc["obs_begin"].astype(np.int32).copy()
    uv = sc["uv"].copy()
    li = sc["lm_index"].astype(np.int32).copy()
    for f in range(pb.size - 1):
        a, b = pb[f], pb[f + 1]
        perm = rng.permutation(b - a) + a
        uv[a:b] = uv[perm]; li[a:b] = li[perm]
    uv[:, 0] = np.clip(uv[:, 0], 0.0, sc["width"] - 1e-6); uv[:, 1] = np.clip(uv[:, 1], 0.0, sc["height"] - 1e-6)
    return pb, uv, li


@pytest.mark.parametrize("order,cell,cam", [(6, 30, "euroc"), (4, 20, "fpv"), (6, 47, "fov")])
def test_selection_matches_oracle(pkg, scene, oracle, order, cell, cam):
    rng = np.random.default_rng(order * 100 + cell)
    sc = scene.make_scene(order=order, duration_s=6.0, n_landmarks=900, track_len=14, imu_hz=0, gp_hz=10, cam=cam, seed=cell, depth=(0.2, 6.0))
    # some landmarks behind / too close to the camera so that the 30 cm / z < 0 rule fires
    lm = sc["landmarks"].copy()
    k = rng.choice(lm.shape[0], 60, replace=False)
    lm[k] = sc["gt_knot_p"][rng.integers(0, sc["gt_knot_p"].shape[0], 60)] + rng.normal(size=(60, 3)) * 0.15
    sc["landmarks"] = lm
    pb, uv, li = _candidates(sc, rng)
    # frames before the spline start and beyond its end are skipped without touching the grid memory
    t = sc["frame_t_ns"].copy()
    t[3] = sc["start_t_ns"] - 5
    t[-2:] += 10 * sc["dt_ns"] * 1000
    prod = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=-1)
    orc = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    for h in (prod, orc):
        h.setCamera(sc["cam_model"], sc["intr"], sc["dist"], sc["q_bc"], sc["p_bc"], sc["t_offset_cam_imu"])
        h.setKnots(sc["knot_q"], sc["knot_p"])
        h.setGlobals(1.07, sc["p_bp"], sc["g"], sc["t_s_gp"])
        h.set3DPoints(sc["landmarks"])
    vp, np_ = prod.computeValidReprojections(t, pb, uv, li, sc["width"], sc["height"], cell)
    vo, no = orc.computeValidReprojections(t, pb, uv, li, sc["width"], sc["height"], cell)
    assert (vp == vo).all() and (np_ == no).all()
    assert np_[3] == 0 and (np_[-2:] == 0).all() and not vp[pb[3]:pb[4]].any()
    assert 0 < vp.sum() < vp.size                      # the grid thins the features out
    assert (np.add.reduceat(vp.astype(np.int64), pb[:-1])[np.diff(pb) > 0] == np_[np.diff(pb) > 0]).all()   # n_valid counts the flags
    ncol = sc["width"] // cell + (sc["width"] % cell > 0)
    # a tracked point keeps being selected while its landmark stays visible: selected landmarks of frame k that reappear in k+1 are selected again
    # unless the depth rule rejects them there
    for f in range(5, 12):
        a, b, c = pb[f], pb[f + 1], pb[f + 2]
        sel = set(li[a:b][vp[a:b]].tolist())
        nxt = dict(zip(li[b:c].tolist(), range(b, c)))
        carried = [nxt[l] for l in sel if l in nxt]
        assert carried
        rejected = [q for q in carried if not vp[q]]
        assert len(rejected) <= len(carried) // 4          # only the depth rule can drop a carried point


def test_selection_feeds_add_frames(pkg, scene):
    """the selected points are what addImagePointsMeasurement consumes (frames with n_valid == 0 are skipped, bspline_so3r3.h:967)"""
    rng = np.random.default_rng(3)
    sc = scene.make_scene(order=6, duration_s=4.0, n_landmarks=500, track_len=10, imu_hz=0, gp_hz=10, seed=8)
    pb, uv, li = _candidates(sc, rng)
    h = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=-1)
    pkg.load_scene(h, sc, with_frames=False)
    v, nv = h.computeValidReprojections(sc["frame_t_ns"], pb, uv, li, sc["width"], sc["height"], 30)
    ob = np.concatenate([[0], np.cumsum(nv)]).astype(np.int32)
    added = h.addFrames(sc["frame_t_ns"], ob, uv[v], li[v], 1.0)
    assert added == int((nv > 0).sum()) and h.counts()["obs"] == int(v.sum())
    with pytest.raises(pkg.CtbaError):
        h.computeValidReprojections(sc["frame_t_ns"], pb, uv, li, sc["width"], sc["height"], 1)   # "Not working when cell_size_px = 1"


@pytest.mark.gpu
def test_evaluate_errors_match_unit_weight_residual_norms(pkg, scene, oracle):
    """evaluate*Errors (bspline_so3r3.h:1229-1579): residual norms with inv_std = 1, against the oracle's residuals."""
    sc = scene.make_scene(order=6, duration_s=3.0, n_landmarks=150, track_len=10, imu_hz=100, gp_hz=10, dt_s_bias=0.25, outlier_frac=0.05, seed=4)
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(gpu, sc); pkg.load_scene(cpu, sc)
    for kind, nres, std in ((pkg.CTBA_REPROJ, 2, sc["sigma_reproj_px"]), (pkg.CTBA_ACCEL, 3, sc["sigma_acc_c"]), (pkg.CTBA_GYRO, 3, sc["sigma_omega_c"]),
                            (pkg.CTBA_GP, 3, sc["gp_std"])):
        e = gpu.evaluateErrors(kind)
        r = cpu.evaluate(kind, jac=False)[0].reshape(-1, nres)
        assert e.size == r.shape[0] > 0
        np.testing.assert_allclose(e, np.linalg.norm(r, axis=1) * std, rtol=1e-10, atol=1e-13)
    assert (gpu.evaluateErrors(pkg.CTBA_REPROJ) == 0).any()      # gated observations (|e|^2 > 100 px^2) report 0 like the reference


@pytest.mark.gpu
def test_evaluate_errors_follow_the_folded_time_offsets(pkg, scene, oracle):
    """evaluate{Accel,Gyro,GlobalPos}Errors re-derive every sample's window from the CURRENT time offsets (bspline_so3r3.h:1239-1256,
    1346-1363, 1692-1707): after relinearize*TimeOffset folded a non-zero correction, the errors must be those of a problem that is built
    afresh with the folded offsets (skip rules re-applied) — not those of the work items the solver keeps."""
    sc = scene.make_scene(order=6, duration_s=4.0, n_landmarks=60, track_len=8, imu_hz=100, gp_hz=20, dt_s_bias=0.25, seed=41,
                          flags=(False, True, True, True, True))
    g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    pkg.load_scene(g, sc)
    corr = np.zeros(10); corr[4] = 1.0; corr[8] = 0.0437; corr[9] = -0.0312     # moves samples across knot boundaries and across the spline's ends
    g.setCorrections(corr)
    before = [g.evaluateErrors(k) for k in (1, 2, 3)]
    g.relinearize()
    after = [g.evaluateErrors(k) for k in (1, 2, 3)]
    sc2 = dict(sc); sc2["t_offset_cam_imu"] = sc["t_offset_cam_imu"] + corr[8]; sc2["t_s_gp"] = sc["t_s_gp"] + corr[9]
    o = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(o, sc2)
    for k, sigma, e in ((1, sc["sigma_acc_c"], after[0]), (2, sc["sigma_omega_c"], after[1]), (3, sc["gp_std"], after[2])):
        r, _, _ = o.evaluate(k, jac=False)
        ref = np.linalg.norm(r.reshape(-1, 3), axis=1) * sigma
        assert e.shape == ref.shape
        np.testing.assert_allclose(e, ref, rtol=1e-9, atol=1e-12)
    assert len(after[0]) != len(before[0]) or len(after[2]) != len(before[2])      # the fold moved samples over the spline's ends: skip rules re-applied
    # the solver's own work items are untouched: cost and counts as before the error pass
    assert g.counts()["accel"] == len(before[0])
