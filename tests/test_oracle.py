"""CPU suite: the oracle against the golden vectors (independent mpmath / OpenCV / LAPACK / scipy
results committed under tests/golden/) and against its own invariants."""
import numpy as np
import pytest

from oracle import pyoracle as po
from sfm_b200 import synth
from tests.util import jacobian_err, kat_problem, rel, residual_err


def test_functor_matches_mpmath_and_opencv(golden_dir):
    g = np.load(f"{golden_dir}/functor_kat.npz")
    p = kat_problem(g)
    r, J, cost = po.evaluate(p)
    assert residual_err(r, g["r"], p.obs) < 1e-12
    # 50-digit derivative of the exact Rodrigues formula; |w| >= 1e-4 in the fixture (the band below
    # is ill-conditioned in the reference itself) and w == 0 (Taylor branch, exact there)
    assert jacobian_err(J, g["J"]) < 1e-11
    # OpenCV's analytic d(u,v)/d(rvec,tvec) — cv2.projectPoints has no skew term, so only the
    # skew == 0 cases are comparable (the reference's k[1] term, CTracker.cpp:597, is pinned by mpmath above)
    ns = g["K"][:, 1] == 0.0
    assert ns.sum() >= 30
    assert jacobian_err(J[ns][:, :, :6], g["J_cv2"][ns]) < 1e-11
    assert cost == pytest.approx(0.5 * (g["r"] ** 2).sum(), rel=1e-13)


def test_w_zero_branch_matches_survey_probe():
    # J_R at w=0, X=(0.1,-0.2,3), K of main.cpp:47-51
    cam = np.zeros((1, 6))
    K = np.array([[synth.FX, 0, synth.CX, 0, synth.FY, synth.CY, 0, 0, 1.0]])
    p = synth.Problem(obs=np.zeros((1, 2)), cam_idx=np.zeros(1, np.int32), pt_idx=np.zeros(1, np.int32), K=K, cam=cam,
                      pts=np.array([[0.1, -0.2, 3.0]]), cam_true=cam, pts_true=None)
    _, J, _ = po.evaluate(p)
    expect = np.array([[2.3836, 1073.798, 71.507], [-1071.941, -2.3716, 35.573]])
    assert np.allclose(J[0, :, :3], expect, rtol=2e-4)


def test_schur_matches_lapack(golden_dir):
    g = np.load(f"{golden_dir}/normal_eq_kat.npz")
    from tests.golden.make_golden import small_problem

    p = small_problem()
    o = po.schur(p, radius=float(g["radius"]))
    nc6 = 6 * p.n_cams
    assert rel(o["scale"], g["scale"]) < 1e-12
    assert rel(o["S"], g["S"]) < 1e-9
    assert rel(o["rhs"], g["rhs"]) < 1e-9
    # first LM step of the oracle == LAPACK solve of the full damped normal equations
    _, _, s, tr = po.solve(p, opts=po.default_options(max_num_iterations=1))
    assert tr[1]["model_cost_change"] == pytest.approx(float(g["model_cost_change"]), rel=1e-8)
    assert s["initial_cost"] == pytest.approx(float(g["cost"]), rel=1e-12)
    yc = np.linalg.solve(o["S"], o["rhs"])
    assert rel(yc, g["y"][:nc6]) < 1e-7


def test_lm_converges_to_scipy_cost(golden_dir):
    g = np.load(f"{golden_dir}/lm_final.npz")
    from tests.golden.make_golden import small_problem

    p = small_problem()
    _, _, s, _ = po.solve(p, opts=po.default_options(function_tolerance=1e-14, parameter_tolerance=1e-14, max_num_iterations=200))
    assert s["final_cost"] == pytest.approx(float(g["cost"]), rel=1e-6)


def test_lm_c1_shape_matches_survey_probe():
    # C1 shape converges by function tolerance in 3 iterations, RMS ~0.68 px
    p = synth.make_config("c1")
    _, _, s, tr = po.solve(p)
    assert s["termination_type"] == 0 and s["num_iterations"] == 3 and s["num_successful_steps"] == 2
    assert 0.6 < s["rms_px"] < 0.75
    assert [round(t["trust_region_radius"]) for t in tr[:3]] == [10000, 30000, 90000]


def test_pcg_equals_dense_on_ring():
    p = synth.make_ring(10, 200, 6, seed=3)
    _, _, a, _ = po.solve(p, opts=po.default_options(linear_solver=1))
    _, _, b, _ = po.solve(p, opts=po.default_options(linear_solver=2, pcg_tolerance=1e-12))
    assert a["num_iterations"] == b["num_iterations"]
    assert b["final_cost"] == pytest.approx(a["final_cost"], rel=1e-9)


def test_cholesky_kat():
    rng = np.random.default_rng(0)
    for n in (1, 7, 96, 200, 301):
        A = rng.normal(size=(n, n))
        A = A @ A.T + n * np.eye(n)
        rc, L = po.cholesky(A, threads=2)
        assert rc == 0
        assert rel(L, np.linalg.cholesky(A)) < 1e-12
    bad = -np.eye(3)
    assert po.cholesky(bad)[0] == 1


def test_modes_and_invalid_mode():
    p = synth.make_ring(8, 100, 5, seed=5)
    for mode in (0, 1):
        cam, pts, s, _ = po.solve(p, mode=mode)
        assert s["final_cost"] < s["initial_cost"]
        if mode == 0:
            assert np.array_equal(cam, p.cam) and not np.array_equal(pts, p.pts)
        else:
            assert np.array_equal(pts, p.pts) and not np.array_equal(cam, p.cam)
    cam, pts, s, _ = po.solve(p, mode=7)  # CTracker.cpp:692-693: default branch adds nothing
    assert np.array_equal(cam, p.cam) and np.array_equal(pts, p.pts)


def test_distortion_extension_reduces_to_reference_at_zero():
    p = synth.make_ring(8, 100, 5, seed=6)
    r0, J0, c0 = po.evaluate(p)
    p.dist = np.zeros((p.n_cams, 2))
    r1, J1, c1 = po.evaluate(p)
    assert np.array_equal(r0, r1) and np.array_equal(J0, J1) and c0 == c1


def test_frame_error_statistics_loop_equals_grouped_numpy():
    """The restated loop of CSfM.cpp:350-382 against a vectorised grouping of the same residuals."""
    p = synth.make_ring(9, 120, 5, seed=3, name="stats")
    rms, mx = po.frame_error_stats(p)
    r, _, _ = po.evaluate(p, want_jac=False)
    d2 = (r * r).sum(axis=1)
    for c in range(p.n_cams):
        sel = d2[p.cam_idx == c]
        assert rms[c] == pytest.approx(np.sqrt(sel.mean()), rel=1e-13) and mx[c] == pytest.approx(np.sqrt(sel.max()), rel=1e-15)


def _scipy_cost_with_fixed(p, cam_fixed, pt_fixed):
    """Converged cost of the same problem with the masked blocks held constant, from an independent optimiser
    (scipy trust-region-reflective over the FREE parameters only, numpy projection)."""
    from scipy.optimize import least_squares

    cf, pf = np.asarray(cam_fixed, bool), np.asarray(pt_fixed, bool)
    nc, npf = int((~cf).sum()), int((~pf).sum())

    def unpack(x):
        cam, pts = p.cam.copy(), p.pts.copy()
        cam[~cf] = x[:6 * nc].reshape(nc, 6)
        pts[~pf] = x[6 * nc:].reshape(npf, 3)
        return cam, pts

    def fun(x):
        cam, pts = unpack(x)
        uv, _ = synth.project(cam, p.K, pts, p.cam_idx, p.pt_idx)
        return (uv - p.obs).ravel()

    x0 = np.concatenate([p.cam[~cf].ravel(), p.pts[~pf].ravel()])
    sol = least_squares(fun, x0, method="trf", xtol=1e-14, ftol=1e-14, gtol=1e-12, max_nfev=300)
    return 0.5 * float((sol.fun ** 2).sum())


def test_constant_blocks_gauge_and_window():
    """parameter blocks held constant.  Fixed blocks keep their value bit for bit; the free ones
    converge to the optimum an independent optimiser finds for the reduced problem; an all-constant residual
    contributes only Summary::fixed_cost."""
    p = synth.make_ring(8, 90, 5, seed=21, name="gauge")
    tight = po.default_options(function_tolerance=1e-14, parameter_tolerance=1e-14, max_num_iterations=200)
    cf = np.zeros(p.n_cams, np.uint8); cf[0] = 1            # gauge: oldest keyframe fixed
    pf = np.zeros(p.n_pts, np.uint8)
    cam, pts, s, _ = po.solve(p, opts=tight, cam_fixed=cf)
    assert np.array_equal(cam[0], p.cam[0]) and not np.array_equal(cam[1:], p.cam[1:])
    assert s["final_cost"] == pytest.approx(_scipy_cost_with_fixed(p, cf, pf), rel=1e-6)
    assert s["fixed_cost"] == 0.0
    # window: only the last three keyframes and the points they see are free
    cf[:] = 1; cf[-3:] = 0
    seen = np.zeros(p.n_pts, bool); seen[p.pt_idx[np.isin(p.cam_idx, np.nonzero(cf == 0)[0])]] = True
    pf = (~seen).astype(np.uint8)
    cam, pts, s, _ = po.solve(p, opts=tight, cam_fixed=cf, pt_fixed=pf)
    assert np.array_equal(cam[cf == 1], p.cam[cf == 1]) and np.array_equal(pts[pf == 1], p.pts[pf == 1])
    assert s["final_cost"] == pytest.approx(_scipy_cost_with_fixed(p, cf, pf), rel=1e-6)
    # residuals between a fixed camera and a fixed point are not part of the reduced program
    r, _, _ = po.evaluate(p, want_jac=False)
    both = (cf[p.cam_idx] == 1) & (pf[p.pt_idx] == 1)
    assert s["fixed_cost"] == pytest.approx(0.5 * float((r[both] ** 2).sum()), rel=1e-12)
    # no mask == null mask == the unmasked entry point
    a = po.solve(p, opts=po.default_options())
    b = po.solve(p, opts=po.default_options(), cam_fixed=np.zeros(p.n_cams, np.uint8), pt_fixed=np.zeros(p.n_pts, np.uint8))
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[3] == b[3]
