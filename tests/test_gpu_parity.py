"""GPU suite (-m gpu): the CUDA path, called through the C ABI of libhsfm_ba.so, against the CPU
oracle on identical seeded inputs and against the committed golden vectors.

Bars (BASELINE.json north_star): per-observation residuals and Jacobians within 1e-10 relative
(definitions in tests/util.py), final cost / RMS within 1e-6 relative after the SAME iteration count."""
import numpy as np
import pytest

from oracle import pyoracle as po
from sfm_b200 import api, synth
from tests.util import jacobian_err, kat_problem, rel, residual_err

pytestmark = pytest.mark.gpu

TOL_RJ = 1e-10
TOL_COST = 1e-6


def _with_dist(p, seed=2):
    rng = np.random.default_rng(seed)
    p.dist = np.ascontiguousarray(rng.normal(0, [0.05, 0.1], (p.n_cams, 2)))
    return p


@pytest.mark.parametrize("make", [
    lambda: synth.make_config("c1"),
    lambda: synth.make_config("c1", skew=0.5),
    lambda: _with_dist(synth.make_ring(12, 500, 7, seed=31, skew=0.2)),
    lambda: synth.make_config("c2", scale=0.1),
    lambda: synth.make_config("c3", scale=0.05),
], ids=["c1", "c1-skew", "ring-distortion", "c2-small", "c3-small"])
def test_residuals_and_jacobians(ctx, make):
    p = make()
    r, J, cost = ctx.evaluate(p)
    r0, J0, cost0 = po.evaluate(p)
    assert residual_err(r, r0, p.obs) < TOL_RJ
    assert jacobian_err(J, J0) < TOL_RJ
    assert cost == pytest.approx(cost0, rel=1e-12)
    # camera 0 sits at w = 0 exactly: Taylor branch of AngleAxisRotatePoint
    m = p.cam_idx == 0
    assert m.any() and jacobian_err(J[m], J0[m]) < TOL_RJ


def test_golden_functor_vectors(ctx, golden_dir):
    g = np.load(f"{golden_dir}/functor_kat.npz")
    p = kat_problem(g)
    r, J, _ = ctx.evaluate(p)
    assert residual_err(r, g["r"], p.obs) < 1e-12
    assert jacobian_err(J, g["J"]) < 1e-11  # mpmath, 50 digits
    ns = g["K"][:, 1] == 0.0
    assert jacobian_err(J[ns][:, :, :6], g["J_cv2"][ns]) < 1e-11  # OpenCV (no skew term there)


def test_distortion_zero_is_bit_identical(ctx):
    p = synth.make_ring(10, 300, 6, seed=8, skew=0.1)
    r0, J0, c0 = ctx.evaluate(p)
    p.dist = np.zeros((p.n_cams, 2))
    r1, J1, c1 = ctx.evaluate(p)
    assert np.array_equal(r0, r1) and np.array_equal(J0, J1) and c0 == c1


def test_observation_order_does_not_matter(ctx):
    p = synth.make_config("c3", scale=0.05)
    r0, J0, c0 = ctx.evaluate(p)
    perm = np.random.default_rng(0).permutation(p.n_obs)
    q = synth.Problem(obs=np.ascontiguousarray(p.obs[perm]), cam_idx=np.ascontiguousarray(p.cam_idx[perm]),
                      pt_idx=np.ascontiguousarray(p.pt_idx[perm]), K=p.K, cam=p.cam, pts=p.pts, cam_true=p.cam_true, pts_true=p.pts_true)
    r1, J1, c1 = ctx.evaluate(q)
    assert np.array_equal(r1, r0[perm]) and np.array_equal(J1, J0[perm])  # per-observation work is order independent
    assert c1 == pytest.approx(c0, rel=1e-13)


def test_edge_cases(ctx):
    # empty problem
    e = synth.make_ring(4, 10, 3, seed=1)
    empty = synth.Problem(obs=np.zeros((0, 2)), cam_idx=np.zeros(0, np.int32), pt_idx=np.zeros(0, np.int32), K=e.K, cam=e.cam, pts=e.pts,
                          cam_true=e.cam, pts_true=e.pts)
    r, J, c = ctx.evaluate(empty)
    assert r.shape == (0, 2) and c == 0.0
    cam, pts, s = ctx.solve_packed(empty)
    assert s["num_iterations"] == 0 and np.array_equal(cam, e.cam) and np.array_equal(pts, e.pts)
    # ragged: points with a single observation, a camera and a point without observations, tracks crossing warps/tiles
    rng = np.random.default_rng(5)
    base = synth.make_ring(9, 400, 9, seed=12)
    keep = np.ones(base.n_obs, bool)
    keep[base.pt_idx == 7] = False          # point 7: no observations
    keep[base.cam_idx == 3] = False         # camera 3: no observations
    one = np.nonzero(base.pt_idx == 11)[0]
    keep[one[1:]] = False                   # point 11: a single observation
    q = synth.Problem(obs=np.ascontiguousarray(base.obs[keep]), cam_idx=np.ascontiguousarray(base.cam_idx[keep]),
                      pt_idx=np.ascontiguousarray(base.pt_idx[keep]), K=base.K, cam=base.cam, pts=base.pts, cam_true=base.cam_true,
                      pts_true=base.pts_true)
    r, J, c = ctx.evaluate(q)
    r0, J0, c0 = po.evaluate(q)
    assert residual_err(r, r0, q.obs) < TOL_RJ and jacobian_err(J, J0) < TOL_RJ
    cam, pts, s = ctx.solve_packed(q, opts=api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR))
    cam0, pts0, so, _ = po.solve(q)
    assert s["num_iterations"] == so["num_iterations"]
    assert s["final_cost"] == pytest.approx(so["final_cost"], rel=TOL_COST)
    assert np.array_equal(pts[7], base.pts[7]) and np.array_equal(cam[3], base.cam[3])  # untouched blocks
    # long tracks: every point seen by 200 cameras (segments span many warps of a tile)
    lt = synth.make_ring(200, 40, 200, seed=13)
    r, J, _ = ctx.evaluate(lt)
    r0, J0, _ = po.evaluate(lt)
    assert jacobian_err(J, J0) < TOL_RJ
    g, o = ctx.schur(lt), po.schur(lt)
    assert rel(g["Hpp"], o["Hpp"]) < 1e-11 and rel(g["rhs"], o["rhs"]) < 1e-9


def test_errors_leave_parameters_untouched(ctx):
    p = synth.make_ring(6, 50, 4, seed=2)
    bad = p.copy()
    bad.cam_idx = p.cam_idx.copy()
    bad.cam_idx[5] = 99
    with pytest.raises(api.BAError) as ei:
        ctx.solve_packed(bad)
    assert ei.value.code == -1
    # a point observed more than 256 times is outside the tile layout: refused, not mis-solved
    big = synth.make_ring(300, 4, 300, seed=3)
    with pytest.raises(api.BAError) as ei:
        ctx.solve_packed(big)
    assert ei.value.code == -4


@pytest.mark.parametrize("make", [lambda: synth.make_ring(6, 40, 4, seed=99, skew=0.3), lambda: synth.make_config("c1"),
                                  lambda: synth.make_config("c3", scale=0.05)], ids=["kat", "c1", "c3-small"])
def test_normal_equation_blocks(ctx, make):
    p = make()
    g, o = ctx.schur(p, radius=1e4), po.schur(p, radius=1e4)
    for k, tol in [("scale", 1e-12), ("Hpp", 1e-11), ("gp", 1e-10), ("Hcc", 1e-11), ("gc", 1e-10), ("rhs", 1e-9), ("S", 1e-9)]:
        assert rel(g[k], o[k]) < tol, k
    assert np.array_equal(g["S"], g["S"].T)


def test_golden_normal_equations_lapack(ctx, golden_dir):
    g = np.load(f"{golden_dir}/normal_eq_kat.npz")
    from tests.golden.make_golden import small_problem

    p = small_problem()
    d = ctx.schur(p, radius=float(g["radius"]))
    assert rel(d["scale"], g["scale"]) < 1e-12 and rel(d["S"], g["S"]) < 1e-9 and rel(d["rhs"], g["rhs"]) < 1e-9
    cam, pts, s = ctx.solve_packed(p, opts=api.default_options(max_num_iterations=1, linear_solver=api.LINEAR_DENSE_SCHUR))
    assert s["trace"][1]["model_cost_change"] == pytest.approx(float(g["model_cost_change"]), rel=1e-8)


def _trajectory(n_cams=420, window=12, n_pts=6000, seed=5, closure=True):
    """Scan trajectory: every point is seen by cameras of a short window of consecutive keyframes; a loop closure
    ties the last keyframes to the first.  The reduced camera matrix is a band plus a few long rows — the case the
    envelope factorisation exists for (n = 2520: 20 panels of 128)."""
    p = synth.make_ring(n_cams, n_pts, 4, seed=seed, skew=0.05)
    rng = np.random.default_rng(seed)
    start = rng.integers(0, n_cams - window, n_pts)
    offs = np.sort(np.argsort(rng.random((n_pts, window)), axis=1)[:, :4], axis=1)  # four distinct keyframes of the window
    cams = start[:, None] + offs
    if closure:
        k = n_pts // 50
        cams[:k] = (n_cams - window) + offs[:k]
        cams[:k, 0] = rng.integers(0, 4, k)  # ... and one of the first keyframes
    cam_idx = cams.reshape(-1).astype(np.int32)
    pt_idx = np.repeat(np.arange(n_pts, dtype=np.int32), 4)
    obs = synth.project(p.cam_true, p.K, p.pts_true, cam_idx, pt_idx)[0] + rng.normal(0.0, 0.5, (cam_idx.size, 2))
    return synth.Problem(obs=np.ascontiguousarray(obs), cam_idx=cam_idx, pt_idx=pt_idx, K=p.K, cam=p.cam, pts=p.pts,
                         cam_true=p.cam_true, pts_true=p.pts_true, name="trajectory")


@pytest.mark.parametrize("closure", [False, True], ids=["band", "band+loop-closure"])
def test_envelope_factorisation_equals_dense(ctx, closure):
    """The envelope Cholesky skips only blocks that are exactly zero: same iterates as the full dense factorisation
    of the same matrix (dense_flags bit 0), and as the unsorted formation of S (bit 1)."""
    p = _trajectory(closure=closure)
    runs = {}
    for name, flags in [("envelope", 0), ("dense", 1), ("unsorted", 2)]:
        o = api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR, max_num_iterations=6)
        o.dense_flags = flags
        runs[name] = ctx.solve_packed(p, opts=o)
    ce, pe, se = runs["envelope"]
    for other in ("dense", "unsorted"):
        co, po_, so = runs[other]
        assert se["num_iterations"] == so["num_iterations"] and se["num_successful_steps"] == so["num_successful_steps"]
        assert se["final_cost"] == pytest.approx(so["final_cost"], rel=1e-10)
        assert np.abs(ce - co).max() < 1e-8 and np.abs(pe - po_).max() < 1e-8
    assert se["final_cost"] < 0.05 * se["initial_cost"]
    so = po.solve(p, opts=po.default_options(num_threads=4, max_num_iterations=6))[2]
    assert se["final_cost"] == pytest.approx(so["final_cost"], rel=TOL_COST)


def test_auto_uses_the_envelope_beyond_dense_max_cams(ctx):
    """AUTO: more cameras than dense_max_cams still get the (envelope) Cholesky when the reduced camera matrix is a
    narrow band; a camera graph without that structure falls back to PCG.  Both agree with the PCG solution."""
    p = _trajectory(n_cams=1680, n_pts=16000, closure=True)  # 79 panels; the envelope holds ~10 % of the triangle
    o = api.default_options(max_num_iterations=5)
    o.dense_max_cams = 200
    sa = ctx.solve_packed(p, opts=o)[2]
    assert sa["linear_solver_used"] == api.LINEAR_DENSE_SCHUR
    sp = ctx.solve_packed(p, opts=api.default_options(max_num_iterations=5, linear_solver=api.LINEAR_PCG, pcg_max_iterations=3000))[2]
    assert sa["final_cost"] == pytest.approx(sp["final_cost"], rel=TOL_COST)
    q = synth.make_ring(300, 3000, 2, seed=77, skew=0.1)  # two-view tracks over random camera pairs: S is full
    sq = ctx.solve_packed(q, opts=o)[2]
    assert sq["linear_solver_used"] == api.LINEAR_PCG


def test_tiles_with_many_distinct_cameras(ctx):
    """Two-view tracks over 300 cameras: a 256-observation tile touches well over a hundred distinct cameras, so
    the shared-memory staging of camera rows needs more than the 48 KB default (opt-in dynamic shared memory)."""
    p = synth.make_ring(300, 3000, 2, seed=77, skew=0.1)
    r, J, cost = ctx.evaluate(p)
    r0, J0, cost0 = po.evaluate(p)
    assert residual_err(r, r0, p.obs) < TOL_RJ and jacobian_err(J, J0) < TOL_RJ
    assert cost == pytest.approx(cost0, rel=1e-12)
    g, o = ctx.schur(p, radius=1e4), po.schur(p, radius=1e4)
    for k, tol in [("Hpp", 1e-11), ("gp", 1e-10), ("Hcc", 1e-11), ("gc", 1e-10), ("rhs", 1e-9), ("S", 1e-9)]:
        assert rel(g[k], o[k]) < tol, k
    sd = ctx.solve_packed(p, opts=api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR, max_num_iterations=3))[2]
    sp = ctx.solve_packed(p, opts=api.default_options(linear_solver=api.LINEAR_PCG, max_num_iterations=3, pcg_max_iterations=3000))[2]
    so = po.solve(p, opts=po.default_options(num_threads=4, max_num_iterations=3))[2]
    assert sd["final_cost"] == pytest.approx(so["final_cost"], rel=TOL_COST)
    assert sp["final_cost"] == pytest.approx(so["final_cost"], rel=TOL_COST)


@pytest.mark.parametrize("linear", [api.LINEAR_DENSE_SCHUR, api.LINEAR_PCG], ids=["dense", "pcg"])
@pytest.mark.parametrize("make", [lambda: synth.make_config("c1"), lambda: synth.make_config("c2", scale=0.1),
                                  lambda: synth.make_config("c3", scale=0.05), lambda: _with_dist(synth.make_ring(12, 500, 7, seed=31))],
                         ids=["c1", "c2-small", "c3-small", "distortion"])
def test_lm_solve_matches_oracle(ctx, make, linear):
    p = make()
    cam, pts, s = ctx.solve_packed(p, opts=api.default_options(linear_solver=linear, pcg_max_iterations=3000))
    cam0, pts0, so, tro = po.solve(p, opts=po.default_options(num_threads=4))
    assert s["num_iterations"] == so["num_iterations"] and s["num_successful_steps"] == so["num_successful_steps"]
    assert s["termination_type"] == so["termination_type"]
    assert s["final_cost"] == pytest.approx(so["final_cost"], rel=TOL_COST)
    assert s["rms_px"] == pytest.approx(so["rms_px"], rel=TOL_COST)
    for a, b in zip(s["trace"], tro):  # same trajectory, not only the same end point
        assert a["cost"] == pytest.approx(b["cost"], rel=TOL_COST)
        assert a["trust_region_radius"] == pytest.approx(b["trust_region_radius"], rel=1e-6)
        assert a["step_is_successful"] == b["step_is_successful"]
    assert np.abs(cam - cam0).max() < 1e-6 and np.abs(pts - pts0).max() < 1e-6
    assert s["gpu_launches"] > 0 and s["linear_solver_used"] == linear


@pytest.mark.parametrize("mode", [api.STRUCT_ONLY, api.POSE_ONLY])
def test_struct_only_and_pose_only(ctx, mode):
    p = synth.make_config("c1")
    cam, pts, s = ctx.solve_packed(p, mode=mode)
    cam0, pts0, so, _ = po.solve(p, mode=mode)
    assert s["num_iterations"] == so["num_iterations"]
    assert s["final_cost"] == pytest.approx(so["final_cost"], rel=TOL_COST)
    if mode == api.STRUCT_ONLY:
        assert np.array_equal(cam, p.cam) and np.abs(pts - pts0).max() < 1e-7
    else:
        assert np.array_equal(pts, p.pts) and np.abs(cam - cam0).max() < 1e-7


def test_ctracker_pointer_interface(ctx):
    """The reference call site: vector<double*> R, t, pts3D with pointer identity as point id
    (CSfM.cpp:321-343).  Storage mimics CFrame::_rot/_t and CMap::_pts3D (sparse ids)."""
    p = synth.make_config("c1")
    frames_rot = [p.cam[c, :3].copy() for c in range(p.n_cams)]
    frames_t = [p.cam[c, 3:].copy() for c in range(p.n_cams)]
    store = np.full((2 * p.n_pts + 5, 3), 123.0)  # map storage with culled (unused) slots in between
    slot = 2 * np.arange(p.n_pts) + 1
    store[slot] = p.pts
    pts3D = [store[slot[j]] for j in p.pt_idx]  # the SAME row view object/memory for every observation of a point
    K = p.K.reshape(-1, 3, 3)
    tr = api.CTracker(0)
    ret = tr.bundleAdjustmentStructAndPose(p.obs, p.cam_idx, K, frames_rot, frames_t, pts3D, api.CTracker.STRUCT_AND_POSE)
    assert ret is None
    cam0, pts0, so, _ = po.solve(p)
    assert tr.last_summary["num_iterations"] == so["num_iterations"]
    assert tr.last_summary["final_cost"] == pytest.approx(so["final_cost"], rel=TOL_COST)
    assert np.abs(np.array(frames_rot) - cam0[:, :3]).max() < 1e-6 and np.abs(np.array(frames_t) - cam0[:, 3:]).max() < 1e-6
    assert np.abs(store[slot] - pts0).max() < 1e-6
    assert np.all(store[slot - 1] == 123.0)  # culled slots untouched
    # default branch of the switch (CTracker.cpp:692-693): nothing happens
    before = store.copy()
    tr.bundleAdjustmentStructAndPose(p.obs, p.cam_idx, K, frames_rot, frames_t, pts3D, 5)
    assert np.array_equal(store, before)


def test_resident_problem_repeatable(ctx):
    p = synth.make_config("c2", scale=0.1)
    ctx.upload(p)
    c0 = ctx.eval_resident()
    assert c0 == pytest.approx(po.evaluate(p, want_jac=False)[2], rel=1e-12)
    a = ctx.solve_resident(api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR))
    cam_a, pts_a = ctx.download()
    ctx.reset()
    assert ctx.eval_resident() == c0
    b = ctx.solve_resident(api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR))
    assert a["num_iterations"] == b["num_iterations"] and b["final_cost"] == pytest.approx(a["final_cost"], rel=1e-12)
    assert ctx.algorithmic_bytes("rj") == 184 * p.n_obs + 88 * p.n_cams + 24 * p.n_pts
    assert ctx.time_kernel("rj", 3) > 0


def test_frame_error_statistics(ctx):
    """per-keyframe (rms, max) reprojection distance, before and after the adjustment, against
    the loop of CSfM.cpp:350-382 over the oracle's residuals.  1e-10 relative to the pixel scale (residuals cancel)."""
    p = synth.make_ring(14, 400, 9, seed=11, name="stats")
    ctx.upload(p)
    rms, mx = ctx.frame_stats()
    rms0, mx0 = po.frame_error_stats(p)
    assert np.abs(rms - rms0).max() <= 1e-10 * max(1.0, rms0.max()) and np.abs(mx - mx0).max() <= 1e-10 * max(1.0, mx0.max())
    ctx.solve_resident(api.default_options(linear_solver=api.LINEAR_DENSE_SCHUR))
    cam, pts = ctx.download()
    rms, mx = ctx.frame_stats()
    rms1, mx1 = po.frame_error_stats(p, cam, pts)
    assert np.abs(rms - rms1).max() <= 1e-9 and np.abs(mx - mx1).max() <= 1e-9
    assert rms.max() < rms0.min()  # the adjustment brought every keyframe down
    # a camera without observations reports 0
    import dataclasses
    keep = p.cam_idx != 3
    q = dataclasses.replace(p, obs=np.ascontiguousarray(p.obs[keep]), cam_idx=np.ascontiguousarray(p.cam_idx[keep]),
                            pt_idx=np.ascontiguousarray(p.pt_idx[keep]), name="nocam3")
    ctx.upload(q)
    rms, mx = ctx.frame_stats()
    assert rms[3] == 0.0 and mx[3] == 0.0 and (rms[np.arange(p.n_cams) != 3] > 0).all()
    rms2, mx2 = po.frame_error_stats(q)
    assert np.abs(rms - rms2).max() <= 1e-10 * rms2.max() and np.abs(mx - mx2).max() <= 1e-10 * mx2.max()


def test_full_size_properties_c4(ctx):
    """BASELINE configs[3] at full size (1778 cameras x 994k points x 5M observations): properties that
    need no CPU solve — consistency of the device-resident loop with an independent evaluation,
    shard additivity (checksum of checksums), and agreement of the two linear solvers."""
    p = synth.make_config("c4")
    ctx.upload(p)
    c_init = ctx.eval_resident()
    s = ctx.solve_resident(api.default_options(linear_solver=api.LINEAR_PCG, max_num_iterations=3, pcg_tolerance=1e-3, pcg_max_iterations=200))
    assert s["initial_cost"] == pytest.approx(c_init, rel=1e-12)
    assert s["final_cost"] < 1e-3 * s["initial_cost"]
    assert ctx.eval_resident() == pytest.approx(s["final_cost"], rel=1e-12)  # the loop's cost == a fresh evaluation at x
    cam, pts = ctx.download()
    # shard additivity of the cost at the solution
    shards = synth.shard_by_point(p, 3)
    total = 0.0
    for sh, gids in shards:
        sh.cam, sh.pts = cam, np.ascontiguousarray(pts[gids])
        ctx.upload(sh)
        total += ctx.eval_resident()
    assert total == pytest.approx(s["final_cost"], rel=1e-12)
