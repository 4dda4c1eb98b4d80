"""oracle/_ref: the reference's OWN functor and problem-assembly text (CTracker.h:64-67,98-151 and
CTracker.cpp:571-702, extracted verbatim by oracle/make_ref.py and compiled against the stand-in Ceres /
OpenCV headers) against the oracle's restatement — bit for bit.  The extracted text never enters the
repository; on a machine without /root/reference and without a prebuilt oracle/_ref these tests skip."""
import numpy as np
import pytest

from oracle import pyoracle as po
from sfm_b200 import synth
from tests.util import kat_problem

pytestmark = pytest.mark.skipif(po.ref_lib() is None, reason="oracle/_ref not built and /root/reference absent")


def _kat(golden_dir):
    g = np.load(f"{golden_dir}/functor_kat.npz")
    return g, kat_problem(g)


def test_struct_and_pose_functor_is_the_oracles(golden_dir):
    g, p = _kat(golden_dir)
    r0, J0, _ = po.evaluate(p)
    r1, J1 = po.ref_evaluate(2, p.obs, p.K, p.cam[:, :3], p.cam[:, 3:], p.pts)
    assert np.array_equal(r0, r1)      # same operations in the same order: bit-identical, not merely close
    assert np.array_equal(J0, J1)
    # and through it the reference text itself meets the independent 50-digit values
    assert np.abs(r1 - g["r"]).max() < 1e-9


def test_pose_and_struct_functors_are_the_oracles(golden_dir):
    _, p = _kat(golden_dir)
    _, J0, _ = po.evaluate(p)
    r0, _, _ = po.evaluate(p, want_jac=False)
    r1, J1 = po.ref_evaluate(1, p.obs, p.K, p.cam[:, :3], p.cam[:, 3:], p.pts)  # BAPoseFunctor: point constant
    assert np.array_equal(J0[:, :, :6], J1[:, :, :6]) and not J1[:, :, 6:].any()
    r2, J2 = po.ref_evaluate(0, p.obs, p.K, p.cam[:, :3], p.cam[:, 3:], p.pts)  # BAStructFunctor: pose constant
    assert np.array_equal(J0[:, :, 6:], J2[:, :, 6:]) and not J2[:, :, :6].any()
    r_j, _, _ = po.evaluate(p)
    assert np.array_equal(r_j, r1) and np.array_equal(r_j, r2)
    # the plain-double instantiation (what Ceres evaluates for a candidate point): also bit-identical.  It is NOT
    # bit-identical to the value part of the Jet pass (Jet division multiplies by the reciprocal, jet.h) — both
    # instantiations are therefore pinned separately.
    for which in (2, 1, 0):
        rd, _ = po.ref_evaluate(which, p.obs, p.K, p.cam[:, :3], p.cam[:, 3:], p.pts, want_jac=False)
        assert np.array_equal(rd, r0)


@pytest.mark.parametrize("mode", [2, 1, 0])
def test_reference_problem_assembly_runs_the_same_trajectory(mode):
    """CTracker::bundleAdjustmentStructAndPose verbatim (AddResidualBlock per observation, pointer identity) with the
    stand-in ceres::Solve == the oracle on the packed arrays: same parameter blocks, same LM trajectory, same result."""
    p = synth.make_ring(9, 120, 5, seed=11, name="ref-assembly")
    o = po.default_options(max_num_iterations=6)
    c0, x0, s0, t0 = po.solve(p, mode=mode, opts=o)
    c1, x1, s1, t1 = po.ref_solve(p, mode=mode, opts=o)
    want_blocks = (2 * p.n_cams if mode != 0 else 0) + (p.n_pts if mode != 1 else 0)
    assert s1["n_parameter_blocks"] == want_blocks
    assert s1["num_iterations"] == s0["num_iterations"] and s1["num_successful_steps"] == s0["num_successful_steps"]
    assert [it["cost"] for it in t1] == [it["cost"] for it in t0]
    assert [it["trust_region_radius"] for it in t1] == [it["trust_region_radius"] for it in t0]
    assert np.array_equal(c0, c1) and np.array_equal(x0, x1)


def test_unknown_mode_adds_nothing():
    # CTracker.cpp:692-693: default branch adds no residual block; Solve on an empty problem changes nothing
    p = synth.make_ring(5, 30, 3, seed=3)
    c1, x1, s1, _ = po.ref_solve(p, mode=7)
    assert np.array_equal(c1, p.cam) and np.array_equal(x1, p.pts) and s1["num_iterations"] == 0
