"""CPU suite: the PRODUCT's per-observation arithmetic (sfm_b200/csrc/ba_math.cuh is
__host__ __device__) compiled with g++ into a tiny harness and compared with the oracle.  This
checks the analytic Jacobian's algebra without a GPU; the -m gpu suite repeats it through the kernels."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import pyoracle as po
from sfm_b200 import synth
from tests.util import jacobian_err, kat_problem, residual_err

HERE = os.path.dirname(os.path.abspath(__file__))
_dp, _ip = C.POINTER(C.c_double), C.POINTER(C.c_int32)


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("h") / "libmath_host.so")
    subprocess.run(["/usr/bin/g++", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(HERE, "harness", "math_host.cpp")], check=True)
    return C.CDLL(so)


def run(lib, p, scale_c=None, scale_p=None):
    n = p.n_obs
    r, J, ro = np.empty((n, 2)), np.empty((n, 2, 9)), np.empty((n, 2))
    d = lambda a: None if a is None else a.ctypes.data_as(_dp)  # noqa: E731
    lib.math_host_eval(C.c_int64(n), d(p.obs), p.cam_idx.ctypes.data_as(_ip), p.pt_idx.ctypes.data_as(_ip), C.c_int32(p.n_cams), d(p.K),
                       d(p.dist), d(p.cam), d(p.pts), d(scale_c), d(scale_p), d(r), d(J), d(ro))
    return r, J, ro


@pytest.mark.parametrize("name,kw", [("c1", {}), ("c1", {"skew": 0.5}), ("c3", {"scale": 0.02})])
def test_analytic_jacobian_equals_autodiff(harness, name, kw):
    p = synth.make_config(name, **kw)
    r, J, ro = run(harness, p)
    r0, J0, _ = po.evaluate(p)
    assert residual_err(r, r0, p.obs) < 1e-13 and residual_err(ro, r0, p.obs) < 1e-13
    assert jacobian_err(J, J0) < 1e-12


def test_distortion_and_scaling(harness):
    p = synth.make_ring(9, 120, 5, seed=11, skew=0.2)
    rng = np.random.default_rng(2)
    p.dist = np.ascontiguousarray(rng.normal(0, [0.05, 0.1], (p.n_cams, 2)))
    r, J, _ = run(harness, p)
    r0, J0, _ = po.evaluate(p)
    assert residual_err(r, r0, p.obs) < 1e-13 and jacobian_err(J, J0) < 1e-12
    sc, sp = rng.uniform(0.1, 1, (p.n_cams, 6)), rng.uniform(0.1, 1, (p.n_pts, 3))
    _, Js, _ = run(harness, p, sc, sp)
    expect = J0.copy()
    expect[:, :, :6] *= sc[p.cam_idx][:, None, :]
    expect[:, :, 6:] *= sp[p.pt_idx][:, None, :]
    assert jacobian_err(Js, expect) < 1e-12
    # k1 = k2 = 0 must reproduce the undistorted functor bit for bit
    q = synth.make_ring(9, 120, 5, seed=11, skew=0.2)
    ra, Ja, _ = run(harness, q)
    q.dist = np.zeros((q.n_cams, 2))
    rb, Jb, _ = run(harness, q)
    assert np.array_equal(ra, rb) and np.array_equal(Ja, Jb)


def test_golden_functor(harness, golden_dir):
    g = np.load(f"{golden_dir}/functor_kat.npz")
    p = kat_problem(g)
    r, J, _ = run(harness, p)
    assert residual_err(r, g["r"], p.obs) < 1e-12 and jacobian_err(J, g["J"]) < 1e-11


def test_block_inverses(harness):
    rng = np.random.default_rng(4)
    for _ in range(20):
        A = rng.normal(size=(3, 3)); A = A @ A.T + 0.1 * np.eye(3)
        B = rng.normal(size=(6, 6)); B = B @ B.T + 0.1 * np.eye(6)
        H6 = np.array([A[0, 0], A[0, 1], A[0, 2], A[1, 1], A[1, 2], A[2, 2]])
        Hi6 = np.zeros(6); B36 = B.reshape(-1).copy()
        rc = harness.math_host_inv(H6.ctypes.data_as(_dp), Hi6.ctypes.data_as(_dp), B36.ctypes.data_as(_dp))
        assert rc == 3
        Ai = np.linalg.inv(A)
        assert np.allclose(Hi6, [Ai[0, 0], Ai[0, 1], Ai[0, 2], Ai[1, 1], Ai[1, 2], Ai[2, 2]], rtol=1e-9)
        assert np.allclose(B36.reshape(6, 6), np.linalg.inv(B), rtol=1e-8)
