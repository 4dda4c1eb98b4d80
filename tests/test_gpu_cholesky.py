"""GPU suite: the blocked DMMA Cholesky + triangular solves (sfm_b200/csrc/ba_chol.cu) through the
hsfm_ba_dense_spd_solve tap, against numpy/LAPACK.  Sizes cover one partial panel, exact panel
multiples, ragged tails and the BASELINE reduced-system sizes of C1 (180) and C2 (1200)."""
import numpy as np
import pytest

from sfm_b200 import api

pytestmark = pytest.mark.gpu


def spd(n, seed, cond_boost=0.0):
    rng = np.random.default_rng(seed)
    A = rng.normal(size=(n, n))
    A = A @ A.T / n + (1.0 + cond_boost) * np.eye(n)
    return A


@pytest.mark.parametrize("n", [2, 6, 52, 128, 130, 180, 256, 300, 1200, 2050])
def test_factor_and_solve_match_lapack(ctx, n):
    A = spd(n, n)
    b = np.random.default_rng(1).normal(size=n)
    x, L = ctx.dense_spd_solve(A, b, want_factor=True)
    L0 = np.linalg.cholesky(A)
    assert np.abs(L - L0).max() / np.abs(L0).max() < 1e-12
    x0 = np.linalg.solve(A, b)
    assert np.abs(x - x0).max() / np.abs(x0).max() < 1e-10
    assert np.abs(A @ x - b).max() < 1e-10 * max(1.0, np.abs(b).max()) * n


def test_ill_conditioned_reduced_system_shape(ctx):
    # graded spectrum like a weakly damped reduced camera matrix (cond ~ 1e10)
    rng = np.random.default_rng(3)
    n = 600
    Q, _ = np.linalg.qr(rng.normal(size=(n, n)))
    A = (Q * np.logspace(0, -10, n)) @ Q.T
    A = 0.5 * (A + A.T)
    b = A @ rng.normal(size=n)
    x = ctx.dense_spd_solve(A, b)
    assert np.abs(A @ x - b).max() / np.abs(b).max() < 1e-8


def test_not_positive_definite_is_reported(ctx):
    A = spd(200, 5)
    A[150, 150] = -1.0
    with pytest.raises(api.BAError) as ei:
        ctx.dense_spd_solve(A, np.ones(200))
    assert ei.value.code == -1 and 1 <= ei.value.pivot <= 151
    with pytest.raises(api.BAError):
        ctx.dense_spd_solve(np.eye(3), np.ones(3))  # odd n is outside the contract (6*n_cams is even)


@pytest.mark.parametrize("where", [(3, 3), (140, 7), (399, 399)])
def test_non_finite_input_is_reported_not_propagated(ctx, where):
    """NaN anywhere in the lower triangle fails a pivot test (NaN > 0 is false) or the finiteness test of the panel
    operands; the call returns an error instead of a factor full of NaN."""
    A = spd(400, 9)
    A[where] = np.nan
    A[where[::-1]] = np.nan
    with pytest.raises(api.BAError) as ei:
        ctx.dense_spd_solve(A, np.ones(400))
    assert ei.value.code == -1
