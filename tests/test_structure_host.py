"""Host side of the structured Cholesky (sfm_b200/csrc/ba_order.cpp + CholPlan::compute), no GPU needed: the camera
order chosen for a banded camera graph, the closed block structure, and the two-chain schedule — checked against a
dense numpy factorisation and against a block-level emulation of the schedule the CUDA launcher follows."""
import numpy as np
import pytest

from sfm_b200 import api

NB = 128


def band_first(n, lo, hi, seed, closures=()):
    rng = np.random.default_rng(seed)
    first = np.maximum(0, np.arange(n) - rng.integers(lo, hi, n)).astype(np.int32)
    for c, f in closures:
        first[c] = f
    return first


def random_spd(first, seed):
    """Symmetric positive definite matrix of order 6 n with the camera-level pattern of the envelope (caller order)."""
    rng = np.random.default_rng(seed)
    n = len(first)
    A = np.zeros((6 * n, 6 * n))
    for c in range(n):
        f = int(first[c])
        A[6 * c:6 * c + 6, 6 * f:6 * c + 6] = rng.normal(size=(6, 6 * (c - f + 1)))
    A = np.tril(A)
    A = A + A.T
    A += np.eye(6 * n) * (np.abs(A).sum(axis=1).max() + 1.0)
    return A


def permuted(A, new_of_old):
    n = len(new_of_old)
    idx = np.empty(6 * n, dtype=np.int64)
    for c in range(n):
        idx[6 * new_of_old[c]:6 * new_of_old[c] + 6] = np.arange(6 * c, 6 * c + 6)
    return A[np.ix_(idx, idx)]


def emulate(A, st, split_a, split_t0):
    """Right-looking block factorisation over the structure, as launch_cholesky_struct schedules it: chain A = panels
    [0, a) whose updates into block columns >= t0 go to a side matrix T, chain B = [a, t0) updating S directly, then
    S_tail += T and the tail [t0, P).  Only structural blocks are ever touched."""
    n = A.shape[0]
    P = st.shape[0]
    S = np.tril(A).copy()
    T = np.zeros_like(S)
    blk = lambda I: slice(I * NB, min(n, (I + 1) * NB))  # noqa: E731
    col = [[I for I in range(K + 1, P) if st[I, K]] for K in range(P)]

    def panel(K, own_end, side):
        S[blk(K), blk(K)] = np.linalg.cholesky(S[blk(K), blk(K)] + np.tril(S[blk(K), blk(K)], -1).T)
        Linv = np.linalg.inv(S[blk(K), blk(K)])
        for I in col[K]:
            S[blk(I), blk(K)] = S[blk(I), blk(K)] @ Linv.T
        for bi, J in enumerate(col[K]):
            C = side if J >= own_end else S
            for I in col[K][bi:]:
                assert st[I, J], "update outside the structure"
                C[blk(I), blk(J)] -= S[blk(I), blk(K)] @ S[blk(J), blk(K)].T

    if split_a > 0:
        for K in range(split_a):
            assert not [I for I in col[K] if split_a <= I < split_t0], "chain A touches chain B's private panels"
            panel(K, split_t0, T)
        for K in range(split_a, split_t0):
            panel(K, P, None)
        S += T
        rest = range(split_t0, P)
    else:
        rest = range(P)
    for K in rest:
        panel(K, P, None)
    return np.tril(S)


@pytest.mark.parametrize("case", ["band", "closure", "ring"])
def test_two_ended_order_structure_and_schedule(case):
    n = 448
    closures = {"band": (), "closure": ((350, 30),), "ring": tuple((c, 0) for c in range(n - 20, n))}[case]
    first = band_first(n, 15, 30, seed=5, closures=closures)
    d = api.debug_structure(first)
    nat = api.debug_structure(first, natural=True)
    assert sorted(d["new_of_old"]) == list(range(n))
    assert nat["structured"] and not nat["reordered"] and nat["split_a"] == -1
    assert d["reordered"] and d["structured"]
    a, t0, P = d["split_a"], d["split_t0"], d["panels"]
    assert 0 < a < t0 < P
    # the point of the exercise: the serial chain (in panels) is much shorter than one chain over all panels ...
    crit = max(a, t0 - a) + (P - t0)
    assert crit <= 0.75 * P, (a, t0, P)
    # ... at (nearly) no extra arithmetic: the structure holds about as many blocks as the envelope of the natural order
    assert d["structure"].sum() <= 1.2 * nat["structure"].sum()
    A = random_spd(first, seed=9)
    Ap = permuted(A, d["new_of_old"])
    L = np.linalg.cholesky(Ap)
    # nothing of L lies outside the structure (exact zeros: no fill is ever computed there)
    mask = np.kron(d["structure"], np.ones((NB, NB)))[: 6 * n, : 6 * n].astype(bool)
    assert np.abs(L[~mask]).max() == 0.0
    # the schedule (side-matrix tail updates included) reproduces the factor
    Le = emulate(Ap, d["structure"], a, t0)
    assert np.abs(Le - L).max() <= 1e-10 * np.abs(L).max()
    # and so does the single chain in the natural order
    Ln = emulate(A, nat["structure"], -1, -1)
    assert np.abs(Ln - np.linalg.cholesky(A)).max() <= 1e-10 * np.abs(L).max()


def test_small_or_dense_graphs_keep_the_callers_order():
    assert not api.debug_structure(np.zeros(200, np.int32))["reordered"]            # everything coupled: dense
    assert not api.debug_structure(band_first(100, 5, 10, 1))["reordered"]          # too few cameras for two chains
    d = api.debug_structure(np.zeros(300, np.int32))
    assert not d["structured"] and d["fill"] == 1.0
