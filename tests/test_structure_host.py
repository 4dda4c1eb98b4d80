"""Host side of the structured Cholesky (sfm_b200/csrc/ba_order.cpp + CholPlan::compute), no GPU needed: the camera
order chosen for a banded camera graph, the closed block structure, and the segment-tree schedule — checked against a
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


def emulate(A, st, segments):
    """Right-looking block factorisation over the structure, as launch_cholesky_struct schedules it: every segment
    (panel chain) factors its panels; updates to blocks of its ancestors' columns go to the segment's own side matrix
    (indexed by its ancestors' panels only), which the parent folds into S (its own columns) and into its own side
    matrix (the rest) before it starts.  Concurrent segments are run here in an order that would expose a missing
    dependency: all leaves first, interleaved panel by panel.  Only structural blocks are ever touched."""
    n = A.shape[0]
    P = st.shape[0]
    S = np.tril(A).copy()
    blk = lambda I: slice(I * NB, min(n, (I + 1) * NB))  # noqa: E731
    col = [[I for I in range(K + 1, P) if st[I, K]] for K in range(P)]
    if not segments:
        segments = [(0, P, -1)]
    G = len(segments)
    anc = []
    for g, (b, e, par) in enumerate(segments):
        a = []
        while par >= 0:
            assert par > g, "parents are numbered after their children"
            a += list(range(segments[par][0], segments[par][1]))
            par = segments[par][2]
        anc.append(sorted(a))
    side = [dict() for _ in range(G)]  # (I, J) -> block, only ancestors' panels are legal keys

    def side_blk(g, I, J):
        assert I in anc[g] and J in anc[g], "side-matrix update outside the segment's ancestors"
        return side[g].setdefault((I, J), np.zeros((NB, NB)))

    def panel(g, K):
        b, e, par = segments[g]
        own_end = e if par >= 0 else P
        S[blk(K), blk(K)] = np.linalg.cholesky(S[blk(K), blk(K)] + np.tril(S[blk(K), blk(K)], -1).T)
        Linv = np.linalg.inv(S[blk(K), blk(K)])
        for I in col[K]:
            assert I < e or I in anc[g], "a panel couples to a panel that is neither its own segment's nor an ancestor's"
            S[blk(I), blk(K)] = S[blk(I), blk(K)] @ Linv.T
        for bi, J in enumerate(col[K]):
            for I in col[K][bi:]:
                assert st[I, J], "update outside the structure"
                upd = S[blk(I), blk(K)] @ S[blk(J), blk(K)].T
                if J >= own_end:
                    t = side_blk(g, I, J)
                    t[: upd.shape[0], : upd.shape[1]] -= upd
                else:
                    S[blk(I), blk(J)] -= upd

    def join(g):
        b, e, par = segments[g]
        for c in [c for c in range(G) if segments[c][2] == g]:
            for (I, J), t in side[c].items():
                if b <= J < e:
                    r, cc = S[blk(I), blk(J)].shape
                    S[blk(I), blk(J)] += t[:r, :cc]
                else:
                    side_blk(g, I, J)[...] += t

    children = [[c for c in range(G) if segments[c][2] == g] for g in range(G)]
    done = [False] * G
    while not all(done):
        ready = [g for g in range(G) if not done[g] and all(done[c] for c in children[g])]
        for g in ready:
            join(g)
        cursors = {g: segments[g][0] for g in ready}
        while cursors:  # interleave the concurrent segments panel by panel
            for g in list(cursors):
                panel(g, cursors[g])
                cursors[g] += 1
                if cursors[g] == segments[g][1]:
                    del cursors[g]
                    done[g] = True
    return np.tril(S)


def check_segments(segments, P):
    assert segments[0][0] == 0 and segments[-1][1] == P
    for (b0, e0, _), (b1, _, _) in zip(segments, segments[1:]):
        assert b0 < e0 == b1
    assert sum(1 for _, _, par in segments if par < 0) == 1  # one root: the last separator


@pytest.mark.parametrize("case", ["band", "closure", "ring"])
def test_two_ended_order_structure_and_schedule(case):
    n = 448
    closures = {"band": (), "closure": ((350, 30),), "ring": tuple((c, 0) for c in range(n - 20, n))}[case]
    first = band_first(n, 15, 30, seed=5, closures=closures)
    d = api.debug_structure(first)
    nat = api.debug_structure(first, natural=True)
    assert sorted(d["new_of_old"]) == list(range(n))
    assert nat["structured"] and not nat["reordered"] and nat["split_a"] == -1 and not nat["segments"]
    assert d["reordered"] and d["structured"]
    a, t0, P = d["split_a"], d["split_t0"], d["panels"]
    assert 0 < a < t0 < P
    assert d["segments"] == [(0, a, 2), (a, t0, 2), (t0, P, -1)]
    # the point of the exercise: the serial chain (in panels) is much shorter than one chain over all panels ...
    crit = max(a, t0 - a) + (P - t0)
    assert crit == d["critical_panels"] and crit <= 0.75 * P, (a, t0, P)
    # ... at (nearly) no extra arithmetic: the structure holds about as many blocks as the envelope of the natural order
    assert d["structure"].sum() <= 1.2 * nat["structure"].sum()
    A = random_spd(first, seed=9)
    Ap = permuted(A, d["new_of_old"])
    L = np.linalg.cholesky(Ap)
    # nothing of L lies outside the structure (exact zeros: no fill is ever computed there)
    mask = np.kron(d["structure"], np.ones((NB, NB)))[: 6 * n, : 6 * n].astype(bool)
    assert np.abs(L[~mask]).max() == 0.0
    # the schedule (side-matrix updates included) reproduces the factor
    Le = emulate(Ap, d["structure"], d["segments"])
    assert np.abs(Le - L).max() <= 1e-10 * np.abs(L).max()
    # and so does the single chain in the natural order
    Ln = emulate(A, nat["structure"], [])
    assert np.abs(Ln - np.linalg.cholesky(A)).max() <= 1e-10 * np.abs(L).max()


@pytest.mark.parametrize("case", ["band", "closures"])
def test_nested_order_runs_four_or_more_chains(case):
    """The band is cut in halves around a separator and each half eliminated from both ends — four concurrent chains,
    inner separators, one root separator (ba_order.cpp: order_band).  The depth is forced: at this size the cost model
    prefers two chains."""
    n = 840
    closures = {"band": (), "closures": ((800, 3), (801, 3), (500, 100))}[case]
    first = band_first(n, 20, 45, seed=11, closures=closures)
    d = api.debug_structure(first, depth=1)
    P, segs = d["panels"], d["segments"]
    assert d["reordered"] and d["structured"] and sorted(d["new_of_old"]) == list(range(n))
    check_segments(segs, P)
    leaves = [g for g in range(len(segs)) if all(par != g for _, _, par in segs)]
    assert len(leaves) >= 4 and len(segs) >= 6, segs
    assert all(par > g or par == -1 for g, (_, _, par) in enumerate(segs))
    assert d["critical_panels"] <= 0.55 * P, (d["critical_panels"], P)
    A = random_spd(first, seed=3)
    Ap = permuted(A, d["new_of_old"])
    L = np.linalg.cholesky(Ap)
    mask = np.kron(d["structure"], np.ones((NB, NB)))[: 6 * n, : 6 * n].astype(bool)
    assert np.abs(L[~mask]).max() == 0.0
    Le = emulate(Ap, d["structure"], segs)
    assert np.abs(Le - L).max() <= 1e-10 * np.abs(L).max()


def test_small_or_dense_graphs_keep_the_callers_order():
    assert not api.debug_structure(np.zeros(200, np.int32))["reordered"]            # everything coupled: dense
    assert not api.debug_structure(band_first(100, 5, 10, 1))["reordered"]          # too few cameras for two chains
    d = api.debug_structure(np.zeros(300, np.int32))
    assert not d["structured"] and d["fill"] == 1.0


def test_depth_two_order_runs_eight_chains():
    """Two levels of dissection: eight leaf chains, four inner separators, two middle ones, one root — the general
    segment tree (side matrices forwarded through two levels of joins)."""
    n = 1300
    first = band_first(n, 10, 22, seed=17, closures=((1250, 5),))
    d = api.debug_structure(first, depth=2)
    P, segs = d["panels"], d["segments"]
    assert d["reordered"] and d["structured"] and sorted(d["new_of_old"]) == list(range(n))
    check_segments(segs, P)
    leaves = [g for g in range(len(segs)) if all(par != g for _, _, par in segs)]
    assert len(leaves) >= 8, segs
    depth_of = lambda g: 0 if segs[g][2] < 0 else 1 + depth_of(segs[g][2])  # noqa: E731
    assert max(depth_of(g) for g in range(len(segs))) >= 3
    assert d["critical_panels"] <= 0.45 * P, (d["critical_panels"], P)
    A = random_spd(first, seed=4)
    Ap = permuted(A, d["new_of_old"])
    L = np.linalg.cholesky(Ap)
    mask = np.kron(d["structure"], np.ones((NB, NB)))[: 6 * n, : 6 * n].astype(bool)
    assert np.abs(L[~mask]).max() == 0.0
    Le = emulate(Ap, d["structure"], segs)
    assert np.abs(Le - L).max() <= 1e-10 * np.abs(L).max()


@pytest.mark.parametrize("case", ["band", "band+closures", "two components"])
def test_reverse_cuthill_mckee_finds_the_band_of_a_shuffled_scan(case):
    """Cameras of a scan delivered in random order: rcm_order (ba_order.cpp) numbers them so that the graph is a band again
    (the fallback of the upload when the delivered order has a full envelope)."""
    rng = np.random.default_rng(7)
    n, bw = 900, 25
    edges = [(c, d) for c in range(n) for d in range(max(0, c - int(rng.integers(5, bw))), c)]
    if case == "band+closures":
        edges += [(800, 3), (801, 4), (450, 20)]
    if case == "two components":
        edges = [(a, b) for a, b in edges if not (a >= 500 > b)]
    shuffle = rng.permutation(n)
    e = shuffle[np.array(edges)]
    pos = api.debug_rcm(n, e)
    assert sorted(pos) == list(range(n))
    width = np.abs(pos[e[:, 0]] - pos[e[:, 1]])
    assert np.abs(shuffle[np.array(edges)][:, 0] - shuffle[np.array(edges)][:, 1]).mean() > 200  # delivered order: no band at all
    if case == "band+closures":
        assert np.sort(width)[:-3].max() <= 3 * bw  # the closures stay long edges; the rest is a band
    else:
        assert width.max() <= 2 * bw
