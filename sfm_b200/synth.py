"""Seeded synthetic bundle-adjustment problems in the argument layout that
``CSfM::bundleAdjustment`` builds for ``CTracker::bundleAdjustmentStructAndPose``
(reference CSfM.cpp:310-341): camera-major observation order, ``cam_idx`` = position in the
frame list, undistorted pinhole intrinsics ``K`` row-major 3x3 (iPhone-6s values of
main/main.cpp:47-51), angle-axis rotation + translation per camera, 3 doubles per point.

Harness code (numpy only): feeds the identical arrays to the CUDA path and to the CPU oracle.
"""
from __future__ import annotations

import dataclasses
import numpy as np

# main/main.cpp:47-51
FX, FY = 1072.606693272117800, 1067.197515608619600
CX, CY = 648.780750477178910, 364.503435962496890
WIDTH, HEIGHT = 1280, 720
SEED0 = 20170519

CONFIGS = {
    # name: (n_cams, n_pts, n_obs, kind)
    "c1": (30, 2_000, 40_000, "ring"),
    "c2": (200, 50_000, 600_000, "ring"),
    "c3": (1_723, 156_502, 678_718, "path"),
    "c4": (1_778, 993_923, 5_001_946, "path"),
    "c5": (13_682, 4_456_117, 28_987_644, "path"),
}
CONFIG_TITLES = {
    "c1": "synthetic small-object scan: 30 keyframes x 2k points (~40k observations)",
    "c2": "synthetic 200 cameras x 50k points (~600k observations)",
    "c3": "synthetic BAL Ladybug-shape: 1723 cameras x 157k points x 680k observations",
    "c4": "synthetic BAL Venice-shape: 1778 cameras x 994k points x 5M observations",
    "c5": "synthetic BAL Final-shape: 13682 cameras x 4.46M points x 29M observations",
}


@dataclasses.dataclass
class Problem:
    """Packed problem.  All float arrays are C-contiguous float64, indices int32."""

    obs: np.ndarray        # [n_obs, 2] observed pixel
    cam_idx: np.ndarray    # [n_obs]
    pt_idx: np.ndarray     # [n_obs]
    K: np.ndarray          # [n_cams, 9] row-major 3x3
    cam: np.ndarray        # [n_cams, 6] initial guess (angle-axis, t)
    pts: np.ndarray        # [n_pts, 3]  initial guess
    cam_true: np.ndarray
    pts_true: np.ndarray
    dist: np.ndarray | None = None  # [n_cams, 2] radial k1,k2 (extension) or None
    name: str = ""

    @property
    def n_obs(self) -> int:
        return int(self.obs.shape[0])

    @property
    def n_cams(self) -> int:
        return int(self.cam.shape[0])

    @property
    def n_pts(self) -> int:
        return int(self.pts.shape[0])

    def copy(self) -> "Problem":
        return dataclasses.replace(self, cam=self.cam.copy(), pts=self.pts.copy())


# ------------------------------------------------------------------------------------------
def rotate(w: np.ndarray, X: np.ndarray) -> np.ndarray:
    """Rodrigues rotation of X[n,3] by angle-axis w[n,3] (vectorised, generator-side only)."""
    th2 = np.einsum("ij,ij->i", w, w)
    th = np.sqrt(th2)
    small = th2 <= np.finfo(np.float64).eps
    ths = np.where(small, 1.0, th)
    u = w / ths[:, None]
    c, s = np.cos(th)[:, None], np.sin(th)[:, None]
    out = X * c + np.cross(u, X) * s + u * (np.einsum("ij,ij->i", u, X)[:, None] * (1.0 - c))
    lin = X + np.cross(w, X)
    return np.where(small[:, None], lin, out)


def project(cam: np.ndarray, K: np.ndarray, pts: np.ndarray, cam_idx, pt_idx, dist=None):
    """Pixels + depth of every (camera, point) pair; pinhole with skew, +z forward."""
    w, t = cam[cam_idx, :3], cam[cam_idx, 3:]
    p = rotate(w, pts[pt_idx]) + t
    xp, yp = p[:, 0] / p[:, 2], p[:, 1] / p[:, 2]
    if dist is not None:
        r2 = xp * xp + yp * yp
        f = 1.0 + dist[cam_idx, 0] * r2 + dist[cam_idx, 1] * r2 * r2
        xp, yp = xp * f, yp * f
    k = K[cam_idx]
    u = k[:, 0] * xp + k[:, 1] * yp + k[:, 2]
    v = k[:, 4] * yp + k[:, 5]
    return np.stack([u, v], axis=1), p[:, 2]


def _rotvec_from_matrix(R: np.ndarray) -> np.ndarray:
    from scipy.spatial.transform import Rotation

    return Rotation.from_matrix(R).as_rotvec()


def _ry(phi: np.ndarray) -> np.ndarray:
    c, s = np.cos(phi), np.sin(phi)
    R = np.zeros((phi.size, 3, 3))
    R[:, 0, 0], R[:, 0, 2], R[:, 1, 1], R[:, 2, 0], R[:, 2, 2] = c, s, 1.0, -s, c
    return R


def _small_rot(rng, n, sigma):
    from scipy.spatial.transform import Rotation

    return Rotation.from_rotvec(rng.normal(0.0, sigma, (n, 3))).as_matrix()


def _intrinsics(n_cams: int, skew: float) -> np.ndarray:
    K = np.zeros((n_cams, 9))
    K[:, 0], K[:, 1], K[:, 2], K[:, 4], K[:, 5], K[:, 8] = FX, skew, CX, FY, CY, 1.0
    return K


def _finish(name, rng, K, cam_true, pts_true, cam_idx, pt_idx, dist, noise_px, sigma_w, sigma_t, sigma_x):
    # camera-major delivery order, as CSfM.cpp:321-341 produces it
    order = np.lexsort((pt_idx, cam_idx))
    cam_idx = np.ascontiguousarray(cam_idx[order], dtype=np.int32)
    pt_idx = np.ascontiguousarray(pt_idx[order], dtype=np.int32)
    uv, _ = project(cam_true, K, pts_true, cam_idx, pt_idx, dist)
    obs = np.ascontiguousarray(uv + rng.normal(0.0, noise_px, uv.shape))
    cam = cam_true.copy()
    cam[:, :3] += rng.normal(0.0, sigma_w, (cam.shape[0], 3))
    cam[:, 3:] += rng.normal(0.0, sigma_t, (cam.shape[0], 3))
    cam[0, :3] = 0.0  # keyframe 0: _rot = (0,0,0) exactly (CFrame.cpp:229-235, CSfM.cpp:813)
    pts = pts_true + rng.normal(0.0, sigma_x, pts_true.shape)
    return Problem(obs=obs, cam_idx=cam_idx, pt_idx=pt_idx, K=np.ascontiguousarray(K), cam=np.ascontiguousarray(cam),
                   pts=np.ascontiguousarray(pts), cam_true=cam_true, pts_true=pts_true, dist=dist, name=name)


def make_ring(n_cams: int, n_pts: int, track: int, seed: int, *, skew: float = 0.0, dist=None, noise_px: float = 0.5,
              sigma_w: float = 0.01, sigma_t: float = 0.02, sigma_x: float = 0.02, name: str = "ring") -> Problem:
    """Small-object scan: cameras on a ring of radius 3 around the object, camera 0 = identity pose
    exactly, points uniform in a unit cube at the ring centre, every point seen by ``track`` distinct
    cameras."""
    rng = np.random.Generator(np.random.Philox(seed))
    centre = np.array([0.0, 0.0, 3.0])
    phi = 2.0 * np.pi * np.arange(n_cams) / n_cams
    phi = np.where(phi > np.pi, phi - 2.0 * np.pi, phi)
    Rc2w = _ry(phi)
    tilt = _small_rot(rng, n_cams, 0.03)
    tilt[0] = np.eye(3)
    Rc2w = Rc2w @ tilt
    C = centre + np.einsum("nij,j->ni", _ry(phi), np.array([0.0, 0.0, -3.0]))
    C[1:] += rng.normal(0.0, 0.05, (n_cams - 1, 3))
    C[0] = 0.0
    Rw2c = np.transpose(Rc2w, (0, 2, 1))
    cam_true = np.zeros((n_cams, 6))
    cam_true[:, :3] = _rotvec_from_matrix(Rw2c)
    cam_true[0, :3] = 0.0
    cam_true[:, 3:] = -np.einsum("nij,nj->ni", Rw2c, C)
    pts_true = centre + rng.uniform(-0.5, 0.5, (n_pts, 3))
    track = min(track, n_cams)
    keys = rng.random((n_pts, n_cams))
    cams = np.argsort(keys, axis=1)[:, :track]
    pt_idx = np.repeat(np.arange(n_pts), track)
    cam_idx = cams.reshape(-1)
    K = _intrinsics(n_cams, skew)
    return _finish(name, rng, K, cam_true, pts_true, cam_idx, pt_idx, dist, noise_px, sigma_w, sigma_t, sigma_x)


def make_path(n_cams: int, n_pts: int, n_obs: int, seed: int, *, skew: float = 0.0, dist=None, noise_px: float = 0.5,
              sigma_w: float = 0.01, sigma_t: float = 0.02, sigma_x: float = 0.02, max_track: int = 96,
              name: str = "path") -> Problem:
    """BAL-shaped problem: cameras along a closed path looking outwards, every point draws a centre
    camera and a truncated-geometric track length (min 2), observers sampled from a window of width
    3L around the centre camera => block-banded reduced camera matrix.  Exact n_pts / n_obs."""
    rng = np.random.Generator(np.random.Philox(seed))
    spacing = 0.05
    rho = n_cams * spacing / (2.0 * np.pi)
    phi = 2.0 * np.pi * np.arange(n_cams) / n_cams
    phi = np.where(phi > np.pi, phi - 2.0 * np.pi, phi)
    Rc2w = _ry(phi)
    tilt = _small_rot(rng, n_cams, 0.03)
    tilt[0] = np.eye(3)
    Rc2w = Rc2w @ tilt
    C = rho * np.stack([np.sin(phi), np.zeros(n_cams), np.cos(phi) - 1.0], axis=1)
    C[1:] += rng.normal(0.0, 0.01, (n_cams - 1, 3))
    Rw2c = np.transpose(Rc2w, (0, 2, 1))
    cam_true = np.zeros((n_cams, 6))
    cam_true[:, :3] = _rotvec_from_matrix(Rw2c)
    cam_true[0, :3] = 0.0
    cam_true[:, 3:] = -np.einsum("nij,nj->ni", Rw2c, C)
    K = _intrinsics(n_cams, skew)

    mean_track = n_obs / n_pts
    max_track = min(max_track, n_cams)
    n_gen = int(n_pts * 1.15) + 64
    factor = 1.06
    for _attempt in range(8):
        p_geo = 1.0 / max(mean_track * factor - 1.0, 1.0 + 1e-9)
        L = np.minimum(1 + rng.geometric(min(p_geo, 1.0), n_gen), max_track).astype(np.int64)
        centre_cam = rng.integers(0, n_cams, n_gen)
        # point in the centre camera's frame, then to world
        depth = rng.uniform(6.0, 14.0, n_gen)
        xc = np.stack([rng.uniform(-0.25, 0.25, n_gen) * depth, rng.uniform(-0.25, 0.25, n_gen) * depth, depth], axis=1)
        pts_all = np.einsum("nij,nj->ni", Rc2w[centre_cam], xc) + C[centre_cam]
        cam_l, pt_l = [], []
        for ell in np.unique(L):
            sel = np.nonzero(L == ell)[0]
            win = min(3 * int(ell), n_cams)
            keys = rng.random((sel.size, win))
            offs = np.argsort(keys, axis=1)[:, : int(ell)] - win // 2
            cams = (centre_cam[sel, None] + offs) % n_cams
            cam_l.append(cams.reshape(-1))
            pt_l.append(np.repeat(sel, int(ell)))
        cam_idx = np.concatenate(cam_l)
        pt_idx = np.concatenate(pt_l)
        uv, z = project(cam_true, K, pts_all, cam_idx, pt_idx, dist)
        ok = (z > 0.1) & (uv[:, 0] >= 0) & (uv[:, 0] < WIDTH) & (uv[:, 1] >= 0) & (uv[:, 1] < HEIGHT)
        cam_idx, pt_idx = cam_idx[ok], pt_idx[ok]
        cnt = np.bincount(pt_idx, minlength=n_gen)
        good = np.nonzero(cnt >= 2)[0]
        got = int(cnt[good[:n_pts]].sum()) if good.size >= n_pts else 0
        if good.size >= n_pts and got >= n_obs:
            break
        if good.size < n_pts:
            n_gen = int(n_gen * 1.3)
        else:
            factor *= max(1.04, 1.03 * n_obs / max(got, 1))
    else:
        raise RuntimeError(f"generator could not reach {n_pts} points / {n_obs} observations")
    # choose n_pts points whose total is >= n_obs and as close as possible: take the first n_pts,
    # then trim single observations off tracks longer than 2 / or swap in longer tracks
    good = good[:n_pts]
    keep_pt = np.zeros(n_gen, dtype=bool)
    keep_pt[good] = True
    m = keep_pt[pt_idx]
    cam_idx, pt_idx = cam_idx[m], pt_idx[m]
    order = np.argsort(pt_idx, kind="stable")
    cam_idx, pt_idx = cam_idx[order], pt_idx[order]
    excess = cam_idx.size - n_obs
    if excess < 0:
        raise RuntimeError(f"generator produced {cam_idx.size} observations, need {n_obs}; raise the mean")
    keep = np.ones(cam_idx.size, dtype=bool)
    cnt = np.bincount(pt_idx, minlength=n_gen)
    start = np.concatenate([[0], np.cumsum(cnt)])[:-1]
    while excess > 0:
        cand = np.nonzero(cnt > 2)[0]
        take = cand[rng.permutation(cand.size)[:excess]]
        keep[start[take] + cnt[take] - 1] = False
        cnt[take] -= 1
        excess -= take.size
    cam_idx, pt_idx = cam_idx[keep], pt_idx[keep]
    remap = np.full(n_gen, -1, dtype=np.int64)
    remap[good] = np.arange(n_pts)
    pt_idx = remap[pt_idx]
    pts_true = pts_all[good]
    assert cam_idx.size == n_obs and pt_idx.max() == n_pts - 1
    return _finish(name, rng, K, cam_true, pts_true, cam_idx, pt_idx, dist, noise_px, sigma_w, sigma_t, sigma_x)


def make_config(name: str, *, skew: float = 0.0, scale: float = 1.0, **kw) -> Problem:
    """The five BASELINE.json configs (c1..c5), seed = 20170519 + config number.  ``scale`` < 1
    shrinks points/observations proportionally (parity-test sizes)."""
    n_cams, n_pts, n_obs, kind = CONFIGS[name]
    seed = SEED0 + int(name[1:])
    if kind == "ring":
        track = n_obs // n_pts
        n_pts = max(8, int(round(n_pts * scale)))
        return make_ring(n_cams, n_pts, track, seed, skew=skew, name=name, **kw)
    mean = n_obs / n_pts
    if scale != 1.0:
        n_pts = max(64, int(round(n_pts * scale)))
        n_obs = int(round(n_pts * mean))
        n_cams = max(16, int(round(n_cams * min(1.0, scale * 4))))
    return make_path(n_cams, n_pts, n_obs, seed, skew=skew, name=name, **kw)


def shard_by_point(prob: Problem, n_ranks: int):
    """Point-sharded split: every point with all its observations goes to exactly
    one rank, balanced by observation count (greedy, longest tracks first); cameras are replicated.
    Returns a list of (Problem shard, global point ids of the shard)."""
    cnt = np.bincount(prob.pt_idx, minlength=prob.n_pts)
    order = np.argsort(-cnt, kind="stable")
    # greedy in blocks: deal sorted points round-robin in a serpentine pattern (near-optimal balance)
    owner = np.empty(prob.n_pts, dtype=np.int32)
    k = np.arange(prob.n_pts)
    rr = k % (2 * n_ranks)
    owner[order] = np.where(rr < n_ranks, rr, 2 * n_ranks - 1 - rr)
    shards = []
    for r in range(n_ranks):
        gids = np.nonzero(owner == r)[0]
        remap = np.full(prob.n_pts, -1, dtype=np.int64)
        remap[gids] = np.arange(gids.size)
        m = owner[prob.pt_idx] == r
        sh = Problem(obs=np.ascontiguousarray(prob.obs[m]), cam_idx=np.ascontiguousarray(prob.cam_idx[m]),
                     pt_idx=np.ascontiguousarray(remap[prob.pt_idx[m]].astype(np.int32)), K=prob.K, cam=prob.cam.copy(),
                     pts=np.ascontiguousarray(prob.pts[gids]), cam_true=prob.cam_true, pts_true=prob.pts_true[gids],
                     dist=prob.dist, name=f"{prob.name}[{r}/{n_ranks}]")
        shards.append((sh, gids))
    return shards


def shuffle_cameras(prob: Problem, seed: int):
    """The same problem with its cameras renumbered at random (a scan whose keyframes arrive in no particular order:
    the camera graph is still a band in SOME order, but not in the delivered one).  Returns (Problem, new_of_old)."""
    rng = np.random.default_rng(seed)
    new_of_old = rng.permutation(prob.n_cams).astype(np.int32)
    old_of_new = np.empty_like(new_of_old)
    old_of_new[new_of_old] = np.arange(prob.n_cams, dtype=np.int32)
    take = lambda a: None if a is None else np.ascontiguousarray(a[old_of_new])  # noqa: E731
    return dataclasses.replace(prob, cam_idx=np.ascontiguousarray(new_of_old[prob.cam_idx]), K=take(prob.K), cam=take(prob.cam),
                               cam_true=take(prob.cam_true), dist=take(prob.dist), name=prob.name + "-shuffled"), new_of_old


def keyframe_stream(prob: Problem, k0: int):
    """Replay ``prob`` as a scan (what CSfM::mapping produces keyframe by keyframe, CSfM.cpp:252-259): a start problem
    over the first ``k0`` keyframes and one append per further keyframe.  A map point enters the stream when its second
    observation arrives (a point is triangulated from two views) and brings its earlier observation with it; point ids
    of the stream are in order of entry.  Returns (start Problem, [dict(K, cam, pts, obs, cam_idx, pt_idx)], final
    Problem = the concatenation, numbered as the stream numbers it)."""
    order = np.lexsort((prob.pt_idx, prob.cam_idx))
    cam_idx, pt_idx, obs = prob.cam_idx[order], prob.pt_idx[order], prob.obs[order]
    n_seen = np.zeros(prob.n_pts, dtype=np.int64)
    new_id = np.full(prob.n_pts, -1, dtype=np.int64)
    pending = {}  # original point id -> index (into the sorted arrays) of its single observation so far
    stages, next_id = [], 0
    bounds = np.searchsorted(cam_idx, np.arange(prob.n_cams + 1))
    cur = dict(obs=[], cam=[], pt=[], new_pts=[])

    def close(c_lo, c_hi):
        nonlocal cur
        st = dict(K=prob.K[c_lo:c_hi].copy(), cam=prob.cam[c_lo:c_hi].copy(), cam_true=prob.cam_true[c_lo:c_hi],
                  pts=prob.pts[cur["new_pts"]].copy() if cur["new_pts"] else np.zeros((0, 3)),
                  pts_true=prob.pts_true[cur["new_pts"]] if cur["new_pts"] else np.zeros((0, 3)),
                  obs=np.ascontiguousarray(np.array(cur["obs"]).reshape(-1, 2)), cam_idx=np.array(cur["cam"], dtype=np.int32),
                  pt_idx=np.array(cur["pt"], dtype=np.int32))
        stages.append(st)
        cur = dict(obs=[], cam=[], pt=[], new_pts=[])

    for c in range(prob.n_cams):
        for j in range(bounds[c], bounds[c + 1]):
            p = int(pt_idx[j])
            n_seen[p] += 1
            if n_seen[p] == 1:
                pending[p] = j
                continue
            if n_seen[p] == 2:
                new_id[p] = next_id
                next_id += 1
                cur["new_pts"].append(p)
                j0 = pending.pop(p)
                cur["obs"].append(obs[j0]); cur["cam"].append(int(cam_idx[j0])); cur["pt"].append(int(new_id[p]))
            cur["obs"].append(obs[j]); cur["cam"].append(c); cur["pt"].append(int(new_id[p]))
        if c == k0 - 1:
            close(0, k0)
        elif c >= k0:
            close(c, c + 1)
    first = stages[0]
    start = Problem(obs=first["obs"], cam_idx=first["cam_idx"], pt_idx=first["pt_idx"], K=np.ascontiguousarray(first["K"]),
                    cam=np.ascontiguousarray(first["cam"]), pts=np.ascontiguousarray(first["pts"]), cam_true=first["cam_true"],
                    pts_true=first["pts_true"], name=prob.name + "[start]")
    cat = lambda k: np.ascontiguousarray(np.concatenate([s[k] for s in stages]))  # noqa: E731
    final = Problem(obs=cat("obs"), cam_idx=cat("cam_idx"), pt_idx=cat("pt_idx"), K=cat("K"), cam=cat("cam"), pts=cat("pts"),
                    cam_true=cat("cam_true"), pts_true=cat("pts_true"), name=prob.name + "[stream]")
    return start, stages[1:], final
