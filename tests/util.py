import numpy as np


def residual_err(r, r_ref, obs):
    """max over observations of |r - r_ref| / max(1, |predicted pixel|)   (SURVEY.md §8d definition:
    r = predicted - observed cancels, so the error is taken relative to the predicted pixel)."""
    pred = r_ref + obs
    return float((np.abs(r - r_ref).max(axis=1) / np.maximum(1.0, np.abs(pred).max(axis=1))).max()) if len(r) else 0.0


def jacobian_err(J, J_ref):
    """max over observations of |J - J_ref|_max / |J_ref|_max (per-observation max-abs)."""
    if len(J) == 0:
        return 0.0
    a = np.abs(J - J_ref).reshape(len(J), -1).max(axis=1)
    b = np.abs(J_ref).reshape(len(J), -1).max(axis=1)
    return float((a / b).max())


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def kat_problem(g):
    """functor_kat.npz as a packed Problem: one camera and one point per case."""
    from sfm_b200 import synth

    n = len(g["w"])
    cam = np.ascontiguousarray(np.concatenate([g["w"], g["t"]], axis=1))
    idx = np.arange(n, dtype=np.int32)
    return synth.Problem(obs=np.ascontiguousarray(g["obs"]), cam_idx=idx, pt_idx=idx.copy(), K=np.ascontiguousarray(g["K"]), cam=cam,
                         pts=np.ascontiguousarray(g["X"]), cam_true=cam, pts_true=g["X"], name="kat")
