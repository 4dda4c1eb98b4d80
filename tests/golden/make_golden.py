"""Generates the golden vectors under tests/golden/ (run in the BUILD container; the GPU box only
reads the committed .npz files).

The reference (hulop/SfM) has no tests or fixtures for its bundle adjustment and cannot be
compiled or imported here, so these vectors come from INDEPENDENT
implementations of the same mathematics, not from the oracle under test:

  functor_kat.npz    the reprojection functor of CTracker.cpp:585-604 evaluated with mpmath at 50
                     digits (value and all 9 partial derivatives by 50-digit central differences of
                     the exact Rodrigues formula), plus OpenCV's analytic Jacobian from
                     cv2.projectPoints (d(u,v)/d(rvec,tvec)) for the six camera columns.
  normal_eq_kat.npz  a 6-camera / 40-point problem: dense J^T J + D^2, the Schur complement and the
                     full Newton step computed with numpy/LAPACK (np.linalg) from the oracle-independent
                     mpmath-verified Jacobian layout.
  lm_final.npz       converged cost of the same small problem from scipy.optimize.least_squares
                     (trust-region reflective; a different algorithm, so only the converged cost is comparable).

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import mpmath as mp  # noqa: E402

mp.mp.dps = 50


def project_mp(w, t, X, k, obs):
    """Exact pinhole projection, mpmath.  Rodrigues formula (the theta^2>eps branch of
    ceres::AngleAxisRotatePoint is the exact rotation; its Taylor branch is an approximation of it)."""
    w = [mp.mpf(v) for v in w]
    t = [mp.mpf(v) for v in t]
    X = [mp.mpf(v) for v in X]
    th = mp.sqrt(w[0] ** 2 + w[1] ** 2 + w[2] ** 2)
    if th == 0:
        p = list(X)
    else:
        u = [v / th for v in w]
        c, s = mp.cos(th), mp.sin(th)
        cr = [u[1] * X[2] - u[2] * X[1], u[2] * X[0] - u[0] * X[2], u[0] * X[1] - u[1] * X[0]]
        d = u[0] * X[0] + u[1] * X[1] + u[2] * X[2]
        p = [X[i] * c + cr[i] * s + u[i] * d * (1 - c) for i in range(3)]
    p = [p[i] + t[i] for i in range(3)]
    xp, yp = p[0] / p[2], p[1] / p[2]
    return [mp.mpf(k[0]) * xp + mp.mpf(k[1]) * yp + mp.mpf(k[2]) - mp.mpf(obs[0]), mp.mpf(k[4]) * yp + mp.mpf(k[5]) - mp.mpf(obs[1])]


def functor_kat(rng):
    import cv2

    cases = []
    norms = [0.0, 1e-4, 1e-3, 1e-2, 0.1, 0.5, 1.0, 2.0, 3.0, 3.1]
    for n in norms:
        for rep in range(6):
            d = rng.normal(size=3)
            d /= np.linalg.norm(d)
            w = d * n
            X = rng.uniform(-0.5, 0.5, 3) + np.array([0, 0, 3.0])
            t = rng.normal(0, 0.2, 3)
            skew = 0.0 if rep % 2 == 0 else 0.5 * rng.normal()
            k = np.array([1072.6066932721178, skew, 648.7807504771789, 0, 1067.1975156086196, 364.5034359624969, 0, 0, 1.0])
            obs = rng.uniform(0, 700, 2)
            cases.append((w, t, X, k, obs))
    W = np.array([c[0] for c in cases]); T = np.array([c[1] for c in cases]); XX = np.array([c[2] for c in cases])
    K = np.array([c[3] for c in cases]); O = np.array([c[4] for c in cases])
    R = np.zeros((len(cases), 2)); J = np.zeros((len(cases), 2, 9)); Jcv = np.zeros((len(cases), 2, 6))
    h = mp.mpf(10) ** -20
    for i, (w, t, X, k, obs) in enumerate(cases):
        r = project_mp(w, t, X, k, obs)
        R[i] = [float(r[0]), float(r[1])]
        x0 = [mp.mpf(v) for v in list(w) + list(t) + list(X)]
        for j in range(9):
            xp = list(x0); xm = list(x0)
            xp[j] += h; xm[j] -= h
            rp = project_mp(xp[:3], xp[3:6], xp[6:], k, obs)
            rm = project_mp(xm[:3], xm[3:6], xm[6:], k, obs)
            J[i, 0, j] = float((rp[0] - rm[0]) / (2 * h))
            J[i, 1, j] = float((rp[1] - rm[1]) / (2 * h))
        Kmat = k.reshape(3, 3).copy()
        _, jac = cv2.projectPoints(X.reshape(1, 3), w.reshape(3, 1), t.reshape(3, 1), Kmat, None)
        Jcv[i] = jac[:, :6]
    np.savez(os.path.join(HERE, "functor_kat.npz"), w=W, t=T, X=XX, K=K, obs=O, r=R, J=J, J_cv2=Jcv)
    print("functor_kat:", len(cases), "cases")


def small_problem():
    from sfm_b200 import synth

    return synth.make_ring(6, 40, 4, seed=99, skew=0.3, name="kat")


def normal_eq_kat():
    """numpy/LAPACK normal equations of the small problem, Jacobian by mpmath central differences."""
    p = small_problem()
    n, nc, npt = p.n_obs, p.n_cams, p.n_pts
    J = np.zeros((2 * n, 6 * nc + 3 * npt)); r = np.zeros(2 * n)
    h = mp.mpf(10) ** -20
    for i in range(n):
        c, q = int(p.cam_idx[i]), int(p.pt_idx[i])
        x0 = [mp.mpf(float(v)) for v in list(p.cam[c]) + list(p.pts[q])]
        rr = project_mp(x0[:3], x0[3:6], x0[6:], p.K[c], p.obs[i])
        r[2 * i], r[2 * i + 1] = float(rr[0]), float(rr[1])
        for j in range(9):
            xp = list(x0); xm = list(x0)
            xp[j] += h; xm[j] -= h
            rp = project_mp(xp[:3], xp[3:6], xp[6:], p.K[c], p.obs[i])
            rm = project_mp(xm[:3], xm[3:6], xm[6:], p.K[c], p.obs[i])
            col = 6 * c + j if j < 6 else 6 * nc + 3 * q + (j - 6)
            J[2 * i, col] = float((rp[0] - rm[0]) / (2 * h))
            J[2 * i + 1, col] = float((rp[1] - rm[1]) / (2 * h))
    # cam 0 has w = 0 exactly: the reference's autodiff takes the Taylor branch there, whose derivative
    # w.r.t. w is e_k x X — equal to the exact derivative at w = 0, so J above is still the reference's J.
    scale = 1.0 / (1.0 + np.sqrt((J * J).sum(axis=0)))
    Js = J * scale
    radius = 1e4
    diag = np.clip((Js * Js).sum(axis=0), 1e-6, 1e32)
    D2 = np.sqrt(diag / radius) ** 2
    H = Js.T @ Js + np.diag(D2)
    g = Js.T @ r
    y = np.linalg.solve(H, g)
    nc6 = 6 * nc
    Hcc, Hcp, Hpp = H[:nc6, :nc6], H[:nc6, nc6:], H[nc6:, nc6:]
    Hpp_inv = np.linalg.inv(Hpp)
    S = Hcc - Hcp @ Hpp_inv @ Hcp.T
    rhs = g[:nc6] - Hcp @ Hpp_inv @ g[nc6:]
    cost = 0.5 * float(r @ r)
    step = -y
    m = Js @ step
    model_cost_change = -float(m @ (r + m / 2))
    np.savez(os.path.join(HERE, "normal_eq_kat.npz"), r=r, J=J, scale=scale, S=S, rhs=rhs, y=y, cost=cost, radius=radius,
             model_cost_change=model_cost_change, g=g)
    print("normal_eq_kat: n_obs", n, "unknowns", J.shape[1], "cost", cost)


def lm_final():
    from scipy.optimize import least_squares

    from sfm_b200 import synth

    p = small_problem()
    nc, npt = p.n_cams, p.n_pts

    def fun(x):
        cam = x[: 6 * nc].reshape(nc, 6)
        pts = x[6 * nc:].reshape(npt, 3)
        uv, _ = synth.project(cam, p.K, pts, p.cam_idx, p.pt_idx)
        return (uv - p.obs).reshape(-1)

    x0 = np.concatenate([p.cam.reshape(-1), p.pts.reshape(-1)])
    sol = least_squares(fun, x0, method="trf", xtol=1e-14, ftol=1e-14, gtol=1e-12, max_nfev=400)
    np.savez(os.path.join(HERE, "lm_final.npz"), cost=sol.cost, x=sol.x)
    print("lm_final: scipy converged cost", sol.cost, "nfev", sol.nfev)


if __name__ == "__main__":
    rng = np.random.default_rng(20170519)
    functor_kat(rng)
    normal_eq_kat()
    lm_final()
