"""Python host side of the drop-in boundary: ctypes over the C ABI of ``libhsfm_ba.so``
(include/hsfm_ba.h) plus a mirror of the reference interface

    void CTracker::bundleAdjustmentStructAndPose(observations, camIdx, K, R, t, pts3D, isStructOrPose)

(reference CTracker.h:65, CTracker.cpp:670-702; same names, argument meaning and — void return,
parameters optimised in place through the passed arrays — error behaviour).

There is NO CPU fallback here: if the CUDA library cannot be loaded or no sm_100 device is
present, construction raises.  The CPU oracle lives under ``oracle/`` and is never imported
from this package.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from ._cdefs import (CONVERGENCE, FAILURE, LINEAR_AUTO, LINEAR_DENSE_SCHUR, LINEAR_PCG, NO_CONVERGENCE, POSE_ONLY,
                     STRUCT_AND_POSE, STRUCT_ONLY, Iteration, Options, Summary)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HSFM_BA_LIB", os.path.join(_HERE, "libhsfm_ba.so"))  # override: instrumented experiment builds

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_pp = C.POINTER(C.c_void_p)
_bp = C.POINTER(C.c_uint8)

# name -> (restype, argtypes): every symbol include/hsfm_ba.h declares
SYMBOLS = {
    "hsfm_ba_abi_version": (C.c_int, []),
    "hsfm_ba_error_string": (C.c_char_p, [C.c_int]),
    "hsfm_ba_default_options": (None, [C.POINTER(Options)]),
    "hsfm_ba_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "hsfm_ba_destroy": (None, [C.c_void_p]),
    "hsfm_ba_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "hsfm_ba_set_ordering": (C.c_int, [C.c_void_p, C.c_int]),
    "hsfm_ba_debug_structure": (C.c_int, [C.c_int32, _ip, C.c_int32, _ip, _ip, _dp, C.POINTER(C.c_uint8)]),
    "hsfm_ba_debug_schedule": (C.c_int, [C.c_int32, _ip, C.c_int32, _ip]),
    "hsfm_ba_debug_rcm": (C.c_int, [C.c_int32, C.c_int64, _ip, _ip]),
    "hsfm_ba_comm_unique_id": (C.c_int, [C.c_void_p]),
    "hsfm_ba_comm_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "hsfm_ba_solve_ptr": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _dp, _ip, C.c_int32, _dp, _pp, _pp, _pp,
                                    C.POINTER(Options), C.POINTER(Summary)]),
    "hsfm_ba_solve_packed": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64,
                                       _dp, C.POINTER(Options), C.POINTER(Summary)]),
    "hsfm_ba_solve_packed_ex": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64,
                                          _dp, _bp, _bp, C.POINTER(Options), C.POINTER(Summary)]),
    "hsfm_ba_problem_set_constant": (C.c_int, [C.c_void_p, _bp, _bp]),
    "hsfm_ba_scan_begin": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp,
                                     C.POINTER(C.c_int64)]),
    "hsfm_ba_scan_append_keyframe": (C.c_int, [C.c_void_p, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp, C.c_int64, _dp, _ip, _ip,
                                               C.POINTER(C.c_int64)]),
    "hsfm_ba_eval": (C.c_int, [C.c_void_p, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp, _dp, _dp, _dp]),
    "hsfm_ba_schur": (C.c_int, [C.c_void_p, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp,
                                C.POINTER(Options), C.c_double, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "hsfm_ba_dense_spd_solve": (C.c_int, [C.c_void_p, C.c_int32, _dp, _dp, _dp, _dp, _ip]),
    "hsfm_ba_bench_cholesky": (C.c_int, [C.c_void_p, C.c_int32, C.c_int, _dp, _dp, _dp]),
    "hsfm_ba_problem_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp]),
    "hsfm_ba_problem_reset": (C.c_int, [C.c_void_p]),
    "hsfm_ba_problem_solve": (C.c_int, [C.c_void_p, C.POINTER(Options), C.POINTER(Summary)]),
    "hsfm_ba_problem_eval": (C.c_int, [C.c_void_p, _dp]),
    "hsfm_ba_problem_frame_stats": (C.c_int, [C.c_void_p, _dp, _dp]),
    "hsfm_ba_problem_download": (C.c_int, [C.c_void_p, _dp, _dp]),
    "hsfm_ba_problem_algorithmic_bytes": (C.c_int64, [C.c_void_p, C.c_char_p]),
    "hsfm_ba_problem_time_kernel": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, _dp]),
    # device pointers (the caller's buffers), passed as integers
    "hsfm_ba_cost_noalloc": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                       C.c_int64, C.c_void_p, C.c_void_p]),
}

NOALLOC_MAX_OBS = 1 << 22  # HSFM_BA_NOALLOC_MAX_OBS

_lib = None


class BAError(RuntimeError):
    def __init__(self, code: int, what: str):
        self.code = code
        try:
            msg = load_library().hsfm_ba_error_string(code).decode()
        except Exception:  # pragma: no cover
            msg = "?"
        super().__init__(f"{what} failed: {code} ({msg})")


def load_library():
    """dlopen libhsfm_ba.so and bind every symbol of the header.  Raises if the library is missing —
    the product never substitutes another implementation."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} is not built; run `python -m sfm_b200.build` (needs nvcc)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _mask(a, n):
    if a is None:
        return None, None
    a = np.ascontiguousarray(np.asarray(a).astype(bool), dtype=np.uint8)
    if a.shape != (n,):
        raise ValueError("mask must have one entry per block")
    return a, a.ctypes.data_as(_bp)


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def default_options(**kw) -> Options:
    o = Options()
    load_library().hsfm_ba_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


class _SummaryBox:
    def __init__(self, trace_capacity=64):
        self.trace_buf = (Iteration * trace_capacity)()
        self.s = Summary()
        self.s.trace = C.cast(self.trace_buf, C.POINTER(Iteration))
        self.s.trace_capacity = trace_capacity

    def as_dict(self):
        d = {f: getattr(self.s, f) for f, _ in self.s._fields_ if f not in ("trace", "phase_ms")}
        d["phase_ms"] = [float(v) for v in self.s.phase_ms]
        d["trace"] = [self.trace_buf[i].as_dict() for i in range(self.s.trace_count)]
        return d


def debug_structure(cam_first, natural=False, depth=None):
    """Host-only: camera order, block structure of the Cholesky factor and chain schedule for an envelope
    (hsfm_ba_debug_structure).  depth: force that dissection depth of the nested two-ended order.  Returns a dict."""
    first = np.ascontiguousarray(cam_first, dtype=np.int32)
    n = first.shape[0]
    perm = np.empty(n, dtype=np.int32)
    info = np.zeros(8, dtype=np.int32)
    fill = C.c_double()
    panels = (6 * n + 127) // 128
    st = np.zeros((panels, panels), dtype=np.uint8)
    natural = -1 - int(depth) if depth is not None else int(bool(natural))
    rc = load_library().hsfm_ba_debug_structure(n, _i(first), natural, _i(perm), _i(info), C.byref(fill),
                                                st.ctypes.data_as(C.POINTER(C.c_uint8)))
    if rc:
        raise BAError(rc, "hsfm_ba_debug_structure")
    seg3 = np.full(3 * 32, -1, dtype=np.int32)
    rc = load_library().hsfm_ba_debug_schedule(n, _i(first), natural, _i(seg3))
    if rc:
        raise BAError(rc, "hsfm_ba_debug_schedule")
    segments = [tuple(int(v) for v in seg3[3 * g:3 * g + 3]) for g in range(int(info[6]))]  # (first panel, end panel, parent)
    return dict(new_of_old=perm, panels=int(info[0]), split_a=int(info[1]), split_t0=int(info[2]), structured=bool(info[3]),
                reordered=bool(info[4]), split_cams=int(info[5]), fill=fill.value, structure=st, segments=segments,
                critical_panels=int(info[7]))


def debug_rcm(n_cams, edges):
    """Host-only: reverse Cuthill-McKee position of every camera for an undirected edge list [n_edges, 2] (hsfm_ba_debug_rcm)."""
    e = np.ascontiguousarray(edges, dtype=np.int32).reshape(-1, 2)
    pos = np.empty(n_cams, dtype=np.int32)
    rc = load_library().hsfm_ba_debug_rcm(n_cams, e.shape[0], _i(e), _i(pos))
    if rc:
        raise BAError(rc, "hsfm_ba_debug_rcm")
    return pos


class Context:
    """One hsfm_ba_ctx (one GPU)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.hsfm_ba_create(C.byref(h), device)
        if rc:
            raise BAError(rc, "hsfm_ba_create")
        self.h = h
        self._keep = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.hsfm_ba_destroy(self.h)
            self.h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- plumbing --
    def set_stream(self, cuda_stream_ptr: int | None):
        rc = self.lib.hsfm_ba_set_stream(self.h, C.c_void_p(cuda_stream_ptr or 0))
        if rc:
            raise BAError(rc, "hsfm_ba_set_stream")

    def set_ordering(self, ordering: int):
        """0 = library-chosen internal camera order (default), 1 = keep the caller's order."""
        rc = self.lib.hsfm_ba_set_ordering(self.h, ordering)
        if rc:
            raise BAError(rc, "hsfm_ba_set_ordering")

    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        rc = load_library().hsfm_ba_comm_unique_id(buf)
        if rc:
            raise BAError(rc, "hsfm_ba_comm_unique_id")
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, n_ranks: int):
        buf = C.create_string_buffer(unique_id, 128)
        rc = self.lib.hsfm_ba_comm_init(self.h, buf, rank, n_ranks)
        if rc:
            raise BAError(rc, "hsfm_ba_comm_init")

    # -- one-shot calls on host buffers --
    def solve_packed(self, prob, mode=STRUCT_AND_POSE, opts=None, trace_capacity=64, inplace=False, cam_fixed=None, pt_fixed=None):
        """Solves a sfm_b200.synth.Problem-like object; returns (cam, pts, summary dict).  ``inplace=True`` is the
        reference's own contract (CTracker.cpp:670-702 updates the caller's pose/point storage): prob.cam / prob.pts
        are read as the initial guess and overwritten with the result, no copies on the Python side."""
        cam, pts = (prob.cam, prob.pts) if inplace else (prob.cam.copy(), prob.pts.copy())
        box = _SummaryBox(trace_capacity)
        cf, cfp = _mask(cam_fixed, prob.n_cams)
        pf, pfp = _mask(pt_fixed, prob.n_pts)
        rc = self.lib.hsfm_ba_solve_packed_ex(self.h, mode, prob.n_obs, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), prob.n_cams,
                                              _d(prob.K), _d(prob.dist), _d(cam), prob.n_pts, _d(pts), cfp, pfp,
                                              C.byref(opts) if opts is not None else None, C.byref(box.s))
        if rc:
            raise BAError(rc, "hsfm_ba_solve_packed")
        return cam, pts, box.as_dict()

    def solve_ptr(self, obs, cam_idx, K9, R_ptrs, t_ptrs, pt_ptrs, mode=STRUCT_AND_POSE, opts=None, trace_capacity=64):
        """hsfm_ba_solve_ptr with the pointer arrays given as uint64 numpy arrays of host addresses (R_ptrs, t_ptrs: one per
        camera; pt_ptrs: one per observation, equal addresses = the same map point).  Optimises in place through them."""
        obs = _f64(obs, (-1, 2))
        cam_idx = np.ascontiguousarray(cam_idx, dtype=np.int32)
        K9 = _f64(K9, (-1, 9))
        R_ptrs, t_ptrs, pt_ptrs = (np.ascontiguousarray(a, dtype=np.uint64) for a in (R_ptrs, t_ptrs, pt_ptrs))
        box = _SummaryBox(trace_capacity)
        rc = self.lib.hsfm_ba_solve_ptr(self.h, int(mode), obs.shape[0], _d(obs), _i(cam_idx), K9.shape[0], _d(K9),
                                        R_ptrs.ctypes.data_as(_pp), t_ptrs.ctypes.data_as(_pp), pt_ptrs.ctypes.data_as(_pp),
                                        C.byref(opts) if opts is not None else None, C.byref(box.s))
        if rc:
            raise BAError(rc, "hsfm_ba_solve_ptr")
        return box.as_dict()

    def evaluate(self, prob, cam=None, pts=None, want_jac=True):
        cam = prob.cam if cam is None else cam
        pts = prob.pts if pts is None else pts
        n = prob.n_obs
        r = np.empty((n, 2))
        J = np.empty((n, 2, 9)) if want_jac else None
        cost = C.c_double()
        rc = self.lib.hsfm_ba_eval(self.h, n, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), prob.n_cams, _d(prob.K), _d(prob.dist),
                                   _d(cam), prob.n_pts, _d(pts), _d(r), _d(J), C.byref(cost))
        if rc:
            raise BAError(rc, "hsfm_ba_eval")
        return r, J, cost.value

    def schur(self, prob, radius=1e4, opts=None, cam=None, pts=None):
        cam = prob.cam if cam is None else cam
        pts = prob.pts if pts is None else pts
        opts = opts or default_options()
        nc, npt = prob.n_cams, prob.n_pts
        out = dict(S=np.zeros((6 * nc, 6 * nc)), rhs=np.zeros(6 * nc), Hcc=np.zeros((nc, 6, 6)), gc=np.zeros((nc, 6)),
                   Hpp=np.zeros((npt, 3, 3)), gp=np.zeros((npt, 3)), scale=np.zeros(6 * nc + 3 * npt))
        rc = self.lib.hsfm_ba_schur(self.h, prob.n_obs, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), nc, _d(prob.K), _d(prob.dist),
                                    _d(cam), npt, _d(pts), C.byref(opts), radius, _d(out["S"]), _d(out["rhs"]), _d(out["Hcc"]),
                                    _d(out["gc"]), _d(out["Hpp"]), _d(out["gp"]), _d(out["scale"]))
        if rc:
            raise BAError(rc, "hsfm_ba_schur")
        return out

    def dense_spd_solve(self, A, b, want_factor=False):
        """x = A^-1 b through the library's dense Cholesky (parity tap)."""
        A = _f64(A); b = _f64(b)
        n = A.shape[0]
        x = np.empty(n)
        L = np.zeros((n, n)) if want_factor else None
        piv = C.c_int32(0)
        rc = self.lib.hsfm_ba_dense_spd_solve(self.h, n, _d(A), _d(b), _d(x), _d(L), C.byref(piv))
        if rc:
            e = BAError(rc, "hsfm_ba_dense_spd_solve")
            e.pivot = piv.value
            raise e
        return (x, np.tril(L)) if want_factor else x

    def bench_cholesky(self, n: int, reps: int = 3):
        """(factor ms, solve ms, relative residual) of the dense Cholesky on a synthetic SPD matrix of order n."""
        f, s, r = C.c_double(), C.c_double(), C.c_double()
        rc = self.lib.hsfm_ba_bench_cholesky(self.h, n, reps, C.byref(f), C.byref(s), C.byref(r))
        if rc:
            raise BAError(rc, "hsfm_ba_bench_cholesky")
        return f.value, s.value, r.value

    # -- resident problem --
    def upload(self, prob, mode=STRUCT_AND_POSE):
        self._keep = prob
        rc = self.lib.hsfm_ba_problem_upload(self.h, mode, prob.n_obs, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), prob.n_cams,
                                             _d(prob.K), _d(prob.dist), _d(prob.cam), prob.n_pts, _d(prob.pts))
        if rc:
            raise BAError(rc, "hsfm_ba_problem_upload")

    def scan_begin(self, prob, mode=STRUCT_AND_POSE):
        """Upload ``prob`` as the start of a scan (device-resident observation table); returns the bytes sent."""
        import dataclasses

        sent = C.c_int64()
        rc = self.lib.hsfm_ba_scan_begin(self.h, mode, prob.n_obs, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), prob.n_cams,
                                         _d(prob.K), _d(prob.dist), _d(prob.cam), prob.n_pts, _d(prob.pts), C.byref(sent))
        if rc:
            raise BAError(rc, "hsfm_ba_scan_begin")
        self._keep = dataclasses.replace(prob)
        return sent.value

    def scan_append_keyframe(self, K_new, cam_new, pts_new, obs, cam_idx, pt_idx, dist_new=None):
        """Append cameras (rows of K_new [k,9] / cam_new [k,6]), points (pts_new [m,3]) and observations (global camera /
        point ids) to the scan; returns the bytes sent.  Only these rows are uploaded."""
        K_new, cam_new, pts_new = _f64(K_new, (-1, 9)), _f64(cam_new, (-1, 6)), _f64(pts_new, (-1, 3))
        obs = _f64(obs, (-1, 2))
        cam_idx = np.ascontiguousarray(cam_idx, dtype=np.int32)
        pt_idx = np.ascontiguousarray(pt_idx, dtype=np.int32)
        dist_new = None if dist_new is None else _f64(dist_new, (-1, 2))
        sent = C.c_int64()
        rc = self.lib.hsfm_ba_scan_append_keyframe(self.h, K_new.shape[0], _d(K_new), _d(dist_new), _d(cam_new), pts_new.shape[0], _d(pts_new),
                                                   obs.shape[0], _d(obs), _i(cam_idx), _i(pt_idx), C.byref(sent))
        if rc:
            raise BAError(rc, "hsfm_ba_scan_append_keyframe")
        k = self._keep
        k.cam = np.concatenate([k.cam, cam_new]); k.pts = np.concatenate([k.pts, pts_new])  # sizes for download()
        return sent.value

    def set_constant(self, cam_fixed=None, pt_fixed=None):
        """Hold cameras / points of the resident problem constant (boolean masks, caller order); None, None clears."""
        cf, cfp = _mask(cam_fixed, self._keep.n_cams)
        pf, pfp = _mask(pt_fixed, self._keep.n_pts)
        rc = self.lib.hsfm_ba_problem_set_constant(self.h, cfp, pfp)
        if rc:
            raise BAError(rc, "hsfm_ba_problem_set_constant")

    def reset(self):
        rc = self.lib.hsfm_ba_problem_reset(self.h)
        if rc:
            raise BAError(rc, "hsfm_ba_problem_reset")

    def solve_resident(self, opts=None, trace_capacity=64):
        opts = opts or default_options()
        box = _SummaryBox(trace_capacity)
        rc = self.lib.hsfm_ba_problem_solve(self.h, C.byref(opts), C.byref(box.s))
        if rc:
            raise BAError(rc, "hsfm_ba_problem_solve")
        return box.as_dict()

    def eval_resident(self, want_cost=True):
        cost = C.c_double()
        rc = self.lib.hsfm_ba_problem_eval(self.h, C.byref(cost) if want_cost else None)
        if rc:
            raise BAError(rc, "hsfm_ba_problem_eval")
        return cost.value if want_cost else None

    def frame_stats(self):
        """Per-camera (rms, max) reprojection distance in pixels at the current parameters
        (the statistics of CSfM.cpp:350-382)."""
        n = self._keep.n_cams
        rms, mx = np.empty(n), np.empty(n)
        rc = self.lib.hsfm_ba_problem_frame_stats(self.h, _d(rms), _d(mx))
        if rc:
            raise BAError(rc, "hsfm_ba_problem_frame_stats")
        return rms, mx

    def download(self):
        p = self._keep
        cam, pts = np.empty((p.n_cams, 6)), np.empty((p.n_pts, 3))
        rc = self.lib.hsfm_ba_problem_download(self.h, _d(cam), _d(pts))
        if rc:
            raise BAError(rc, "hsfm_ba_problem_download")
        return cam, pts

    def cost_noalloc(self, n_obs, d_obs, d_cam_idx, d_pt_idx, n_cams, d_K9, d_cam_rt, n_pts, d_pts, d_cost):
        """hsfm_ba_cost_noalloc: all d_* are DEVICE addresses (ints) of caller-owned buffers; allocates nothing; the
        result lands in d_cost once the context's stream has been synchronised."""
        rc = self.lib.hsfm_ba_cost_noalloc(self.h, n_obs, d_obs, d_cam_idx, d_pt_idx, n_cams, d_K9, d_cam_rt, n_pts, d_pts, d_cost)
        if rc:
            raise BAError(rc, "hsfm_ba_cost_noalloc")

    def algorithmic_bytes(self, kernel: str) -> int:
        return int(self.lib.hsfm_ba_problem_algorithmic_bytes(self.h, kernel.encode()))

    def time_kernel(self, kernel: str, reps: int = 20) -> float:
        ms = C.c_double()
        rc = self.lib.hsfm_ba_problem_time_kernel(self.h, kernel.encode(), reps, C.byref(ms))
        if rc:
            raise BAError(rc, "hsfm_ba_problem_time_kernel")
        return ms.value


class CTracker:
    """Mirror of the BA facade of the reference's ``CTracker`` (CTracker.h:64-67)."""

    STRUCT_ONLY, POSE_ONLY, STRUCT_AND_POSE = STRUCT_ONLY, POSE_ONLY, STRUCT_AND_POSE  # enum BA_TYPE, CTracker.h:67

    def __init__(self, device: int = 0):
        self._ctx = Context(device)
        # initialiseBAOptions(), CTracker.cpp:571-577: DENSE_SCHUR, silent, everything else default
        self._opts = default_options(linear_solver=LINEAR_DENSE_SCHUR)
        self.last_summary = None

    def bundleAdjustmentStructAndPose(self, observations, camIdx, K, R, t, pts3D, isStructOrPose):
        """Same arguments as the reference:
        observations  [n,2] float64 (vector<Point2d>), camIdx [n] int, K [n_cams,3,3] (vector<Matx33d>),
        R, t          sequences of n_cams float64 arrays of 3 (the ``double*`` into CFrame::_rot/_t),
        pts3D         sequence of n float64 arrays of 3; entries that alias the same memory are the same
                      map point (pointer identity, CSfM.cpp:331 / CMap.cpp:215).
        Optimises R, t, pts3D in place; returns None (the reference returns void and drops the summary)."""
        obs = _f64(observations, (-1, 2))
        cam_idx = np.ascontiguousarray(camIdx, dtype=np.int32)
        n, n_cams = obs.shape[0], len(R)
        K9 = _f64(np.asarray(K), (n_cams, 9)) if n_cams else np.zeros((0, 9))
        if len(pts3D) != n or cam_idx.shape[0] != n or len(t) != n_cams:
            raise ValueError("observations, camIdx and pts3D must have one entry per observation; R, t, K one per camera")
        for a in list(R) + list(t) + list(pts3D):
            if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.size == 3 and a.flags.c_contiguous and a.flags.writeable):
                raise TypeError("R, t, pts3D entries must be writable contiguous float64 arrays (views) of 3 elements")
        Rp = (C.c_void_p * max(1, n_cams))(*[a.ctypes.data for a in R])
        tp = (C.c_void_p * max(1, n_cams))(*[a.ctypes.data for a in t])
        Pp = (C.c_void_p * max(1, n))(*[a.ctypes.data for a in pts3D])
        box = _SummaryBox()
        rc = self._ctx.lib.hsfm_ba_solve_ptr(self._ctx.h, int(isStructOrPose), n, _d(obs), _i(cam_idx), n_cams, _d(K9), Rp, tp, Pp,
                                             C.byref(self._opts), C.byref(box.s))
        self.last_summary = box.as_dict()
        if rc:
            raise BAError(rc, "hsfm_ba_solve_ptr")
        return None
