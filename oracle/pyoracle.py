"""ctypes binding of the CPU ORACLE (oracle/libba_oracle.so).

TEST INFRASTRUCTURE.  Import only from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs — never from the sfm_b200 package."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libba_oracle.so")


class OOptions(C.Structure):
    _fields_ = [
        ("max_num_iterations", C.c_int32), ("max_consecutive_invalid_steps", C.c_int32), ("jacobi_scaling", C.c_int32),
        ("linear_solver", C.c_int32), ("dense_max_cams", C.c_int32), ("pcg_max_iterations", C.c_int32),
        ("verbose", C.c_int32), ("num_threads", C.c_int32),
        ("initial_trust_region_radius", C.c_double), ("max_trust_region_radius", C.c_double),
        ("min_trust_region_radius", C.c_double), ("min_relative_decrease", C.c_double), ("min_lm_diagonal", C.c_double),
        ("max_lm_diagonal", C.c_double), ("function_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
        ("parameter_tolerance", C.c_double), ("pcg_tolerance", C.c_double),
    ]


class OIteration(C.Structure):
    _fields_ = [
        ("cost", C.c_double), ("cost_change", C.c_double), ("gradient_max_norm", C.c_double), ("step_norm", C.c_double),
        ("relative_decrease", C.c_double), ("trust_region_radius", C.c_double), ("model_cost_change", C.c_double),
        ("step_is_valid", C.c_int32), ("step_is_successful", C.c_int32), ("linear_iterations", C.c_int32),
        ("reserved0", C.c_int32),
    ]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_ if f != "reserved0"}


class OSummary(C.Structure):
    _fields_ = [
        ("initial_cost", C.c_double), ("final_cost", C.c_double), ("rms_px", C.c_double),
        ("num_iterations", C.c_int32), ("num_successful_steps", C.c_int32), ("termination_type", C.c_int32),
        ("linear_solver_used", C.c_int32), ("total_linear_iterations", C.c_int32), ("n_threads", C.c_int32),
        ("eval_seconds", C.c_double), ("linear_seconds", C.c_double), ("total_seconds", C.c_double),
        ("trace", C.POINTER(OIteration)), ("trace_capacity", C.c_int32), ("trace_count", C.c_int32),
        ("fixed_cost", C.c_double), ("setup_seconds", C.c_double),
    ]


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO) or any(
        os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_SO)
        for f in ("ba_oracle.cpp", "ba_oracle_chol.cpp", "ba_oracle.h", "ba_oracle_jet.h", "Makefile")
    ):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


def build_ref(force: bool = False):
    """oracle/_ref/libba_ref.so: the reference's own functor + problem-assembly text (CTracker.h:64-67,98-151,
    CTracker.cpp:571-702) compiled from where it lies under /root/reference against the stand-in headers
    (oracle/make_ref.py).  Returns None when neither the reference tree nor a prebuilt file is present."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("_make_ref", os.path.join(_HERE, "make_ref.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.build(force=force)


_lib = None
_ref = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_bp = C.POINTER(C.c_uint8)
_EVAL_FN = C.c_void_p


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.ba_oracle_default_options.argtypes = [C.POINTER(OOptions)]
        _lib.ba_oracle_eval.restype = C.c_int
        _lib.ba_oracle_eval.argtypes = [C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp, _dp, _dp, _dp, C.c_int]
        _lib.ba_oracle_schur.restype = C.c_int
        _lib.ba_oracle_schur.argtypes = [C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp,
                                         C.POINTER(OOptions), C.c_double, _dp, _dp, _dp, _dp, _dp, _dp, _dp]
        _lib.ba_oracle_solve.restype = C.c_int
        _lib.ba_oracle_solve.argtypes = [C.c_int, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp,
                                         C.POINTER(OOptions), C.POINTER(OSummary)]
        _lib.ba_oracle_solve_ex.restype = C.c_int
        _lib.ba_oracle_solve_ex.argtypes = [C.c_int, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, _dp, C.c_int64, _dp, _bp, _bp,
                                            C.POINTER(OOptions), C.POINTER(OSummary), _EVAL_FN, C.c_void_p]
        _lib.ba_oracle_cholesky.restype = C.c_int
        _lib.ba_oracle_cholesky.argtypes = [_dp, C.c_int64, C.c_int]
    return _lib


def ref_lib():
    """libba_ref.so or None (reference tree absent and nothing prebuilt)."""
    global _ref
    if _ref is None:
        so = build_ref()
        if so is None or not os.path.exists(so):
            return None
        _ref = C.CDLL(so)
        _ref.ba_ref_eval.restype = C.c_int
        _ref.ba_ref_eval.argtypes = [C.c_int, C.c_int64, _dp, _dp, _dp, _dp, _dp, _dp, _dp]
        _ref.ba_ref_solve.restype = C.c_int
        _ref.ba_ref_solve.argtypes = [C.c_int, C.c_int64, _dp, _ip, _ip, C.c_int32, _dp, _dp, C.c_int64, _dp, C.POINTER(OOptions),
                                      C.POINTER(OSummary)]
    return _ref


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _b(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint8).ctypes.data_as(_bp)


def _i(a):
    return a.ctypes.data_as(_ip)


def default_options(**kw) -> OOptions:
    o = OOptions()
    lib().ba_oracle_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def evaluate(prob, cam=None, pts=None, want_jac=True, threads=1):
    """r [n,2], J [n,2,9], cost by dual numbers (reference functor CTracker.cpp:585-604)."""
    cam = prob.cam if cam is None else cam
    pts = prob.pts if pts is None else pts
    n = prob.n_obs
    r = np.empty((n, 2))
    J = np.empty((n, 2, 9)) if want_jac else None
    cost = C.c_double()
    rc = lib().ba_oracle_eval(n, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), prob.n_cams, _d(prob.K), _d(prob.dist),
                              _d(cam), prob.n_pts, _d(pts), _d(r), _d(J), C.byref(cost), threads)
    if rc:
        raise RuntimeError(f"ba_oracle_eval -> {rc}")
    return r, J, cost.value


def frame_error_stats(prob, cam=None, pts=None):
    """Per-keyframe reprojection statistics as the reference computes them in the block it keeps commented out
    (CSfM.cpp:350-382): walking the observations frame by frame, dist = squared pixel distance between the
    reprojected point and the observation (CSfM.cpp:376), meanDist /= count (:381), and the two numbers handed
    to updateFrameErrorStatistics are sqrt(meanDist), sqrt(maxDist) (:382).  Plain loop over the oracle's
    residuals (r = projected - observed, CTracker.cpp:601-602)."""
    r, _, _ = evaluate(prob, cam, pts, want_jac=False)
    mean = np.zeros(prob.n_cams)
    mx = np.zeros(prob.n_cams)
    count = np.zeros(prob.n_cams, dtype=np.int64)
    for i in range(prob.n_obs):
        c = int(prob.cam_idx[i])
        dist = r[i, 0] * r[i, 0] + r[i, 1] * r[i, 1]
        mx[c] = dist if dist > mx[c] else mx[c]
        mean[c] += dist
        count[c] += 1
    mean = np.where(count > 0, mean / np.maximum(count, 1), 0.0)
    return np.sqrt(mean), np.sqrt(mx)


def schur(prob, radius=1e4, opts=None, cam=None, pts=None):
    cam = prob.cam if cam is None else cam
    pts = prob.pts if pts is None else pts
    opts = opts or default_options()
    nc, npt = prob.n_cams, prob.n_pts
    out = dict(S=np.zeros((6 * nc, 6 * nc)), rhs=np.zeros(6 * nc), Hcc=np.zeros((nc, 6, 6)), gc=np.zeros((nc, 6)),
               Hpp=np.zeros((npt, 3, 3)), gp=np.zeros((npt, 3)), scale=np.zeros(6 * nc + 3 * npt))
    rc = lib().ba_oracle_schur(prob.n_obs, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), nc, _d(prob.K), _d(prob.dist),
                               _d(cam), npt, _d(pts), C.byref(opts), radius, _d(out["S"]), _d(out["rhs"]), _d(out["Hcc"]),
                               _d(out["gc"]), _d(out["Hpp"]), _d(out["gp"]), _d(out["scale"]))
    if rc:
        raise RuntimeError(f"ba_oracle_schur -> {rc}")
    return out


def solve(prob, mode=2, opts=None, trace_capacity=64, cam_fixed=None, pt_fixed=None):
    """In-place Ceres-semantics LM on copies; returns (cam, pts, summary dict, trace list).  cam_fixed / pt_fixed:
    optional boolean masks of parameter blocks held constant (ceres::Problem::SetParameterBlockConstant)."""
    opts = opts or default_options()
    cam, pts = prob.cam.copy(), prob.pts.copy()
    trace = (OIteration * trace_capacity)()
    s = OSummary()
    s.trace = C.cast(trace, C.POINTER(OIteration))
    s.trace_capacity = trace_capacity
    cf = None if cam_fixed is None else np.ascontiguousarray(cam_fixed, dtype=np.uint8)
    pf = None if pt_fixed is None else np.ascontiguousarray(pt_fixed, dtype=np.uint8)
    rc = lib().ba_oracle_solve_ex(mode, prob.n_obs, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), prob.n_cams, _d(prob.K),
                                  _d(prob.dist), _d(cam), prob.n_pts, _d(pts), _b(cf), _b(pf), C.byref(opts), C.byref(s), None, None)
    if rc:
        raise RuntimeError(f"ba_oracle_solve -> {rc}")
    summ = {f: getattr(s, f) for f, _ in s._fields_ if f != "trace"}
    return cam, pts, summ, [trace[i].as_dict() for i in range(s.trace_count)]


def ref_evaluate(which, obs, K9, R, t, X, want_jac=True):
    """One observation per row through the REFERENCE's own functor text (oracle/_ref): which = 2
    BAStructAndPoseFunctor (CTracker.cpp:585-604), 1 BAPoseFunctor (:615-636), 0 BAStructFunctor (:647-668).
    Returns (r [n,2], J [n,2,9]); Jacobian columns of the functor's constant arguments are zero."""
    n = len(obs)
    a = [np.ascontiguousarray(v, dtype=np.float64) for v in (obs, K9, R, t, X)]
    r = np.empty((n, 2))
    J = np.zeros((n, 2, 9)) if want_jac else None
    rc = ref_lib().ba_ref_eval(which, n, *[_d(v) for v in a], _d(r), _d(J))
    if rc:
        raise RuntimeError(f"ba_ref_eval -> {rc}")
    return r, J


def ref_solve(prob, mode=2, opts=None, trace_capacity=64):
    """The reference call itself: CTracker::bundleAdjustmentStructAndPose (CTracker.cpp:670-702, verbatim in oracle/_ref)
    on arguments laid out as CSfM.cpp:311-341 builds them; the stand-in ceres::Solve runs the oracle's trust-region loop
    over the reference's own cost functions.  Returns (cam, pts, summary dict, trace list); summary["n_parameter_blocks"]
    = the parameter blocks found by pointer identity."""
    opts = opts or default_options()
    cam, pts = prob.cam.copy(), prob.pts.copy()
    trace = (OIteration * trace_capacity)()
    s = OSummary()
    s.trace = C.cast(trace, C.POINTER(OIteration))
    s.trace_capacity = trace_capacity
    rc = ref_lib().ba_ref_solve(mode, prob.n_obs, _d(prob.obs), _i(prob.cam_idx), _i(prob.pt_idx), prob.n_cams, _d(prob.K), _d(cam),
                                prob.n_pts, _d(pts), C.byref(opts), C.byref(s))
    if rc:
        raise RuntimeError(f"ba_ref_solve -> {rc}")
    summ = {f: getattr(s, f) for f, _ in s._fields_ if f != "trace"}
    summ["n_parameter_blocks"] = summ.pop("total_linear_iterations")
    return cam, pts, summ, [trace[i].as_dict() for i in range(s.trace_count)]


def cholesky(A, threads=1):
    L = np.array(A, dtype=np.float64, order="C")
    rc = lib().ba_oracle_cholesky(_d(L), L.shape[0], threads)
    return rc, np.tril(L)
