"""ctypes mirrors of the PODs in include/hsfm_ba.h (kept field-for-field in sync with the header;
tests/test_abi.py checks sizes against the compiled library)."""
import ctypes as C


class Options(C.Structure):
    _fields_ = [
        ("max_num_iterations", C.c_int32),
        ("max_consecutive_invalid_steps", C.c_int32),
        ("jacobi_scaling", C.c_int32),
        ("linear_solver", C.c_int32),
        ("dense_max_cams", C.c_int32),
        ("pcg_max_iterations", C.c_int32),
        ("verbose", C.c_int32),
        ("dense_flags", C.c_int32),  # same slot as the oracle's num_threads
        ("initial_trust_region_radius", C.c_double),
        ("max_trust_region_radius", C.c_double),
        ("min_trust_region_radius", C.c_double),
        ("min_relative_decrease", C.c_double),
        ("min_lm_diagonal", C.c_double),
        ("max_lm_diagonal", C.c_double),
        ("function_tolerance", C.c_double),
        ("gradient_tolerance", C.c_double),
        ("parameter_tolerance", C.c_double),
        ("pcg_tolerance", C.c_double),
    ]


class Iteration(C.Structure):
    _fields_ = [
        ("cost", C.c_double),
        ("cost_change", C.c_double),
        ("gradient_max_norm", C.c_double),
        ("step_norm", C.c_double),
        ("relative_decrease", C.c_double),
        ("trust_region_radius", C.c_double),
        ("model_cost_change", C.c_double),
        ("step_is_valid", C.c_int32),
        ("step_is_successful", C.c_int32),
        ("linear_iterations", C.c_int32),
        ("reserved0", C.c_int32),
    ]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_ if f != "reserved0"}


class Summary(C.Structure):
    _fields_ = [
        ("initial_cost", C.c_double),
        ("final_cost", C.c_double),
        ("rms_px", C.c_double),
        ("num_iterations", C.c_int32),
        ("num_successful_steps", C.c_int32),
        ("termination_type", C.c_int32),
        ("linear_solver_used", C.c_int32),
        ("total_linear_iterations", C.c_int32),
        ("n_ranks", C.c_int32),
        ("gpu_launches", C.c_int64),
        ("setup_ms", C.c_double),
        ("h2d_ms", C.c_double),
        ("solve_ms", C.c_double),
        ("d2h_ms", C.c_double),
        ("trace", C.POINTER(Iteration)),
        ("trace_capacity", C.c_int32),
        ("trace_count", C.c_int32),
        ("phase_ms", C.c_double * 8),
        ("fixed_cost", C.c_double),
        ("chol_panels", C.c_int32),
        ("chol_critical_panels", C.c_int32),
        ("cameras_reordered", C.c_int32),
        ("chol_chains", C.c_int32),
        ("chol_flops", C.c_double),
    ]


STRUCT_ONLY, POSE_ONLY, STRUCT_AND_POSE = 0, 1, 2
LINEAR_AUTO, LINEAR_DENSE_SCHUR, LINEAR_PCG = 0, 1, 2
CONVERGENCE, NO_CONVERGENCE, FAILURE = 0, 1, 2
