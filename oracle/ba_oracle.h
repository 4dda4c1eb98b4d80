/*
 * ba_oracle.h — CPU ORACLE for the bundle-adjustment hot path of hulop/SfM.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product library
 * (sfm_b200/csrc → libhsfm_ba.so) never links, loads or calls anything in this directory.
 *
 * PARITY UNPINNED: the reference ships no tests, fixtures or golden vectors for this path
 *  and its arithmetic lives in Ceres Solver ("3.12.0" sic = 1.12.0,
 * README.md:30), which is neither vendored in /root/reference nor installable here.  The
 * oracle therefore restates
 *   - the reference's own functor, operation for operation (CTracker.cpp:585-604, 615-636,
 *     647-668), instantiated on a dual-number type that follows ceres::Jet's published rules,
 *   - the problem structure of CTracker.cpp:670-702 (no loss, no constant blocks, R / t / point
 *     parameter blocks, block order R,t,X of CTracker.h:108-110),
 *   - ceres::AngleAxisRotatePoint, the Levenberg–Marquardt trust-region loop, Jacobi scaling and
 *     DENSE_SCHUR elimination as published for Ceres 1.12,
 * and is cross-checked against independent implementations available in the build container
 * (cv2.projectPoints' analytic Jacobian, mpmath 50-digit derivatives, a dense numpy
 * normal-equation solve): tests/golden/ holds those vectors with the generating script.
 */
#ifndef BA_ORACLE_H_
#define BA_ORACLE_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Field-for-field the same layout as hsfm_ba_options / hsfm_ba_iteration (include/hsfm_ba.h) so
 * one ctypes.Structure serves both; duplicated on purpose — the oracle includes nothing from
 * the product. */
typedef struct ba_oracle_options {
  int32_t max_num_iterations, max_consecutive_invalid_steps, jacobi_scaling, linear_solver;
  int32_t dense_max_cams, pcg_max_iterations, verbose, num_threads /* reserved0 in product */;
  double initial_trust_region_radius, max_trust_region_radius, min_trust_region_radius;
  double min_relative_decrease, min_lm_diagonal, max_lm_diagonal;
  double function_tolerance, gradient_tolerance, parameter_tolerance, pcg_tolerance;
} ba_oracle_options;

typedef struct ba_oracle_iteration {
  double cost, cost_change, gradient_max_norm, step_norm, relative_decrease, trust_region_radius,
      model_cost_change;
  int32_t step_is_valid, step_is_successful, linear_iterations, reserved0;
} ba_oracle_iteration;

typedef struct ba_oracle_summary {
  double initial_cost, final_cost, rms_px;
  int32_t num_iterations, num_successful_steps, termination_type, linear_solver_used;
  int32_t total_linear_iterations, n_threads;
  double eval_seconds;   /* time inside residual+Jacobian evaluation */
  double linear_seconds; /* Schur elimination + reduced solve + back-substitution */
  double total_seconds;
  ba_oracle_iteration* trace;
  int32_t trace_capacity, trace_count;
  double fixed_cost;     /* 1/2 sum r^2 over residuals whose parameter blocks are all constant (Ceres Summary::fixed_cost) */
  double setup_seconds;  /* index build + Jacobi scaling + the initial linearisation: everything before the first trust-region attempt */
} ba_oracle_summary;

void ba_oracle_default_options(ba_oracle_options* o);

/* r [n_obs][2], J [n_obs][2][9] (dR dt dX, row-major) by dual numbers; cost = 1/2 sum r^2. */
int ba_oracle_eval(int64_t n_obs, const double* obs_xy, const int32_t* cam_idx,
                   const int32_t* pt_idx, int32_t n_cams, const double* K9, const double* dist,
                   const double* cam_rt, int64_t n_pts, const double* pts, double* residuals,
                   double* jac, double* cost, int num_threads);

/* Same outputs as hsfm_ba_schur. */
int ba_oracle_schur(int64_t n_obs, const double* obs_xy, const int32_t* cam_idx,
                    const int32_t* pt_idx, int32_t n_cams, const double* K9, const double* dist,
                    const double* cam_rt, int64_t n_pts, const double* pts,
                    const ba_oracle_options* opts, double radius, double* S_dense, double* rhs,
                    double* Hcc, double* gc, double* Hpp, double* gp, double* scale);

/* Ceres-semantics LM in place on cam_rt / pts. */
int ba_oracle_solve(int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx,
                    const int32_t* pt_idx, int32_t n_cams, const double* K9, const double* dist,
                    double* cam_rt, int64_t n_pts, double* pts, const ba_oracle_options* opts,
                    ba_oracle_summary* summary);

/* Residual + Jacobian callback for ba_oracle_solve_ex: fills r [n_obs][2] and, when Jc != NULL, the UNSCALED
 * Jc [n_obs][12] = rows (du/d[R t], dv/d[R t]) and Jp [n_obs][6] = rows (du/dX, dv/dX) at (cam [n_cams][6], pts [n_pts][3]);
 * returns non-zero on failure.  oracle/_ref supplies one that calls the reference's own functors. */
typedef int (*ba_oracle_eval_fn)(void* user, const double* cam, const double* pts, double* r, double* Jc, double* Jp);

/* ba_oracle_solve with
 *   cam_fixed [n_cams] / pt_fixed [n_pts] (NULL = none): parameter blocks held constant, as
 *     ceres::Problem::SetParameterBlockConstant would (gauge fixing / windowed BA, SURVEY.md 8f row 2;
 *     reference intent CMap.cpp:80-112,317-330, CSfM.cpp:123,254);
 *   eval_fn (NULL = the built-in restatement of CTracker.cpp:585-604): external residual/Jacobian evaluation;
 *     obs_xy / K9 / dist are then unused and may be NULL. */
int ba_oracle_solve_ex(int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx,
                       const int32_t* pt_idx, int32_t n_cams, const double* K9, const double* dist,
                       double* cam_rt, int64_t n_pts, double* pts, const uint8_t* cam_fixed,
                       const uint8_t* pt_fixed, const ba_oracle_options* opts,
                       ba_oracle_summary* summary, ba_oracle_eval_fn eval_fn, void* eval_user);

/* In-place lower Cholesky of a dense row-major n×n SPD matrix (blocked, OpenMP); returns 0 or
 * the 1-based index of the first non-positive pivot.  Exposed for the Cholesky KATs. */
int ba_oracle_cholesky(double* A, int64_t n, int num_threads);

#ifdef __cplusplus
}
#endif
#endif
