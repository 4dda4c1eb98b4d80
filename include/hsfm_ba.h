/*
 * hsfm_ba.h — C ABI of the B200-native bundle-adjustment solver that replaces the body of
 *
 *     void CTracker::bundleAdjustmentStructAndPose(const vector<Point2d>& observations,
 *             const vector<int>& camIdx, const vector<Matx33d>& K, vector<double*>& R,
 *             vector<double*>& t, vector<double*>& pts3D, int isStructOrPose)
 *
 * of hulop/SfM (declaration CTracker.h:65, definition CTracker.cpp:670-702; the only caller is
 * CSfM::bundleAdjustment, CSfM.cpp:310-348).  In the reference that body builds a ceres::Problem
 * (one AutoDiffCostFunction<…,2,3,3,3> per observation, CTracker.h:108-110) and calls
 * ceres::Solve with DENSE_SCHUR and otherwise default options (CTracker.cpp:571-577,700-701).
 *
 * Everything here is plain C: pointers, sizes, PODs.  No torch / STL / OpenCV types cross this
 * boundary.  All floating point is IEEE double.  The library has NO CPU fallback: every entry
 * point that computes returns HSFM_BA_ERR_NO_DEVICE when no sm_100 device is usable.
 *
 * Return convention: 0 = ok, <0 = invalid argument / state (HSFM_BA_ERR_*), >0 = CUDA (1000+e)
 * or NCCL (2000+e) error code.  Nothing throws across the ABI.  On any non-zero return the
 * caller's parameter arrays are left untouched (Ceres leaves the best-so-far x; for a failed
 * first step that is the input — CTracker.cpp:700-701 drops the summary, so this is the only
 * observable behaviour).
 */
#ifndef HSFM_BA_H_
#define HSFM_BA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HSFM_BA_ABI_VERSION 2

/* enum BA_TYPE, CTracker.h:67.  Any other value: the reference adds no residual block
 * (CTracker.cpp:692-693) and Solve() returns immediately; we return 0 and touch nothing. */
#define HSFM_BA_STRUCT_ONLY 0     /* cameras constant, points free  (BAStructFunctor, CTracker.cpp:647-668) */
#define HSFM_BA_POSE_ONLY 1       /* points constant, cameras free  (BAPoseFunctor,   CTracker.cpp:615-636) */
#define HSFM_BA_STRUCT_AND_POSE 2 /* both free                      (BAStructAndPoseFunctor, :585-604)     */

/* error codes (<0) */
#define HSFM_BA_OK 0
#define HSFM_BA_ERR_ARG (-1)         /* null pointer, negative count, index out of range            */
#define HSFM_BA_ERR_NO_DEVICE (-2)   /* no CUDA device / not sm_100: there is no CPU path           */
#define HSFM_BA_ERR_STATE (-3)       /* call order (e.g. solve before upload)                       */
#define HSFM_BA_ERR_UNSUPPORTED (-4) /* e.g. dense Schur requested above the dense camera limit     */
#define HSFM_BA_ERR_NOMEM (-5)
#define HSFM_BA_ERR_COMM (-6)        /* multi-rank requested but NCCL could not be loaded/initialised */

/* termination types, same order as ceres::TerminationType */
#define HSFM_BA_CONVERGENCE 0
#define HSFM_BA_NO_CONVERGENCE 1
#define HSFM_BA_FAILURE 2

/* linear solver for the reduced camera system */
#define HSFM_BA_LINEAR_AUTO 0        /* Cholesky when n_cams <= dense_max_cams, or up to 16x that when the reduced camera
                                        matrix is a narrow envelope (< 15 % of the triangle) and fits in memory; else PCG */
#define HSFM_BA_LINEAR_DENSE_SCHUR 1 /* what the reference selects, CTracker.cpp:574                 */
#define HSFM_BA_LINEAR_PCG 2         /* implicit Schur complement, block-Jacobi preconditioned CG    */

/* Options.  hsfm_ba_default_options() fills in ceres::Solver::Options defaults (Ceres 1.12; the
 * reference overrides only linear_solver_type / logging, CTracker.cpp:571-577). */
typedef struct hsfm_ba_options {
  int32_t max_num_iterations;            /* 50   */
  int32_t max_consecutive_invalid_steps; /* 5    */
  int32_t jacobi_scaling;                /* 1    */
  int32_t linear_solver;                 /* HSFM_BA_LINEAR_AUTO */
  int32_t dense_max_cams;                /* 2048: largest camera count solved by dense Cholesky under AUTO */
  int32_t pcg_max_iterations;            /* 1000 */
  int32_t verbose;                       /* 0; 1 prints a Ceres-like progress table to stderr */
  int32_t dense_flags;                   /* 0; bit 0: factor S as a full dense matrix (ignore its envelope), bit 1: form S
                                            from unsorted observation pairs.  Same result either way; for tests and benches */
  double initial_trust_region_radius;    /* 1e4  */
  double max_trust_region_radius;        /* 1e16 */
  double min_trust_region_radius;        /* 1e-32*/
  double min_relative_decrease;          /* 1e-3 */
  double min_lm_diagonal;                /* 1e-6 */
  double max_lm_diagonal;                /* 1e32 */
  double function_tolerance;             /* 1e-6 */
  double gradient_tolerance;             /* 1e-10*/
  double parameter_tolerance;            /* 1e-8 */
  double pcg_tolerance;                  /* 1e-10: stop when |r|_{M^-1} <= tol * |b|_{M^-1}  */
} hsfm_ba_options;

/* One row per trust-region iteration (row 0 = the initial evaluation), the columns Ceres'
 * IterationSummary carries. */
typedef struct hsfm_ba_iteration {
  double cost;                /* 1/2 sum r^2 at the accepted x after this iteration */
  double cost_change;         /* x_cost - candidate_cost (0 for invalid steps)      */
  double gradient_max_norm;   /* max |J^T r| (unscaled J)                           */
  double step_norm;           /* |delta| in parameter space                          */
  double relative_decrease;   /* rho                                                 */
  double trust_region_radius; /* radius AFTER this iteration's update                */
  double model_cost_change;
  int32_t step_is_valid;
  int32_t step_is_successful;
  int32_t linear_iterations; /* PCG iterations (0 for dense Cholesky) */
  int32_t reserved0;
} hsfm_ba_iteration;

typedef struct hsfm_ba_summary {
  double initial_cost;
  double final_cost;
  double rms_px;               /* sqrt(2*final_cost/n_obs) */
  int32_t num_iterations;      /* trust-region attempts = Ceres iterations.size()-1 */
  int32_t num_successful_steps;
  int32_t termination_type;    /* HSFM_BA_CONVERGENCE / NO_CONVERGENCE / FAILURE */
  int32_t linear_solver_used;  /* HSFM_BA_LINEAR_DENSE_SCHUR or _PCG */
  int32_t total_linear_iterations;
  int32_t n_ranks;
  int64_t gpu_launches;        /* kernels of this library launched by the call */
  /* wall/device times of the call, milliseconds */
  double setup_ms;             /* index build: sort by point / by camera, tiles (device) */
  double h2d_ms;
  double solve_ms;             /* LM loop, device events */
  double d2h_ms;
  /* optional per-iteration trace: caller supplies the array, we fill min(n, capacity) rows */
  hsfm_ba_iteration* trace;
  int32_t trace_capacity;
  int32_t trace_count;
  /* device time of the solve by phase (CUDA events on the solve stream), milliseconds:
   * [0] linearise (residual+Jacobian, point and camera blocks, gradient norm)
   * [1] reduced-system formation (point inverses, explicit S or block-Jacobi blocks, reduced rhs)
   * [2] dense Cholesky factorisation      [3] triangular solves or PCG iterations
   * [4] step, back-substitution, model cost change, candidate cost
   * [5] scalar read-back + host decision   [6],[7] reserved */
  double phase_ms[8];
  /* ---- ABI version 2 ---- */
  double fixed_cost;            /* 1/2 sum r^2 of residuals whose parameter blocks are all constant (ceres Summary::fixed_cost);
                                   included in initial_cost / final_cost, not in the per-iteration costs */
  int32_t chol_panels;          /* 128-column panels of the reduced camera system (0: no Cholesky in this solve) */
  int32_t chol_critical_panels; /* panels on the longest serial chain of the factorisation: = chol_panels with one chain, fewer with concurrent chains */
  int32_t cameras_reordered;    /* 1: the problem is held in an internal camera order (nested two-ended order and / or reverse
                                   Cuthill-McKee, hsfm_ba_set_ordering); results come back in the caller's order */
  int32_t chol_chains;           /* panel chains (segments of the block elimination tree) the factorisation ran as; 1 = one chain */
  double chol_flops;            /* FP64 flops of ONE factorisation of the reduced camera system over its block structure:
                                   sum over panels of 128^3/3 (diagonal block) + |R_K| 128^3 (panel solve) + |R_K|(|R_K|+1) 128^3
                                   (trailing update); n^3/3 when factored as a dense matrix */
} hsfm_ba_summary;

typedef struct hsfm_ba_ctx hsfm_ba_ctx;

/* ---- context ---------------------------------------------------------------------------- */

/* One context per process per GPU.  device_id: CUDA ordinal.  The context owns a stream and
 * caches device buffers across the per-keyframe calls of CSfM::mapping (CSfM.cpp:252-259). */
int hsfm_ba_create(hsfm_ba_ctx** out, int device_id);
void hsfm_ba_destroy(hsfm_ba_ctx* ctx);
int hsfm_ba_abi_version(void);
const char* hsfm_ba_error_string(int code);
void hsfm_ba_default_options(hsfm_ba_options* opts);

/* Run all subsequent work of this context on an externally owned cudaStream_t (passed as
 * void*), e.g. a harness' timing stream.  NULL restores the context's own stream. */
int hsfm_ba_set_stream(hsfm_ba_ctx* ctx, void* cuda_stream);

/* Internal camera order of subsequently uploaded problems.  0 (default): chosen by the library — a banded camera
 * graph (a scan trajectory) is cut into two or four pieces, each numbered from both ends towards its middle, so that
 * the Cholesky of the reduced camera system runs as concurrent panel chains that meet in small separator blocks;
 * results are returned in the caller's order either way.  1: keep the caller's order (one chain).  The solution is the
 * same up to rounding. */
int hsfm_ba_set_ordering(hsfm_ba_ctx* ctx, int ordering);

/* Host-only tap (no device needed): the camera order, the 128-row block structure of the Cholesky factor and the
 * panel-chain schedule the library derives from an envelope cam_first[c] = smallest camera sharing a point with
 * camera c (caller order).  natural > 0: keep the caller's order; natural = -1 - d: the nested two-ended order of
 * dissection depth d (2, 4, 8 chains for d = 0, 1, 2) whatever the cost model prefers.  new_of_old [n_cams] (optional) internal position
 * of every camera; info [8] = {panels, first panel of the second chain or -1, first common panel or -1 (both set only
 * for the plain two-chain schedule), structured factorisation used (0/1), reordered (0/1), first camera of the second
 * chain or -1, panel chains (segments; 0 = one chain), panels on the critical path}; fill (optional) = structural
 * blocks / triangle blocks; structure (optional) [panels][panels] bytes, lower triangle, 1 = block of L may be non-zero;
 * seg3 (optional) [3 * info[6]] = {first panel, end panel, parent segment or -1} of every panel chain, children before
 * parents: chains that are not ancestor / descendant run concurrently (size it for 3 * 32 ints). */
int hsfm_ba_debug_structure(int32_t n_cams, const int32_t* cam_first, int32_t natural, int32_t* new_of_old, int32_t* info,
                            double* fill, uint8_t* structure);
int hsfm_ba_debug_schedule(int32_t n_cams, const int32_t* cam_first, int32_t natural, int32_t* seg3);

/* Host-only tap: the reverse Cuthill-McKee camera order the upload falls back to when the cameras do NOT arrive in a
 * banded order (envelope of the delivered order > 35 % of the triangle; the reference's dense LLT is indifferent to the
 * order, CTracker.cpp:574).  edges [2 * n_edges]: undirected camera pairs (each once); position [n_cams] receives the
 * new position of every camera.  On the device the graph is one edge per observation (camera, smallest camera of its
 * point), and the envelope of the new order is measured before it is adopted. */
int hsfm_ba_debug_rcm(int32_t n_cams, int64_t n_edges, const int32_t* edges, int32_t* position);

/* Multi-GPU (one process per GPU, observations sharded by point, cameras replicated).
 * unique_id: the 128-byte ncclUniqueId produced by hsfm_ba_comm_unique_id() on rank 0 and
 * broadcast by the host (torch.distributed / MPI / a file).  After this call every solve on
 * the context is collective over the n_ranks contexts. */
int hsfm_ba_comm_unique_id(void* unique_id_128);
int hsfm_ba_comm_init(hsfm_ba_ctx* ctx, const void* unique_id_128, int rank, int n_ranks);

/* ---- one-shot solves on HOST buffers (what the reference call site binds) ----------------- */

/* Pointer form — argument-for-argument the reference signature (CTracker.h:65) in C:
 *   obs_xy   = observations.data()   (cv::Point2d = 2 doubles)            [n_obs][2]
 *   cam_idx  = camIdx.data()         position in the frame list, CSfM.cpp:336-340
 *   K9       = K.data()              (cv::Matx33d row-major; k0,k1,k2,k4,k5 used, CTracker.cpp:597-598)
 *   R, t     = R.data(), t.data()    n_cams pointers to 3 doubles each (angle-axis, translation)
 *   pts3d    = pts3D.data()          n_obs pointers to 3 doubles; POINTER IDENTITY is the point id
 *                                    (CSfM.cpp:331, CMap.cpp:215)
 * Optimises in place through the pointers.  summary may be NULL. */
int hsfm_ba_solve_ptr(hsfm_ba_ctx* ctx, int mode, int64_t n_obs, const double* obs_xy,
                      const int32_t* cam_idx, int32_t n_cams, const double* K9,
                      double* const* R, double* const* t, double* const* pts3d,
                      const hsfm_ba_options* opts, hsfm_ba_summary* summary);

/* Packed form — same problem with explicit point indices.
 *   cam_rt [n_cams][6] = (angle-axis, translation) per camera, updated in place
 *   pts    [n_pts][3]  updated in place
 *   dist   [n_cams][2] radial distortion k1,k2 (x_d = x_n (1 + k1 r^2 + k2 r^4)) or NULL.
 *          Extension named by the north star; NULL / zeros reproduce the reference functor
 *          bit for bit. */
int hsfm_ba_solve_packed(hsfm_ba_ctx* ctx, int mode, int64_t n_obs, const double* obs_xy,
                         const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
                         const double* K9, const double* dist, double* cam_rt, int64_t n_pts,
                         double* pts, const hsfm_ba_options* opts, hsfm_ba_summary* summary);

/* Packed form with CONSTANT parameter blocks (gauge fixing / windowed BA; the reference intends a
 * covisibility-limited camera set with the oldest keyframe fixed — CMap.cpp:80-112,317-330, CSfM.cpp:123,254 — and
 * would express it with ceres::Problem::SetParameterBlockConstant on R[c], t[c] / pts3D[i]).
 *   cam_fixed [n_cams], pt_fixed [n_pts]: non-zero = the camera (both its R and t block) / the point keeps its value
 *   bit for bit; either may be NULL (= none).  Residuals whose blocks are all constant leave the reduced program;
 *   their cost is reported as summary->fixed_cost and included in initial_cost / final_cost, as Ceres does. */
int hsfm_ba_solve_packed_ex(hsfm_ba_ctx* ctx, int mode, int64_t n_obs, const double* obs_xy,
                            const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams, const double* K9,
                            const double* dist, double* cam_rt, int64_t n_pts, double* pts,
                            const uint8_t* cam_fixed, const uint8_t* pt_fixed, const hsfm_ba_options* opts,
                            hsfm_ba_summary* summary);

/* ---- parity taps on HOST buffers ---------------------------------------------------------- */

/* Residuals and Jacobians exactly as ceres::AutoDiffCostFunction<…,2,3,3,3> lays them out:
 *   residuals [n_obs][2];  jac [n_obs][2][9] row-major, columns = dR(3) dt(3) dX(3);
 *   cost = 1/2 sum r^2.  Output order = input observation order.  Any output may be NULL. */
int hsfm_ba_eval(hsfm_ba_ctx* ctx, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx,
                 const int32_t* pt_idx, int32_t n_cams, const double* K9, const double* dist,
                 const double* cam_rt, int64_t n_pts, const double* pts, double* residuals,
                 double* jac, double* cost);

/* Normal-equation pieces at the given x with LM diagonal D^2 = clamp(diag(J^T J))/radius
 * (jacobi_scaling as in opts; all in the scaled space Ceres solves in):
 *   S_dense [6 n_cams][6 n_cams] row-major reduced camera matrix (both triangles),
 *   rhs     [6 n_cams]           reduced right-hand side  g_c - W Hpp^-1 g_p,
 *   Hcc     [n_cams][36]         camera blocks incl. D_c^2,   gc [n_cams][6],
 *   Hpp     [n_pts][9]           point blocks incl. D_p^2,    gp [n_pts][3],
 *   scale   [6 n_cams + 3 n_pts] Jacobi column scaling 1/(1+|col|).
 * Any output may be NULL. */
int hsfm_ba_schur(hsfm_ba_ctx* ctx, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx,
                  const int32_t* pt_idx, int32_t n_cams, const double* K9, const double* dist,
                  const double* cam_rt, int64_t n_pts, const double* pts,
                  const hsfm_ba_options* opts, double radius, double* S_dense, double* rhs,
                  double* Hcc, double* gc, double* Hpp, double* gp, double* scale);

/* Dense SPD solve tap: the blocked DMMA Cholesky + triangular solves that stand in for the Eigen LLT
 * of Ceres' DENSE_SCHUR (CTracker.cpp:574), on a caller-supplied row-major n×n matrix (lower
 * triangle read; n even, as 6*n_cams always is).  x [n] = A^-1 b; L_out (optional) [n][n] receives the
 * lower factor.  Returns HSFM_BA_ERR_ARG with *pivot (optional) = 1-based failing column when A is
 * not positive definite. */
int hsfm_ba_dense_spd_solve(hsfm_ba_ctx* ctx, int32_t n, const double* A, const double* b, double* x,
                            double* L_out, int32_t* pivot);

/* Measurement tap for the same kernels: builds a synthetic SPD matrix of order n on the device
 * (n even), factors it `reps` times (after one warm-up) and solves one right-hand side each time;
 * returns the average device time of the factorisation and of the two triangular solves in ms
 * (CUDA events on the context's stream) and the max-norm residual |A x - b| / |b| of the last solve. */
int hsfm_ba_bench_cholesky(hsfm_ba_ctx* ctx, int32_t n, int reps, double* factor_ms, double* solve_ms,
                           double* rel_residual);

/* ---- device-resident problem (bench / repeated solves; inputs stay in HBM) ----------------- */

/* Upload + index build (sort by point, pack tiles, camera-major permutation).  Under a
 * multi-rank communicator each rank passes ITS shard of the observations (a disjoint set of
 * points with all their observations; pt_idx local to the rank, cameras global). */
int hsfm_ba_problem_upload(hsfm_ba_ctx* ctx, int mode, int64_t n_obs, const double* obs_xy,
                           const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
                           const double* K9, const double* dist, const double* cam_rt,
                           int64_t n_pts, const double* pts);
/* ---- scan: a resident problem that grows by a keyframe per call ----------------------
 * The reference re-gathers and re-sends its whole map at every keyframe (CSfM::mapping -> CSfM::bundleAdjustment,
 * CSfM.cpp:252-259, 310-348; two heap objects and a hash insert per observation in CTracker.cpp:676-696).  A scan keeps
 * the raw arrays of the problem in device memory between the calls:
 *   hsfm_ba_scan_begin            = hsfm_ba_problem_upload that also starts the device-resident table;
 *   hsfm_ba_scan_append_keyframe  adds n_new_cams cameras (ids n_cams, n_cams+1, ...), n_new_pts points (ids n_pts, ...) and
 *                                 n_new_obs observations between ANY cameras / points of the grown problem (a new map point
 *                                 brings its observations in older keyframes with it).  Only these rows cross PCIe; the
 *                                 poses / points refined by earlier solves are taken from the device as they are.
 * After either call the problem is resident and indexed exactly as a one-shot upload of the concatenated arrays would
 * be: hsfm_ba_problem_solve / _download / _frame_stats / _set_constant apply (constant-block masks are cleared by an
 * append).  bytes_uploaded (optional) receives the host->device bytes of the call.  Invalid new indices return
 * HSFM_BA_ERR_ARG and leave the scan as it was. */
int hsfm_ba_scan_begin(hsfm_ba_ctx* ctx, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx,
                       const int32_t* pt_idx, int32_t n_cams, const double* K9, const double* dist,
                       const double* cam_rt, int64_t n_pts, const double* pts, int64_t* bytes_uploaded);
int hsfm_ba_scan_append_keyframe(hsfm_ba_ctx* ctx, int32_t n_new_cams, const double* K9, const double* dist,
                                 const double* cam_rt, int64_t n_new_pts, const double* pts, int64_t n_new_obs,
                                 const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                                 int64_t* bytes_uploaded);

/* Constant-block masks of the resident problem (see hsfm_ba_solve_packed_ex), caller's camera / point order; both NULL
 * clears them.  They stay in force for every later solve of this problem (a new upload clears them). */
int hsfm_ba_problem_set_constant(hsfm_ba_ctx* ctx, const uint8_t* cam_fixed, const uint8_t* pt_fixed);
/* Restore the uploaded initial parameters on the device (no host traffic). */
int hsfm_ba_problem_reset(hsfm_ba_ctx* ctx);
/* LM solve on the resident problem; parameters stay on the device. */
int hsfm_ba_problem_solve(hsfm_ba_ctx* ctx, const hsfm_ba_options* opts, hsfm_ba_summary* summary);
/* One residual + Jacobian pass (kernel RJ) on the resident problem; cost (host double) may be
 * NULL, in which case nothing is copied back and the call is fully asynchronous. */
int hsfm_ba_problem_eval(hsfm_ba_ctx* ctx, double* cost);
/* Copy the current parameters back. */
int hsfm_ba_problem_download(hsfm_ba_ctx* ctx, double* cam_rt, double* pts);
/* Per-keyframe reprojection statistics at the current parameters: for camera c over its
 * observations, rms_px[c] = sqrt(mean d^2) and max_px[c] = sqrt(max d^2), d = reprojection distance in pixels —
 * the two numbers CSfM::bundleAdjustment hands to CKeyFrame::updateFrameErrorStatistics in the block the reference
 * keeps commented out (CSfM.cpp:350-382; TODO at CSfM.cpp:24-27).  Either array ([n_cams]) may be NULL; cameras
 * without observations get 0.  With several ranks every rank receives the statistics over all shards. */
int hsfm_ba_problem_frame_stats(hsfm_ba_ctx* ctx, double* rms_px, double* max_px);
/* Algorithmic byte counts of one launch of the named kernel on the resident problem
 * ("rj", "schur_matvec", "cam_form", "backsub"); 0 if unknown.  DESIGN.md derives them. */
int64_t hsfm_ba_problem_algorithmic_bytes(hsfm_ba_ctx* ctx, const char* kernel);
/* Time `reps` back-to-back launches of the named kernel on the resident problem with CUDA
 * events on the context's stream; returns average ms per launch in *ms. */
int hsfm_ba_problem_time_kernel(hsfm_ba_ctx* ctx, const char* kernel, int reps, double* ms);

/* ---- allocation-free entry point ---------------------------------------------------------------- */

/* No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * Reprojection cost 1/2 sum |r|^2 of n_obs observations at the given parameters — the candidate-cost evaluation of
 * the trust-region loop (functor of CTracker.cpp:585-604, residuals only) — with every input and the output in
 * DEVICE memory owned by the caller:
 *   d_obs_xy [n_obs][2], d_cam_idx [n_obs], d_pt_idx [n_obs], d_K9 [n_cams][9], d_cam_rt [n_cams][6], d_pts [n_pts][3],
 *   d_cost   [1] (written).
 * The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine): its scratch memory, one
 * partial sum per 256 observations, is a static device array of HSFM_BA_NOALLOC_MAX_OBS / 256 doubles (128 KB) that
 * every GPU using the library holds until exit.  n_obs > HSFM_BA_NOALLOC_MAX_OBS returns HSFM_BA_ERR_UNSUPPORTED and
 * launches nothing.  The work is enqueued on the context's stream (hsfm_ba_set_stream) and NOT waited for: d_cost is
 * valid once that stream has been synchronised.  An out-of-range camera / point index makes the cost NaN.
 * All calls on one GPU share the scratch array (it keeps the last call's contents; nothing in it has to start at
 * zero): calls that overlap on the GPU — other streams or other contexts of the same device — race on it. */
#define HSFM_BA_NOALLOC_MAX_OBS 4194304 /* 2^22 observations */
int hsfm_ba_cost_noalloc(hsfm_ba_ctx* ctx, int64_t n_obs, const double* d_obs_xy, const int32_t* d_cam_idx,
                         const int32_t* d_pt_idx, int32_t n_cams, const double* d_K9, const double* d_cam_rt,
                         int64_t n_pts, const double* d_pts, double* d_cost);

#ifdef __cplusplus
}
#endif
#endif /* HSFM_BA_H_ */
