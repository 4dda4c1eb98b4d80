// ba_internal.h — internal declarations shared by the translation units of libhsfm_ba.so.
// Device-resident problem state, kernel launch wrappers, the LM driver.  Not installed.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/hsfm_ba.h"
#include "ba_math.cuh"

namespace hsfm {

constexpr int TILE = 256;        // observations per point-major tile == threads per CTA
constexpr int CHUNK = 256;       // observations per camera-major warp chunk
constexpr int MAX_TRACK = TILE;  // a point's observations must fit one tile (checked at upload)

#define HSFM_CUDA(call)                                         \
  do {                                                          \
    cudaError_t e__ = (call);                                   \
    if (e__ != cudaSuccess) return 1000 + (int)e__;             \
  } while (0)

// Plain device pointers of one resident problem (passed to kernels by value).
struct Dev {
  int n_obs, n_cams, n_pts, n_tiles, n_chunks;
  // static, point-major order (sorted by point; ties keep input order)
  const double2* obs;    // [n_obs] observed pixel
  const int* cam;        // [n_obs]
  const int* pt;         // [n_obs]
  const int* pt_start;   // [n_pts+1] CSR over point-major observations
  const int* tile_start; // [n_tiles+1] first observation of each tile (tiles hold whole points)
  const int4* tile_hdr;  // [n_tiles][2] {first obs, end obs, first point, end point}, {distinct cameras, 0, 0, 0}: one
                         // 32-byte record so that a CTA learns everything about its tile in a single memory round trip
  const int* tile_cams;  // [n_tiles][TILE] distinct cameras of each tile, ascending (first hdr.ncam entries valid)
  const unsigned char* slot;  // [n_obs] position of the observation's camera in its tile's list
  const int* pair_off;        // [n_tiles+1] first entry of each tile in pair_list (null: more than 2^31 pairs)
  const unsigned* pair_list;  // per tile: the (i,j) observation pairs of its points SORTED by destination block of S,
                              // (slot_i << 24 | slot_j << 16 | i << 8 | j); built lazily by the first dense solve
  // camera-keyed sums without atomics and without a per-observation round trip: inside a tile the observations'
  // contributions are summed per camera slot in shared memory (srank = position in slot order, slot_begin = first
  // position of every slot), one partial per (tile, slot) ENTRY goes to HBM, and a camera's sum is the sum of its entries
  const unsigned char* srank;        // [n_obs]
  const unsigned char* sorder;       // [n_obs] inverse within the tile: tile-local index of the observation at slot-order position j
  const unsigned short* slot_begin;  // [n_tiles][TILE] (first hdr.ncam entries valid)
  const int* ent_off;                // [n_tiles+1] first entry of each tile; entry = ent_off[tile] + slot
  const int* cs_start;               // [n_cams+1] CSR over cs_ent
  const int* cs_ent;                 // [n_entries] entries of every camera, in tile order
  int n_entries;
  // static, camera-major order
  const double2* obs_cs; // [n_obs]
  const int* pt_cs;      // [n_obs]
  const int* pos_cs;     // [n_obs] point-major position of the j-th camera-major observation
  const int* chunk_cam;  // [n_chunks]
  const int* chunk_b;    // [n_chunks]
  const int* chunk_e;    // [n_chunks]
  const int* cam_chunk_start;  // [n_cams+1]
  // linearisation, point-major SoA planes of double2 = (d u, d v)
  double2* r;            // [n_obs]
  double2* Jc;           // [6][n_obs]
  double2* Jp;           // [3][n_obs]
};

// Pageable host memory -> device at link speed.  cudaMemcpyAsync from pageable memory goes through the driver's own
// staging at ~10 GB/s and blocks the caller; the reference's containers (std::vector, CSfM.cpp:311-343) are pageable.
// Here worker threads copy slices of the source into pinned rings and the DMA of one chunk runs while the next one is
// being copied: ~4 threads keep up with the link.  Sources that are pinned or device memory never come here.
struct HostStager {
  static constexpr int LANES = 4, CHUNKS = 3;
  static constexpr size_t CHUNK = (size_t)2 << 20, MIN_BYTES = (size_t)4 << 20;
  char* pinned = nullptr;  // [LANES][CHUNKS][CHUNK], allocated on first use
  cudaEvent_t ev[LANES][CHUNKS] = {};
  bool used[LANES][CHUNKS] = {};
  struct Job { void* dst; const void* src; size_t bytes; };
  static bool pageable(const void* p);  // plain host memory (neither registered / pinned nor device)
  int init();
  void destroy();
  int lane_copy(int lane, void* dst, const void* src, size_t bytes, cudaStream_t st);
  // lanes first_lane .. LANES-1 each take their slice of every job; on the calling thread (blocking) ...
  int copy(const std::vector<Job>& jobs, cudaStream_t st, int device);
  // ... or on background threads of their own: join() them before relying on the stream
  void start(const std::vector<Job>& jobs, cudaStream_t st, int device, std::vector<std::thread>& threads, int* rc);
};

// Everything the solver keeps on the device for one problem.
struct Problem {
  int mode = 2;
  int n_obs = 0, n_cams = 0, n_pts = 0, n_tiles = 0, n_chunks = 0;
  int max_track = 0;
  int n_sm = 148;         // SM count of the context's device (set by hsfm_ba_create; survives release())
  int n_entries = 0;      // (tile, slot) pairs: sum over tiles of their distinct cameras
  double* ent_part = nullptr;  // [n_entries][6] per-entry partial sums (grow-only, outside the arena like S)
  size_t ent_part_cap = 0;
  int max_tile_cams = 0;  // largest number of distinct cameras in one tile (sizes the shared-memory row staging)
  std::vector<int> cam_first;  // [n_cams] CALLER order: smallest camera sharing a point with camera c (all ranks) — envelope of S
  // internal camera order (ba_order.cpp): null / empty = the caller's order
  std::vector<int> h_cam_new_of_old;      // caller's camera -> internal camera
  std::vector<int> h_struct_new_of_base;  // base order (the caller's, or reverse Cuthill-McKee: what cam_first is in) -> internal; empty = identity
  int* cam_new_of_old = nullptr;  // device [n_cams]
  int* cam_old_of_new = nullptr;  // device [n_cams]
  int split_cams = -1;            // first internal camera of the second chain (two-ended order) or -1
  int tail_cams = -1;             // first internal camera of the tail (cameras coupled to both chains) or -1
  double* cam_io = nullptr;       // [n_cams][9] staging in the caller's camera order (upload / download / taps)
  HostStager stager;                   // pageable uploads (survives release())
  cudaStream_t copy_stream = nullptr;  // upload of the observations' coordinates and the points beside the index build (survives release())
  cudaEvent_t ev_copy_go = nullptr, ev_copy_done = nullptr;
  // The last ordering decision (survives release()): the per-keyframe calls of a scan and repeated solves mostly bring
  // the camera graph of the previous call back, and the host-side choice (a few ms of symbolic work per candidate
  // order at Venice size) is then skipped.
  struct OrderCache {
    bool valid = false, reordered = false;
    std::vector<int> first, new_of_old;  // envelope in base order -> base position to internal position
    int split_cams = -1, tail_cams = -1;
    bool rcm_valid = false;
    std::vector<int> first_caller, base_of_caller;  // envelope in the caller's order -> reverse Cuthill-McKee position of every camera
  } order_cache;
  Dev d{};
  // Device arena: ONE grow-only allocation per context that every per-problem buffer is carved from, so the
  // per-keyframe calls of CSfM::mapping (CSfM.cpp:252-259) pay for cudaMalloc/cudaFree only when the problem
  // outgrows the arena.  release() rewinds the arena; destroy() frees it (together with S and the pinned mirrors).
  char* arena = nullptr;
  size_t arena_cap = 0, arena_off = 0;
  int arena_reserve(size_t bytes);  // grow (never shrink); invalidates previous carve-outs
  void* arena_take(size_t bytes);   // 256-byte aligned; nullptr when exhausted
  void destroy();
  // input order → point-major position (for the eval tap) and static copies
  int* order_pm = nullptr;  // [n_obs] input index of the j-th point-major observation
  int* pt_old_of_new = nullptr;  // [n_pts] caller's point id of internal point k (internal order: by smallest observing camera)
  int* pt_new_of_old = nullptr;  // [n_pts] inverse
  double* pts_io = nullptr;      // [n_pts][3] staging in the caller's point order (upload / download / taps)
  double* K9 = nullptr;     // [n_cams][9]
  double* dist = nullptr;   // [n_cams][2]
  // parameters
  double *cam_x = nullptr, *pts_x = nullptr;        // current x   [n_cams][6], [n_pts][3]
  double *cam_c = nullptr, *pts_c = nullptr;        // candidate
  double *cam_x0 = nullptr, *pts_x0 = nullptr;      // uploaded initial values (problem_reset)
  CamRow *rows_x = nullptr, *rows_c = nullptr;      // prepared camera tables for x / candidate
  double *scale_c = nullptr, *scale_p = nullptr;    // Jacobi scaling [6 n_cams], [3 n_pts]
  int *cam_nobs = nullptr, *pt_nobs_mask = nullptr; // global observation count per camera (all ranks)
  // constant parameter blocks (hsfm_ba_problem_set_constant), internal camera / point order
  unsigned char *cam_fixed = nullptr, *pt_fixed = nullptr;  // [n_cams], [n_pts]; valid when has_fixed
  bool has_fixed = false;
  // normal-equation blocks
  double *Hpp = nullptr, *gp = nullptr, *Hpp_inv = nullptr, *zp = nullptr;  // [n_pts][6],[3],[6],[3]
  double *Hcc = nullptr, *gc = nullptr;             // [n_cams][21] upper-tri row-major, [n_cams][6]
  double *chunk_part = nullptr;                     // [n_chunks][27]
  double *tobs = nullptr;                           // [n_obs][6] per-observation camera-space vectors (point-major), gathered camera-major
  double *Dc2 = nullptr;                            // [6 n_cams] squared LM diagonal of the camera block
  double *Minv = nullptr;                           // [n_cams][36] block-Jacobi preconditioner / pose-only inverse
  double *E = nullptr;                              // [n_cams][21] sum_i Jc^T Jp Hpp^-1 Jp^T Jc
  double *rhs = nullptr, *yc = nullptr, *yp = nullptr;
  double *pcg_x = nullptr, *pcg_r = nullptr, *pcg_z = nullptr, *pcg_p = nullptr, *pcg_q = nullptr;
  double *S = nullptr;                              // dense reduced camera matrix (lazily allocated)
  size_t S_capacity = 0;
  unsigned* pair_list = nullptr;                    // sorted pair lists of dense_schur (lazily built, persists like S)
  size_t pair_capacity = 0;
  long long total_pairs = 0;
  bool pairs_ready = false;
  // scalar plumbing
  double *part = nullptr;    // [PART_SLOTS][part_stride] per-block partial sums
  int part_stride = 0;
  double *scal = nullptr;    // [32] reduced scalars (device)
  double *scal_host = nullptr;  // pinned mirror
  int *flags = nullptr;      // [8] device flags: 0 = not-PD / invalid, 1 = pcg done, 2 = pcg iterations
  int *flags_host = nullptr; // pinned mirror
  int *idx_host = nullptr;   // pinned staging of the index build (point CSR offsets), grow-only
  size_t idx_host_cap = 0;
  int64_t launches = 0;
  void release();
};

enum PartSlot { P_COST = 0, P_MCC, P_DN_P, P_XN_P, P_DN_C, P_XN_C, P_GMAX_C, P_GMAX_P, PART_SLOTS };
// slots 0..4 are per-rank partial sums (all-reduced with SUM under a communicator), S_GMAX_P with MAX;
// the camera-side slots are computed from replicated data and are identical on every rank.
enum ScalSlot { S_COST = 0, S_MCC, S_DN_P, S_XN_P, S_CAND_COST, S_GMAX_P, S_DN_C, S_XN_C, S_GMAX_C, S_PCG_RHO0, S_PCG_RHO, S_FIXED_COST, SCAL_SLOTS = 32 };
enum FlagSlot { F_INVALID = 0, F_PCG_DONE = 1, F_PCG_ITERS = 2, F_CHOL_FAIL = 3, FLAG_SLOTS = 8 };

// ---- communicator (NCCL, loaded at run time; ba_comm.cu) ----
struct Comm {
  void* handle = nullptr;  // ncclComm_t
  int rank = 0, n_ranks = 1;
  int allreduce_sum(double* buf, size_t count, cudaStream_t s);
  int allreduce_max(double* buf, size_t count, cudaStream_t s);
  int allreduce_sum_i32(int* buf, size_t count, cudaStream_t s);
  int allreduce_min_i32(int* buf, size_t count, cudaStream_t s);
  void destroy();
};
int comm_unique_id(void* out128);
int comm_init(Comm& c, const void* id128, int rank, int n_ranks);

// ---- index build (ba_index.cu) ----
int build_problem(Problem& P, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                  int32_t n_cams, const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts,
                  Comm& comm, cudaStream_t s, double* setup_ms, double* h2d_ms, bool natural_order = false);
void launch_set_fixed(Problem& P, const unsigned char* cam_caller, const unsigned char* pt_caller, cudaStream_t s);  // device staging -> internal order
// camera-indexed rows between the caller's camera order and the internal one
void launch_cam_rows_to_caller(Problem& P, int w, const double* internal_order, double* caller_order, cudaStream_t s);

// point-indexed rows between the caller's point order and the internal one (ba_index.cu)
void launch_rows_to_internal(Problem& P, int w, const double* caller_order, double* internal_order, cudaStream_t s);
void launch_rows_to_caller(Problem& P, int w, const double* internal_order, double* caller_order, cudaStream_t s);

// ---- kernel launchers (ba_kernels.cu); every launcher bumps P.launches ----
int init_kernel_attributes();  // per-device dynamic shared-memory opt-ins (call with the context's device current)
void launch_cam_prep(Problem& P, const double* cam, const double* scale_c /*null = 1*/, CamRow* rows, cudaStream_t s);
// residual + Jacobian (+ point blocks Hpp,gp when form) ; scale_p null = 1
void launch_rj(Problem& P, const CamRow* rows, const double* pts, const double* scale_p, bool write_j, bool form, cudaStream_t s);
void launch_cost(Problem& P, const CamRow* rows, const double* pts, int part_slot, cudaStream_t s);
// camera-major passes: mode 0 → Hcc,gc ; mode 1 → E (needs Hpp_inv)
void launch_cam_major(Problem& P, const CamRow* rows, const double* pts, const double* scale_p, int what, cudaStream_t s);
void launch_jacobi_scale(Problem& P, bool enabled, cudaStream_t s);  // from Hcc / Hpp diagonals (unscaled pass); 0 for constant blocks
void launch_fixed_cost(Problem& P, const CamRow* rows, const double* pts, cudaStream_t s);  // -> scal[S_FIXED_COST]
void launch_point_invert(Problem& P, double radius, double min_diag, double max_diag, cudaStream_t s);
void launch_cam_diag(Problem& P, double radius, double min_diag, double max_diag, cudaStream_t s);  // Dc2
void launch_precond_invert(Problem& P, bool with_E, cudaStream_t s);  // Minv = (Hcc + Dc2 - E)^-1
void launch_schur_rhs(Problem& P, bool preload_gc, cudaStream_t s);  // rhs = [gc] - sum Jc^T Jp zp
void launch_schur_matvec(Problem& P, const double* pvec, double* q, cudaStream_t s);  // q += (Jc^T (I - Jp Hpp^-1 Jp^T) Jc) p ; honours F_PCG_DONE
void launch_backsub(Problem& P, cudaStream_t s);        // yp, candidate points, model-cost-change and norm partials
void launch_cam_step(Problem& P, cudaStream_t s);       // candidate cameras, norm partials
void launch_block_solve_cams(Problem& P, cudaStream_t s);  // pose-only: yc = Minv gc
void launch_grad_norm(Problem& P, cudaStream_t s);      // max |g/scale| partials
void launch_finalize_scalars(Problem& P, cudaStream_t s);  // partials -> scal[]
void launch_frame_stats(Problem& P, const CamRow* rows, const double* pts, double* out3 /*[3 n_cams]: sum d^2, count, max d^2*/, cudaStream_t s);
void launch_pcg_init(Problem& P, cudaStream_t s);
void launch_pcg_step(Problem& P, double tol, int max_it, cudaStream_t s);
struct CholPlan;
void launch_dense_init(Problem& P, bool with_diag, cudaStream_t s, const CholPlan* plan = nullptr);  // S = 0 (+ blockdiag(Hcc + D^2)); plan: zero the envelope only
void launch_dense_schur(Problem& P, cudaStream_t s, bool sorted_pairs = true);    // S -= sum_p W Hpp^-1 W^T (lower block triangle)
void launch_scatter_eval(Problem& P, double* r_out, double* jac_out, cudaStream_t s);  // back to input order, Ceres layout
void launch_export_blocks(Problem& P, double* Hcc36, double* Hpp9, cudaStream_t s);  // parity tap (adds LM diagonal)
void launch_mirror_lower(Problem& P, cudaStream_t s);   // S upper <- lower (parity tap)

// ---- dense Cholesky (ba_chol.cu) ----
// In-place lower Cholesky of row-major n×n S (lower triangle read/written); *fail_flag set to
// non-zero on a non-positive pivot.  Then solve S y = b in place in b.
// auxiliary high-priority stream + events for the panel look-ahead of the factorisation
constexpr int MAX_CHOL_SEGMENTS = 32;  // more panel chains than this: the factorisation runs as ONE chain
struct CholChainSet {  // streams + events one panel chain (segment) of the structured factorisation runs on
  cudaStream_t chain = nullptr;  // high priority: diagonal blocks and next-diagonal-tile kernels
  cudaStream_t bulk = nullptr;   // panel solves and trailing updates, one step behind
  cudaStream_t tail = nullptr;   // side-matrix updates: wanted only when the parent segment starts
  cudaEvent_t ev_panel = nullptr, ev_next = nullptr, ev_rest = nullptr, ev_trsm = nullptr, ev_tail = nullptr, ev_joined = nullptr;
};
constexpr int CHOL_STREAM_SETS = 8;  // concurrent segments beyond this share a set (still correct: stream order)
struct CholAux {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_panel = nullptr, ev_rest = nullptr, ev_join = nullptr, ev_next = nullptr, ev_zero = nullptr;
  std::vector<CholChainSet> sets;     // segment tree schedule (CholPlan::segs)
  std::vector<cudaEvent_t> ev_seg;    // [MAX_CHOL_SEGMENTS] segment g complete (all its streams)
  // captured launch sequence of the envelope factorisation (static per problem: same S, workspace and plan)
  cudaGraphExec_t graph = nullptr;
  const void* g_S = nullptr; const void* g_ws = nullptr; const void* g_fail = nullptr; const void* g_stream = nullptr;
  int g_n = 0, g_version = -1;
  int64_t g_launches = 0;
  int n_sm = 0, coop = 0;  // of the context's device
  int init();              // also opts the factorisation kernels into their dynamic shared memory ON THE CURRENT DEVICE
  void destroy();
};
// Block structure of the reduced camera matrix and of its Cholesky factor (static per problem), at the granularity
// of the factorisation's 128-row panels.  Block (I,J) of S is non-zero only when a camera of row block I shares a
// point with a camera of column block J; closing that pattern under elimination (ba_order.cpp) gives the structure
// of L.  Panel K then only has to touch the row blocks R_K = { I > K : L(I,K) != 0 } — everything else is an exact
// zero that is never computed, so the result is the factor a dense LLT (CTracker.cpp:574) would produce.  For a scan
// trajectory (keyframes see the points of nearby keyframes, plus a few loop closures) the structure is a band, or —
// in the two-ended camera order chosen at upload — two bands meeting in a tail block: the factorisation of the
// Venice-shaped problem costs 2-3 % of the dense flops.
struct CholPlan {
  int n = 0, panels = 0;
  int version = 0;             // bumped when the structure changes: invalidates captured launch graphs
  bool use = false;            // false: treat the matrix as dense (the structure is most of the triangle)
  double fill = 1.0;           // structural blocks / lower-triangle blocks
  std::vector<int> off, cnt;   // per panel K: list[off[K] .. off[K]+cnt[K]) = R_K ascending  (column structure)
  std::vector<int> list;
  std::vector<int> row_off, row_list;  // per row block I: row_list[row_off[I] .. row_off[I+1]) = { J <= I : L(I,J) != 0 } ascending (ends with I)
  // Panel chains (ba_order.cpp: panel_segments): the paths of the block elimination tree.  Segments that are not
  // ancestor / descendant of each other are factored concurrently.  A segment with a parent must not write the blocks of
  // its ancestors' columns in S (its siblings and cousins update the same blocks at the same time): those updates go to
  // the segment's SIDE MATRIX, a dense square over its ancestors' panels anc(s) (ascending), which the parent folds into
  // S (its own columns) and into its own side matrix (the columns of ITS ancestors) when it starts.  One segment: none.
  struct Seg {
    int begin = 0, end = 0, parent = -1;
    int slot = 0;                // stream set the segment runs on (CholAux)
    std::vector<int> children;   // ascending
    std::vector<int> anc;        // panels of all ancestor segments, ascending
    int anc_off = 0;             // position of anc in d_anc
    long long side_off = 0;      // doubles from the start of the side-matrix area of the workspace
  };
  std::vector<Seg> segs;         // empty or one entry: a single chain over all panels
  std::vector<int> side;         // parallel to list: position of list[q] in anc(segment of K), or -1 inside K's own segment
  std::vector<int> anc_all, map_all;  // anc lists of all segments; per segment [panels]: panel -> position in its anc (or -1)
  long long side_total = 0;      // doubles of all side matrices
  int critical = 0;              // panels on the longest root-to-leaf path
  int split_a = -1, split_t0 = -1;  // set when the schedule is the plain two-chain one (chains [0,a) and [a,t0), tail [t0,P)); reporting only
  int* d_ints = nullptr;       // device copy: list | row_off | row_list | col_off | col_cnt | side | anc_all | map_all
  int* d_list = nullptr; int* d_row_off = nullptr; int* d_row_list = nullptr; int* d_col_off = nullptr; int* d_col_cnt = nullptr;
  int* d_side = nullptr; int* d_anc = nullptr; int* d_map = nullptr; int* d_col_slot = nullptr; int* d_sweep = nullptr;
  std::vector<int> sweep;      // columns by height in the block elimination tree: the order the triangular solves deal them to CTAs
  std::vector<int> col_slot;   // parallel to list: position of block (list[q], K) in row_list (slot of the forward triangular solve)
  size_t d_cap = 0;
  // packed structure (multi-GPU: only these entries of S are summed across ranks): row r of block I occupies
  // packed[blk_off[I] + (r - 128 I) * width(I)) .. + width(I)), width(I) = 128 * (blocks in row I), blocks in row_list order
  std::vector<long long> blk_off;  // [panels + 1]
  long long* d_blk_off = nullptr;
  size_t d_blk_cap = 0;
  std::vector<uint8_t> key;    // the pattern the plan was built from: an unchanged structure keeps plan, version and graph
  long long packed_doubles() const { return blk_off.empty() ? 0 : blk_off.back(); }
  // side matrices of the segments, after the L11^-1 blocks in the workspace
  long long tail_doubles() const { return use ? side_total : 0; }
  // pattern: P x P bytes, lower triangle, NOT yet closed under elimination (nullptr = dense)
  int build(int n_cams, const uint8_t* pattern, int P, cudaStream_t s);
  bool compute(int n_cams, const uint8_t* pattern, int P);  // host part only (no device); false: unchanged
  void destroy();
};
// host-side structure code (ba_order.cpp)
struct PanelSeg { int begin, end, parent; };  // panels [begin, end) form one chain; parent: index of the segment that continues it, or -1
void structure_lists(int P, const std::vector<uint8_t>& st, std::vector<int>& off, std::vector<int>& cnt, std::vector<int>& list);
void panel_segments(int P, const std::vector<int>& off, const std::vector<int>& cnt, const std::vector<int>& list, std::vector<PanelSeg>& segs,
                    std::vector<int>& seg_of);
int critical_panels(const std::vector<PanelSeg>& segs);
bool choose_camera_order(int n_cams, const int* first, std::vector<int>& new_of_old, int* split_cams, int* tail_cams = nullptr,
                         int force_depth = -1);
void rcm_order(int n, const std::vector<std::pair<int, int>>& edges, std::vector<int>& pos_of_cam);  // reverse Cuthill-McKee
void block_pattern_from_envelope(int n_cams, const int* first, const int* new_of_old, std::vector<uint8_t>& pattern, int* panels);
void symbolic_block_factor(int P, std::vector<uint8_t>& pattern);
size_t cholesky_workspace_doubles(int n);  // L11^-1 blocks (one 128×128 per panel) + flags of the cooperative triangular solve
void launch_cholesky(double* S, int n, double* ws, int* fail_flag, cudaStream_t s, CholAux* aux, int64_t* launches,
                     const CholPlan* plan = nullptr);
void launch_chol_solve(const double* L, int n, double* ws, double* b, cudaStream_t s, int64_t* launches, const CholAux* aux,
                       const CholPlan* plan = nullptr);

// S envelope <-> packed buffer (to_packed: gather; else scatter)
void launch_envelope_pack(const CholPlan& pl, double* S, double* packed, bool to_packed, cudaStream_t s);
void launch_structure_zero(const CholPlan& pl, double* S, cudaStream_t s);  // zero the structural blocks of S only
void launch_spd_fill(double* A, int n, double* b /*null = matrix only*/, cudaStream_t s);  // bench tap: synthetic SPD matrix, b = A 1
void launch_spd_err(int n, const double* x, double* err, cudaStream_t s);                 // max |x - 1|

// ---- allocation-free cost evaluation (ba_noalloc.cu) ----
int noalloc_setup(int device);  // once per device, with that device current: device address of the static scratch array
int launch_cost_noalloc(int device, int n_obs, const double* obs_xy, const int* cam_idx, const int* pt_idx, int n_cams, const double* K9,
                        const double* cam_rt, int n_pts, const double* pts, double* cost, cudaStream_t s);

// ---- LM driver (ba_solver.cu) ----
// Persistent device-resident observation table of a scan: the raw arrays of the problem in the
// CALLER's order — what CSfM::bundleAdjustment re-gathers from the map and re-sends at every keyframe
// (CSfM.cpp:310-348, CTracker.cpp:676-696) — kept in grow-only device buffers outside the per-problem arena, so that a
// new keyframe uploads only its own rows (ba_scan.cu).
struct ScanStore {
  bool active = false;
  int mode = 2;
  int64_t n_obs = 0, n_pts = 0, cap_obs = 0, cap_pts = 0;
  int n_cams = 0, cap_cams = 0;
  bool has_dist = false;
  double* obs = nullptr;      // [cap_obs][2]
  int *cam_idx = nullptr, *pt_idx = nullptr;  // [cap_obs]
  double *K9 = nullptr, *dist = nullptr, *cam = nullptr;  // [cap_cams][9], [2], [6]
  double* pts = nullptr;      // [cap_pts][3]
  int64_t bytes_uploaded = 0; // host->device bytes of the last append
  void destroy();
};

// Pointer identity -> compact point ids ON THE DEVICE (hsfm_ba_solve_ptr; CSfM.cpp:331 / CMap.cpp:215: the same
// double* repeated for every observation of a map point).  Grow-only scratch, kept across the per-keyframe calls.
struct PtrScratch {
  char* buf = nullptr;
  size_t cap = 0;
  unsigned long long* uniq_host = nullptr;  // pinned [uniq_cap]
  size_t uniq_cap = 0;
  double* pts_host = nullptr;               // pinned [pts_cap]: the distinct points, packed, on their way to / from the device
  size_t pts_cap = 0;
  void destroy();
};
// d_pt_idx_out: device array [n] of compact ids (valid until the next call), ids in ascending pointer order;
// uniq_out [n_uniq_out]: the distinct pointers by id (pinned mirror, valid until the next call); W.pts_host is sized for
// them.  HSFM_BA_ERR_ARG when a pointer is NULL.
int dedupe_pointers(PtrScratch& W, int64_t n, double* const* host_ptrs, int** d_pt_idx_out, double* const** uniq_out, int64_t* n_uniq_out,
                    cudaStream_t s, HostStager* stager = nullptr);

struct Ctx {
  int device = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  Comm comm;
  CholAux chol_aux;
  CholPlan chol_plan;     // envelope plan of the resident problem (built on first dense solve)
  bool chol_plan_valid = false;
  std::vector<int> plan_first, plan_new_of_old;  // the envelope and camera order chol_plan was built from
  Problem prob;
  bool has_problem = false;
  int ordering = 0;         // 0 = automatic camera order, 1 = the caller's (hsfm_ba_set_ordering)
  bool tap_upload = false;  // upload on behalf of a parity tap: caller's camera order
  std::vector<cudaEvent_t> phase_events;  // pool reused across solves (phase timing of lm_solve)
  ScanStore scan;
  PtrScratch ptr_ws;
};
// ba_scan.cu
int scan_begin(Ctx& c, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
               const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts);
int scan_append(Ctx& c, int32_t n_new_cams, const double* K9, const double* dist, const double* cam_rt, int64_t n_new_pts, const double* pts,
                int64_t n_new_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, double* setup_ms, double* h2d_ms);
int lm_solve(Ctx& ctx, const hsfm_ba_options& o, hsfm_ba_summary* summary);

}  // namespace hsfm
