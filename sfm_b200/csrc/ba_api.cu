// ba_api.cu — the extern "C" boundary declared in include/hsfm_ba.h.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

#include "ba_internal.h"

using namespace hsfm;

struct hsfm_ba_ctx {
  Ctx c;
};

namespace {

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int check_device(int device_id) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) { cudaGetLastError(); return HSFM_BA_ERR_NO_DEVICE; }
  if (device_id < 0 || device_id >= count) return HSFM_BA_ERR_ARG;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device_id) != cudaSuccess) return HSFM_BA_ERR_NO_DEVICE;
  if (prop.major != 10) return HSFM_BA_ERR_NO_DEVICE;  // built for sm_100a only; no fallback path
  return 0;
}

int valid_mode(int mode) { return mode >= 0 && mode <= 2; }

// fn(begin, end) over [0, n) on a few host threads (plain copies through pointers; small inputs stay on the caller's thread)
template <class F>
void parallel_rows(int64_t n, F fn) {
  const int T = n < (1 << 16) ? 1 : 4;
  if (T == 1) { fn(0, n); return; }
  std::vector<std::thread> th;
  const int64_t step = (n + T - 1) / T;
  for (int k = 1; k < T; ++k) th.emplace_back([=] { fn(std::min(n, k * step), std::min(n, (k + 1) * step)); });
  fn(0, std::min(n, step));
  for (auto& t : th) t.join();
}

int64_t global_obs(Ctx& c) {
  // sum of the (already all-reduced) per-camera observation counts
  std::vector<int> h((size_t)c.prob.n_cams);
  if (c.prob.n_cams == 0) return 0;
  cudaMemcpyAsync(h.data(), c.prob.cam_nobs, sizeof(int) * h.size(), cudaMemcpyDeviceToHost, c.stream);
  cudaStreamSynchronize(c.stream);
  int64_t t = 0;
  for (int v : h) t += v;
  return t;
}

}  // namespace

extern "C" {

int hsfm_ba_abi_version(void) { return HSFM_BA_ABI_VERSION; }

const char* hsfm_ba_error_string(int code) {
  static thread_local char buf[96];
  switch (code) {
    case HSFM_BA_OK: return "ok";
    case HSFM_BA_ERR_ARG: return "invalid argument";
    case HSFM_BA_ERR_NO_DEVICE: return "no usable sm_100 CUDA device (this library has no CPU path)";
    case HSFM_BA_ERR_STATE: return "invalid call order / no resident problem";
    case HSFM_BA_ERR_UNSUPPORTED:
      return "unsupported problem (a point with more than 256 observations, sizes beyond int32, or more observations than "
             "HSFM_BA_NOALLOC_MAX_OBS in hsfm_ba_cost_noalloc)";
    case HSFM_BA_ERR_NOMEM: return "out of device memory";
    case HSFM_BA_ERR_COMM: return "NCCL unavailable or communicator initialisation failed";
    default: break;
  }
  if (code >= 2000) { std::snprintf(buf, sizeof(buf), "NCCL error %d", code - 2000); return buf; }
  if (code >= 1000) { std::snprintf(buf, sizeof(buf), "CUDA error %d: %s", code - 1000, cudaGetErrorString((cudaError_t)(code - 1000))); return buf; }
  return "unknown error";
}

void hsfm_ba_default_options(hsfm_ba_options* o) {
  if (!o) return;
  std::memset(o, 0, sizeof(*o));
  o->max_num_iterations = 50;
  o->max_consecutive_invalid_steps = 5;
  o->jacobi_scaling = 1;
  o->linear_solver = HSFM_BA_LINEAR_AUTO;
  o->dense_max_cams = 2048;
  o->pcg_max_iterations = 1000;
  o->verbose = 0;
  o->dense_flags = 0;
  o->initial_trust_region_radius = 1e4;
  o->max_trust_region_radius = 1e16;
  o->min_trust_region_radius = 1e-32;
  o->min_relative_decrease = 1e-3;
  o->min_lm_diagonal = 1e-6;
  o->max_lm_diagonal = 1e32;
  o->function_tolerance = 1e-6;
  o->gradient_tolerance = 1e-10;
  o->parameter_tolerance = 1e-8;
  o->pcg_tolerance = 1e-10;
}

int hsfm_ba_create(hsfm_ba_ctx** out, int device_id) {
  if (!out) return HSFM_BA_ERR_ARG;
  *out = nullptr;
  int rc = check_device(device_id);
  if (rc) return rc;
  HSFM_CUDA(cudaSetDevice(device_id));
  hsfm_ba_ctx* h = new (std::nothrow) hsfm_ba_ctx();
  if (!h) return HSFM_BA_ERR_NOMEM;
  h->c.device = device_id;
  cudaError_t e = cudaStreamCreateWithFlags(&h->c.own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete h; return 1000 + (int)e; }
  h->c.stream = h->c.own_stream;
  { int rc2 = h->c.chol_aux.init(); if (rc2) { cudaStreamDestroy(h->c.own_stream); delete h; return rc2; } }
  { int rc2 = init_kernel_attributes(); if (rc2) { h->c.chol_aux.destroy(); cudaStreamDestroy(h->c.own_stream); delete h; return rc2; } }
  h->c.prob.n_sm = h->c.chol_aux.n_sm;
  { int rc2 = noalloc_setup(device_id); if (rc2) { h->c.chol_aux.destroy(); cudaStreamDestroy(h->c.own_stream); delete h; return rc2; } }
  *out = h;
  return 0;
}

void hsfm_ba_destroy(hsfm_ba_ctx* h) {
  if (!h) return;
  cudaSetDevice(h->c.device);
  cudaStreamSynchronize(h->c.stream);
  h->c.prob.destroy();
  h->c.scan.destroy();
  h->c.ptr_ws.destroy();
  h->c.comm.destroy();
  h->c.chol_aux.destroy();
  h->c.chol_plan.destroy();
  for (cudaEvent_t e : h->c.phase_events) cudaEventDestroy(e);
  if (h->c.own_stream) cudaStreamDestroy(h->c.own_stream);
  delete h;
}

int hsfm_ba_set_stream(hsfm_ba_ctx* h, void* cuda_stream) {
  if (!h) return HSFM_BA_ERR_ARG;
  cudaStreamSynchronize(h->c.stream);
  h->c.stream = cuda_stream ? (cudaStream_t)cuda_stream : h->c.own_stream;
  return 0;
}

int hsfm_ba_set_ordering(hsfm_ba_ctx* h, int ordering) {
  if (!h || ordering < 0 || ordering > 1) return HSFM_BA_ERR_ARG;
  h->c.ordering = ordering;
  return 0;
}

static int debug_plan(int32_t n_cams, const int32_t* cam_first, int32_t natural, std::vector<int>& perm, bool* re, CholPlan& pl, int* split_cams) {
  if (n_cams <= 0 || !cam_first) return HSFM_BA_ERR_ARG;
  for (int c = 0; c < n_cams; ++c)
    if (cam_first[c] < 0 || cam_first[c] > c) return HSFM_BA_ERR_ARG;
  *split_cams = -1;
  *re = natural <= 0 && choose_camera_order(n_cams, cam_first, perm, split_cams, nullptr, natural < 0 ? -natural - 1 : -1);
  if (!*re) { perm.clear(); *split_cams = -1; }
  std::vector<uint8_t> pattern;
  int panels = 0;
  block_pattern_from_envelope(n_cams, cam_first, *re ? perm.data() : nullptr, pattern, &panels);
  pl.compute(n_cams, pattern.data(), panels);
  return 0;
}

int hsfm_ba_debug_structure(int32_t n_cams, const int32_t* cam_first, int32_t natural, int32_t* new_of_old, int32_t* info, double* fill,
                            uint8_t* structure) {
  if (!info) return HSFM_BA_ERR_ARG;
  std::vector<int> perm;
  CholPlan pl;
  bool re = false;
  int split_cams = -1;
  const int rc = debug_plan(n_cams, cam_first, natural, perm, &re, pl, &split_cams);
  if (rc) return rc;
  if (new_of_old)
    for (int c = 0; c < n_cams; ++c) new_of_old[c] = re ? perm[(size_t)c] : c;
  const int panels = pl.panels;
  info[0] = panels; info[1] = pl.split_a; info[2] = pl.split_t0; info[3] = pl.use ? 1 : 0; info[4] = re ? 1 : 0; info[5] = split_cams;
  info[6] = (int)pl.segs.size(); info[7] = pl.critical;
  if (fill) *fill = pl.fill;
  if (structure) {
    std::memset(structure, 0, (size_t)panels * panels);
    if (!pl.row_list.empty())
      for (int I = 0; I < panels; ++I)
        for (int q = pl.row_off[(size_t)I]; q < pl.row_off[(size_t)I + 1]; ++q) structure[(size_t)I * panels + pl.row_list[(size_t)q]] = 1;
  }
  return 0;
}

int hsfm_ba_debug_schedule(int32_t n_cams, const int32_t* cam_first, int32_t natural, int32_t* seg3) {
  if (!seg3) return HSFM_BA_ERR_ARG;
  std::vector<int> perm;
  CholPlan pl;
  bool re = false;
  int split_cams = -1;
  const int rc = debug_plan(n_cams, cam_first, natural, perm, &re, pl, &split_cams);
  if (rc) return rc;
  for (size_t g = 0; g < pl.segs.size(); ++g) { seg3[3 * g] = pl.segs[g].begin; seg3[3 * g + 1] = pl.segs[g].end; seg3[3 * g + 2] = pl.segs[g].parent; }
  return 0;
}

int hsfm_ba_debug_rcm(int32_t n_cams, int64_t n_edges, const int32_t* edges, int32_t* position) {
  if (n_cams <= 0 || n_edges < 0 || (n_edges > 0 && !edges) || !position) return HSFM_BA_ERR_ARG;
  std::vector<std::pair<int, int>> e;
  e.reserve((size_t)n_edges);
  for (int64_t k = 0; k < n_edges; ++k) {
    const int a = edges[2 * k], b = edges[2 * k + 1];
    if (a < 0 || b < 0 || a >= n_cams || b >= n_cams) return HSFM_BA_ERR_ARG;
    if (a != b) e.emplace_back(std::min(a, b), std::max(a, b));
  }
  std::vector<int> pos;
  rcm_order(n_cams, e, pos);
  for (int c = 0; c < n_cams; ++c) position[c] = pos[(size_t)c];
  return 0;
}

int hsfm_ba_comm_unique_id(void* unique_id_128) {
  if (!unique_id_128) return HSFM_BA_ERR_ARG;
  return comm_unique_id(unique_id_128);
}

int hsfm_ba_comm_init(hsfm_ba_ctx* h, const void* unique_id_128, int rank, int n_ranks) {
  if (!h || n_ranks < 1 || rank < 0 || rank >= n_ranks) return HSFM_BA_ERR_ARG;
  if (n_ranks > 1 && !unique_id_128) return HSFM_BA_ERR_ARG;
  cudaSetDevice(h->c.device);
  h->c.comm.destroy();
  return comm_init(h->c.comm, unique_id_128, rank, n_ranks);
}

// ---- resident problem -------------------------------------------------------------------------

int hsfm_ba_problem_upload(hsfm_ba_ctx* h, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                           int32_t n_cams, const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts) {
  if (!h || !valid_mode(mode) || n_obs < 0 || n_cams < 0 || n_pts < 0) return HSFM_BA_ERR_ARG;
  if (n_obs > 0 && (!obs_xy || !cam_idx || !pt_idx)) return HSFM_BA_ERR_ARG;
  if (n_cams > 0 && (!K9 || !cam_rt)) return HSFM_BA_ERR_ARG;
  if (n_pts > 0 && !pts) return HSFM_BA_ERR_ARG;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  h->c.has_problem = false;
  h->c.chol_plan_valid = false;
  h->c.scan.active = false;  // the resident problem no longer is the scan store's
  double setup = 0, h2d = 0;
  const bool natural = h->c.ordering == 1 || h->c.tap_upload;
  int rc = build_problem(h->c.prob, mode, n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, dist, cam_rt, n_pts, pts, h->c.comm, h->c.stream, &setup, &h2d, natural);
  if (rc) { h->c.prob.release(); return rc; }
  launch_jacobi_scale(h->c.prob, false, h->c.stream);  // scale = 1 until a solve computes it
  h->c.prob.scal_host[20] = setup;
  h->c.prob.scal_host[21] = h2d;
  h->c.has_problem = true;
  return 0;
}

int hsfm_ba_scan_begin(hsfm_ba_ctx* h, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                       int32_t n_cams, const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts,
                       int64_t* bytes_uploaded) {
  if (!h || !valid_mode(mode) || n_obs < 0 || n_cams < 0 || n_pts < 0) return HSFM_BA_ERR_ARG;
  if (n_obs > 0 && (!obs_xy || !cam_idx || !pt_idx)) return HSFM_BA_ERR_ARG;
  if (n_cams > 0 && (!K9 || !cam_rt)) return HSFM_BA_ERR_ARG;
  if (n_pts > 0 && !pts) return HSFM_BA_ERR_ARG;
  if (n_obs >= ((int64_t)1 << 31) || n_pts >= ((int64_t)1 << 30)) return HSFM_BA_ERR_UNSUPPORTED;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  const int rc = scan_begin(h->c, mode, n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, dist, cam_rt, n_pts, pts);
  if (bytes_uploaded) *bytes_uploaded = rc ? 0 : h->c.scan.bytes_uploaded;
  return rc;
}

int hsfm_ba_scan_append_keyframe(hsfm_ba_ctx* h, int32_t n_new_cams, const double* K9, const double* dist, const double* cam_rt,
                                 int64_t n_new_pts, const double* pts, int64_t n_new_obs, const double* obs_xy, const int32_t* cam_idx,
                                 const int32_t* pt_idx, int64_t* bytes_uploaded) {
  if (!h || n_new_cams < 0 || n_new_pts < 0 || n_new_obs < 0) return HSFM_BA_ERR_ARG;
  if (n_new_obs > 0 && (!obs_xy || !cam_idx || !pt_idx)) return HSFM_BA_ERR_ARG;
  if (n_new_cams > 0 && (!K9 || !cam_rt)) return HSFM_BA_ERR_ARG;
  if (n_new_pts > 0 && !pts) return HSFM_BA_ERR_ARG;
  if (!h->c.scan.active) return HSFM_BA_ERR_STATE;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  const int rc = scan_append(h->c, n_new_cams, K9, dist, cam_rt, n_new_pts, pts, n_new_obs, obs_xy, cam_idx, pt_idx, nullptr, nullptr);
  if (bytes_uploaded) *bytes_uploaded = rc ? 0 : h->c.scan.bytes_uploaded;
  return rc;
}

int hsfm_ba_problem_set_constant(hsfm_ba_ctx* h, const uint8_t* cam_fixed, const uint8_t* pt_fixed) {
  if (!h) return HSFM_BA_ERR_ARG;
  if (!h->c.has_problem) return HSFM_BA_ERR_STATE;
  Problem& P = h->c.prob;
  cudaStream_t s = h->c.stream;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  unsigned char* stage = nullptr;
  const size_t nc = (size_t)P.n_cams, np = (size_t)P.n_pts;
  if (cam_fixed || pt_fixed) {
    if (cudaMalloc(&stage, std::max<size_t>(1, nc + np)) != cudaSuccess) { cudaGetLastError(); return HSFM_BA_ERR_NOMEM; }
    if (cam_fixed && nc) cudaMemcpyAsync(stage, cam_fixed, nc, cudaMemcpyHostToDevice, s);
    if (pt_fixed && np) cudaMemcpyAsync(stage + nc, pt_fixed, np, cudaMemcpyHostToDevice, s);
  }
  launch_set_fixed(P, cam_fixed ? stage : nullptr, pt_fixed ? stage + nc : nullptr, s);
  const cudaError_t e = cudaStreamSynchronize(s);
  if (stage) cudaFree(stage);
  return e == cudaSuccess ? 0 : 1000 + (int)e;
}

int hsfm_ba_problem_reset(hsfm_ba_ctx* h) {
  if (!h || !h->c.has_problem) return HSFM_BA_ERR_STATE;
  Problem& P = h->c.prob;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  HSFM_CUDA(cudaMemcpyAsync(P.cam_x, P.cam_x0, sizeof(double) * 6 * P.n_cams, cudaMemcpyDeviceToDevice, h->c.stream));
  HSFM_CUDA(cudaMemcpyAsync(P.pts_x, P.pts_x0, sizeof(double) * 3 * (size_t)P.n_pts, cudaMemcpyDeviceToDevice, h->c.stream));
  HSFM_CUDA(cudaMemcpyAsync(P.pts_c, P.pts_x0, sizeof(double) * 3 * (size_t)P.n_pts, cudaMemcpyDeviceToDevice, h->c.stream));
  return 0;
}

int hsfm_ba_problem_solve(hsfm_ba_ctx* h, const hsfm_ba_options* opts, hsfm_ba_summary* summary) {
  if (!h || !opts) return HSFM_BA_ERR_ARG;
  if (!h->c.has_problem) return HSFM_BA_ERR_STATE;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  if (summary) { summary->setup_ms = h->c.prob.scal_host[20]; summary->h2d_ms = h->c.prob.scal_host[21]; }
  int rc = lm_solve(h->c, *opts, summary);
  if (rc) return rc;
  if (summary) {
    const int64_t n = global_obs(h->c);
    summary->rms_px = n > 0 ? std::sqrt(2.0 * summary->final_cost / (double)n) : 0.0;
  }
  return 0;
}

int hsfm_ba_problem_eval(hsfm_ba_ctx* h, double* cost) {
  if (!h || !h->c.has_problem) return HSFM_BA_ERR_STATE;
  Problem& P = h->c.prob;
  cudaStream_t s = h->c.stream;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  launch_cam_prep(P, P.cam_x, nullptr, P.rows_x, s);
  launch_rj(P, P.rows_x, P.pts_x, nullptr, true, true, s);
  launch_finalize_scalars(P, s);
  if (h->c.comm.n_ranks > 1) { int rc = h->c.comm.allreduce_sum(P.scal + S_COST, 1, s); if (rc) return rc; }
  if (cost) {
    HSFM_CUDA(cudaMemcpyAsync(P.scal_host, P.scal, sizeof(double) * 16, cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    *cost = P.scal_host[S_COST];
  }
  return 0;
}

int hsfm_ba_problem_frame_stats(hsfm_ba_ctx* h, double* rms_px, double* max_px) {
  if (!h || !h->c.has_problem) return HSFM_BA_ERR_STATE;
  if (!rms_px && !max_px) return HSFM_BA_ERR_ARG;
  Problem& P = h->c.prob;
  cudaStream_t s = h->c.stream;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  const size_t nc = (size_t)P.n_cams;
  if (nc == 0) return 0;
  launch_cam_prep(P, P.cam_x, nullptr, P.rows_x, s);
  double* out3 = P.rhs;  // 6 n_cams doubles of per-solve scratch: [sum d^2 | count | max d^2]
  launch_frame_stats(P, P.rows_x, P.pts_x, out3, s);
  if (h->c.comm.n_ranks > 1) {  // observations are sharded by point: a camera's observations live on every rank
    int rc = h->c.comm.allreduce_sum(out3, 2 * nc, s);
    if (rc) return rc;
    rc = h->c.comm.allreduce_max(out3 + 2 * nc, nc, s);
    if (rc) return rc;
  }
  std::vector<double> host(3 * nc);
  HSFM_CUDA(cudaMemcpyAsync(host.data(), out3, sizeof(double) * 3 * nc, cudaMemcpyDeviceToHost, s));
  HSFM_CUDA(cudaStreamSynchronize(s));
  for (size_t c = 0; c < nc; ++c) {
    const size_t k = P.h_cam_new_of_old.empty() ? c : (size_t)P.h_cam_new_of_old[c];  // internal row of the caller's camera c
    const double cnt = host[nc + k];
    if (rms_px) rms_px[c] = cnt > 0.0 ? std::sqrt(host[k] / cnt) : 0.0;
    if (max_px) max_px[c] = std::sqrt(host[2 * nc + k]);
  }
  return 0;
}

int hsfm_ba_problem_download(hsfm_ba_ctx* h, double* cam_rt, double* pts) {
  if (!h || !h->c.has_problem) return HSFM_BA_ERR_STATE;
  Problem& P = h->c.prob;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  if (cam_rt && P.cam_new_of_old) {
    launch_cam_rows_to_caller(P, 6, P.cam_x, P.cam_io, h->c.stream);
    HSFM_CUDA(cudaMemcpyAsync(cam_rt, P.cam_io, sizeof(double) * 6 * P.n_cams, cudaMemcpyDeviceToHost, h->c.stream));
  } else if (cam_rt) {
    HSFM_CUDA(cudaMemcpyAsync(cam_rt, P.cam_x, sizeof(double) * 6 * P.n_cams, cudaMemcpyDeviceToHost, h->c.stream));
  }
  if (pts) {
    launch_rows_to_caller(P, 3, P.pts_x, P.pts_io, h->c.stream);
    HSFM_CUDA(cudaMemcpyAsync(pts, P.pts_io, sizeof(double) * 3 * (size_t)P.n_pts, cudaMemcpyDeviceToHost, h->c.stream));
  }
  HSFM_CUDA(cudaStreamSynchronize(h->c.stream));
  return 0;
}

int64_t hsfm_ba_problem_algorithmic_bytes(hsfm_ba_ctx* h, const char* kernel) {
  if (!h || !h->c.has_problem || !kernel) return 0;
  const Problem& P = h->c.prob;
  const int64_t N = P.n_obs, C = P.n_cams, Q = P.n_pts;
  // SURVEY.md §8(d) / DESIGN.md §5
  if (!std::strcmp(kernel, "rj")) return 184 * N + 88 * C + 24 * Q;              // obs 24 + r,J 160 per obs; unique parameters
  if (!std::strcmp(kernel, "rj_form")) return 184 * N + 88 * C + 24 * Q + 72 * Q;  // + Hpp (48) and g_p (24) per point
  if (!std::strcmp(kernel, "schur_matvec")) return 150 * N + 48 * Q + 96 * C;      // Jc 96 + Jp 48 + pt 4 + slot 1 + rank 1 ; Hpp^-1 ; p, q
  if (!std::strcmp(kernel, "backsub")) return 168 * N + 120 * Q + 48 * C;          // + r 16 ; Hpp^-1 48, z_p 24, y_p 24, cand 24
  if (!std::strcmp(kernel, "cam_form")) return 20 * N + 24 * Q + 88 * C + 216 * C; // obs_cs 16 + pt 4 ; point gather ; rows ; Hcc+gc out
  if (!std::strcmp(kernel, "cost")) return 24 * N + 24 * Q + 88 * C;
  if (!std::strcmp(kernel, "schur_rhs")) return 149 * N + 24 * Q + 96 * C;         // Jc 96 + Jp 48 + pt 4 + rank 1 ; z_p ; g_c in, rhs out
  if (!std::strcmp(kernel, "dense_schur")) return 152 * N + 48 * Q;                // Jc, Jp, cam, pt ; Hpp^-1 (S blocks stay in L2)
  return 0;
}

int hsfm_ba_problem_time_kernel(hsfm_ba_ctx* h, const char* kernel, int reps, double* ms) {
  if (!h || !h->c.has_problem || !kernel || reps <= 0 || !ms) return HSFM_BA_ERR_ARG;
  Problem& P = h->c.prob;
  cudaStream_t s = h->c.stream;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  cudaEvent_t e0, e1;
  HSFM_CUDA(cudaEventCreate(&e0));
  HSFM_CUDA(cudaEventCreate(&e1));
  auto run = [&]() -> int {
    if (!std::strcmp(kernel, "rj")) launch_rj(P, P.rows_x, P.pts_x, P.scale_p, true, false, s);
    else if (!std::strcmp(kernel, "rj_form")) launch_rj(P, P.rows_x, P.pts_x, P.scale_p, true, true, s);
    else if (!std::strcmp(kernel, "schur_matvec")) launch_schur_matvec(P, P.pcg_p, P.pcg_q, s);
    else if (!std::strcmp(kernel, "backsub")) launch_backsub(P, s);
    else if (!std::strcmp(kernel, "cam_form")) launch_cam_major(P, P.rows_x, P.pts_x, P.scale_p, 0, s);
    else if (!std::strcmp(kernel, "cost")) launch_cost(P, P.rows_x, P.pts_x, P_COST, s);
    else if (!std::strcmp(kernel, "schur_rhs")) launch_schur_rhs(P, true, s);
    else if (!std::strcmp(kernel, "dense_schur")) {
      if (!P.S) return HSFM_BA_ERR_ARG;  // needs a dense solve first (S is allocated lazily)
      launch_dense_schur(P, s);
    }
    else return HSFM_BA_ERR_ARG;
    return 0;
  };
  // make sure the flag that silences the mat-vec is clear and a linearisation exists
  HSFM_CUDA(cudaMemsetAsync(P.flags, 0, sizeof(int) * FLAG_SLOTS, s));
  int rc = run();
  if (rc) return rc;
  HSFM_CUDA(cudaEventRecord(e0, s));
  for (int i = 0; i < reps; ++i) run();
  HSFM_CUDA(cudaEventRecord(e1, s));
  HSFM_CUDA(cudaEventSynchronize(e1));
  float t = 0.f;
  cudaEventElapsedTime(&t, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  launch_finalize_scalars(P, s);  // drain pending reduction tasks queued by the launchers
  *ms = (double)t / reps;
  return 0;
}

// ---- one-shot solves on host buffers ------------------------------------------------------------

int hsfm_ba_solve_packed(hsfm_ba_ctx* h, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                         int32_t n_cams, const double* K9, const double* dist, double* cam_rt, int64_t n_pts, double* pts,
                         const hsfm_ba_options* opts, hsfm_ba_summary* summary) {
  return hsfm_ba_solve_packed_ex(h, mode, n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, dist, cam_rt, n_pts, pts, nullptr, nullptr, opts, summary);
}

int hsfm_ba_solve_packed_ex(hsfm_ba_ctx* h, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                            int32_t n_cams, const double* K9, const double* dist, double* cam_rt, int64_t n_pts, double* pts,
                            const uint8_t* cam_fixed, const uint8_t* pt_fixed, const hsfm_ba_options* opts, hsfm_ba_summary* summary) {
  if (!h) return HSFM_BA_ERR_ARG;
  if (!valid_mode(mode)) return HSFM_BA_OK;  // CTracker.cpp:692-693: nothing is added, Solve() is a no-op
  hsfm_ba_options def;
  if (!opts) { hsfm_ba_default_options(&def); opts = &def; }
  int rc = hsfm_ba_problem_upload(h, mode, n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, dist, cam_rt, n_pts, pts);
  if (rc) return rc;
  if (cam_fixed || pt_fixed) {
    rc = hsfm_ba_problem_set_constant(h, cam_fixed, pt_fixed);
    if (rc) return rc;
  }
  rc = hsfm_ba_problem_solve(h, opts, summary);
  if (rc) return rc;
  const double t0 = now_ms();
  rc = hsfm_ba_problem_download(h, cam_rt, pts);
  if (summary) summary->d2h_ms = now_ms() - t0;
  return rc;
}

int hsfm_ba_solve_ptr(hsfm_ba_ctx* h, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, int32_t n_cams,
                      const double* K9, double* const* R, double* const* t, double* const* pts3d, const hsfm_ba_options* opts,
                      hsfm_ba_summary* summary) {
  if (!h || n_obs < 0 || n_cams < 0) return HSFM_BA_ERR_ARG;
  if (!valid_mode(mode)) return HSFM_BA_OK;
  if (n_obs > 0 && (!obs_xy || !cam_idx || !pts3d)) return HSFM_BA_ERR_ARG;
  if (n_cams > 0 && (!K9 || !R || !t)) return HSFM_BA_ERR_ARG;
  // pointer identity -> compact point ids (CSfM.cpp:331, CMap.cpp:215), on the device: the 8-byte pointers are uploaded,
  // radix-sorted and numbered there; the host only touches the distinct points (not every observation)
  HSFM_CUDA(cudaSetDevice(h->c.device));
  static const bool timing = getenv("HSFM_PTR_TIMING") != nullptr;  // experiment knob: where the call's host time goes
  const double tt0 = now_ms();
  double* const* uniq = nullptr;
  int64_t n_pts = 0;
  int* d_pt_idx = nullptr;
  {
    const int rc0 = dedupe_pointers(h->c.ptr_ws, n_obs, pts3d, &d_pt_idx, &uniq, &n_pts, h->c.stream, &h->c.prob.stager);
    if (rc0) return rc0;
  }
  const double tt1 = now_ms();
  std::vector<double> cam((size_t)n_cams * 6);
  double* const pts = h->c.ptr_ws.pts_host;  // pinned, sized by dedupe_pointers
  for (int c = 0; c < n_cams; ++c) {
    if (!R[c] || !t[c]) return HSFM_BA_ERR_ARG;
    for (int k = 0; k < 3; ++k) { cam[6 * (size_t)c + k] = R[c][k]; cam[6 * (size_t)c + 3 + k] = t[c][k]; }
  }
  // the distinct points, through their pointers (ids are in address order: a mostly sequential walk of the map's storage)
  parallel_rows(n_pts, [&](int64_t b, int64_t e) {
    for (int64_t p = b; p < e; ++p)
      for (int k = 0; k < 3; ++k) pts[3 * (size_t)p + k] = uniq[(size_t)p][k];
  });
  const double tt2 = now_ms();
  // pt_idx is a DEVICE array: the upload copies with cudaMemcpyDefault
  int rc = hsfm_ba_solve_packed(h, mode, n_obs, obs_xy, cam_idx, d_pt_idx, n_cams, K9, nullptr, cam.data(), n_pts, pts, opts, summary);
  if (rc) return rc;  // parameters untouched on failure
  const double tt3 = now_ms();
  if (mode != HSFM_BA_STRUCT_ONLY)
    for (int c = 0; c < n_cams; ++c)
      for (int k = 0; k < 3; ++k) { R[c][k] = cam[6 * (size_t)c + k]; t[c][k] = cam[6 * (size_t)c + 3 + k]; }
  if (mode != HSFM_BA_POSE_ONLY)
    parallel_rows(n_pts, [&](int64_t b, int64_t e) {
      for (int64_t p = b; p < e; ++p)
        for (int k = 0; k < 3; ++k) uniq[(size_t)p][k] = pts[3 * (size_t)p + k];
    });
  if (timing && summary)
    std::fprintf(stderr, "[hsfm_ba_solve_ptr] dedupe %.2f ms, gather %.2f, packed call %.2f (h2d %.2f setup %.2f solve %.2f d2h %.2f), scatter %.2f\n", tt1 - tt0,
                 tt2 - tt1, tt3 - tt2, summary->h2d_ms, summary->setup_ms, summary->solve_ms, summary->d2h_ms, now_ms() - tt3);
  return 0;
}

// ---- allocation-free entry point (ba_noalloc.cu) ---------------------------------------------------
int hsfm_ba_cost_noalloc(hsfm_ba_ctx* h, int64_t n_obs, const double* d_obs_xy, const int32_t* d_cam_idx, const int32_t* d_pt_idx,
                         int32_t n_cams, const double* d_K9, const double* d_cam_rt, int64_t n_pts, const double* d_pts, double* d_cost) {
  if (!h || n_obs < 0 || n_cams < 0 || n_pts < 0 || !d_cost) return HSFM_BA_ERR_ARG;
  if (n_obs > 0 && (!d_obs_xy || !d_cam_idx || !d_pt_idx || !d_K9 || !d_cam_rt || !d_pts)) return HSFM_BA_ERR_ARG;
  if (n_obs > HSFM_BA_NOALLOC_MAX_OBS || n_pts >= ((int64_t)1 << 31)) return HSFM_BA_ERR_UNSUPPORTED;  // beyond the static scratch array
  HSFM_CUDA(cudaSetDevice(h->c.device));
  return launch_cost_noalloc(h->c.device, (int)n_obs, d_obs_xy, d_cam_idx, d_pt_idx, n_cams, d_K9, d_cam_rt, (int)n_pts, d_pts, d_cost,
                             h->c.stream);
}

// ---- parity taps --------------------------------------------------------------------------------

int hsfm_ba_dense_spd_solve(hsfm_ba_ctx* h, int32_t n, const double* A, const double* b, double* x, double* L_out, int32_t* pivot) {
  if (!h || n <= 0 || (n & 1) || !A || !b || !x) return HSFM_BA_ERR_ARG;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  cudaStream_t s = h->c.stream;
  const size_t nn = (size_t)n * n;
  double *dA = nullptr, *db = nullptr;
  int* dflag = nullptr;
  if (cudaMalloc(&dA, sizeof(double) * (nn + cholesky_workspace_doubles(n))) != cudaSuccess) { cudaGetLastError(); return HSFM_BA_ERR_NOMEM; }
  if (cudaMalloc(&db, sizeof(double) * 3 * n) != cudaSuccess || cudaMalloc(&dflag, sizeof(int)) != cudaSuccess) {
    cudaGetLastError(); cudaFree(dA); if (db) cudaFree(db); return HSFM_BA_ERR_NOMEM;
  }
  int rc = 0, flag = 0;
  int64_t launches = 0;
  auto run = [&]() -> int {
    HSFM_CUDA(cudaMemcpyAsync(dA, A, sizeof(double) * nn, cudaMemcpyHostToDevice, s));
    HSFM_CUDA(cudaMemcpyAsync(db, b, sizeof(double) * n, cudaMemcpyHostToDevice, s));
    HSFM_CUDA(cudaMemsetAsync(dflag, 0, sizeof(int), s));
    launch_cholesky(dA, n, dA + nn, dflag, s, &h->c.chol_aux, &launches);
    launch_chol_solve(dA, n, dA + nn, db, s, &launches, &h->c.chol_aux);
    HSFM_CUDA(cudaMemcpyAsync(&flag, dflag, sizeof(int), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaMemcpyAsync(x, db, sizeof(double) * n, cudaMemcpyDeviceToHost, s));
    if (L_out) HSFM_CUDA(cudaMemcpyAsync(L_out, dA, sizeof(double) * nn, cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    HSFM_CUDA(cudaGetLastError());
    return 0;
  };
  rc = run();
  cudaFree(dA); cudaFree(db); cudaFree(dflag);
  if (rc) return rc;
  if (pivot) *pivot = flag;
  return flag ? HSFM_BA_ERR_ARG : 0;
}

int hsfm_ba_bench_cholesky(hsfm_ba_ctx* h, int32_t n, int reps, double* factor_ms, double* solve_ms, double* rel_residual) {
  if (!h || n <= 0 || (n & 1) || reps <= 0) return HSFM_BA_ERR_ARG;
  HSFM_CUDA(cudaSetDevice(h->c.device));
  cudaStream_t s = h->c.stream;
  const size_t nn = (size_t)n * n;
  double *dA = nullptr, *db = nullptr, *derr = nullptr;
  int* dflag = nullptr;
  if (cudaMalloc(&dA, sizeof(double) * (nn + cholesky_workspace_doubles(n))) != cudaSuccess) { cudaGetLastError(); return HSFM_BA_ERR_NOMEM; }
  if (cudaMalloc(&db, sizeof(double) * 4 * n) != cudaSuccess || cudaMalloc(&dflag, sizeof(int)) != cudaSuccess ||
      cudaMalloc(&derr, sizeof(double)) != cudaSuccess) {
    cudaGetLastError(); cudaFree(dA); if (db) cudaFree(db); if (dflag) cudaFree(dflag); return HSFM_BA_ERR_NOMEM;
  }
  cudaEvent_t e0, e1, e2;
  cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
  int64_t launches = 0;
  float tf = 0.f, ts = 0.f;
  int flag = 0;
  double err = 0.0;
  auto run = [&]() -> int {
    for (int it = 0; it <= reps; ++it) {
      launch_spd_fill(dA, n, it == 0 ? db + 3 * n : nullptr, s);
      HSFM_CUDA(cudaMemcpyAsync(db, db + 3 * n, sizeof(double) * n, cudaMemcpyDeviceToDevice, s));
      HSFM_CUDA(cudaMemsetAsync(dflag, 0, sizeof(int), s));
      HSFM_CUDA(cudaEventRecord(e0, s));
      launch_cholesky(dA, n, dA + nn, dflag, s, &h->c.chol_aux, &launches);
      HSFM_CUDA(cudaEventRecord(e1, s));
      launch_chol_solve(dA, n, dA + nn, db, s, &launches, &h->c.chol_aux);
      HSFM_CUDA(cudaEventRecord(e2, s));
      HSFM_CUDA(cudaEventSynchronize(e2));
      if (it > 0) {
        float a = 0.f, b = 0.f;
        cudaEventElapsedTime(&a, e0, e1);
        cudaEventElapsedTime(&b, e1, e2);
        tf += a; ts += b;
      }
    }
    launch_spd_err(n, db, derr, s);
    HSFM_CUDA(cudaMemcpyAsync(&flag, dflag, sizeof(int), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaMemcpyAsync(&err, derr, sizeof(double), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    HSFM_CUDA(cudaGetLastError());
    return 0;
  };
  const int rc = run();
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
  cudaFree(dA); cudaFree(db); cudaFree(dflag); cudaFree(derr);
  if (rc) return rc;
  if (flag) return HSFM_BA_ERR_STATE;
  if (factor_ms) *factor_ms = tf / reps;
  if (solve_ms) *solve_ms = ts / reps;
  if (rel_residual) *rel_residual = err;
  return 0;
}

int hsfm_ba_eval(hsfm_ba_ctx* h, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
                 const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts, double* residuals,
                 double* jac, double* cost) {
  if (!h) return HSFM_BA_ERR_ARG;
  h->c.tap_upload = true;  // parity taps report per-camera quantities: keep the caller's camera order
  int rc = hsfm_ba_problem_upload(h, 2, n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, dist, cam_rt, n_pts, pts);
  h->c.tap_upload = false;
  if (rc) return rc;
  Problem& P = h->c.prob;
  cudaStream_t s = h->c.stream;
  rc = hsfm_ba_problem_eval(h, cost);
  if (rc) return rc;
  if (n_obs == 0) { if (cost) *cost = 0.0; return 0; }
  double *d_r = nullptr, *d_j = nullptr;
  if (residuals) HSFM_CUDA(cudaMalloc(&d_r, sizeof(double) * 2 * (size_t)n_obs));
  if (jac) HSFM_CUDA(cudaMalloc(&d_j, sizeof(double) * 18 * (size_t)n_obs));
  launch_scatter_eval(P, d_r, d_j, s);
  if (residuals) HSFM_CUDA(cudaMemcpyAsync(residuals, d_r, sizeof(double) * 2 * (size_t)n_obs, cudaMemcpyDeviceToHost, s));
  if (jac) HSFM_CUDA(cudaMemcpyAsync(jac, d_j, sizeof(double) * 18 * (size_t)n_obs, cudaMemcpyDeviceToHost, s));
  HSFM_CUDA(cudaStreamSynchronize(s));
  if (d_r) cudaFree(d_r);
  if (d_j) cudaFree(d_j);
  return 0;
}

int hsfm_ba_schur(hsfm_ba_ctx* h, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
                  const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts, const hsfm_ba_options* opts,
                  double radius, double* S_dense, double* rhs, double* Hcc, double* gc, double* Hpp, double* gp, double* scale) {
  if (!h || !opts || !(radius > 0.0)) return HSFM_BA_ERR_ARG;
  h->c.tap_upload = true;
  int rc = hsfm_ba_problem_upload(h, 2, n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, dist, cam_rt, n_pts, pts);
  h->c.tap_upload = false;
  if (rc) return rc;
  Problem& P = h->c.prob;
  cudaStream_t s = h->c.stream;
  const size_t n6 = 6 * (size_t)n_cams;
  {
    const size_t need = std::max<size_t>(1, n6 * n6) + cholesky_workspace_doubles((int)n6);
    if (P.S_capacity < need) {
      if (P.S) cudaFree(P.S);
      P.S = nullptr; P.S_capacity = 0;
      if (cudaMalloc(&P.S, sizeof(double) * need) != cudaSuccess) { cudaGetLastError(); P.S = nullptr; return HSFM_BA_ERR_NOMEM; }
      P.S_capacity = need;
    }
  }
  // scaling pass, linearisation, blocks
  launch_cam_prep(P, P.cam_x, nullptr, P.rows_x, s);
  launch_rj(P, P.rows_x, P.pts_x, nullptr, false, true, s);
  launch_cam_major(P, P.rows_x, P.pts_x, nullptr, 0, s);
  launch_jacobi_scale(P, opts->jacobi_scaling != 0, s);
  launch_cam_prep(P, P.cam_x, P.scale_c, P.rows_x, s);
  launch_rj(P, P.rows_x, P.pts_x, P.scale_p, true, true, s);
  launch_cam_major(P, P.rows_x, P.pts_x, P.scale_p, 0, s);
  launch_point_invert(P, radius, opts->min_lm_diagonal, opts->max_lm_diagonal, s);
  launch_cam_diag(P, radius, opts->min_lm_diagonal, opts->max_lm_diagonal, s);
  launch_dense_init(P, true, s);
  launch_dense_schur(P, s);
  launch_schur_rhs(P, true, s);
  launch_mirror_lower(P, s);
  launch_finalize_scalars(P, s);
  double *d_Hcc = nullptr, *d_Hpp = nullptr, *d_tmp = nullptr;  // point rows go back in the caller's point order
  if (Hcc) HSFM_CUDA(cudaMalloc(&d_Hcc, sizeof(double) * 36 * std::max(1, (int)n_cams)));
  if (Hpp) HSFM_CUDA(cudaMalloc(&d_Hpp, sizeof(double) * 9 * std::max<size_t>(1, (size_t)n_pts)));
  HSFM_CUDA(cudaMalloc(&d_tmp, sizeof(double) * 9 * std::max<size_t>(1, (size_t)n_pts)));
  P.scal_host[31] = radius; P.scal_host[30] = opts->min_lm_diagonal; P.scal_host[29] = opts->max_lm_diagonal;
  launch_export_blocks(P, d_Hcc, d_Hpp, s);
  if (S_dense) HSFM_CUDA(cudaMemcpyAsync(S_dense, P.S, sizeof(double) * n6 * n6, cudaMemcpyDeviceToHost, s));
  if (rhs) HSFM_CUDA(cudaMemcpyAsync(rhs, P.rhs, sizeof(double) * n6, cudaMemcpyDeviceToHost, s));
  if (Hcc) HSFM_CUDA(cudaMemcpyAsync(Hcc, d_Hcc, sizeof(double) * 36 * n_cams, cudaMemcpyDeviceToHost, s));
  if (gc) HSFM_CUDA(cudaMemcpyAsync(gc, P.gc, sizeof(double) * n6, cudaMemcpyDeviceToHost, s));
  if (Hpp) {
    launch_rows_to_caller(P, 9, d_Hpp, d_tmp, s);
    HSFM_CUDA(cudaMemcpyAsync(Hpp, d_tmp, sizeof(double) * 9 * (size_t)n_pts, cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
  }
  if (gp) {
    launch_rows_to_caller(P, 3, P.gp, d_tmp, s);
    HSFM_CUDA(cudaMemcpyAsync(gp, d_tmp, sizeof(double) * 3 * (size_t)n_pts, cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
  }
  if (scale) {
    HSFM_CUDA(cudaMemcpyAsync(scale, P.scale_c, sizeof(double) * n6, cudaMemcpyDeviceToHost, s));
    launch_rows_to_caller(P, 3, P.scale_p, d_tmp, s);
    HSFM_CUDA(cudaMemcpyAsync(scale + n6, d_tmp, sizeof(double) * 3 * (size_t)n_pts, cudaMemcpyDeviceToHost, s));
  }
  HSFM_CUDA(cudaStreamSynchronize(s));
  if (d_Hcc) cudaFree(d_Hcc);
  if (d_Hpp) cudaFree(d_Hpp);
  cudaFree(d_tmp);
  return 0;
}

}  // extern "C"
