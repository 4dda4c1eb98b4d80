// ba_scan.cu — the persistent, incrementally updated device-resident observation table.
//
// The reference rebuilds its whole problem at every keyframe: CSfM::bundleAdjustment walks the map's hash tables to
// gather every observation of every keyframe (CSfM.cpp:321-341 through CMap.cpp:206-223), and CTracker allocates two
// heap objects and does a hash insert per observation (CTracker.cpp:676-696), called from CSfM::mapping at every
// keyframe (CSfM.cpp:252-259).  Here the raw arrays of the problem stay in HBM between the calls (ScanStore,
// ba_internal.h): a new keyframe sends ONLY its own camera row, the map points it created and the observations that
// were added — O(new) bytes over PCIe instead of O(scan) — and the refined poses/points of the previous adjustment
// are taken from the device, not re-sent.  The index (point-major tiles, camera-major chunks, camera order) is then
// rebuilt ON THE DEVICE from the resident arrays by the same code path as a one-shot upload, so it is identical to the
// index a one-shot upload of the concatenated arrays would build and the solve follows the same trajectory.  (A
// pure merge of the sorted tables is not possible in general: internal point ids are ordered by smallest observing
// camera, and camera order is chosen from the whole graph, so a new keyframe can renumber everything; the radix sorts
// cost ~2 ms at 5 M observations, the PCIe transfer they replace ~10 ms.)
#include <algorithm>
#include <chrono>

#include "ba_internal.h"

namespace hsfm {

void ScanStore::destroy() {
  for (void* p : {(void*)obs, (void*)cam_idx, (void*)pt_idx, (void*)K9, (void*)dist, (void*)cam, (void*)pts})
    if (p) cudaFree(p);
  *this = ScanStore();
}

namespace {

// grow-only device array of `width` T per row; keeps the first `used` rows
template <typename T>
int grow(T** buf, int64_t used, int64_t need, int64_t* cap_io, int64_t cap_new, int width, cudaStream_t s) {
  (void)need;
  if (cap_new <= *cap_io && *buf) return 0;
  T* nb = nullptr;
  if (cudaMalloc(&nb, sizeof(T) * (size_t)std::max<int64_t>(1, cap_new) * width) != cudaSuccess) { cudaGetLastError(); return HSFM_BA_ERR_NOMEM; }
  if (*buf && used > 0) {
    if (cudaMemcpyAsync(nb, *buf, sizeof(T) * (size_t)used * width, cudaMemcpyDeviceToDevice, s) != cudaSuccess) { cudaFree(nb); return HSFM_BA_ERR_NOMEM; }
    cudaStreamSynchronize(s);
  }
  if (*buf) cudaFree(*buf);
  *buf = nb;
  return 0;
}

int reserve(ScanStore& S, int64_t n_obs, int n_cams, int64_t n_pts, cudaStream_t s) {
  auto want = [](int64_t need, int64_t cap) { return need <= cap ? cap : std::max<int64_t>(need + need / 2, 1024); };  // 1.5x head-room
  const int64_t co = want(n_obs, S.cap_obs), cc = want(n_cams, S.cap_cams), cp = want(n_pts, S.cap_pts);
  int rc = 0;
  int64_t dummy;
  if (co != S.cap_obs || !S.obs) {
    dummy = S.cap_obs; if ((rc = grow(&S.obs, S.n_obs, n_obs, &dummy, co, 2, s))) return rc;
    dummy = S.cap_obs; if ((rc = grow(&S.cam_idx, S.n_obs, n_obs, &dummy, co, 1, s))) return rc;
    dummy = S.cap_obs; if ((rc = grow(&S.pt_idx, S.n_obs, n_obs, &dummy, co, 1, s))) return rc;
    S.cap_obs = co;
  }
  if (cc != S.cap_cams || !S.cam) {
    dummy = S.cap_cams; if ((rc = grow(&S.K9, (int64_t)S.n_cams, n_cams, &dummy, cc, 9, s))) return rc;
    dummy = S.cap_cams; if ((rc = grow(&S.dist, (int64_t)S.n_cams, n_cams, &dummy, cc, 2, s))) return rc;
    dummy = S.cap_cams; if ((rc = grow(&S.cam, (int64_t)S.n_cams, n_cams, &dummy, cc, 6, s))) return rc;
    S.cap_cams = (int)cc;
  }
  if (cp != S.cap_pts || !S.pts) {
    dummy = S.cap_pts; if ((rc = grow(&S.pts, S.n_pts, n_pts, &dummy, cp, 3, s))) return rc;
    S.cap_pts = cp;
  }
  return 0;
}

__global__ void validate_new_idx_kernel(int n, const int* __restrict__ cam, const int* __restrict__ pt, int n_cams, long long n_pts,
                                        int* __restrict__ bad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && ((unsigned)cam[i] >= (unsigned)n_cams || pt[i] < 0 || (long long)pt[i] >= n_pts)) *bad = 1;
}

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// append rows to the store (host sources); returns bytes sent
int64_t send(ScanStore& S, int n_new_cams, const double* K9, const double* dist, const double* cam_rt, int64_t n_new_pts, const double* pts,
             int64_t n_new_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, cudaStream_t s) {
  int64_t bytes = 0;
  auto put = [&](void* dst, const void* src, size_t b) { if (b) { cudaMemcpyAsync(dst, src, b, cudaMemcpyHostToDevice, s); bytes += (int64_t)b; } };
  put(S.obs + 2 * S.n_obs, obs_xy, sizeof(double) * 2 * (size_t)n_new_obs);
  put(S.cam_idx + S.n_obs, cam_idx, sizeof(int) * (size_t)n_new_obs);
  put(S.pt_idx + S.n_obs, pt_idx, sizeof(int) * (size_t)n_new_obs);
  put(S.K9 + 9 * (size_t)S.n_cams, K9, sizeof(double) * 9 * (size_t)n_new_cams);
  if (dist) put(S.dist + 2 * (size_t)S.n_cams, dist, sizeof(double) * 2 * (size_t)n_new_cams);
  else if (n_new_cams) cudaMemsetAsync(S.dist + 2 * (size_t)S.n_cams, 0, sizeof(double) * 2 * (size_t)n_new_cams, s);
  put(S.cam + 6 * (size_t)S.n_cams, cam_rt, sizeof(double) * 6 * (size_t)n_new_cams);
  put(S.pts + 3 * (size_t)S.n_pts, pts, sizeof(double) * 3 * (size_t)n_new_pts);
  if (dist) S.has_dist = true;
  return bytes;
}

int reindex(Ctx& c, double* setup_ms, double* h2d_ms) {
  ScanStore& S = c.scan;
  c.has_problem = false;
  c.chol_plan_valid = false;
  double setup = 0, h2d = 0;
  int rc = build_problem(c.prob, S.mode, S.n_obs, S.obs, S.cam_idx, S.pt_idx, S.n_cams, S.K9, S.has_dist ? S.dist : nullptr, S.cam, S.n_pts, S.pts,
                         c.comm, c.stream, &setup, &h2d, c.ordering == 1);
  if (rc) { c.prob.release(); return rc; }
  launch_jacobi_scale(c.prob, false, c.stream);
  c.prob.scal_host[20] = setup + h2d;  // on-device copies + index build
  c.prob.scal_host[21] = 0.0;
  c.has_problem = true;
  if (setup_ms) *setup_ms = setup + h2d;
  if (h2d_ms) *h2d_ms = 0.0;
  return 0;
}

}  // namespace

int scan_begin(Ctx& c, int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
               const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts) {
  ScanStore& S = c.scan;
  S.n_obs = 0; S.n_cams = 0; S.n_pts = 0; S.has_dist = false; S.mode = mode; S.active = false;
  int rc = reserve(S, n_obs, n_cams, n_pts, c.stream);
  if (rc) return rc;
  S.bytes_uploaded = send(S, n_cams, K9, dist, cam_rt, n_pts, pts, n_obs, obs_xy, cam_idx, pt_idx, c.stream);
  S.n_obs = n_obs; S.n_cams = n_cams; S.n_pts = n_pts;
  rc = reindex(c, nullptr, nullptr);
  if (rc) return rc;
  S.active = true;
  return 0;
}

int scan_append(Ctx& c, int32_t n_new_cams, const double* K9, const double* dist, const double* cam_rt, int64_t n_new_pts, const double* pts,
                int64_t n_new_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, double* setup_ms, double* h2d_ms) {
  ScanStore& S = c.scan;
  if (!S.active || !c.has_problem) return HSFM_BA_ERR_STATE;
  cudaStream_t s = c.stream;
  Problem& P = c.prob;
  if (S.n_obs + n_new_obs >= ((int64_t)1 << 31) || S.n_pts + n_new_pts >= ((int64_t)1 << 30)) return HSFM_BA_ERR_UNSUPPORTED;
  // the refined parameters of the previous adjustment go back into the store (caller order) — device to device
  if (S.n_cams > 0) {
    if (P.cam_new_of_old) launch_cam_rows_to_caller(P, 6, P.cam_x, S.cam, s);
    else HSFM_CUDA(cudaMemcpyAsync(S.cam, P.cam_x, sizeof(double) * 6 * (size_t)S.n_cams, cudaMemcpyDeviceToDevice, s));
  }
  if (S.n_pts > 0) launch_rows_to_caller(P, 3, P.pts_x, S.pts, s);
  HSFM_CUDA(cudaStreamSynchronize(s));  // reserve() may move the store
  int rc = reserve(S, S.n_obs + n_new_obs, S.n_cams + n_new_cams, S.n_pts + n_new_pts, s);
  if (rc) return rc;
  const double t0 = now_ms();
  S.bytes_uploaded = send(S, n_new_cams, K9, dist, cam_rt, n_new_pts, pts, n_new_obs, obs_xy, cam_idx, pt_idx, s);
  // the new rows may reference any camera / point of the grown problem; validate them before anything uses them as addresses
  if (n_new_obs > 0) {
    int* d_bad = P.flags + F_INVALID;  // any device int: the flag block is reset by the next solve
    int h_bad = 0;
    HSFM_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), s));
    validate_new_idx_kernel<<<(int)((n_new_obs + 255) / 256), 256, 0, s>>>((int)n_new_obs, S.cam_idx + S.n_obs, S.pt_idx + S.n_obs,
                                                                          S.n_cams + n_new_cams, (long long)(S.n_pts + n_new_pts), d_bad);
    HSFM_CUDA(cudaMemcpyAsync(&h_bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    if (h_bad) return HSFM_BA_ERR_ARG;  // the store keeps its old extent: nothing was appended
  } else {
    HSFM_CUDA(cudaStreamSynchronize(s));
  }
  const double t1 = now_ms();
  S.n_obs += n_new_obs; S.n_cams += n_new_cams; S.n_pts += n_new_pts;
  double setup = 0;
  rc = reindex(c, &setup, nullptr);
  if (rc) return rc;
  c.prob.scal_host[21] = t1 - t0;
  if (setup_ms) *setup_ms = setup;
  if (h2d_ms) *h2d_ms = t1 - t0;
  return 0;
}

}  // namespace hsfm
