// ba_noalloc.cu — hsfm_ba_cost_noalloc: the reprojection cost 1/2 sum |r|^2 (the candidate-cost evaluation of the
// trust-region loop) for callers that must not allocate device memory — code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// A variant of cost_kernel (ba_kernels.cu) whose scratch memory — one partial sum per block of 256 observations — is a
// file-scope static __device__ array instead of the context's arena: every GPU that loads the library holds its
// HSFM_BA_NOALLOC_MAX_OBS / 256 doubles (128 KB) until exit.  The caller passes DEVICE pointers for all inputs and the
// output.  Nothing here calls cudaMalloc / cudaFree or an allocating library routine.
//
// The array's name is a host-side stand-in in host code, not a device address: the address the kernels are given is
// fetched with cudaGetSymbolAddress once per device, with that device current (noalloc_setup, called by
// hsfm_ba_create).  Every call on a device shares the array; it keeps the last call's contents, which are never read
// (a call writes partials [0, blocks) and folds exactly those), so nothing has to be cleared — but two calls that
// overlap on the GPU (different streams / contexts of one device) race on it.
#include <mutex>

#include "ba_internal.h"

namespace hsfm {
namespace {

constexpr int NA_BLOCK = 256;
constexpr int NA_MAX_BLOCKS = HSFM_BA_NOALLOC_MAX_OBS / NA_BLOCK;

static __device__ __align__(256) double g_noalloc_scratch[NA_MAX_BLOCKS];

constexpr int NA_MAX_DEVICES = 64;
double* g_scratch_addr[NA_MAX_DEVICES];  // device address of g_noalloc_scratch on each device (null: not set up)
std::mutex g_scratch_mutex;

// One thread per observation, in the caller's observation order: camera row prepared on the fly (the same
// cam_prepare / reproject_r arithmetic as the solver's candidate-cost pass), |r|^2 summed per block in a fixed order.
// An out-of-range camera or point index poisons the block's partial (the cost comes out NaN) instead of being read.
__global__ void __launch_bounds__(NA_BLOCK) cost_noalloc_kernel(int n_obs, const double2* __restrict__ obs, const int* __restrict__ cam_idx,
                                                                 const int* __restrict__ pt_idx, int n_cams, const double* __restrict__ K9,
                                                                 const double* __restrict__ cam_rt, int n_pts, const double* __restrict__ pts,
                                                                 double* __restrict__ partial) {
  __shared__ double sm[NA_BLOCK / 32];
  const int i = blockIdx.x * NA_BLOCK + threadIdx.x;
  double e = 0.0;
  if (i < n_obs) {
    const int c = cam_idx[i], p = pt_idx[i];
    if ((unsigned)c >= (unsigned)n_cams || (unsigned)p >= (unsigned)n_pts) {
      e = __longlong_as_double(0x7ff8000000000000ll);
    } else {
      CamRow C;
      cam_prepare(cam_rt + 6 * (size_t)c, K9 + 9 * (size_t)c, nullptr, nullptr, C);
      const double2 o = obs[i];
      double r0, r1;
      reproject_r(C, pts[3 * (size_t)p], pts[3 * (size_t)p + 1], pts[3 * (size_t)p + 2], o.x, o.y, r0, r1);
      e = r0 * r0 + r1 * r1;
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) e += __shfl_down_sync(0xffffffffu, e, d);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = e;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < NA_BLOCK / 32; ++w) t += sm[w];
    partial[blockIdx.x] = t;
  }
}

// cost = 1/2 sum of the first n partials, one CTA, fixed order.
__global__ void __launch_bounds__(1024) cost_noalloc_fold_kernel(int n, const double* __restrict__ partial, double* __restrict__ cost) {
  __shared__ double sm[32];
  double v = 0.0;
  for (int k = threadIdx.x; k < n; k += 1024) v += partial[k];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 32; ++w) s += sm[w];
    *cost = 0.5 * s;
  }
}

}  // namespace

// Once per device, with that device current (hsfm_ba_create): the device address of the static scratch array.
int noalloc_setup(int device) {
  if (device < 0 || device >= NA_MAX_DEVICES) return 0;  // the entry point then reports HSFM_BA_ERR_UNSUPPORTED
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_addr[device]) return 0;
  double* ptr = nullptr;
  const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
  if (e != cudaSuccess) return 1000 + (int)e;
  g_scratch_addr[device] = ptr;
  return 0;
}

int launch_cost_noalloc(int device, int n_obs, const double* obs_xy, const int* cam_idx, const int* pt_idx, int n_cams, const double* K9,
                        const double* cam_rt, int n_pts, const double* pts, double* cost, cudaStream_t s) {
  if (device < 0 || device >= NA_MAX_DEVICES || !g_scratch_addr[device]) return HSFM_BA_ERR_UNSUPPORTED;
  double* scratch = g_scratch_addr[device];
  const int blocks = (n_obs + NA_BLOCK - 1) / NA_BLOCK;
  if (blocks > 0)
    cost_noalloc_kernel<<<blocks, NA_BLOCK, 0, s>>>(n_obs, reinterpret_cast<const double2*>(obs_xy), cam_idx, pt_idx, n_cams, K9, cam_rt, n_pts,
                                                    pts, scratch);
  cost_noalloc_fold_kernel<<<1, 1024, 0, s>>>(blocks, scratch, cost);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : 1000 + (int)e;
}

}  // namespace hsfm
