// ba_chol.cu — dense FP64 Cholesky + triangular solves for the reduced camera system
// (the Eigen LLT that Ceres' DENSE_SCHUR calls, CTracker.cpp:574).
// Right-looking blocked factorisation of the lower triangle of a row-major n×n matrix, all
// launches enqueued back to back on one stream (no host synchronisation; a non-positive pivot
// raises a device flag and later kernels become no-ops on NaNs that the LM driver rejects).
#include "ba_internal.h"

namespace hsfm {
namespace {

constexpr int NB = 64;

// Factor the kb×kb diagonal block at (k0,k0) in shared memory.
__global__ void __launch_bounds__(256) chol_diag_kernel(double* __restrict__ A, int n, int k0, int kb, int* __restrict__ fail) {
  __shared__ double s[NB][NB + 1];
  const int tid = threadIdx.x;
  for (int e = tid; e < kb * kb; e += 256) {
    const int i = e / kb, j = e % kb;
    s[i][j] = (j <= i) ? A[(size_t)(k0 + i) * n + k0 + j] : 0.0;
  }
  __syncthreads();
  for (int j = 0; j < kb; ++j) {
    const double djj = s[j][j];
    if (!(djj > 0.0) || !isfinite(djj)) {
      if (tid == 0) atomicExch(fail, k0 + j + 1);
      return;  // uniform: every thread reads the same s[j][j]
    }
    const double d = sqrt(djj);
    const double inv = 1.0 / d;
    __syncthreads();
    if (tid == 0) s[j][j] = d;
    for (int i = j + 1 + tid; i < kb; i += 256) s[i][j] *= inv;
    __syncthreads();
    // trailing update of the lower triangle: s[i][k] -= s[i][j] s[k][j], j < k <= i
    const int m = kb - j - 1;
    for (int e = tid; e < m * m; e += 256) {
      const int i = j + 1 + e / m, k = j + 1 + e % m;
      if (k <= i) s[i][k] -= s[i][j] * s[k][j];
    }
    __syncthreads();
  }
  for (int e = tid; e < kb * kb; e += 256) {
    const int i = e / kb, j = e % kb;
    if (j <= i) A[(size_t)(k0 + i) * n + k0 + j] = s[i][j];
  }
}

// Panel below the diagonal block: X L11^T = A21  =>  row-wise forward substitution.
__global__ void __launch_bounds__(NB) chol_trsm_kernel(double* __restrict__ A, int n, int k0, int kb, const int* __restrict__ fail) {
  extern __shared__ double dyn_sm[];
  double (*sL)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(dyn_sm);
  double (*sA)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(dyn_sm + NB * (NB + 1));
  if (*fail) return;
  const int tid = threadIdx.x;
  const int r0 = k0 + kb + blockIdx.x * NB;
  const int rows = min(NB, n - r0);
  for (int e = tid; e < kb * kb; e += NB) {
    const int i = e / kb, j = e % kb;
    sL[i][j] = A[(size_t)(k0 + i) * n + k0 + j];
  }
  for (int e = tid; e < rows * kb; e += NB) {
    const int i = e / kb, j = e % kb;
    sA[i][j] = A[(size_t)(r0 + i) * n + k0 + j];
  }
  __syncthreads();
  if (tid < rows) {
    for (int j = 0; j < kb; ++j) {
      double sacc = sA[tid][j];
      for (int k = 0; k < j; ++k) sacc -= sA[tid][k] * sL[j][k];
      sA[tid][j] = sacc / sL[j][j];
    }
  }
  __syncthreads();
  for (int e = tid; e < rows * kb; e += NB) {
    const int i = e / kb, j = e % kb;
    A[(size_t)(r0 + i) * n + k0 + j] = sA[i][j];
  }
}

// Trailing update C[bi,bj] -= P[bi] P[bj]^T over lower tiles (bi >= bj) of 64×64, panel width kb.
__global__ void __launch_bounds__(256) chol_syrk_kernel(double* __restrict__ A, int n, int k0, int kb, const int* __restrict__ fail) {
  __shared__ double sA[32][NB + 1];  // [k][row of tile bi]
  __shared__ double sB[32][NB + 1];  // [k][row of tile bj]
  if (*fail) return;
  // decode linear lower-triangular tile index
  const int t = blockIdx.x;
  int bi = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
  while ((bi + 1) * (bi + 2) / 2 <= t) ++bi;
  while (bi * (bi + 1) / 2 > t) --bi;
  const int bj = t - bi * (bi + 1) / 2;
  const int base = k0 + kb;
  const int ri = base + bi * NB, rj = base + bj * NB;
  const int ni = min(NB, n - ri), nj = min(NB, n - rj);
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  for (int kk = 0; kk < kb; kk += 32) {
    const int kc = min(32, kb - kk);
    for (int e = threadIdx.x; e < NB * 32; e += 256) {
      const int row = e / 32, k = e % 32;
      sA[k][row] = (row < ni && k < kc) ? A[(size_t)(ri + row) * n + k0 + kk + k] : 0.0;
      sB[k][row] = (row < nj && k < kc) ? A[(size_t)(rj + row) * n + k0 + kk + k] : 0.0;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      double a[4], b[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { a[q] = sA[k][ty * 4 + q]; b[q] = sB[k][tx * 4 + q]; }
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] += a[p] * b[q];
    }
    __syncthreads();
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int i = ty * 4 + p;
    if (i >= ni) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int j = tx * 4 + q;
      if (j >= nj) continue;
      if (bi == bj && j > i) continue;
      A[(size_t)(ri + i) * n + rj + j] -= acc[p][q];
    }
  }
}

// Forward substitution step for block k0: every CTA solves the kb×kb diagonal system from the
// (read-only in this launch) right-hand side, CTA 0 publishes x_k, CTA r>=1 updates 64 rows below.
__global__ void __launch_bounds__(NB) trsv_fwd_kernel(const double* __restrict__ L, int n, int k0, int kb, double* __restrict__ b,
                                                     double* __restrict__ x) {
  __shared__ double sL[NB][NB + 1];
  __shared__ double sx[NB];
  const int tid = threadIdx.x;
  for (int e = tid; e < kb * kb; e += NB) {
    const int i = e / kb, j = e % kb;
    sL[i][j] = L[(size_t)(k0 + i) * n + k0 + j];
  }
  if (tid < kb) sx[tid] = b[k0 + tid];
  __syncthreads();
  for (int j = 0; j < kb; ++j) {
    if (tid == j) sx[j] = sx[j] / sL[j][j];
    __syncthreads();
    if (tid > j && tid < kb) sx[tid] -= sL[tid][j] * sx[j];
    __syncthreads();
  }
  if (blockIdx.x == 0) {
    if (tid < kb) x[k0 + tid] = sx[tid];
    return;
  }
  const int row = k0 + kb + (blockIdx.x - 1) * NB + tid;
  if (row < n) {
    const double* Lr = L + (size_t)row * n + k0;
    double s = 0.0;
    for (int k = 0; k < kb; ++k) s += Lr[k] * sx[k];
    b[row] -= s;
  }
}

// Backward substitution step (L^T x = y) for block k0, processed from the last block upwards.
__global__ void __launch_bounds__(NB) trsv_bwd_kernel(const double* __restrict__ L, int n, int k0, int kb, double* __restrict__ y,
                                                     double* __restrict__ x) {
  __shared__ double sL[NB][NB + 1];
  __shared__ double sx[NB];
  const int tid = threadIdx.x;
  for (int e = tid; e < kb * kb; e += NB) {
    const int i = e / kb, j = e % kb;
    sL[i][j] = L[(size_t)(k0 + i) * n + k0 + j];
  }
  if (tid < kb) sx[tid] = y[k0 + tid];
  __syncthreads();
  for (int j = kb - 1; j >= 0; --j) {
    if (tid == j) sx[j] = sx[j] / sL[j][j];
    __syncthreads();
    if (tid < j) sx[tid] -= sL[j][tid] * sx[j];
    __syncthreads();
  }
  if (blockIdx.x == 0) {
    if (tid < kb) x[k0 + tid] = sx[tid];
    return;
  }
  const int col = (blockIdx.x - 1) * NB + tid;  // rows above the block: j < k0
  if (col < k0) {
    double s = 0.0;
    for (int i = 0; i < kb; ++i) s += L[(size_t)(k0 + i) * n + col] * sx[i];
    y[col] -= s;
  }
}

}  // namespace

void launch_cholesky(double* S, int n, int* fail_flag, cudaStream_t s, int64_t* launches) {
  constexpr int kTrsmSmem = 2 * NB * (NB + 1) * (int)sizeof(double);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(chol_trsm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTrsmSmem);
    attr_set = true;
  }
  for (int k0 = 0; k0 < n; k0 += NB) {
    const int kb = min(NB, n - k0);
    chol_diag_kernel<<<1, 256, 0, s>>>(S, n, k0, kb, fail_flag);
    ++*launches;
    const int m = n - (k0 + kb);
    if (m <= 0) break;
    const int rb = (m + NB - 1) / NB;
    chol_trsm_kernel<<<rb, NB, kTrsmSmem, s>>>(S, n, k0, kb, fail_flag);
    chol_syrk_kernel<<<rb * (rb + 1) / 2, 256, 0, s>>>(S, n, k0, kb, fail_flag);
    *launches += 2;
  }
}

// Solves S y = b with the factor from launch_cholesky; b is overwritten by the solution.
void launch_chol_solve(const double* L, int n, double* b, cudaStream_t s, int64_t* launches) {
  // forward: L w = b  (w written into b's storage through x == b is unsafe; use in-place carefully:
  // x_k is published into b[k0..] by CTA 0 only after it has been read into shared memory by every
  // CTA of the same launch?  No such guarantee — so publish into a scratch tail kept after b.)
  // The caller provides b with room for 2n doubles: [0,n) rhs/solution, [n,2n) scratch.
  double* x = b + n;
  for (int k0 = 0; k0 < n; k0 += NB) {
    const int kb = min(NB, n - k0);
    const int m = n - (k0 + kb);
    trsv_fwd_kernel<<<1 + (m + NB - 1) / NB, NB, 0, s>>>(L, n, k0, kb, b, x);
    ++*launches;
  }
  // backward: L^T y = w, w lives in x; result published back into b
  const int last = ((n - 1) / NB) * NB;
  for (int k0 = last; k0 >= 0; k0 -= NB) {
    const int kb = min(NB, n - k0);
    trsv_bwd_kernel<<<1 + (k0 + NB - 1) / NB, NB, 0, s>>>(L, n, k0, kb, x, b);
    ++*launches;
  }
}

}  // namespace hsfm
