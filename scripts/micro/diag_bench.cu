// diag_bench.cu — latency of the Cholesky chain kernels in isolation (one 128x128 diagonal block:
// factor + inverse), with the clock64 phase marks of HSFM_DIAG_TIMING.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I sfm_b200/csrc -I include scripts/micro/diag_bench.cu -o scripts/micro/diag_bench
#define HSFM_DIAG_TIMING
#include "../../sfm_b200/csrc/ba_chol.cu"
#include <vector>
#include <cmath>

int main() {
  using namespace hsfm;
  const int n = 1024, k0 = 256;
  std::vector<double> h((size_t)n * n, 0.0);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) h[(size_t)i * n + j] = (i == j ? 4.0 + 0.01 * (i % 7) : 1.0 / (1.0 + std::abs(i - j)));
  double *A, *A0, *Linv;
  int* fail;
  cudaMalloc(&A, sizeof(double) * n * n);
  cudaMalloc(&A0, sizeof(double) * n * n);
  cudaMalloc(&Linv, sizeof(double) * NB * NB);
  cudaMalloc(&fail, 4);
  cudaMemset(fail, 0, 4);
  cudaMemcpy(A0, h.data(), sizeof(double) * n * n, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDiagSmem);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int it = 0; it < 10; ++it) {
    cudaMemcpy(A, A0, sizeof(double) * n * n, cudaMemcpyDeviceToDevice);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    chol_diag_kernel<<<1, 256, kDiagSmem>>>(A, n, k0, NB, Linv, fail);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    best = std::min(best, ms);
  }
  long long c[32];
  cudaMemcpyFromSymbol(c, g_diag_clk, sizeof(c));
  std::printf("diag kernel: best %.2f us (event), err=%s\n", best * 1e3, cudaGetErrorString(cudaGetLastError()));
  std::printf("clk: load %lld | ", c[1] - c[0]);
  long long prev = c[1];
  for (int kq = 0; kq < 4; ++kq) {
    std::printf("kq%d factor+inv %lld", kq, c[8 + 3 * kq] - prev);
    if (kq < 3) { std::printf(" chain step %lld | ", c[10 + 3 * kq] - c[8 + 3 * kq]); prev = c[10 + 3 * kq]; }
  }
  std::printf("\n     tail %lld total %lld clk\n", c[6] - c[17], c[6] - c[0]);
  for (int kq = 1; kq < 4; ++kq) std::printf("kq%d: factor done +%lld (scale+inv +%lld), shadow done +%lld (from phase start)\n", kq, c[20 + kq] - c[10 + 3 * (kq - 1)], c[8 + 3 * kq] - c[20 + kq], c[24 + kq] - c[10 + 3 * (kq - 1)]);
  std::printf("shadow<0>: a0 %lld a %lld b %lld c1 %lld d %lld c2 %lld (start +%lld after phase start)\n", c[29]-c[28], c[30]-c[29], c[31]-c[30], c[32]-c[31], c[33]-c[32], c[34]-c[33], c[28]-c[10]);
#ifdef HSFM_STEP_TIMING
  long long sc[40];
  cudaMemcpyFromSymbol(sc, g_step_clk, sizeof(sc));
  std::printf("warp_chol32 (last sub-block): prologue %lld, steps:", sc[0] - sc[34]);
  for (int j = 0; j < 32; ++j) std::printf(" %lld", sc[j + 1] - sc[j]);
  std::printf("\n");
  unsigned mk[40];
  cudaMemcpyFromSymbol(mk, g_mask, sizeof(mk));
  std::printf("active masks: pro %08x steps", mk[34]);
  for (int j = 0; j < 34; ++j) std::printf(" %08x", mk[j]);
  std::printf("\n");
  cudaMemcpyFromSymbol(mk, g_mask2, sizeof(mk));
  std::printf("phase masks:");
  for (int j = 0; j < 28; ++j) std::printf(" %d:%08x", j, mk[j]);
  std::printf("\n");
#endif
  // check: L L^T == A block, Linv L == I
  std::vector<double> L((size_t)n * n), Li(NB * NB);
  cudaMemcpy(L.data(), A, sizeof(double) * n * n, cudaMemcpyDeviceToHost);
  cudaMemcpy(Li.data(), Linv, sizeof(double) * NB * NB, cudaMemcpyDeviceToHost);
  double e_f = 0, e_i = 0;
  for (int i = 0; i < NB; ++i)
    for (int j = 0; j <= i; ++j) {
      double s = 0, t = 0;
      for (int k = 0; k <= j; ++k) s += L[(size_t)(k0 + i) * n + k0 + k] * L[(size_t)(k0 + j) * n + k0 + k];
      for (int k = j; k <= i; ++k) t += Li[i * NB + k] * L[(size_t)(k0 + k) * n + k0 + j];
      e_f = std::max(e_f, std::abs(s - h[(size_t)(k0 + i) * n + k0 + j]));
      e_i = std::max(e_i, std::abs(t - (i == j ? 1.0 : 0.0)));
    }
  std::printf("max |LL^T - A| %.2e   max |Linv L - I| %.2e\n", e_f, e_i);
  return 0;
}
