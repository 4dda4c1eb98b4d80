// chain_tile_test.cu — chol_chain_tile_kernel against a host computation: tile(K+1,K+1) - L21 L21^T, L21 = A21 Linv^T.
#include "../../sfm_b200/csrc/ba_chol.cu"
#include <vector>
#include <cmath>
int main() {
  using namespace hsfm;
  const int n = 640, k0 = 128, blk = 2;  // panel K = 1, block K+1 = 2
  std::vector<double> h((size_t)n * n), li((size_t)NB * NB, 0.0);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) h[(size_t)i * n + j] = std::sin(0.37 * i + 0.11 * j) + (i == j ? 3.0 : 0.0);
  for (int i = 0; i < NB; ++i)
    for (int j = 0; j <= i; ++j) li[(size_t)i * NB + j] = std::cos(0.05 * i - 0.13 * j) / (1.0 + 0.1 * (i - j)) + (i == j ? 1.0 : 0.0);
  double *S, *Linv; int *fail, *list;
  cudaMalloc(&S, sizeof(double) * n * n); cudaMalloc(&Linv, sizeof(double) * NB * NB); cudaMalloc(&fail, 4); cudaMalloc(&list, 4);
  cudaMemset(fail, 0, 4);
  cudaMemcpy(list, &blk, 4, cudaMemcpyHostToDevice);
  cudaMemcpy(S, h.data(), sizeof(double) * n * n, cudaMemcpyHostToDevice);
  cudaMemcpy(Linv, li.data(), sizeof(double) * NB * NB, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(chol_chain_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kChainTileSmem);
  chol_chain_tile_kernel<<<32, 256, kChainTileSmem>>>(S, n, k0, Linv, list, fail);
  std::vector<double> out((size_t)n * n);
  cudaMemcpy(out.data(), S, sizeof(double) * n * n, cudaMemcpyDeviceToHost);
  std::printf("launch: %s\n", cudaGetErrorString(cudaGetLastError()));
  const int b0 = blk * NB;
  std::vector<double> L21((size_t)NB * NB);
  for (int r = 0; r < NB; ++r)
    for (int c = 0; c < NB; ++c) {
      double s = 0;
      for (int k = 0; k <= c; ++k) s += h[(size_t)(b0 + r) * n + k0 + k] * li[(size_t)c * NB + k];
      L21[(size_t)r * NB + c] = s;
    }
  double err = 0, other = 0; int wr = -1, wc = -1;
  for (int r = 0; r < NB; ++r)
    for (int c = 0; c < NB; ++c) {
      double s = h[(size_t)(b0 + r) * n + b0 + c];
      for (int k = 0; k < NB; ++k) s -= L21[(size_t)r * NB + k] * L21[(size_t)c * NB + k];
      const double got = out[(size_t)(b0 + r) * n + b0 + c];
      if (c / 8 <= r / 8) { const double e = std::abs(got - s); if (e > err) { err = e; wr = r; wc = c; } }
    }
  for (size_t i = 0; i < (size_t)n * n; ++i) {
    const int r = (int)(i / n), c = (int)(i % n);
    if (!(r >= b0 && r < b0 + NB && c >= b0 && c < b0 + NB)) other = std::max(other, std::abs(out[i] - h[i]));
  }
  std::printf("max |tile - ref| over the lower 8x8-tile triangle: %.3e at (%d,%d); change outside the tile: %.3e\n", err, wr, wc, other);
  return 0;
}
