// How fast can one CTA-per-tile kernel write the linearisation (10 planes of double2, 160 B/observation)?
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a plane_store.cu -o plane_store
#include <cstdio>
#include <cuda_runtime.h>
template <int NP, int MODE>
__global__ void __launch_bounds__(256) store_k(double2* __restrict__ out, size_t n, int obs_per_tile, int spin) {
  const size_t i = (size_t)blockIdx.x * obs_per_tile + threadIdx.x;
  if (threadIdx.x >= obs_per_tile || i >= n) return;
  double a = (double)i, b = 1.0;
  for (int k = 0; k < spin; ++k) { a = fma(a, 1.0000001, b); b = fma(b, 0.9999999, a); }  // dependent FP64 chain
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    double2 v = make_double2(a + p, b - p);
    if (MODE == 0) out[p * n + i] = v;
    if (MODE == 1) __stcs(out + p * n + i, v);
    if (MODE == 2) __stwt(out + p * n + i, v);
    if (MODE == 3) out[i * NP + p] = v;  // AoS: 160 contiguous bytes per observation
  }
}
template <int NP, int MODE>
void run(const char* name, double2* buf, size_t n, int opt, int spin) {
  const int grid = (int)((n + opt - 1) / opt);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int w = 0; w < 2; ++w) store_k<NP, MODE><<<grid, 256>>>(buf, n, opt, spin);
  cudaEventRecord(e0);
  const int reps = 10;
  for (int r = 0; r < reps; ++r) store_k<NP, MODE><<<grid, 256>>>(buf, n, opt, spin);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= reps;
  printf("%-28s planes %2d obs/tile %3d spin %4d : %7.1f us  %7.1f GB/s\n", name, NP, opt, spin, ms * 1e3, NP * 16.0 * n / ms / 1e6);
}
int main() {
  const size_t n = 5001831;
  double2* buf; cudaMalloc(&buf, sizeof(double2) * 10 * n + 4096);
  for (int spin : {0, 100, 300}) {
    run<10, 0>("st.global", buf, n, 256, spin);
    run<10, 0>("st.global ragged 253", buf, n, 253, spin);
    run<10, 1>("st.cs ragged 253", buf, n, 253, spin);
    run<10, 2>("st.wt ragged 253", buf, n, 253, spin);
    run<10, 3>("AoS ragged 253", buf, n, 253, spin);
    run<5, 0>("5 planes ragged 253", buf, n, 253, spin);
  }
  return 0;
}
