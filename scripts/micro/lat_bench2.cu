// lat_bench2.cu — shared-memory broadcast load issue rates and named-barrier round trips (one CTA).
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ double lds64(unsigned a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ double2 lds128(unsigned a) { double2 v; asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a)); return v; }
__global__ void k(double* out, long long* clk, int off) {
  __shared__ __align__(16) double sm[4096];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = 1.0 + i * 1e-9;
  __syncthreads();
  const unsigned base = (unsigned)__cvta_generic_to_shared(sm) + off * 16;
  double a[32], x = 1.0 + lane * 1e-9;
  long long t0, t1;
#pragma unroll
  for (int i = 0; i < 32; ++i) a[i] = x + i;
  if (warp == 0) {
    // (a) 32 LDS.64 broadcast + 32 DFMA, x8
    t0 = clock64();
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 32; ++i) a[i] = fma(a[i], lds64(base + (r * 40 + i) * 8), x);
    t1 = clock64();
    if (lane == 0) clk[0] = t1 - t0;
    // (b) 16 LDS.128 + 32 DFMA, x8
    t0 = clock64();
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 32; i += 2) { const double2 v = lds128(base + (r * 40 + i) * 8); a[i] = fma(a[i], v.x, x); a[i + 1] = fma(a[i + 1], v.y, x); }
    t1 = clock64();
    if (lane == 0) clk[1] = t1 - t0;
    // (c) LDS.64 only x256 (sum to keep)
    t0 = clock64();
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma unroll
    for (int i = 0; i < 256; i += 4) { s0 += lds64(base + i * 8); s1 += lds64(base + i * 8 + 8); s2 += lds64(base + i * 8 + 16); s3 += lds64(base + i * 8 + 24); }
    t1 = clock64();
    x += s0 + s1 + s2 + s3;
    if (lane == 0) clk[2] = t1 - t0;
    // (d) 256 DFMA independent (32 accumulators), operands in registers
    t0 = clock64();
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 32; ++i) a[i] = fma(a[i], x, a[(i + 7) & 31] );
    t1 = clock64();
    if (lane == 0) clk[3] = t1 - t0;
    // (e) per-lane (non-broadcast) LDS.64 x 32 + DFMA
    t0 = clock64();
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 32; ++i) a[i] = fma(a[i], lds64(base + (r * 40 + i) * 32 * 8 % 16384 + lane * 8), x);
    t1 = clock64();
    if (lane == 0) clk[4] = t1 - t0;
  }
  __syncthreads();
  // (f) STS -> bar.sync(all warps of CTA via named barrier 1) -> LDS, chained 32 times; 4 warps
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 32; ++i) {
    if (warp == (i & 3)) sm[lane] = x;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    x += sm[(i + 1) & 31];
  }
  t1 = clock64();
  if (threadIdx.x == 0) clk[5] = (t1 - t0) / 32;
  // (g) same with 2 warps participating (64)
  if (warp < 2) {
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < 32; ++i) {
      if (warp == (i & 1)) sm[lane] = x;
      asm volatile("bar.sync 2, 64;" ::: "memory");
      x += sm[(i + 1) & 31];
    }
    t1 = clock64();
    if (threadIdx.x == 0) clk[6] = (t1 - t0) / 32;
  }
  // (h) flag hand-off through volatile shared memory: warp 0 writes data+flag, warp 1 polls (ping-pong)
  __syncthreads();
  volatile double* vsm = sm;
  volatile int* flag = reinterpret_cast<volatile int*>(sm + 2048);
  if (threadIdx.x == 0) { flag[0] = 0; }
  __syncthreads();
  if (warp < 2) {
    t0 = clock64();
#pragma unroll 1
    for (int i = 1; i <= 32; ++i) {
      if ((i & 1) == warp) {  // my turn to publish
        vsm[lane] = x + i;
        __threadfence_block();
        __syncwarp();
        if (lane == 0) flag[0] = i;
      } else {
        while (flag[0] < i) { }
        x += vsm[(i + 1) & 31];
      }
    }
    t1 = clock64();
    if (threadIdx.x == 0) clk[7] = (t1 - t0) / 32;
  }
#pragma unroll
  for (int i = 0; i < 32; ++i) x += a[i];
  out[threadIdx.x] = x;
}
int main() {
  double* out; long long* clk;
  cudaMalloc(&out, 8 * 256); cudaMalloc(&clk, 8 * 16);
  for (int rep = 0; rep < 2; ++rep) k<<<1, 128>>>(out, clk, 1);
  long long h[16];
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  printf("256x(LDS.64 bcast + DFMA) %lld | 128 LDS.128 + 256 DFMA %lld | 256 LDS.64 %lld | 256 DFMA %lld | 256x(LDS.64 per-lane + DFMA) %lld | STS-bar128-LDS %lld | STS-bar64-LDS %lld | flag hand-off %lld (%s)\n",
         h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7], cudaGetErrorString(cudaGetLastError()));
  return 0;
}
