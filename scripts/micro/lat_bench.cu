// lat_bench.cu — dependent-issue latencies of the instructions on the Cholesky pivot chain (one warp, one CTA).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* clk, double seed) {
  __shared__ double sm[128];
  const int lane = threadIdx.x & 31;
  double x = seed + lane * 1e-9, y = 1.0000001;
  long long t0, t1;
  // DFMA chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; ++i) x = fma(x, y, 1e-9);
  t1 = clock64();
  if (threadIdx.x == 0) clk[0] = (t1 - t0) / 64;
  // DMUL chain
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 64; ++i) x = x * y;
  t1 = clock64();
  if (threadIdx.x == 0) clk[1] = (t1 - t0) / 64;
  // MUFU.RCP64H chain (rcp.approx.ftz.f64)
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; ++i) asm volatile("rcp.approx.ftz.f64 %0, %0;" : "+d"(x));
  t1 = clock64();
  if (threadIdx.x == 0) clk[2] = (t1 - t0) / 32;
  // shuffle of a double (2 SHFL)
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; ++i) x = __shfl_sync(0xffffffffu, x, (i + 1) & 31);
  t1 = clock64();
  if (threadIdx.x == 0) clk[3] = (t1 - t0) / 32;
  // STS -> syncwarp -> LDS round trip
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    sm[lane] = x;
    __syncwarp();
    x = sm[(i + 1) & 31];
    __syncwarp();
  }
  t1 = clock64();
  if (threadIdx.x == 0) clk[4] = (t1 - t0) / 32;
  // 32 independent DFMA (issue rate of one warp)
  double a[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) a[i] = x + i;
  t0 = clock64();
#pragma unroll
  for (int r = 0; r < 8; ++r)
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = fma(a[i], y, x);
  t1 = clock64();
  if (threadIdx.x == 0) clk[5] = (t1 - t0);  // 256 DFMA
#pragma unroll
  for (int i = 0; i < 32; ++i) x += a[i];
  // 32 independent LDS.64 broadcast + DFMA
  t0 = clock64();
#pragma unroll
  for (int r = 0; r < 8; ++r)
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = fma(a[i], sm[(r * 7 + i) & 127], x);
  t1 = clock64();
  if (threadIdx.x == 0) clk[6] = (t1 - t0);  // 256 LDS+DFMA
#pragma unroll
  for (int i = 0; i < 32; ++i) x += a[i];
  // full fast_rcp3
  t0 = clock64();
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    double yy;
    asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(yy) : "d"(x));
    const double e = fma(-x, yy, 1.0);
    const double y1 = fma(yy, e, yy);
    x = fma(y1, e * e, y1);
  }
  t1 = clock64();
  if (threadIdx.x == 0) clk[7] = (t1 - t0) / 16;
  // bar.sync among 64 threads (named barrier) round trip
  out[threadIdx.x] = x;
}
int main() {
  double* out; long long* clk;
  cudaMalloc(&out, 8 * 256); cudaMalloc(&clk, 8 * 16);
  for (int w = 1; w <= 4; w *= 4) {
    k<<<1, 32 * w>>>(out, clk, 1.5);
    k<<<1, 32 * w>>>(out, clk, 1.5);
    long long h[16];
    cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    printf("warps %d: DFMA dep %lld  DMUL dep %lld  MUFU.RCP64H dep %lld  SHFL.f64 %lld  STS-sync-LDS-sync %lld  256 indep DFMA %lld  256 LDS+DFMA %lld  fast_rcp3 %lld  (%s)\n", w, h[0], h[1], h[2],
           h[3], h[4], h[5], h[6], h[7], cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
