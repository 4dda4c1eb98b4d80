// fp64_peaks.cu — microbenchmarks that set the denominators for the dense-Cholesky kernels:
//   (a) DMMA  (mma.sync.m8n8k4.f64) issue rate per SM,   (b) DFMA rate per SM,
//   (c) RED.ADD.F64 throughput for the two address patterns a Schur-complement scatter can use.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peaks fp64_peaks.cu ; run on a B200.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int NACC>
__global__ void dmma_kernel(int iters, double* out) {
  double acc[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { acc[i][0] = threadIdx.x; acc[i][1] = 1.0; }
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma884(acc[i][0], acc[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += acc[i][0] + acc[i][1];
  if (s == 12345.678) out[0] = s;
}

template <int NACC>
__global__ void dfma_kernel(int iters, double* out) {
  double acc[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i] = threadIdx.x + i;
  double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += acc[i];
  if (s == 12345.678) out[0] = s;
}

// pattern 0: every lane adds 36 values of its own random 6x6 block, one element per instruction (lanes -> 32 different blocks)
// pattern 1: a group of 36 "virtual lanes" covers one 6x6 block per step: lanes write contiguous rows (6 doubles = 48 B)
__global__ void red_kernel(int pattern, int n_blocks_per_thread, int nb /*blocks per row of S*/, double* S, size_t ld, unsigned seed) {
  const unsigned tid = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned st = tid * 2654435761u + seed;
  if (pattern == 0) {
    for (int k = 0; k < n_blocks_per_thread; ++k) {
      st = st * 1664525u + 1013904223u; const unsigned bi = (st >> 8) % nb;
      st = st * 1664525u + 1013904223u; const unsigned bj = (st >> 8) % nb;
      double* dst = S + (size_t)(6 * bi) * ld + 6 * bj;
#pragma unroll
      for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int b = 0; b < 6; ++b) atomicAdd(dst + a * ld + b, 1.0);
    }
  } else {
    // warp handles 32 blocks in sequence; for each block, lanes 0..31 cover elements 0..31, then lanes 0..3 cover 32..35
    const int lane = threadIdx.x & 31;
    unsigned wst = (tid >> 5) * 2654435761u + seed;
    for (int k = 0; k < n_blocks_per_thread * 32; ++k) {
      wst = wst * 1664525u + 1013904223u; const unsigned bi = (wst >> 8) % nb;
      wst = wst * 1664525u + 1013904223u; const unsigned bj = (wst >> 8) % nb;
      double* dst = S + (size_t)(6 * bi) * ld + 6 * bj;
      { const int e = lane; atomicAdd(dst + (e / 6) * ld + (e % 6), 1.0); }
      if (lane < 4) { const int e = 32 + lane; atomicAdd(dst + (e / 6) * ld + (e % 6), 1.0); }
    }
  }
}

template <typename F>
float time_ms(F f, int reps = 5) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms / reps;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  printf("device %s, %d SMs, %d kHz\n", p.name, sms, p.clockRate);
  double* out; cudaMalloc(&out, 8);
  const int iters = 20000;
  for (int warps : {4, 8, 16}) {
    float ms = time_ms([&] { dmma_kernel<8><<<sms, warps * 32>>>(iters, out); });
    double flop = (double)sms * warps * iters * 8 * 512.0;
    printf("DMMA m8n8k4 x8 acc, %2d warps/SM: %.3f ms  %.2f TFLOP/s\n", warps, ms, flop / ms * 1e-9);
    ms = time_ms([&] { dmma_kernel<16><<<sms, warps * 32>>>(iters, out); });
    flop = (double)sms * warps * iters * 16 * 512.0;
    printf("DMMA m8n8k4 x16 acc, %2d warps/SM: %.3f ms  %.2f TFLOP/s\n", warps, ms, flop / ms * 1e-9);
  }
  for (int warps : {4, 8, 16, 32}) {
    float ms = time_ms([&] { dfma_kernel<16><<<sms, warps * 32>>>(iters, out); });
    double flop = (double)sms * warps * 32 * iters * 16 * 2.0;
    printf("DFMA x16 acc, %2d warps/SM: %.3f ms  %.2f TFLOP/s\n", warps, ms, flop / ms * 1e-9);
  }
  // atomics: S of 10668^2 doubles (910 MB), 1778 block rows
  const int nb = 1778; const size_t ld = 6 * (size_t)nb;
  double* S; cudaMalloc(&S, ld * ld * 8); cudaMemset(S, 0, ld * ld * 8);
  const int threads = 256, grid = sms * 16, per = 64;
  for (int pat = 0; pat < 2; ++pat) {
    float ms = time_ms([&] { red_kernel<<<grid, threads>>>(pat, per, nb, S, ld, 17u); }, 3);
    double atoms = (double)grid * threads * per * 36.0;
    printf("RED.F64 pattern %d (full matrix): %.3f ms  %.1f G atomics/s  (%.1f M blocks)\n", pat, ms, atoms / ms * 1e-6, atoms / 36e6);
  }
  // band-limited (288 block band -> L2-resident-ish region ~ 1778*288*36*8 = 147 MB)
  for (int pat = 0; pat < 2; ++pat) {
    float ms = time_ms([&] { red_kernel<<<grid, threads>>>(pat, per, 400, S, ld, 17u); }, 3);
    double atoms = (double)grid * threads * per * 36.0;
    printf("RED.F64 pattern %d (400x400 blocks = 46 MB footprint): %.3f ms  %.1f G atomics/s\n", pat, ms, atoms / ms * 1e-6);
  }
  return 0;
}
