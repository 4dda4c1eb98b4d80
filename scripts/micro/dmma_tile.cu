// dmma_tile.cu — how fast can a warp tile of DMMA m8n8k4 fragments run when the fragments come from
// shared memory (the inner loop of the Cholesky SYRK/TRSM tiles), as a function of the warp-tile shape,
// warps per CTA and CTAs per SM.  No global traffic: isolates the LDS + DMMA issue pattern.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
constexpr int KC = 32, LD = KC + 4;

// warp grid WR x WC, warp tile (8*MA) x (8*NB)
template <int MA, int NB, int WR, int WC, bool SYNC>
__global__ void __launch_bounds__(WR * WC * 32) tile_kernel(int chunks, double* out) {
  extern __shared__ double sm[];
  constexpr int ROWS = WR * MA * 8, COLS = WC * NB * 8;
  double (*a)[LD] = reinterpret_cast<double (*)[LD]>(sm);
  double (*b)[LD] = reinterpret_cast<double (*)[LD]>(sm + ROWS * LD);
  for (int e = threadIdx.x; e < (ROWS + COLS) * LD; e += blockDim.x) sm[e] = 1.0 + e * 1e-7;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wr = warp / WC, wc = warp % WC, fr = lane >> 2, fk = lane & 3;
  double acc[MA][NB][2];
#pragma unroll
  for (int i = 0; i < MA; ++i)
#pragma unroll
    for (int j = 0; j < NB; ++j) { acc[i][j][0] = 0; acc[i][j][1] = 0; }
#pragma unroll 1
  for (int ch = 0; ch < chunks; ++ch) {
#pragma unroll
    for (int ks = 0; ks < KC / 4; ++ks) {
      double af[MA], bf[NB];
#pragma unroll
      for (int i = 0; i < MA; ++i) af[i] = a[wr * MA * 8 + i * 8 + fr][ks * 4 + fk];
#pragma unroll
      for (int j = 0; j < NB; ++j) bf[j] = b[wc * NB * 8 + j * 8 + fr][ks * 4 + fk];
#pragma unroll
      for (int i = 0; i < MA; ++i)
#pragma unroll
        for (int j = 0; j < NB; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
    }
    if (SYNC) __syncthreads();
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < MA; ++i)
#pragma unroll
    for (int j = 0; j < NB; ++j) s += acc[i][j][0] + acc[i][j][1];
  if (s == 1234.5) out[0] = s;
}

template <int MA, int NB, int WR, int WC, bool SYNC>
void run(const char* name, int ctas_per_sm, int sms, double* out) {
  constexpr int ROWS = WR * MA * 8, COLS = WC * NB * 8;
  const int smem = (ROWS + COLS) * LD * 8;
  auto kern = tile_kernel<MA, NB, WR, WC, SYNC>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WR * WC * 32, smem);
  const int chunks = 2000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  kern<<<sms * ctas_per_sm, WR * WC * 32, smem>>>(10, out);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  kern<<<sms * ctas_per_sm, WR * WC * 32, smem>>>(chunks, out);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double flop = (double)sms * ctas_per_sm * WR * WC * chunks * (KC / 4) * MA * NB * 512.0;
  printf("%-34s warp tile %3dx%-3d CTA %3dx%-3d x%d CTA/SM (occ %d) sync %d: %7.3f ms %6.2f TFLOP/s  [%s]\n", name, MA * 8, NB * 8, ROWS, COLS,
         ctas_per_sm, occ, (int)SYNC, ms, flop / ms * 1e-9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  double* out; cudaMalloc(&out, 8);
  run<8, 4, 2, 2, true>("current syrk (4 warps)", 2, sms, out);
  run<8, 4, 2, 2, true>("current syrk (4 warps)", 1, sms, out);
  run<8, 4, 2, 2, false>("current syrk nosync", 2, sms, out);
  run<4, 8, 2, 2, true>("32x64 warp", 2, sms, out);
  run<4, 4, 4, 2, true>("32x32 warp, 8 warps", 1, sms, out);
  run<4, 4, 4, 2, true>("32x32 warp, 8 warps", 2, sms, out);
  run<4, 4, 4, 4, true>("32x32 warp, 16 warps", 1, sms, out);
  run<8, 4, 2, 4, true>("64x32 warp, 8 warps (128x128)", 1, sms, out);
  run<4, 8, 4, 2, true>("32x64 warp, 8 warps (128x128)", 1, sms, out);
  run<8, 8, 2, 2, true>("64x64 warp, 4 warps (128x128)", 1, sms, out);
  run<2, 4, 4, 4, true>("16x32 warp, 16 warps (64x128)", 1, sms, out);
  run<8, 2, 2, 4, true>("64x16 warp, 8 warps", 1, sms, out);
  run<8, 1, 2, 4, true>("64x8 warp, 8 warps", 1, sms, out);
  return 0;
}
