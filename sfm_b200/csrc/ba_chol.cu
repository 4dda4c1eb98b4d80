// ba_chol.cu — dense FP64 Cholesky + triangular solves for the reduced camera system
// (the Eigen LLT that Ceres' DENSE_SCHUR calls, CTracker.cpp:574).
//
// Right-looking blocked factorisation of the lower triangle of a row-major n×n matrix with
// 128-wide panels.  Per panel three launches, enqueued back to back (no host synchronisation):
//   1. chol_diag_kernel   one CTA factors the 128×128 diagonal block in shared memory (4 x 32 sub-blocks:
//                         one warp runs the pivot chain, six work in its shadow) and also forms L11^-1,
//                         so that
//   2. chol_gemm<TRSM>    the panel solve  L21 = A21 L11^-T  is a GEMM, and
//   3. chol_gemm<SYRK>    the trailing update  A22 -= L21 L21^T  (lower tiles only)
// both run on the FP64 tensor path: 128×128 CTA tiles, K staged through shared memory with
// cp.async double buffering, mma.sync.m8n8k4.f64 (DMMA) fragments.  tcgen05 has no FP64 kind, so
// DMMA is the tensor-core instruction this workload can use (north_star: "tensor cores only for the
// dense Cholesky").  The stored L11^-1 blocks also turn the two triangular solves into mat-vecs.
// A non-positive pivot raises *fail; later kernels of the same factorisation return immediately.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "ba_internal.h"

namespace hsfm {
namespace {

constexpr int NB = 128;          // panel width / tile edge
constexpr int KC = 32;           // K chunk staged per pipeline stage
constexpr int LDS_ = KC + 4;     // padded shared-memory row stride (doubles): conflict-free DMMA fragment loads

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// Programmatic dependent launch (the chain kernels of the envelope factorisation): a kernel lets its successor in
// the stream be launched at once and waits for its predecessor's results itself, so the hand-over from kernel to
// kernel costs no launch latency.  Both are no-ops in a launch without the attribute.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

// Global loads for the latency-bound chain kernels (volatile: kept where they are written by the front end).
__device__ __forceinline__ int is_nonfinite(double v) { return (__double2hiint(v) & 0x7ff00000) == 0x7ff00000; }  // Inf or NaN
__device__ __forceinline__ double ld_now(const double* p) {
  double v;
  asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ld_now2(const double* p) {
  double2 v;
  asm volatile("ld.global.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// ---------------------------------------------------------------------------------------------
// 1. diagonal block: factor + inverse, one CTA, blocked 4 x 32
// ---------------------------------------------------------------------------------------------
// The 128x128 block lives in shared memory as its ten lower 32x32 sub-blocks (row stride 36 doubles:
// DMMA fragment loads in either orientation are bank-conflict free).  The kernel is a latency chain —
// 128 dependent pivots — so the work is split by whether it is ON that chain:
//   chain  (per sub-block column kq): warp 0 runs the 32 pivots of sub-block (kq,kq) (warp_chol32);
//          all warps scale it to L and put it out (diag_scale_out); warp 0 inverts it in place on DMMA
//          (warp_trinv32_dmma); all warps form L(kq+1,kq) as a product with that inverse and update
//          sub-block (kq+1,kq+1) alone (diag_chain_step) — the next pivots can start;
//   shadow (warps 1,2,3,5,6,7 while warp 0 runs the pivots of kq+1; warp 4 shares warp 0's scheduler and
//          idles): the rest of the sub-panel, the rest of the trailing update, L out to global, and the
//          rows of L11^-1 that are already determined (diag_shadow).
// After the fourth sub-block only the last row of L11^-1 remains (three 32^3 products).
constexpr int SB = 32;             // sub-block edge
constexpr int LB = SB + 4;         // sub-block row stride in shared memory (doubles)
constexpr int SBLK = SB * LB;      // doubles per stored sub-block
__device__ __forceinline__ constexpr int blk_id(int bi, int bj) { return bi * (bi + 1) / 2 + bj; }  // bj <= bi

// 1/d and 1/sqrt(d) from the hardware seed (MUFU, ~2^-22 relative) and two Newton steps of fused multiply-adds:
// full double precision for normal d > 0 without the special-case branches of the library routines.  The
// factorisation checks d > 0 and finiteness itself.
__device__ __forceinline__ double fast_rcp(double d) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  double e = fma(-d, y, 1.0);
  y = fma(y, e, y);
  e = fma(-d, y, 1.0);
  return fma(y, e, y);
}
__device__ __forceinline__ double fast_rsqrt(double d) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  double e = fma(-(d * y), y, 1.0);
  y = fma(0.5 * y, e, y);
  e = fma(-(d * y), y, 1.0);
  return fma(0.5 * y, e, y);
}

#ifdef HSFM_DIAG_TIMING
__device__ long long g_diag_clk[40];
__device__ long long g_step_clk[40];
__device__ unsigned g_mask[40];
__device__ unsigned g_mask2[40];
#define DIAG_TICK(k) do { const unsigned am_ = __activemask(); if (threadIdx.x == 0) { g_diag_clk[k] = clock64(); g_mask2[k] = am_; } __syncwarp(); } while (0)
#define SHADOW_TICK(k) do { if (KQ == 0) { if (tid6 == 0) g_diag_clk[k] = clock64(); __syncwarp(); } } while (0)
#ifdef HSFM_STEP_TIMING
#define STEP_TICK(k) do { const unsigned am_ = __activemask(); if (threadIdx.x == 0) { g_step_clk[k] = clock64(); g_mask[k] = am_; } __syncwarp(); } while (0)
#else
#define STEP_TICK(k) do { } while (0)
#endif
#else
#define DIAG_TICK(k) do { } while (0)
#define STEP_TICK(k) do { } while (0)
#define SHADOW_TICK(k) do { } while (0)
#endif

// 1/d with the second Newton correction taken as e*e (1 - d*y1 = e^2 exactly): three dependent multiply-adds
// after the seed instead of four; <= 2 ulp.
__device__ __forceinline__ double fast_rcp3(double d) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  const double e = fma(-d, y, 1.0);
  const double y1 = fma(y, e, y);
  const double e2 = e * e;
  return fma(y1, e2, y1);
}

// One pivot step of warp_chol32.  PAR = j & 1 (fixes which history entries are 16-byte pairs), W = live window
// width: columns c < W are updated, a[c-1] = a[c] - u h[c].
template <int PAR, int W>
__device__ __forceinline__ void chol32_step(double (&a)[SB], double* __restrict__ hist, int j, int lane, double& r, double& sq, double& g, int& bad) {
  STEP_TICK(j);
  const double* hs = hist + j * SB + j;  // hs[c] = A(j)[j+c][j]
  const double dn = fma(-sq, r, g);      // pivot j+1 (don't-care at j = 31): the serial chain
  const double rn = fast_rcp3(dn);
  const double nu = -(a[0] * r);
  if (PAR == 0) {
    a[0] = fma(nu, hs[1], a[1]);
#pragma unroll
    for (int c = 2; c + 1 < W; c += 2) {
      const double2 h2 = *reinterpret_cast<const double2*>(hs + c);
      a[c - 1] = fma(nu, h2.x, a[c]);
      a[c] = fma(nu, h2.y, a[c + 1]);
    }
  } else {
#pragma unroll
    for (int c = 1; c + 1 < W; c += 2) {
      const double2 h2 = *reinterpret_cast<const double2*>(hs + c);
      a[c - 1] = fma(nu, h2.x, a[c]);
      a[c] = fma(nu, h2.y, a[c + 1]);
    }
    a[W - 2] = fma(nu, hs[W - 1], a[W - 1]);
  }
  hist[(j + 1) * SB + lane] = a[0];
  const double hq = __shfl_sync(0xffffffffu, a[0], (j + 2) & 31), gq = __shfl_sync(0xffffffffu, a[1], (j + 2) & 31);
  sq = hq * hq;
  g = gq;
  if (j + 1 < SB && !bad && !(dn > 0.0 && dn < 1e300)) bad = j + 2;
  r = rn;
  __syncwarp();
}

// Cholesky of a 32x32 sub-block by one warp, lane r = row r in a register window that shifts left one column
// per step.  The serial chain of the whole factorisation is pivot -> 1/pivot -> next pivot; everything else is
// kept off it:
//   * step j publishes the lanes' next column entries A(j+1)[.][j+1] to slot j+1 of a history buffer (the
//     sub-block's own shared memory: its rows are in registers); the next-but-one pivot's two ingredients come
//     by shuffle from the lane that owns them, so the next pivot is d(j+1) = g - h*h*r with h*h ready before r;
//   * the reciprocal square roots that scale L are taken once, after the loop, 32 in parallel;
//   * the window narrows every eight steps (32, 24, 16, 8 columns), history entries are read as 16-byte pairs.
// D (row stride LB) receives L (strict upper zeroed), dinv[r] = 1/L[r][r].  Returns 0 or the 1-based failing column.
__device__ __noinline__ int warp_chol32(double* __restrict__ D, double* __restrict__ dinv, int lane) {
  STEP_TICK(34);
  double a[SB];
#pragma unroll
  for (int c = 0; c < SB; c += 2) {
    const double2 v = *reinterpret_cast<const double2*>(D + lane * LB + c);
    a[c] = v.x; a[c + 1] = v.y;
  }
  __syncwarp();
  double* hist = D;  // slot j, entry i: A(j)[i][j]  (unscaled column j of L); 33 slots of 32 fit in 32 x LB
  hist[lane] = a[0];
  const double d = __shfl_sync(0xffffffffu, a[0], 0);
  int bad = (d > 0.0 && d < 1e300) ? 0 : 1;
  double r = fast_rcp(d);
  const double h1 = __shfl_sync(0xffffffffu, a[0], 1);
  double sq = h1 * h1, g = __shfl_sync(0xffffffffu, a[1], 1);
  __syncwarp();
#pragma unroll 1
  for (int j = 0; j < 8; j += 2) { chol32_step<0, 32>(a, hist, j, lane, r, sq, g, bad); chol32_step<1, 32>(a, hist, j + 1, lane, r, sq, g, bad); }
#pragma unroll 1
  for (int j = 8; j < 16; j += 2) { chol32_step<0, 24>(a, hist, j, lane, r, sq, g, bad); chol32_step<1, 24>(a, hist, j + 1, lane, r, sq, g, bad); }
#pragma unroll 1
  for (int j = 16; j < 24; j += 2) { chol32_step<0, 16>(a, hist, j, lane, r, sq, g, bad); chol32_step<1, 16>(a, hist, j + 1, lane, r, sq, g, bad); }
#pragma unroll 1
  for (int j = 24; j < 32; j += 2) { chol32_step<0, 8>(a, hist, j, lane, r, sq, g, bad); chol32_step<1, 8>(a, hist, j + 1, lane, r, sq, g, bad); }
  STEP_TICK(32);
  return bad;
}

// Barrier of the two chain warps (0 and 4).
__device__ __forceinline__ void chain_sync() { asm volatile("bar.sync 2, 64;" ::: "memory"); }

// After warp_chol32: warps 0 and 4 turn the pivot-column history (slot j, entry i = A(j)[i][j], in the sub-block's
// own storage) into L[i][j] = A(j)[i][j] / sqrt(A(j)[j][j]) in place (row stride LB, strict upper zeroed),
// dinv[j] = 1/L[j][j], and put the lower triangle out to global memory.  Thread = (row lane, sixteen columns).
__device__ __forceinline__ void diag_scale_out(double* __restrict__ D, double* __restrict__ dinv, double* __restrict__ gdst, int n, int rows_valid,
                                               int warp, int lane) {
  const int c16 = warp ? 16 : 0;
  if (warp == 0) dinv[lane] = fast_rsqrt(D[lane * SB + lane]);
  chain_sync();
  double l[16];
#pragma unroll
  for (int q = 0; q < 16; q += 2) {
    const double2 iv = *reinterpret_cast<const double2*>(dinv + c16 + q);
    const int j = c16 + q;
    l[q] = (lane >= j) ? D[j * SB + lane] * iv.x : 0.0;
    l[q + 1] = (lane >= j + 1) ? D[(j + 1) * SB + lane] * iv.y : 0.0;
  }
  chain_sync();
  double* dst = gdst + (size_t)lane * n + c16;
#pragma unroll
  for (int q = 0; q < 16; q += 2) {
    *reinterpret_cast<double2*>(D + lane * LB + c16 + q) = make_double2(l[q], l[q + 1]);
    const int c = c16 + q;
    if (lane < rows_valid) {
      if (c + 1 <= lane) *reinterpret_cast<double2*>(dst + q) = make_double2(l[q], l[q + 1]);
      else if (c == lane) dst[q] = l[q];
    }
  }
  chain_sync();
}

// In-place inverse of the lower-triangular 32x32 sub-block D (row stride LB, strict upper zero) by one warp, by
// recursive doubling on DMMA: the four 8x8 diagonal blocks by substitution in registers (lane = block*8 + column),
// then X21 = -X22 (L21 X11) at block size 8 and 16.  ~1/5 of the latency of a 32-step substitution.  scr: 320 doubles.
__device__ __noinline__ void warp_trinv32_dmma(double* D, const double* __restrict__ dinv, double* scr, int lane) {
  const int fr = lane >> 2, fk = lane & 3;
  {
    const int b = lane >> 3, c = lane & 7;
    const double* Lb = D + (b * 8) * LB + b * 8;
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      double s = (i == c) ? 1.0 : 0.0;
#pragma unroll
      for (int k = 0; k < i; ++k) s = fma(-Lb[i * LB + k], x[k], s);
      x[i] = s * dinv[b * 8 + i];
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 8; ++i) D[(b * 8 + i) * LB + b * 8 + c] = x[i];
    __syncwarp();
  }
  // block size 8: pairs (0,1) and (2,3)
  {
    double t0[2] = {0.0, 0.0}, t1[2] = {0.0, 0.0};
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const double* L21 = D + (16 * p + 8) * LB + 16 * p;
      const double* X11 = D + (16 * p) * LB + 16 * p;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) dmma884(t0[p], t1[p], L21[fr * LB + ks * 4 + fk], X11[(ks * 4 + fk) * LB + fr]);
    }
#pragma unroll
    for (int p = 0; p < 2; ++p) *reinterpret_cast<double2*>(scr + p * 96 + fr * 12 + 2 * fk) = make_double2(t0[p], t1[p]);
    __syncwarp();
    double x0[2] = {0.0, 0.0}, x1[2] = {0.0, 0.0};
#pragma unroll
    for (int p = 0; p < 2; ++p) {
      const double* X22 = D + (16 * p + 8) * LB + 16 * p + 8;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) dmma884(x0[p], x1[p], X22[fr * LB + ks * 4 + fk], scr[p * 96 + (ks * 4 + fk) * 12 + fr]);
    }
#pragma unroll
    for (int p = 0; p < 2; ++p)
      *reinterpret_cast<double2*>(D + (16 * p + 8 + fr) * LB + 16 * p + 2 * fk) = make_double2(-x0[p], -x1[p]);
    __syncwarp();
  }
  // block size 16
  {
    double t0[4], t1[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) { t0[q] = 0.0; t1[q] = 0.0; }
    const double* L21 = D + 16 * LB;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int ti = q >> 1, tj = q & 1;
        dmma884(t0[q], t1[q], L21[(ti * 8 + fr) * LB + ks * 4 + fk], D[(ks * 4 + fk) * LB + tj * 8 + fr]);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int ti = q >> 1, tj = q & 1;
      *reinterpret_cast<double2*>(scr + (ti * 8 + fr) * 20 + tj * 8 + 2 * fk) = make_double2(t0[q], t1[q]);
    }
    __syncwarp();
    double x0[4], x1[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) { x0[q] = 0.0; x1[q] = 0.0; }
    const double* X22 = D + 16 * LB + 16;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int ti = q >> 1, tj = q & 1;
        dmma884(x0[q], x1[q], X22[(ti * 8 + fr) * LB + ks * 4 + fk], scr[(ks * 4 + fk) * 20 + tj * 8 + fr]);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int ti = q >> 1, tj = q & 1;
      *reinterpret_cast<double2*>(D + (16 + ti * 8 + fr) * LB + tj * 8 + 2 * fk) = make_double2(-x0[q], -x1[q]);
    }
    __syncwarp();
  }
}

// Barrier of the six "shadow" warps (1,2,3,5,6,7) that work beside the factoring warp 0; warp 4 shares warp 0's
// scheduler and stays idle.
__device__ __forceinline__ void shadow_sync() { asm volatile("bar.sync 1, 192;" ::: "memory"); }

// One 8x8 output tile of the 32x32 block product A * B over K = 32, accumulated into (c0, c1).  A is [m][k];
// B is [k][n] when B_KN, else [n][k] (i.e. the product is A * B^T).
template <bool B_KN>
__device__ __forceinline__ void tile_acc32(double& c0, double& c1, const double* __restrict__ A, const double* __restrict__ B, int ti, int tj,
                                           int fr, int fk, int ks_lo = 0, int ks_hi = SB / 4) {
#pragma unroll
  for (int ks = 0; ks < SB / 4; ++ks) {
    if (ks >= ks_lo && ks < ks_hi) {  // warp-uniform: triangular operands skip their zero k-steps
      const double a = A[(ti * 8 + fr) * LB + ks * 4 + fk];
      const double b = B_KN ? B[(ks * 4 + fk) * LB + tj * 8 + fr] : B[(tj * 8 + fr) * LB + ks * 4 + fk];
      dmma884(c0, c1, a, b);
    }
  }
}

// T output tiles of a warp advance together through the eight k-steps of a 32^3 block product, so that their DMMA
// dependency chains overlap.  oa/ob: offsets (doubles, from sL) of the lane's A and B fragment at k-step 0; a tile
// takes k-steps lo <= ks < hi only (triangular operands skip their zero k-steps; hi = 0 switches a tile off).
template <int T, bool B_KN>
__device__ __forceinline__ void tiles_mma(double (&c0)[T], double (&c1)[T], const double* __restrict__ sL, const int (&oa)[T], const int (&ob)[T],
                                          const int (&lo)[T], const int (&hi)[T]) {
#pragma unroll
  for (int ks = 0; ks < SB / 4; ++ks) {
#pragma unroll
    for (int q = 0; q < T; ++q)
      if (ks >= lo[q] && ks < hi[q]) dmma884(c0[q], c1[q], sL[oa[q] + ks * 4], sL[ob[q] + (B_KN ? ks * 4 * LB : ks * 4)]);
  }
}
__device__ __forceinline__ int frag_a(int blk, int ti, int fr, int fk) { return blk * SBLK + (ti * 8 + fr) * LB + fk; }       // A [m][k]
__device__ __forceinline__ int frag_bt(int blk, int tj, int fr, int fk) { return blk * SBLK + (tj * 8 + fr) * LB + fk; }      // B as [n][k]
__device__ __forceinline__ int frag_bn(int blk, int tj, int fr, int fk) { return blk * SBLK + fk * LB + tj * 8 + fr; }        // B as [k][n]
__device__ __forceinline__ int frag_c(int blk, int ti, int tj, int fr, int fk) { return blk * SBLK + (ti * 8 + fr) * LB + tj * 8 + 2 * fk; }

// Work of step KQ that is NOT on the chain, done by warps 1,2,3,5,6,7 while warp 0 runs the pivots of sub-block KQ+1:
//   (a0) the sub-panel rows below sub-block (KQ+1,KQ),           (a) the trailing update of every sub-block but (KQ+1,KQ+1),
//   (b) L(.,KQ) below the diagonal -> global,                     (c1) row KQ of the inverse  X(KQ,j) = -X(KQ,KQ) Y(KQ,j),
//   (d) that row -> global,
//   (c2) Y(KQ+1,j) = sum_{k=j..KQ} L(KQ+1,k) X(k,j) over L(KQ+1,j)   (the part of row KQ+1 that needs no X(KQ+1,KQ+1)).
template <int KQ>
__device__ __forceinline__ void diag_shadow(double* sL, double* __restrict__ A, int n, int k0, int kb, double* __restrict__ Linv, int tid6) {
  const int w6 = tid6 >> 5, lane = tid6 & 31, fr = lane >> 2, fk = lane & 3;
  SHADOW_TICK(28);
  // (a0) L(bi,KQ) = P(bi,KQ) X(KQ,KQ)^T, bi >= KQ+2
  if constexpr (KQ < 2) {
    constexpr int NT = (2 - KQ) * 16, T = (NT + 5) / 6;
    double c0[T], c1[T];
    int oa[T], ob[T], oc[T], lo[T], hi[T];
#pragma unroll
    for (int q = 0; q < T; ++q) {
      const int t = w6 + 6 * q, bi = KQ + 2 + (t >> 4), ti = (t >> 2) & 3, tj = t & 3;
      c0[q] = 0.0; c1[q] = 0.0;
      oa[q] = frag_a(blk_id(bi, KQ), ti, fr, fk);
      ob[q] = frag_bt(blk_id(KQ, KQ), tj, fr, fk);
      oc[q] = frag_c(blk_id(bi, KQ), ti, tj, fr, fk);
      lo[q] = 0; hi[q] = t < NT ? 2 * tj + 2 : 0;  // X is lower triangular: X[n][k] = 0 for k > n
    }
    tiles_mma<T, false>(c0, c1, sL, oa, ob, lo, hi);
    shadow_sync();
#pragma unroll
    for (int q = 0; q < T; ++q)
      if (hi[q]) *reinterpret_cast<double2*>(sL + oc[q]) = make_double2(c0[q], c1[q]);
    shadow_sync();
  }
  SHADOW_TICK(29);
  // (a) L(bi,bj) -= L(bi,KQ) L(bj,KQ)^T, in two halves (registers)
  if constexpr (KQ < 2) {
    constexpr int NBLK = KQ == 0 ? 5 : 2;
    constexpr int BI[5] = {KQ == 0 ? 2 : 3, KQ == 0 ? 2 : 3, 3, 3, 3};
    constexpr int BJ[5] = {KQ == 0 ? 1 : 2, KQ == 0 ? 2 : 3, 1, 2, 3};
    constexpr int T = (NBLK * 16 + 11) / 12;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      double c0[T], c1[T];
      int oa[T], ob[T], oc[T], lo[T], hi[T];
#pragma unroll
      for (int q = 0; q < T; ++q) {
        const int t = w6 + 6 * (half * T + q), b = t >> 4, ti = (t >> 2) & 3, tj = t & 3;
        int bi = BI[0], bj = BJ[0];
#pragma unroll
        for (int z = 1; z < 5; ++z)
          if (b == z) { bi = BI[z]; bj = BJ[z]; }
        c0[q] = 0.0; c1[q] = 0.0;
        oa[q] = frag_a(blk_id(bi, KQ), ti, fr, fk);
        ob[q] = frag_bt(blk_id(bj, KQ), tj, fr, fk);
        oc[q] = frag_c(blk_id(bi, bj), ti, tj, fr, fk);
        lo[q] = 0; hi[q] = (t < NBLK * 16 && (bi != bj || ti >= tj)) ? SB / 4 : 0;  // strict upper of a diagonal sub-block: never read
      }
      tiles_mma<T, false>(c0, c1, sL, oa, ob, lo, hi);
#pragma unroll
      for (int q = 0; q < T; ++q)
        if (hi[q]) {
          double2* pd = reinterpret_cast<double2*>(sL + oc[q]);
          double2 v = *pd;
          v.x -= c0[q]; v.y -= c1[q];
          *pd = v;
        }
    }
  }
  SHADOW_TICK(30);
  // (b)
  {
    constexpr int NE = (3 - KQ) * SB * (SB / 2);
    for (int e = tid6; e < NE; e += 192) {
      const int blk = e / (SB * SB / 2), rem = e - blk * (SB * SB / 2), r = rem >> 4, c = (rem & 15) * 2;
      const int i = (KQ + 1 + blk) * SB + r;
      if (i < kb)
        *reinterpret_cast<double2*>(A + (size_t)(k0 + i) * n + k0 + KQ * SB + c) =
            *reinterpret_cast<const double2*>(sL + blk_id(KQ + 1 + blk, KQ) * SBLK + r * LB + c);
    }
  }
  shadow_sync();
  SHADOW_TICK(31);
  // (c1) X(KQ,j) = -X(KQ,KQ) Y(KQ,j), j < KQ
  if constexpr (KQ >= 1) {
    constexpr int NT = KQ * 16, T = (NT + 5) / 6;
    double c0[T], c1[T];
    int oa[T], ob[T], oc[T], lo[T], hi[T];
#pragma unroll
    for (int q = 0; q < T; ++q) {
      const int t = w6 + 6 * q, j = t >> 4, ti = (t >> 2) & 3, tj = t & 3;
      c0[q] = 0.0; c1[q] = 0.0;
      oa[q] = frag_a(blk_id(KQ, KQ), ti, fr, fk);
      ob[q] = frag_bn(blk_id(KQ, j < KQ ? j : 0), tj, fr, fk);
      oc[q] = frag_c(blk_id(KQ, j < KQ ? j : 0), ti, tj, fr, fk);
      lo[q] = 0; hi[q] = t < NT ? 2 * ti + 2 : 0;  // X(KQ,KQ)[m][k] = 0 for k > m
    }
    tiles_mma<T, true>(c0, c1, sL, oa, ob, lo, hi);
    shadow_sync();
#pragma unroll
    for (int q = 0; q < T; ++q)
      if (hi[q]) *reinterpret_cast<double2*>(sL + oc[q]) = make_double2(-c0[q], -c1[q]);
    shadow_sync();
  }
  SHADOW_TICK(32);
  // (d)
  for (int e = tid6; e < SB * (NB / 2); e += 192) {
    const int r = e >> 6, cc = (e & 63) * 2, bj = cc >> 5;
    double2 v = make_double2(0.0, 0.0);
    if (bj <= KQ) v = *reinterpret_cast<const double2*>(sL + blk_id(KQ, bj) * SBLK + r * LB + (cc & 31));
    *reinterpret_cast<double2*>(Linv + (size_t)(KQ * SB + r) * NB + cc) = v;
  }
  SHADOW_TICK(33);
  // (c2) Y(KQ+1,j) = sum_{k=j..KQ} L(KQ+1,k) X(k,j), j <= KQ
  {
    constexpr int NT = (KQ + 1) * 16, T = (NT + 5) / 6;
    double c0[T], c1[T];
    int oc[T];
#pragma unroll
    for (int q = 0; q < T; ++q) { c0[q] = 0.0; c1[q] = 0.0; }
#pragma unroll
    for (int k = 0; k <= KQ; ++k) {
      int oa[T], ob[T], lo[T], hi[T];
#pragma unroll
      for (int q = 0; q < T; ++q) {
        const int t = w6 + 6 * q, j = t >> 4, ti = (t >> 2) & 3, tj = t & 3, jj = j <= k ? j : k;
        oa[q] = frag_a(blk_id(KQ + 1, k), ti, fr, fk);
        ob[q] = frag_bn(blk_id(k, jj), tj, fr, fk);
        oc[q] = frag_c(blk_id(KQ + 1, j <= KQ ? j : 0), ti, tj, fr, fk);
        lo[q] = k == j ? 2 * tj : 0;  // X(j,j)[k][n] = 0 for k < n
        hi[q] = (t < NT && k >= j) ? SB / 4 : 0;
      }
      tiles_mma<T, true>(c0, c1, sL, oa, ob, lo, hi);
    }
    shadow_sync();
#pragma unroll
    for (int q = 0; q < T; ++q)
      if (w6 + 6 * q < NT) *reinterpret_cast<double2*>(sL + oc[q]) = make_double2(c0[q], c1[q]);
  }
  SHADOW_TICK(34);
}

// The part of step KQ that IS on the chain, all eight warps: sub-block (KQ+1,KQ) of the panel as a product with the
// inverse, L(KQ+1,KQ) = P(KQ+1,KQ) X(KQ,KQ)^T, then the trailing update of sub-block (KQ+1,KQ+1) alone.
template <int KQ>
__device__ __forceinline__ void diag_chain_step(double* sL, int warp, int lane) {
  const int fr = lane >> 2, fk = lane & 3, tj = warp & 3, ti0 = warp >> 2;
  {
    double c0[2] = {0.0, 0.0}, c1[2] = {0.0, 0.0};
    const double* X = sL + blk_id(KQ, KQ) * SBLK;
    double* P = sL + blk_id(KQ + 1, KQ) * SBLK;
#pragma unroll
    for (int ks = 0; ks < SB / 4; ++ks) {
      if (ks < 2 * tj + 2) {  // X is lower triangular: X[n][k] = 0 for k > n
        const double b = X[(tj * 8 + fr) * LB + ks * 4 + fk];
#pragma unroll
        for (int q = 0; q < 2; ++q) dmma884(c0[q], c1[q], P[((ti0 + 2 * q) * 8 + fr) * LB + ks * 4 + fk], b);
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 2; ++q) *reinterpret_cast<double2*>(P + ((ti0 + 2 * q) * 8 + fr) * LB + tj * 8 + 2 * fk) = make_double2(c0[q], c1[q]);
    __syncthreads();
  }
  {
    const double* P = sL + blk_id(KQ + 1, KQ) * SBLK;
    double c0[2] = {0.0, 0.0}, c1[2] = {0.0, 0.0};
#pragma unroll
    for (int ks = 0; ks < SB / 4; ++ks) {
      const double b = P[(tj * 8 + fr) * LB + ks * 4 + fk];
#pragma unroll
      for (int q = 0; q < 2; ++q) dmma884(c0[q], c1[q], P[((ti0 + 2 * q) * 8 + fr) * LB + ks * 4 + fk], b);
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      double2* pd = reinterpret_cast<double2*>(sL + blk_id(KQ + 1, KQ + 1) * SBLK + ((ti0 + 2 * q) * 8 + fr) * LB + tj * 8 + 2 * fk);
      double2 v = *pd;
      v.x -= c0[q]; v.y -= c1[q];
      *pd = v;
    }
    __syncthreads();
  }
}

// Step KQ of the diagonal block.  Returns true when a pivot failed (uniform over the CTA).
template <int KQ>
__device__ __forceinline__ bool diag_phase(double* sL, double* sDinv, double* sScr, int* sBad, double* __restrict__ A, int n, int k0,
                                           int kb, double* __restrict__ Linv, int* __restrict__ fail) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* D = sL + blk_id(KQ, KQ) * SBLK;
  if ((warp & 3) == 0) {
    // the chain: warp 0 runs the pivots (warp 4, on the same scheduler, waits), both scale and put L out, warp 0 inverts
    if (warp == 0) {
      const int bad = warp_chol32(D, sDinv + KQ * SB, lane);
      if (bad && lane == 0) { *sBad = 1; atomicExch(fail, k0 + KQ * SB + bad); }
    }
    chain_sync();
    if (!*sBad) {
      diag_scale_out(D, sDinv + KQ * SB, A + (size_t)(k0 + KQ * SB) * n + k0 + KQ * SB, n, kb - KQ * SB, warp, lane);
      if (warp == 0) warp_trinv32_dmma(D, sDinv + KQ * SB, sScr, lane);
    }
  } else if (KQ > 0) {
    diag_shadow<(KQ > 0 ? KQ - 1 : 0)>(sL, A, n, k0, kb, Linv, warp < 4 ? tid - 32 : tid - 64);
#ifdef HSFM_DIAG_TIMING
    if (tid == 32) g_diag_clk[24 + KQ] = clock64();
    __syncwarp();
#endif
  }
  __syncthreads();
  if (*sBad) return true;
  DIAG_TICK(8 + 3 * KQ);
  if (KQ < 3) {
    diag_chain_step<(KQ < 3 ? KQ : 0)>(sL, warp, lane);
    DIAG_TICK(10 + 3 * KQ);
  }
  return false;
}

#ifdef HSFM_CHAIN_TIMING
__device__ unsigned long long g_chain_ts[4096];  // [2 * panel] start, [2 * panel + 1] end of the diagonal-block kernel (globaltimer, ns)
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#endif
// 10 sub-blocks + 1/diag + inverse scratch: 95 KB, so that (with <= 128 registers per thread) the kernel
// fits on an SM NEXT TO a resident trailing-update CTA and never has to wait for a whole SM to drain.
constexpr int kDiagSmem = (10 * SBLK + NB + 320) * (int)sizeof(double) + 16;

__global__ void __launch_bounds__(256, 2) chol_diag_kernel(double* __restrict__ A, int n, int k0, int kb, double* Linv, int* __restrict__ fail) {
  extern __shared__ __align__(16) double dsm[];
  double* sL = dsm;                 // [10][32][LB] lower sub-blocks: the matrix, then L, then (in place) the pieces of L^-1
  double* sDinv = sL + 10 * SBLK;   // [128] 1 / L[i][i]
  double* sScr = sDinv + NB;        // [320] scratch of warp_trinv32_dmma
  int* sBad = reinterpret_cast<int*>(sScr + 320);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_enter();
  const int failed = *fail;
  if (tid == 0) *sBad = 0;
#ifdef HSFM_CHAIN_TIMING
  if (tid == 0) g_chain_ts[2 * ((k0 / NB) & 2047)] = gtime();
#endif
  DIAG_TICK(0);
  // load the lower sub-blocks (identity beyond kb; strict upper of the diagonal sub-blocks zero).  Element
  // e = tid + 256 it lies in sub-block it >> 1: the block decode is a compile-time constant of the unrolled
  // loop and the twenty 16-byte loads of a thread are all in flight before the first store.
  {
    double2 v[20];
#pragma unroll
    for (int it = 0; it < 20; ++it) {
      const int blk = it >> 1, rem = tid + 256 * (it & 1);
      const int bi = blk < 1 ? 0 : blk < 3 ? 1 : blk < 6 ? 2 : 3, bj = blk - blk_id(bi, 0);
      const int r = rem / (SB / 2), c = (rem - r * (SB / 2)) * 2;
      const int i = bi * SB + r, j = bj * SB + c;
      v[it] = make_double2(0.0, 0.0);
      if (i < kb && j + 1 < kb) v[it] = *reinterpret_cast<const double2*>(A + (size_t)(k0 + i) * n + k0 + j);
      else if (i < kb && j < kb) v[it].x = A[(size_t)(k0 + i) * n + k0 + j];
    }
#pragma unroll
    for (int it = 0; it < 20; ++it) {
      const int blk = it >> 1, rem = tid + 256 * (it & 1);
      const int bi = blk < 1 ? 0 : blk < 3 ? 1 : blk < 6 ? 2 : 3, bj = blk - blk_id(bi, 0);
      const int r = rem / (SB / 2), c = (rem - r * (SB / 2)) * 2;
      const int i = bi * SB + r, j = bj * SB + c;
      double2 w = v[it];
      if (j > i) w.x = 0.0;
      if (j + 1 > i) w.y = 0.0;
      if (i >= kb) { w.x = (i == j) ? 1.0 : 0.0; w.y = (i == j + 1) ? 1.0 : 0.0; }
      *reinterpret_cast<double2*>(sL + blk * SBLK + r * LB + c) = w;
    }
  }
  __syncthreads();
  if (failed) return;
  DIAG_TICK(1);
  if (diag_phase<0>(sL, sDinv, sScr, sBad, A, n, k0, kb, Linv, fail)) return;
  if (diag_phase<1>(sL, sDinv, sScr, sBad, A, n, k0, kb, Linv, fail)) return;
  if (diag_phase<2>(sL, sDinv, sScr, sBad, A, n, k0, kb, Linv, fail)) return;
  if (diag_phase<3>(sL, sDinv, sScr, sBad, A, n, k0, kb, Linv, fail)) return;
  // ---- last row of the inverse: X(3,3) and X(3,j) = -X(3,3) Y(3,j), straight to global ----
  {
    const int fr = lane >> 2, fk = lane & 3;
    const double* X33 = sL + blk_id(3, 3) * SBLK;
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      const int t = warp + 8 * q, j = t >> 4, tile = t & 15, ti = tile >> 2, tj = tile & 3;
      double c0 = 0.0, c1 = 0.0;
      const double* Y = sL + blk_id(3, j) * SBLK;
#pragma unroll
      for (int ks = 0; ks < SB / 4; ++ks)
        if (ks < 2 * ti + 2) dmma884(c0, c1, X33[(ti * 8 + fr) * LB + ks * 4 + fk], Y[(ks * 4 + fk) * LB + tj * 8 + fr]);
      *reinterpret_cast<double2*>(Linv + (size_t)(3 * SB + ti * 8 + fr) * NB + j * SB + tj * 8 + 2 * fk) = make_double2(-c0, -c1);
    }
    for (int e = tid; e < SB * (SB / 2); e += 256) {
      const int r = e >> 4, c = (e & 15) * 2;
      *reinterpret_cast<double2*>(Linv + (size_t)(3 * SB + r) * NB + 3 * SB + c) = *reinterpret_cast<const double2*>(X33 + r * LB + c);
    }
  }
  DIAG_TICK(6);
#ifdef HSFM_CHAIN_TIMING
  __syncthreads();
  if (tid == 0) g_chain_ts[2 * ((k0 / NB) & 2047) + 1] = gtime();
#endif
}

// ---------------------------------------------------------------------------------------------
// 3. Trailing update  S[bi,bj] -= L21[bi] L21[bj]^T  on DMMA: 128×64 output tiles, K = 128 (the panel) in four
// chunks of 32 staged through shared memory.  PERSISTENT: the grid is (at most) two CTAs per SM and every CTA
// walks tiles t = blockIdx.x, blockIdx.x + gridDim.x, …; the two-stage cp.async ring runs straight across
// tile boundaries (the first chunk of the next tile streams in while the last chunk of this one is in the
// tensor pipe) and the next tile's C is pulled into L2 a whole tile ahead, so the only exposed latency per
// tile is one L2 round trip for the accumulator preload.  4 warps as 2×2, each 64×32 = 8×4 DMMA tiles
// (m8n8k4): 64 FP64 accumulators per lane that START at C; the loop feeds NEGATED A fragments (free operand
// modifier), so the epilogue is a plain store — no read-modify-write, no arithmetic between the 32
// independent accumulator loads (they stay in flight together).
// Tile sets (64-wide column blocks cb, 128-row tiles bi, relative to the trailing matrix):
//   col_first >= 0 : row tiles row_first.., column blocks {col_first, col_first+1}   (2 tiles per row tile)
//   col_first <  0 : bulk — column blocks c0 = -col_first (even, >= 2) and beyond, lower tiles only
// ---------------------------------------------------------------------------------------------
constexpr int TN = 64;  // tile columns
// TM = tile rows: 128 (8 warps as 4 x 2 of 32 x 32, two CTAs per SM) or 64 (4 warps as 2 x 2, three CTAs per SM: the
// structured factorisation's launches are single partial waves, and half-size tiles spread them evenly over the SMs)
template <int TM>
struct GemmSmem {
  double a[2][TM][LDS_];
  double b[2][TN][LDS_];
};
constexpr int kGemmSmem = (int)sizeof(GemmSmem<NB>), kGemmSmemHalf = (int)sizeof(GemmSmem<64>);

template <int ROWS, int THREADS>
__device__ __forceinline__ void load_tile_async(double (*dst)[LDS_], const double* __restrict__ src, int ld, int row0, int rows_valid,
                                                int col0) {
  // ROWS rows × 32 doubles = ROWS*16 chunks of 16 B over the CTA's threads
#pragma unroll
  for (int q = 0; q < ROWS * 16 / THREADS; ++q) {
    const int e = threadIdx.x + THREADS * q;
    const int row = e >> 4, ch = e & 15;
    double* d = &dst[row][ch * 2];
    if (row < rows_valid) cp_async16(d, src + (size_t)(row0 + row) * ld + col0 + ch * 2);
    else { d[0] = 0.0; d[1] = 0.0; }
  }
}

struct SyrkTile { int ri, ni, rj, nj, oi, oj; };  // operand rows in S; oi, oj: row / column of the output tile in its matrix
// rb_list != nullptr: structured mode — the tiles are pairs (a, b), b <= a, of the panel's ACTIVE 128-row blocks
// rb_list[0 .. cnt) (global block indices, ascending), two 64-wide tiles per pair, enumerated column by column:
// column b holds the rows a = b .. cnt-1.  A launch with fewer tiles than all pairs covers the first block columns
// only (the two-chain schedule defers the others); col_first == -2 skips pair (0, 0), which the chain stream
// updates itself.
template <int TM>
__device__ __forceinline__ SyrkTile syrk_tile(int t, int n, int base, int row_first, int col_first, const int* __restrict__ rb_list, int cnt,
                                              const int* __restrict__ side_list) {
  SyrkTile T;
  if (rb_list) {
    const int half = t & 1;
    const int rh = TM == NB ? 0 : (t >> 1) & 1;  // row half of the pair's block (64-row tiles: four tiles per pair)
    int pp = (t >> (TM == NB ? 1 : 2)) + (col_first == -2 ? 1 : 0);
    int b = 0, rem = cnt;
    while (pp >= rem) { pp -= rem; --rem; ++b; }
    const int a = b + pp;
    T.ri = __ldg(rb_list + a) * NB + rh * TM; T.ni = max(0, min(TM, n - T.ri));
    T.rj = __ldg(rb_list + b) * NB + half * TN; T.nj = max(0, min(TN, n - T.rj));
    T.oi = side_list ? __ldg(side_list + a) * NB + rh * TM : T.ri;
    T.oj = side_list ? __ldg(side_list + b) * NB + half * TN : T.rj;
    return T;
  }
  int bi, bj;
  if (col_first >= 0) {
    bi = row_first + (t >> 1); bj = col_first + (t & 1);
  } else {
    // with h = c0/2 - 1, row tile bi = q + h (q >= 1) owns column blocks c0 .. 2*bi+1 → 2q tiles; before row q: q(q-1)
    const int c0 = -col_first, h = c0 / 2 - 1;
    int q = (int)((1.0f + sqrtf(1.0f + 4.0f * (float)t)) * 0.5f);
    while (q * (q + 1) <= t) ++q;
    while (q * (q - 1) > t) --q;
    bi = q + h;
    bj = c0 + (t - q * (q - 1));
  }
  T.ri = base + bi * NB; T.ni = min(NB, n - T.ri);
  T.rj = base + bj * TN; T.nj = min(TN, n - T.rj);
  T.oi = T.ri; T.oj = T.rj;
  return T;
}

template <int TM>
__global__ void __launch_bounds__(2 * TM, TM == NB ? 2 : 3) chol_syrk_kernel(double* __restrict__ S, int n, int k0, int row_first, int col_first, int n_tiles,
                                                           int stagger, const int* __restrict__ fail, const int* __restrict__ rb_list, int rb_cnt,
                                                           double* __restrict__ Cside, int ldc_side, const int* __restrict__ side_list) {
  extern __shared__ __align__(16) unsigned char gsm_raw[];
  constexpr int THREADS = 2 * TM;
  GemmSmem<TM>& sm = *reinterpret_cast<GemmSmem<TM>*>(gsm_raw);
  int t = blockIdx.x;
  if (t >= n_tiles) return;
  const int failed = *fail;
  const int base = k0 + NB;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wr = warp >> 1, wc = warp & 1;  // (TM / 32) × 2 warps, 32 × 32 each
  const int fr = lane >> 2, fk = lane & 3;

  // All CTAs of a launch start together; left alone, the two CTAs of an SM (and the whole GPU) stay in
  // phase — everyone loads C at once, then everyone shares the DMMA pipe at half speed.  Delaying the second
  // half of the grid by half a tile time puts the co-resident CTAs in anti-phase (a stable state: one streams
  // its C tile while the other owns the tensor pipe), which also spreads the HBM traffic evenly in time.
  if (stagger && n_tiles > 2 * (int)gridDim.x && blockIdx.x >= gridDim.x / 2) {
    const long long t0 = clock64();
    while (clock64() - t0 < 16000) { }
  }
  SyrkTile cur = syrk_tile<TM>(t, n, base, row_first, col_first, rb_list, rb_cnt, side_list);
  load_tile_async<TM, THREADS>(sm.a[0], S, n, cur.ri, cur.ni, k0);
  load_tile_async<TN, THREADS>(sm.b[0], S, n, cur.rj, cur.nj, k0);
  cp_async_commit();
  if (failed) { cp_async_wait<0>(); return; }
  int stage = 0;
#pragma unroll 1
  for (; t < n_tiles; t += gridDim.x) {
    const int t_next = t + gridDim.x;
    const bool has_next = t_next < n_tiles;
    SyrkTile nxt = cur;
    if (has_next) nxt = syrk_tile<TM>(t_next, n, base, row_first, col_first, rb_list, rb_cnt, side_list);
    // C → accumulators.  Full tiles take an unpredicated path: one row pointer per 8-row group, the four
    // column tiles at immediate offsets, every load straight into its accumulator pair — nothing recycles a
    // register, so all 32 loads are in flight at once.
    double acc[4][4][2];
    // the output tile: in place in S, or (Cside) in a side matrix at the block positions side_list gives (CholPlan::Seg)
    double* const cbase = Cside ? Cside : S;
    const size_t ldc = Cside ? (size_t)ldc_side : (size_t)n;
    double* const crow = cbase + (size_t)(cur.oi + wr * 32 + fr) * ldc + cur.oj + wc * 32 + 2 * fk;
    const size_t rstep = 8 * ldc;
    const bool full = (cur.ni == TM) && (cur.nj == TN);
#ifdef HSFM_SYRK_NOLOAD
    if (full) {
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) { acc[a][b][0] = 0.0; acc[a][b][1] = 0.0; }
    } else
#endif
    if (full) {
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const double2* p = reinterpret_cast<const double2*>(crow + a * rstep);
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const double2 c = __ldcs(p + b * 4);
          acc[a][b][0] = c.x; acc[a][b][1] = c.y;
        }
      }
    } else {
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          acc[a][b][0] = 0.0; acc[a][b][1] = 0.0;
          const int row = wr * 32 + a * 8 + fr, col = wc * 32 + b * 8 + 2 * fk;
          if (row < cur.ni && col < cur.nj) {
            const double2 c = __ldcs(reinterpret_cast<const double2*>(crow + a * rstep) + b * 4);
            acc[a][b][0] = c.x; acc[a][b][1] = c.y;
          }
        }
    }
    if (has_next) {
      // next tile's C → L2: 128 rows × 512 B = 4 lines per row, one line per thread per pass
#pragma unroll
      for (int q = 0; q < 4 * TM / THREADS; ++q) {
        const int e = threadIdx.x + THREADS * q;
        const int row = e >> 2, seg = e & 3;
        if (row < nxt.ni && seg * 16 < nxt.nj)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(cbase + (size_t)(nxt.oi + row) * ldc + nxt.oj + seg * 16));
      }
    }
#pragma unroll 1
    for (int ch = 0; ch < NB / KC; ++ch) {
      if (ch + 1 < NB / KC) {
        load_tile_async<TM, THREADS>(sm.a[stage ^ 1], S, n, cur.ri, cur.ni, k0 + (ch + 1) * KC);
        load_tile_async<TN, THREADS>(sm.b[stage ^ 1], S, n, cur.rj, cur.nj, k0 + (ch + 1) * KC);
      } else if (has_next) {
        load_tile_async<TM, THREADS>(sm.a[stage ^ 1], S, n, nxt.ri, nxt.ni, k0);
        load_tile_async<TN, THREADS>(sm.b[stage ^ 1], S, n, nxt.rj, nxt.nj, k0);
      }
      cp_async_commit();
      cp_async_wait<1>();
      __syncthreads();
#pragma unroll
      for (int ks = 0; ks < KC / 4; ++ks) {
        double af[4], bf[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) af[a] = -sm.a[stage][wr * 32 + a * 8 + fr][ks * 4 + fk];
#pragma unroll
        for (int b = 0; b < 4; ++b) bf[b] = sm.b[stage][wc * 32 + b * 8 + fr][ks * 4 + fk];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
      }
      __syncthreads();
      stage ^= 1;
    }
    // epilogue: lane holds C[row = fr][col = 2*fk, 2*fk+1] of each 8×8 tile
#ifdef HSFM_SYRK_NOSTORE
    if (full) {
      if (acc[0][0][0] == 1.2345e300) crow[0] = 1.0;
    } else
#endif
    if (full) {
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        double2* p = reinterpret_cast<double2*>(crow + a * rstep);
#pragma unroll
        for (int b = 0; b < 4; ++b) p[b * 4] = make_double2(acc[a][b][0], acc[a][b][1]);
      }
    } else {
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int row = wr * 32 + a * 8 + fr;
        if (row >= cur.ni) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int col = wc * 32 + b * 8 + 2 * fk;
          if (col >= cur.nj) continue;
          reinterpret_cast<double2*>(crow + a * rstep)[b * 4] = make_double2(acc[a][b][0], acc[a][b][1]);
        }
      }
    }
    cur = nxt;
  }
  cp_async_wait<0>();
}

// Latency variant of the trailing update for the envelope factorisation's critical path: the update of the
// NEXT panel's column (128 columns of block rb_list[0]) against every active row block.  One 128x64 tile per
// CTA would put 8.4 us of DMMA issue on a single SM; here a CTA takes a 64x64 tile (four warps of 32x32), the
// whole K = 128 of both operands arrives in shared memory with ONE cp.async burst, and 4 x cnt CTAs spread over
// the GPU: ~2 us of tensor work per SM.
constexpr int LT_MAX = 64;         // largest tile edge (LT = 64: throughput; LT = 32: the chain's diagonal tile on 10 SMs)
constexpr int LT_LD = NB + 4;      // operand row stride in shared memory (doubles)
constexpr int kSyrkLatSmem = 2 * LT_MAX * LT_LD * (int)sizeof(double);
template <int LT>
__global__ void __launch_bounds__(128) chol_syrk_col_kernel(double* __restrict__ S, int n, int k0, const int* __restrict__ rb_list,
                                                            const int* __restrict__ fail, int a0) {
  pdl_enter();
  extern __shared__ __align__(16) double lsm[];
  double (*sA)[LT_LD] = reinterpret_cast<double (*)[LT_LD]>(lsm);
  double (*sB)[LT_LD] = reinterpret_cast<double (*)[LT_LD]>(lsm + LT * LT_LD);
  constexpr int PER = NB / LT, H = LT / 2, NT = H / 8;  // tiles per block edge, warp tile edge, 8x8 tiles per warp edge
  const int a = a0 + blockIdx.x / (PER * PER), rh = (blockIdx.x / PER) % PER, ch = blockIdx.x % PER;
  if (a == 0 && ch > rh) return;  // strictly upper part of the diagonal block
  const int ri = __ldg(rb_list + a) * NB + rh * LT, rj = __ldg(rb_list) * NB + ch * LT;
  const int ni = min(LT, n - ri), nj = min(LT, n - rj);
  if (ni <= 0 || nj <= 0 || *fail) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, fr = lane >> 2, fk = lane & 3;
  // LT == 32 (the chain's tile): operands by plain loads, issued together with the accumulator preload — one memory
  // round trip (ptxas puts a cp.async wait ahead of the other loads).  LT == 64: cp.async.
  constexpr int NLD = LT == 32 ? LT * (NB / 2) / 128 : 1;
  double2 opa[NLD], opb[NLD];
  if (LT == 32) {
#pragma unroll
    for (int it = 0; it < NLD; ++it) {
      const int e = tid + 128 * it, row = e >> 6, c2 = (e & 63) * 2;
      opa[it] = make_double2(0.0, 0.0); opb[it] = make_double2(0.0, 0.0);
      if (row < ni) opa[it] = ld_now2(S + (size_t)(ri + row) * n + k0 + c2);
      if (row < nj) opb[it] = ld_now2(S + (size_t)(rj + row) * n + k0 + c2);
    }
  } else {
    for (int e = tid; e < LT * (NB / 2); e += 128) {
      const int row = e >> 6, c2 = (e & 63) * 2;
      if (row < ni) cp_async16(&sA[row][c2], S + (size_t)(ri + row) * n + k0 + c2);
      else { sA[row][c2] = 0.0; sA[row][c2 + 1] = 0.0; }
      if (row < nj) cp_async16(&sB[row][c2], S + (size_t)(rj + row) * n + k0 + c2);
      else { sB[row][c2] = 0.0; sB[row][c2 + 1] = 0.0; }
    }
    cp_async_commit();
  }
  const int wr = warp >> 1, wc = warp & 1;  // 2 x 2 warps of H x H
  double acc[NT][NT][2];
  double* const crow = S + (size_t)(ri + wr * H + fr) * n + rj + wc * H + 2 * fk;
#pragma unroll
  for (int p = 0; p < NT; ++p)
#pragma unroll
    for (int q = 0; q < NT; ++q) {
      acc[p][q][0] = 0.0; acc[p][q][1] = 0.0;
      const int row = wr * H + p * 8 + fr, col = wc * H + q * 8 + 2 * fk;
      if (row < ni && col < nj) {
        const double2 c = ld_now2(crow + (size_t)p * 8 * n + q * 8);
        acc[p][q][0] = c.x; acc[p][q][1] = c.y;
      }
    }
  int nonfinite = 0;  // ties the accumulator preload to the barrier (see chol_trsm_strip_kernel)
#pragma unroll
  for (int p = 0; p < NT; ++p)
#pragma unroll
    for (int q = 0; q < NT; ++q) nonfinite |= is_nonfinite(acc[p][q][0]) | is_nonfinite(acc[p][q][1]);
  if (LT == 32) {
#pragma unroll
    for (int it = 0; it < NLD; ++it) {
      const int e = tid + 128 * it, row = e >> 6, c2 = (e & 63) * 2;
      *reinterpret_cast<double2*>(&sA[row][c2]) = opa[it];
      *reinterpret_cast<double2*>(&sB[row][c2]) = opb[it];
    }
  } else {
    cp_async_wait<0>();
  }
  if (__syncthreads_or(nonfinite)) return;
#pragma unroll 4
  for (int ks = 0; ks < NB / 4; ++ks) {
    double af[NT], bf[NT];
#pragma unroll
    for (int p = 0; p < NT; ++p) af[p] = -sA[wr * H + p * 8 + fr][ks * 4 + fk];
#pragma unroll
    for (int q = 0; q < NT; ++q) bf[q] = sB[wc * H + q * 8 + fr][ks * 4 + fk];
#pragma unroll
    for (int p = 0; p < NT; ++p)
#pragma unroll
      for (int q = 0; q < NT; ++q) dmma884(acc[p][q][0], acc[p][q][1], af[p], bf[q]);
  }
#pragma unroll
  for (int p = 0; p < NT; ++p)
#pragma unroll
    for (int q = 0; q < NT; ++q) {
      const int row = wr * H + p * 8 + fr, col = wc * H + q * 8 + 2 * fk;
      if (row < ni && col < nj) *reinterpret_cast<double2*>(crow + (size_t)p * 8 * n + q * 8) = make_double2(acc[p][q][0], acc[p][q][1]);
    }
}

// Panel solve  L21 = A21 L11^-T  as a product with the stored inverse, IN PLACE: every CTA owns a strip of
// rows (all 128 columns), stages the strip in shared memory and writes the result over it.  The B fragments
// (rows of L11^-1, 128 KB shared by every CTA, L2-resident) are read straight from global memory four k-steps
// ahead of their use, so the kernel needs only the strip in shared memory (<= 93 KB) and fits on an SM next
// to a resident trailing-update CTA.  Only k <= column contributes (L11^-1 is lower triangular); warp w takes
// the 8-column tiles w and 15-w, which together need the same 34 k-steps for every warp.
// ---------------------------------------------------------------------------------------------
// Chain kernel of the envelope factorisation: the next diagonal tile straight from the UNSOLVED panel block.
//   S(K+1,K+1) -= L21 L21^T   with   L21 = A21 Linv^T   is   S(K+1,K+1) -= ((A21 Linv^T) Linv) A21^T,
// and in that form every 8-row strip of the tile needs only its own rows of A21 (plus all of A21 and Linv as
// read-only operands): independent CTAs (two per strip, one per half of the block's columns), ONE kernel and one memory round trip between two diagonal-block kernels
// instead of panel solve -> hand-over -> tile update.  A21 is only read here; the panel solve that overwrites it with
// L21 runs afterwards on the main stream together with the other rows of the panel.
// Per CTA (strip r, 8 rows): Linv and A21 arrive by coalesced 16-byte loads (registers -> shared memory, the A21
// copy takes over Linv's buffer once Linv is no longer needed); Z = A_r Linv^T, T = Z Linv (both triangular: half
// the k-steps), D = T A21^T over the column tiles at or left of the strip's diagonal.  DMMA m8n8k4 throughout.
constexpr int CT_LD = NB + 4;
constexpr int kChainTileSmem = (NB + 3 * 8) * CT_LD * (int)sizeof(double);
__global__ void __launch_bounds__(256, 1) chol_chain_tile_kernel(double* __restrict__ S, int n, int k0, const double* __restrict__ Linv,
                                                                 const int* __restrict__ rb_list, const int* __restrict__ fail) {
  extern __shared__ __align__(16) double csm[];
  double (*sM)[CT_LD] = reinterpret_cast<double (*)[CT_LD]>(csm);                      // [128] Linv, later A21
  double (*sA)[CT_LD] = reinterpret_cast<double (*)[CT_LD]>(csm + NB * CT_LD);          // [8] strip of A21
  double (*sZ)[CT_LD] = reinterpret_cast<double (*)[CT_LD]>(csm + (NB + 8) * CT_LD);    // [8] Z
  double (*sT)[CT_LD] = reinterpret_cast<double (*)[CT_LD]>(csm + (NB + 16) * CT_LD);   // [8] T
  pdl_enter();
  const int failed = *fail;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, fr = lane >> 2, fk = lane & 3;
  const int blk = __ldg(rb_list);
  const int b0 = blk * NB, nb = min(NB, n - b0);  // first row and number of valid rows of block K+1
  // two CTAs per strip: each forms Z and T (cheap) and updates the column tiles of its half of the block, so it
  // needs only 64 rows of A21
  const int strip = blockIdx.x >> 1, half = blockIdx.x & 1, r0 = b0 + strip * 8;
  if (strip * 8 >= nb || (half && strip < 8)) return;
  // ---- every operand in one round trip ----
  double2 li[32], am[16], cpre;
#pragma unroll
  for (int it = 0; it < 32; ++it) {
    const int e = tid + 256 * it, row = e >> 6, c2 = (e & 63) * 2;
    li[it] = make_double2(0.0, 0.0);
    if (c2 <= row) li[it] = ld_now2(Linv + (size_t)row * NB + c2);  // lower triangular: the rest is zero
  }
#pragma unroll
  for (int it = 0; it < 16; ++it) {
    const int e = tid + 256 * it, row = half * 64 + (e >> 6), c2 = (e & 63) * 2;
    am[it] = make_double2(0.0, 0.0);
    if (row < nb && row < 8 * (strip + 1)) am[it] = ld_now2(S + (size_t)(b0 + row) * n + k0 + c2);  // rows of the column tiles this CTA updates
  }
  double2 st2[2];
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int e = tid + 256 * it, row = e >> 6, c2 = (e & 63) * 2;
    st2[it] = make_double2(0.0, 0.0);
    if (strip * 8 + row < nb) st2[it] = ld_now2(S + (size_t)(r0 + row) * n + k0 + c2);
  }
  const int ct_lo = warp, ct_hi = 15 - warp;
  const bool row_ok = strip * 8 + fr < nb;
  double* const crow = S + (size_t)(r0 + fr) * n + b0 + 2 * fk;
  const int ct3 = half * 8 + warp;  // the column tile this warp updates in the last product
  const bool do3 = ct3 <= strip;
  cpre = make_double2(0.0, 0.0);
  if (row_ok && do3 && b0 + ct3 * 8 + 2 * fk + 1 < n) cpre = ld_now2(crow + ct3 * 8);
  int nonfinite = 0;
#pragma unroll
  for (int it = 0; it < 32; ++it) nonfinite |= is_nonfinite(li[it].x) | is_nonfinite(li[it].y);
#pragma unroll
  for (int it = 0; it < 16; ++it) nonfinite |= is_nonfinite(am[it].x) | is_nonfinite(am[it].y);
  nonfinite |= is_nonfinite(cpre.x) | is_nonfinite(cpre.y);
#pragma unroll
  for (int it = 0; it < 32; ++it) {
    const int e = tid + 256 * it, row = e >> 6, c2 = (e & 63) * 2;
    if (c2 < 8 * ((row >> 3) + 1)) *reinterpret_cast<double2*>(&sM[row][c2]) = li[it];  // nothing right of the diagonal 8x8 tile is ever read
  }
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int e = tid + 256 * it, row = e >> 6, c2 = (e & 63) * 2;
    *reinterpret_cast<double2*>(&sA[row][c2]) = st2[it];
  }
  nonfinite = __syncthreads_or(nonfinite);
  if (failed || nonfinite) return;
  // ---- Z = A_r Linv^T : Z[r][c] = sum_{k <= c} A[r][k] Linv[c][k] ----
  double z[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
  for (int u = 0; u < 32; ++u) {
    if (u < 2 * (ct_hi + 1)) {
      const double a = sA[fr][u * 4 + fk];
      dmma884(z[1][0], z[1][1], a, sM[ct_hi * 8 + fr][u * 4 + fk]);
      if (u < 2 * (ct_lo + 1)) dmma884(z[0][0], z[0][1], a, sM[ct_lo * 8 + fr][u * 4 + fk]);
    }
  }
  *reinterpret_cast<double2*>(&sZ[fr][ct_lo * 8 + 2 * fk]) = make_double2(z[0][0], z[0][1]);
  *reinterpret_cast<double2*>(&sZ[fr][ct_hi * 8 + 2 * fk]) = make_double2(z[1][0], z[1][1]);
  __syncthreads();
  // ---- T = Z Linv : T[r][c] = sum_{k >= c} Z[r][k] Linv[k][c] ----
  double t[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
  for (int u = 0; u < 32; ++u) {
    if (u >= 2 * ct_lo) {
      const double a = sZ[fr][u * 4 + fk];
      dmma884(t[0][0], t[0][1], a, sM[u * 4 + fk][ct_lo * 8 + fr]);
      if (u >= 2 * ct_hi) dmma884(t[1][0], t[1][1], a, sM[u * 4 + fk][ct_hi * 8 + fr]);
    }
  }
  __syncthreads();  // everybody is done with Linv in sM
  *reinterpret_cast<double2*>(&sT[fr][ct_lo * 8 + 2 * fk]) = make_double2(t[0][0], t[0][1]);
  *reinterpret_cast<double2*>(&sT[fr][ct_hi * 8 + 2 * fk]) = make_double2(t[1][0], t[1][1]);
#pragma unroll
  for (int it = 0; it < 16; ++it) {
    const int e = tid + 256 * it, row = half * 64 + (e >> 6), c2 = (e & 63) * 2;
    *reinterpret_cast<double2*>(&sM[row][c2]) = am[it];
  }
  __syncthreads();
  // ---- tile rows of the strip: C[r][c] -= sum_k T[r][k] A21[c][k], column tiles at or left of the diagonal ----
  if (!do3) return;
  double c0 = cpre.x, c1 = cpre.y;
#pragma unroll 8
  for (int u = 0; u < 32; ++u) dmma884(c0, c1, -sT[fr][u * 4 + fk], sM[ct3 * 8 + fr][u * 4 + fk]);
  if (row_ok && b0 + ct3 * 8 + 2 * fk + 1 < n) *reinterpret_cast<double2*>(crow + ct3 * 8) = make_double2(c0, c1);
}


constexpr int TS_LD = NB + 4;
constexpr int TS_MAXROWS = 88;
__global__ void __launch_bounds__(256) chol_trsm_strip_kernel(double* __restrict__ S, int n, int k0, const double* __restrict__ Linv,
                                                               int rows_per, const int* __restrict__ fail, const int* __restrict__ rb_list,
                                                               int strips_per_block) {
  extern __shared__ __align__(16) double tsm[];
  double (*sA)[TS_LD] = reinterpret_cast<double (*)[TS_LD]>(tsm);
  pdl_enter();
  const int failed = *fail;
  int r0, nr;
  if (rb_list) {  // envelope mode: strips of the active 128-row blocks only
    const int blk = __ldg(rb_list + blockIdx.x / strips_per_block);
    r0 = blk * NB + (blockIdx.x % strips_per_block) * rows_per;
    nr = min(rows_per, min(n, (blk + 1) * NB) - r0);
  } else {
    r0 = k0 + NB + blockIdx.x * rows_per;
    nr = min(rows_per, n - r0);
  }
  const int tid = threadIdx.x;
  // Latency path (<= 16 rows): the strip comes by plain loads into registers, issued together with the B fragments
  // below — ONE memory round trip.  (With cp.async ptxas puts the wait for the copies ahead of the fragment loads.)
  double2 strip[4];
  if (rows_per <= 16) {
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int e = tid + 256 * it, row = e >> 6, c2 = (e & 63) * 2;
      strip[it] = make_double2(0.0, 0.0);
      if (e < rows_per * (NB / 2) && row < nr) strip[it] = ld_now2(S + (size_t)(r0 + row) * n + k0 + c2);
    }
  } else {
    for (int e = tid; e < rows_per * (NB / 2); e += 256) {
      const int row = e >> 6, c2 = (e & 63) * 2;
      if (row < nr) cp_async16(&sA[row][c2], S + (size_t)(r0 + row) * n + k0 + c2);
      else { sA[row][c2] = 0.0; sA[row][c2 + 1] = 0.0; }
    }
    cp_async_commit();
  }
  const int warp = tid >> 5, lane = tid & 31, fr = lane >> 2, fk = lane & 3;
  const int RT = rows_per >> 3;
  const int ct_lo = warp, ct_hi = 15 - warp;
  const int ks_hi = 2 * (ct_hi + 1), ks_lo = 2 * (ct_lo + 1);  // k-steps (of 4) each column tile needs
  const double* __restrict__ Bhi = Linv + (ct_hi * 8 + fr) * NB + fk;
  const double* __restrict__ Blo = Linv + (ct_lo * 8 + fr) * NB + fk;
  double acc[TS_MAXROWS / 8][2][2];
#pragma unroll
  for (int rt = 0; rt < TS_MAXROWS / 8; ++rt) { acc[rt][0][0] = acc[rt][0][1] = acc[rt][1][0] = acc[rt][1][1] = 0.0; }
  if (rows_per <= 16) {
    // Latency path (envelope factorisation: few rows per CTA, the kernel sits on the critical path): all 34
    // B fragments of the warp's two column tiles are fetched in ONE round trip, before the strip lands.
    double bh[32], bl[16];
#pragma unroll
    for (int u = 0; u < 32; ++u) bh[u] = (u < ks_hi) ? ld_now(Bhi + 4 * u) : 0.0;
#pragma unroll
    for (int u = 0; u < 16; ++u) bl[u] = (u < ks_lo) ? ld_now(Blo + 4 * u) : 0.0;
    // The barrier's predicate depends on every fragment (a NaN/Inf check that is worth having anyway), so ptxas cannot
    // sink these loads below it into the multiply loop — which it otherwise does, and which costs a second memory
    // round trip (2.3 us per panel of the chain, measured with scripts/chain_timing.py).
    int nonfinite = 0;
#pragma unroll
    for (int u = 0; u < 32; ++u) nonfinite |= is_nonfinite(bh[u]);
#pragma unroll
    for (int u = 0; u < 16; ++u) nonfinite |= is_nonfinite(bl[u]);
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int e = tid + 256 * it, row = e >> 6, c2 = (e & 63) * 2;
      if (e < rows_per * (NB / 2)) *reinterpret_cast<double2*>(&sA[row][c2]) = strip[it];
    }
    nonfinite = __syncthreads_or(nonfinite);
    if (failed || nr <= 0 || nonfinite) return;
#pragma unroll
    for (int u = 0; u < 32; ++u) {
      if (u < ks_hi) {
#pragma unroll
        for (int rt = 0; rt < 2; ++rt) {
          if (rt < RT) {
            const double a = sA[rt * 8 + fr][u * 4 + fk];
            dmma884(acc[rt][1][0], acc[rt][1][1], a, bh[u]);
            if (u < 16 && u < ks_lo) dmma884(acc[rt][0][0], acc[rt][0][1], a, bl[u < 16 ? u : 0]);
          }
        }
      }
    }
  } else {
    double bh[4], bl[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      bh[u] = (u < ks_hi) ? __ldg(Bhi + 4 * u) : 0.0;
      bl[u] = (u < ks_lo) ? __ldg(Blo + 4 * u) : 0.0;
    }
    cp_async_wait<0>();
    __syncthreads();
    if (failed || nr <= 0) return;
#pragma unroll 1
    for (int ks0 = 0; ks0 < ks_hi; ks0 += 4) {
      double nh[4], nl[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {  // next group of B fragments, in flight while this group is multiplied
        const int ks = ks0 + 4 + u;
        nh[u] = (ks < ks_hi) ? __ldg(Bhi + 4 * ks) : 0.0;
        nl[u] = (ks < ks_lo) ? __ldg(Blo + 4 * ks) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int ks = ks0 + u;
        if (ks < ks_hi) {
          const bool lo = ks < ks_lo;
#pragma unroll
          for (int rt = 0; rt < TS_MAXROWS / 8; ++rt) {
            if (rt < RT) {
              const double a = sA[rt * 8 + fr][ks * 4 + fk];
              dmma884(acc[rt][1][0], acc[rt][1][1], a, bh[u]);
              if (lo) dmma884(acc[rt][0][0], acc[rt][0][1], a, bl[u]);
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) { bh[u] = nh[u]; bl[u] = nl[u]; }
    }
  }
#pragma unroll
  for (int rt = 0; rt < TS_MAXROWS / 8; ++rt) {
    const int row = rt * 8 + fr;
    if (rt < RT && row < nr) {
      double* dst = S + (size_t)(r0 + row) * n + k0 + 2 * fk;
      *reinterpret_cast<double2*>(dst + ct_lo * 8) = make_double2(acc[rt][0][0], acc[rt][0][1]);
      *reinterpret_cast<double2*>(dst + ct_hi * 8) = make_double2(acc[rt][1][0], acc[rt][1][1]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Triangular solves with the stored L11^-1 blocks: per panel one launch; every CTA forms
// x_k = L11^-1 b_k (resp. L11^-T y_k) redundantly — a 128×128 mat-vec — CTA 0 publishes it and
// CTA r >= 1 folds it into 128 rows below (resp. 128 columns to the left).  Loads are issued four
// rows (eight for the column-oriented backward pass) at a time so their latencies overlap.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void warp_rows_dot4(const double* __restrict__ M, size_t ld, int row, int nrows, const double* __restrict__ sv,
                                               int lane, double out[4]) {
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  double v[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q) v[r][q] = (r < nrows) ? M[(size_t)(row + r) * ld + lane + 32 * q] : 0.0;
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[r] += v[r][q] * sv[lane + 32 * q];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], d);
    out[r] = acc[r];
  }
}

__global__ void __launch_bounds__(256) trsv_fwd_kernel(const double* __restrict__ L, int n, int k0, int kb, const double* __restrict__ Linv,
                                                      double* __restrict__ b, double* __restrict__ x) {
  __shared__ double sb[NB];
  __shared__ double sx[NB];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < NB) sb[tid] = (tid < kb) ? b[k0 + tid] : 0.0;
  __syncthreads();
#pragma unroll 1
  for (int g = 0; g < 4; ++g) {
    const int i = warp * 16 + g * 4;
    double o[4];
    warp_rows_dot4(Linv, NB, i, 4, sb, lane, o);
    if (lane < 4) sx[i + lane] = (lane == 0) ? o[0] : (lane == 1) ? o[1] : (lane == 2) ? o[2] : o[3];
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    if (tid < kb) x[k0 + tid] = sx[tid];
    return;
  }
  const int r0 = k0 + NB + (blockIdx.x - 1) * NB;
#pragma unroll 1
  for (int g = 0; g < 4; ++g) {
    const int row = r0 + warp * 16 + g * 4;
    const int nrows = min(4, n - row);
    if (nrows <= 0) break;
    double o[4];
    warp_rows_dot4(L + k0, (size_t)n, row, nrows, sx, lane, o);
    if (lane < nrows) b[row + lane] -= (lane == 0) ? o[0] : (lane == 1) ? o[1] : (lane == 2) ? o[2] : o[3];
  }
}

__global__ void __launch_bounds__(256) trsv_bwd_kernel(const double* __restrict__ L, int n, int k0, int kb, const double* __restrict__ Linv,
                                                      double* __restrict__ y, double* __restrict__ x) {
  __shared__ double sy[NB];
  __shared__ double sx[NB];
  __shared__ double sp[NB];
  const int tid = threadIdx.x, j = tid & (NB - 1), half = tid >> 7;
  if (tid < NB) sy[tid] = (tid < kb) ? y[k0 + tid] : 0.0;
  __syncthreads();
  {
    // x[j] = sum_{i >= j} Linv[i][j] y[i]; rows i split between the two halves, 8 loads in flight
    double acc = 0.0;
    const int i0 = half * 64;
#pragma unroll 1
    for (int ib = i0; ib < i0 + 64; ib += 8) {
      double v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = (ib + q >= j) ? Linv[(size_t)(ib + q) * NB + j] : 0.0;
#pragma unroll
      for (int q = 0; q < 8; ++q) acc += v[q] * sy[ib + q];
    }
    if (half) sp[j] = acc;
    __syncthreads();
    if (!half) sx[j] = acc + sp[j];
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    if (tid < kb) x[k0 + tid] = sx[tid];
    return;
  }
  const int col = (blockIdx.x - 1) * NB + j;
  double acc = 0.0;
  if (col < k0) {
    const int i0 = half * 64;
#pragma unroll 1
    for (int ib = i0; ib < i0 + 64; ib += 8) {
      double v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = (ib + q < kb) ? L[(size_t)(k0 + ib + q) * n + col] : 0.0;
#pragma unroll
      for (int q = 0; q < 8; ++q) acc += v[q] * sx[ib + q];
    }
  }
  if (half) sp[j] = acc;
  __syncthreads();
  if (!half && col < k0) y[col] -= acc + sp[j];
}

// ---------------------------------------------------------------------------------------------
// Both triangular solves in ONE cooperative launch.  CTA c owns the 128-row blocks c, c+G, … .  A block's result is
// handed to its consumers through the data itself: vectors and contribution slots are preset to a sentinel bit pattern
// and every consumer thread polls its own element (poll_value / publish_value below) — no flags.  Blocks only ever wait
// on blocks with a smaller (forward) / larger (backward) index and every CTA walks its blocks in that order, so the
// waits cannot cycle; the cooperative launch guarantees all CTAs are resident.  The dependency chain is two 128x128
// mat-vecs plus one hand-over per block instead of two kernel launches.
// ---------------------------------------------------------------------------------------------

// Block mat-vec pieces of the triangular solves, split into "fetch the matrix block into registers" and "multiply
// with the vector", so that the fetch is issued BEFORE the CTA waits for the flag of the vector's producer: when
// the flag arrives only the vector load and a short multiply-add tree are left on the critical path.
// Row form: thread t owns row t>>1, columns (t&1)*64 .. +63 (32 x 16-byte loads).  Rows >= nrows read as zero.
struct RowFrag { double2 v[32]; };
// 32-byte loads (whole sectors) that do not allocate in L1: a CTA has 8192 of them in flight for one block, and with the
// L11^-1 block in shared memory the L1 that is left cannot track that many lines (the sweep time follows the L1 size)
__device__ __forceinline__ void ld_sector(const double* p, double2& a, double2& b) {
  asm volatile("ld.global.L1::no_allocate.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a.x), "=d"(a.y), "=d"(b.x), "=d"(b.y) : "l"(p));
}
__device__ __forceinline__ void row_fetch(RowFrag& f, const double* __restrict__ M, size_t ld, int nrows, int tid) {
  const int r = tid >> 1, h = tid & 1;
  const double* p = M + (size_t)r * ld + h * 64;
  if (r < nrows && !(ld & 3)) {  // rows 32-byte aligned (n a multiple of 4)
#pragma unroll
    for (int q = 0; q < 16; ++q) ld_sector(p + 4 * q, f.v[2 * q], f.v[2 * q + 1]);
  } else if (r < nrows) {
#pragma unroll
    for (int q = 0; q < 32; ++q) f.v[q] = __ldg(reinterpret_cast<const double2*>(p) + q);
  } else {
#pragma unroll
    for (int q = 0; q < 32; ++q) f.v[q] = make_double2(0.0, 0.0);
  }
}
__device__ __forceinline__ double row_dot(const RowFrag& f, const double* __restrict__ sv, int tid) {  // valid on even threads
  const int h = tid & 1;
  double a[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int q = 0; q < 32; ++q) {
    a[(2 * q) & 7] = fma(f.v[q].x, sv[h * 64 + 2 * q], a[(2 * q) & 7]);
    a[(2 * q + 1) & 7] = fma(f.v[q].y, sv[h * 64 + 2 * q + 1], a[(2 * q + 1) & 7]);
  }
  double s = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  return s;
}
// Column form: thread t owns the column pair 2 (t & 63), 2 (t & 63) + 1 and the rows (t >> 6) * 32 .. +31: 32 loads of
// 16 bytes, a warp's load one contiguous 512-byte piece of a row (whole sectors, nothing for L1 to keep).
struct ColFrag { double2 v[32]; };
__device__ __forceinline__ void col_fetch(ColFrag& f, const double* __restrict__ M, size_t ld, int nrows, int ncols, int tid) {
  const int c = 2 * (tid & 63), r0 = (tid >> 6) * 32;
  const double* p = M + (size_t)r0 * ld + c;
#pragma unroll
  for (int q = 0; q < 32; ++q) {
    double2 v = make_double2(0.0, 0.0);
    if (r0 + q < nrows && c < ncols) {  // ncols is even (6 per camera): a pair is in or out as a whole
      asm volatile("ld.global.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p + (size_t)q * ld));
    }
    f.v[q] = v;
  }
}
__device__ __forceinline__ double2 col_dot(const ColFrag& f, const double* __restrict__ sv, int tid) {  // partial sums of this quarter
  const int r0 = (tid >> 6) * 32;
  double ax[4] = {0.0, 0.0, 0.0, 0.0}, ay[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int q = 0; q < 32; ++q) {
    const double s = sv[r0 + q];
    ax[q & 3] = fma(f.v[q].x, s, ax[q & 3]);
    ay[q & 3] = fma(f.v[q].y, s, ay[q & 3]);
  }
  return make_double2((ax[0] + ax[1]) + (ax[2] + ax[3]), (ay[0] + ay[1]) + (ay[2] + ay[3]));
}

// L11^-1 block in shared memory (row stride 130 doubles: the 16 rows a warp reads together fall in distinct
// 16-byte bank groups), fetched with cp.async while the off-diagonal blocks are folded in.
constexpr int TI_LD = NB + 2;
constexpr int kTrsvSmem = NB * TI_LD * (int)sizeof(double);
__device__ __forceinline__ void inv_stage(double* sInv, const double* __restrict__ Linv, int tid) {
  for (int e = tid; e < NB * (NB / 2); e += 256) {
    const int row = e >> 6, c2 = (e & 63) * 2;
    cp_async16(sInv + row * TI_LD + c2, Linv + row * NB + c2);
  }
  cp_async_commit();
}

// Structure (optional, all null = dense): row_off/row_list = the column blocks of every row block (ascending, the
// diagonal block last), col_off/col_cnt/col_list = the row blocks below the diagonal of every column block.
// Hand-over of a solved 128-vector between CTAs WITHOUT a flag: the vectors w (forward) and x (backward) are preset to a
// sentinel bit pattern (all ones: a NaN no arithmetic produces) and every consumer thread polls ITS element until it
// differs — one L2 round trip instead of flag acquire + vector load, and no release fence on the producer side (an
// aligned 8-byte store is atomic; the value is its own flag).
__device__ __forceinline__ double poll_value(const double* p) {
  unsigned long long v;
  do {
    asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  } while (v == 0xffffffffffffffffull);
  return __longlong_as_double((long long)v);
}
__device__ __forceinline__ void publish_value(double* p, double v) {
  // a genuine all-ones result cannot occur (not a canonical NaN); canonical NaNs pass through and end the consumers' waits
  asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(__double_as_longlong(v)) : "memory");
}
// Forward sweep, FAN-OUT: the CTA that owns block column k waits for the contributions to its rows (one 128-vector per
// structural block of block row k, each in a SLOT of its own: slot = position of the block in row_list, or
// i (i + 1) / 2 + k dense), forms y_k = L11^-1 (b_k - sum of the slots, in list order: the result does not depend on
// arrival order), and then multiplies y_k into the blocks below its diagonal, the first of them — the one the next
// column waits for — from registers it filled BEFORE it started waiting.  Every CTA touches the blocks of its own
// columns only (about as many per column everywhere), where a row-wise gather leaves the loop-closure rows with one
// block per panel of the whole matrix.
__global__ void __launch_bounds__(256) trsv_coop_kernel(const double* __restrict__ L, int n, const double* __restrict__ LinvAll, double* b,
                                                        double* w, double* x, double* slots, const int* __restrict__ row_off,
                                                        const int* __restrict__ col_off, const int* __restrict__ col_cnt,
                                                        const int* __restrict__ col_list, const int* __restrict__ col_slot,
                                                        const int* __restrict__ sweep) {
  extern __shared__ __align__(16) double sInv[];  // [128][TI_LD]
  __shared__ double sv[NB];    // vector of the block being folded in (y_k / x_i)
  __shared__ double sacc[NB];  // running right-hand side of the owned block
  __shared__ __align__(16) double sp[4][NB];  // partial sums of the four row quarters (backward), of the two halves (L11^-T)
  const int tid = threadIdx.x;
  const int P = (n + NB - 1) / NB, G = gridDim.x;
  // Columns are dealt to the CTAs in SWEEP order (sweep[]: by height in the block elimination tree; null = by index):
  // with more columns than CTAs a CTA owns columns of several concurrent chains, and in index order the early columns
  // of a later chain would wait behind the late columns of an earlier one.  Dependencies point to strictly lower (forward)
  // / higher (backward) positions of the sweep order, every CTA walks its columns in that order: no wait can cycle.
  // ---- forward: L y = b, y -> w ----
  for (int idx = blockIdx.x; idx < P; idx += G) {
    const int k = sweep ? __ldg(sweep + idx) : idx;
    const int r0 = k * NB, nr = min(NB, n - r0);
    __syncthreads();
    inv_stage(sInv, LinvAll + (size_t)k * NB * NB, tid);
    const int c_beg = col_off ? __ldg(col_off + k) : 0, c_cnt = col_off ? __ldg(col_cnt + k) : P - 1 - k;
    RowFrag f;
    int i_next = c_cnt > 0 ? (col_off ? __ldg(col_list + c_beg) : k + 1) : -1;
    if (i_next >= 0) row_fetch(f, L + (size_t)i_next * NB * n + (size_t)k * NB, (size_t)n, min(NB, n - i_next * NB), tid);  // before the wait
    if (tid < NB) {
      double a = 0.0;
      if (tid < nr) {
        a = b[r0 + tid];
        const int q_beg = row_off ? __ldg(row_off + k) : k * (k + 1) / 2;
        const int q_end = row_off ? __ldg(row_off + k + 1) - 1 : q_beg + k;  // the diagonal entry (last) has no slot content
        const double* sl = slots + (size_t)q_beg * NB + tid;
        int q = q_beg;
        for (; q + 4 <= q_end; q += 4, sl += 4 * NB) {  // four loads in flight; the sum keeps list order
          unsigned long long v0, v1, v2, v3;
          asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v0) : "l"(sl) : "memory");
          asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v1) : "l"(sl + NB) : "memory");
          asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v2) : "l"(sl + 2 * NB) : "memory");
          asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v3) : "l"(sl + 3 * NB) : "memory");
          a -= (v0 == 0xffffffffffffffffull) ? poll_value(sl) : __longlong_as_double((long long)v0);
          a -= (v1 == 0xffffffffffffffffull) ? poll_value(sl + NB) : __longlong_as_double((long long)v1);
          a -= (v2 == 0xffffffffffffffffull) ? poll_value(sl + 2 * NB) : __longlong_as_double((long long)v2);
          a -= (v3 == 0xffffffffffffffffull) ? poll_value(sl + 3 * NB) : __longlong_as_double((long long)v3);
        }
        for (; q < q_end; ++q, sl += NB) a -= poll_value(sl);
      }
      sacc[tid] = a;
    }
    cp_async_wait<0>();
    __syncthreads();
    {
      // y = L11^-1 acc from shared memory: thread t owns row t>>1, half t&1
      const int r = tid >> 1, h = tid & 1;
      const double* row = sInv + r * TI_LD + h * 64;
      double a[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int q = 0; q < 32; ++q) {
        const double2 m = *reinterpret_cast<const double2*>(row + 2 * q);
        a[(2 * q) & 3] = fma(m.x, sacc[h * 64 + 2 * q], a[(2 * q) & 3]);
        a[(2 * q + 1) & 3] = fma(m.y, sacc[h * 64 + 2 * q + 1], a[(2 * q + 1) & 3]);
      }
      double y = (a[0] + a[1]) + (a[2] + a[3]);
      y += __shfl_xor_sync(0xffffffffu, y, 1);
      if (!h) {
        sv[r] = (r < nr) ? y : 0.0;
        if (r < nr) publish_value(w + r0 + r, y);  // kept for the backward sweep
      }
    }
    __syncthreads();
    for (int q = 0; q < c_cnt; ++q) {
      const int i = i_next;
      const double dd = row_dot(f, sv, tid);
      if (q + 1 < c_cnt) {  // the next block is on its way while this one goes out
        i_next = col_off ? __ldg(col_list + c_beg + q + 1) : k + 2 + q;
        row_fetch(f, L + (size_t)i_next * NB * n + (size_t)k * NB, (size_t)n, min(NB, n - i_next * NB), tid);
      }
      const int slot = col_slot ? __ldg(col_slot + c_beg + q) : i * (i + 1) / 2 + k;
      if (!(tid & 1) && (tid >> 1) < min(NB, n - i * NB)) publish_value(slots + (size_t)slot * NB + (tid >> 1), dd);
    }
  }
  // ---- backward: L^T x = y, x published in its own vector ----
  // (Measured and dropped: the block whose x arrives last kept in shared memory beside a packed lower triangle of
  // L11^-1 — 200 KB of shared memory leave ~28 KB of L1, and the register fetches of the other blocks, 16 / 8 bytes per
  // thread along rows, then miss L1 on every second half of a sector: the sweep took 1.2 ms instead of 0.45 at c4.)
  const int j = tid & (NB - 1), half = tid >> 7;
  for (int idx = P - 1 - (int)blockIdx.x; idx >= 0; idx -= G) {
    const int kk = sweep ? __ldg(sweep + idx) : idx;
    const int c0 = kk * NB, kb = min(NB, n - c0);
    __syncthreads();
    inv_stage(sInv, LinvAll + (size_t)kk * NB * NB, tid);
    if (tid < NB) sacc[tid] = (tid < kb) ? poll_value(w + c0 + tid) : 0.0;  // this block's forward result (another CTA's)
    const int c_beg = col_off ? __ldg(col_off + kk) : 0, c_cnt = col_off ? __ldg(col_cnt + kk) : P - 1 - kk;
    for (int q = c_cnt - 1; q >= 0; --q) {  // structural blocks of column kk, bottom up
      const int i = col_off ? __ldg(col_list + c_beg + q) : kk + 1 + q;
      const int r0 = i * NB, nr = min(NB, n - r0);
      ColFrag f;
      col_fetch(f, L + (size_t)r0 * n + c0, (size_t)n, nr, kb, tid);
      __syncthreads();
      if (tid < NB) sv[tid] = (tid < nr) ? poll_value(x + r0 + tid) : 0.0;
      __syncthreads();
      const double2 acc = col_dot(f, sv, tid);
      *reinterpret_cast<double2*>(&sp[tid >> 6][2 * (tid & 63)]) = acc;
      __syncthreads();
      if (tid < NB) sacc[tid] -= (sp[0][tid] + sp[1][tid]) + (sp[2][tid] + sp[3][tid]);
    }
    cp_async_wait<0>();
    __syncthreads();
    {
      // x_k = L11^-T acc : x[j] = sum_{i >= j} Linv[i][j] acc[i]  (column j of the staged block, this half's rows)
      const int r0h = half * 64;
      double a[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
      for (int q = 0; q < 64; ++q) {
        const int r = r0h + q;
        const double m = (r >= j) ? sInv[r * TI_LD + j] : 0.0;
        a[q & 3] = fma(m, sacc[r], a[q & 3]);
      }
      const double acc = (a[0] + a[1]) + (a[2] + a[3]);
      if (half) sp[0][j] = acc;
      __syncthreads();
      if (!half && j < kb) publish_value(x + c0 + j, acc + sp[0][j]);
    }
  }
}

// synthetic SPD test matrix A_ij = 1/(1+|i-j|) + 2 delta_ij (Polya: convex decreasing kernel), b = A 1
__global__ void spd_fill_kernel(double* __restrict__ A, int n) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j < n) A[(size_t)i * n + j] = 1.0 / (1.0 + fabs((double)(i - j))) + (i == j ? 2.0 : 0.0);
}
__global__ void spd_rowsum_kernel(int n, double* __restrict__ b) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 2.0;
  for (int j = 0; j < n; ++j) s += 1.0 / (1.0 + fabs((double)(i - j)));
  b[i] = s;
}
__global__ void spd_err_kernel(int n, const double* __restrict__ x, double* __restrict__ err) {
  __shared__ double sm[256];
  double m = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) m = fmax(m, fabs(x[i] - 1.0));
  sm[threadIdx.x] = m;
  __syncthreads();
  for (int d = 128; d > 0; d >>= 1) {
    if ((int)threadIdx.x < d) sm[threadIdx.x] = fmax(sm[threadIdx.x], sm[threadIdx.x + d]);
    __syncthreads();
  }
  if (threadIdx.x == 0) *err = sm[0];
}

}  // namespace

void launch_spd_fill(double* A, int n, double* b, cudaStream_t s) {
  dim3 g((n + 255) / 256, n);
  spd_fill_kernel<<<g, 256, 0, s>>>(A, n);
  if (b) spd_rowsum_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, b);
}
void launch_spd_err(int n, const double* x, double* err, cudaStream_t s) {
  spd_err_kernel<<<1, 256, 0, s>>>(n, x, err);
#ifdef HSFM_DIAG_TIMING
  long long h[32];
  cudaStreamSynchronize(s);
  cudaMemcpyFromSymbol(h, g_diag_clk, sizeof(h));
  const char* nm[7] = {"", "load", "factor", "storeL", "trinv", "M", "levels"};
  for (int k = 1; k <= 6; ++k) std::fprintf(stderr, "%s %lld  ", nm[k], h[k] - h[k - 1]);
  std::fprintf(stderr, "total %lld clk\n", h[6] - h[0]);
#endif
}

#ifdef HSFM_CHAIN_TIMING
extern "C" int hsfm_ba_debug_chain_ts(unsigned long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, g_chain_ts, sizeof(unsigned long long) * (size_t)n);
}
#endif
size_t cholesky_workspace_doubles(int n) {
  const size_t panels = (size_t)((n + NB - 1) / NB);
  // one L11^-1 block per panel + 2 x panels ints (even: what follows stays 16-byte aligned) + the contribution slots of the
  // forward triangular solve (one 128-vector per block of the lower block triangle: the dense case)
  return ((panels * NB * NB + panels + 1) & ~(size_t)1) + panels * (panels + 1) / 2 * NB;
}

int CholAux::init() {
  if (stream) return 0;
  // per device (function attributes and the SM count belong to the device of the owning context)
  int dev = 0;
  HSFM_CUDA(cudaGetDevice(&dev));
  cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
  if (n_sm <= 0) n_sm = 148;
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  if (getenv("HSFM_TRSV_LAUNCHES")) coop = 0;
  HSFM_CUDA(cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDiagSmem));
  HSFM_CUDA(cudaFuncSetAttribute(chol_trsm_strip_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_MAXROWS * TS_LD * (int)sizeof(double)));
  HSFM_CUDA(cudaFuncSetAttribute(chol_syrk_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
  HSFM_CUDA(cudaFuncSetAttribute(chol_syrk_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemHalf));
  HSFM_CUDA(cudaFuncSetAttribute(chol_syrk_col_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrkLatSmem));
  HSFM_CUDA(cudaFuncSetAttribute(chol_syrk_col_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSyrkLatSmem / 2));
  HSFM_CUDA(cudaFuncSetAttribute(chol_chain_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kChainTileSmem));
  HSFM_CUDA(cudaFuncSetAttribute(trsv_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTrsvSmem));
  int lo = 0, hi = 0;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  HSFM_CUDA(cudaStreamCreateWithPriority(&stream, cudaStreamNonBlocking, hi));
  HSFM_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
  HSFM_CUDA(cudaEventCreateWithFlags(&ev_panel, cudaEventDisableTiming));
  HSFM_CUDA(cudaEventCreateWithFlags(&ev_rest, cudaEventDisableTiming));
  HSFM_CUDA(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
  HSFM_CUDA(cudaEventCreateWithFlags(&ev_next, cudaEventDisableTiming));
  HSFM_CUDA(cudaEventCreateWithFlags(&ev_zero, cudaEventDisableTiming));
  sets.resize(CHOL_STREAM_SETS);
  for (CholChainSet& c : sets) {
    HSFM_CUDA(cudaStreamCreateWithPriority(&c.chain, cudaStreamNonBlocking, hi));
    HSFM_CUDA(cudaStreamCreateWithPriority(&c.bulk, cudaStreamNonBlocking, lo));
    HSFM_CUDA(cudaStreamCreateWithPriority(&c.tail, cudaStreamNonBlocking, lo));
    for (cudaEvent_t* e : {&c.ev_panel, &c.ev_next, &c.ev_rest, &c.ev_trsm, &c.ev_tail, &c.ev_joined}) HSFM_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  }
  ev_seg.assign(MAX_CHOL_SEGMENTS, nullptr);
  for (cudaEvent_t& e : ev_seg) HSFM_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  return 0;
}
void CholAux::destroy() {
  if (graph) { cudaGraphExecDestroy(graph); graph = nullptr; }
  for (CholChainSet& c : sets) {
    if (c.chain) cudaStreamDestroy(c.chain);
    if (c.bulk) cudaStreamDestroy(c.bulk);
    if (c.tail) cudaStreamDestroy(c.tail);
    for (cudaEvent_t e : {c.ev_panel, c.ev_next, c.ev_rest, c.ev_trsm, c.ev_tail, c.ev_joined})
      if (e) cudaEventDestroy(e);
  }
  sets.clear();
  for (cudaEvent_t e : ev_seg)
    if (e) cudaEventDestroy(e);
  ev_seg.clear();
  if (ev_zero) { cudaEventDestroy(ev_zero); ev_zero = nullptr; }
  if (!stream) return;
  cudaStreamDestroy(stream);
  cudaEventDestroy(ev_fork); cudaEventDestroy(ev_panel); cudaEventDestroy(ev_rest); cudaEventDestroy(ev_join); cudaEventDestroy(ev_next);
  stream = nullptr;
}

// Factorisation with two panels of look-ahead.  Relative to panel k (trailing matrix starts at k0+128,
// 64-wide column blocks cb): cb 0,1 are panel k+1's columns, cb 2,3 panel k+2's, cb >= 4 the bulk.
//   chain stream (high priority):  diag(k) → trsm(k) → [panel k ready] → update cb 0,1  → diag(k+1) …
//   main stream:                   wait panel k → update cb 2,3 → [ahead k ready] → bulk update cb >= 4
// The chain's update of cb 0,1 at step k+1 waits for "ahead k" (those are the columns cb 2,3 of step k), so
// the bulk update of step k has the whole of chain step k+1 to finish before anything waits on it: the
// chain (latency-bound) and the bulk (DMMA-bound) overlap and the slower of the two sets the pace.
// ---------------------------------------------------------------------------------------------
// Envelope plan
// ---------------------------------------------------------------------------------------------
// Host part of build(): structure lists and schedule.  Returns false when the plan already holds this structure.
bool CholPlan::compute(int n_cams, const uint8_t* pattern, int P) {
  const int n_new = 6 * n_cams, panels_new = (n_new + NB - 1) / NB;
  {
    // same structure as the plan already holds (the per-keyframe calls of a scan mostly re-solve an unchanged camera
    // graph, and bench restarts the same problem): keep plan, device lists, version — and the captured launch graph
    std::vector<uint8_t> k;
    if (pattern && P == panels_new) k.assign(pattern, pattern + (size_t)P * P);
    if (n == n_new && version > 0 && k == key) return false;
    key.swap(k);
  }
  ++version;
  n = n_new;
  panels = panels_new;
  use = false; fill = 1.0; split_a = split_t0 = -1;
  off.assign((size_t)panels, 0); cnt.assign((size_t)panels, 0); list.clear();
  row_off.assign((size_t)panels + 1, 0); row_list.clear();
  blk_off.clear();
  segs.clear(); side.clear(); anc_all.clear(); map_all.clear(); side_total = 0; critical = panels;
  if (!pattern || P != panels || panels < 4) return true;
  std::vector<uint8_t> st(pattern, pattern + (size_t)P * P);
  symbolic_block_factor(P, st);
  structure_lists(P, st, off, cnt, list);
  for (int I = 0; I < P; ++I) {
    row_off[I] = (int)row_list.size();
    for (int J = 0; J < I; ++J)
      if (st[(size_t)I * P + J]) row_list.push_back(J);
    row_list.push_back(I);
  }
  row_off[P] = (int)row_list.size();
  fill = (double)row_list.size() / ((double)P * (P + 1) / 2);
  static const bool force_dense = getenv("HSFM_CHOL_DENSE") != nullptr;  // experiment knob
  use = fill < 0.6 && !force_dense;
  if (!use) return true;
  // panel chains: the paths of the block elimination tree
  {
    std::vector<PanelSeg> ps;
    std::vector<int> seg_of;
    panel_segments(P, off, cnt, list, ps, seg_of);
    static const bool one_chain = getenv("HSFM_ONE_CHAIN") != nullptr;  // experiment knob
    const int crit = critical_panels(ps);
    // worth it only when the concurrent chains shorten the critical path for real
    if (!one_chain && ps.size() > 1 && ps.size() <= (size_t)MAX_CHOL_SEGMENTS && crit <= (9 * P) / 10) {
      segs.resize(ps.size());
      int next_slot = 0;
      for (size_t g = 0; g < ps.size(); ++g) { segs[g].begin = ps[g].begin; segs[g].end = ps[g].end; segs[g].parent = ps[g].parent; }
      for (size_t g = 0; g < ps.size(); ++g)
        if (segs[g].parent >= 0) segs[(size_t)segs[g].parent].children.push_back((int)g);
      for (size_t g = 0; g < segs.size(); ++g) {
        Seg& sg = segs[g];
        sg.slot = sg.children.empty() ? next_slot++ : segs[(size_t)sg.children[0]].slot;  // a parent carries on where its first child ran
        for (int a = sg.parent; a >= 0; a = segs[(size_t)a].parent)
          for (int K = segs[(size_t)a].begin; K < segs[(size_t)a].end; ++K) sg.anc.push_back(K);
        sg.anc_off = (int)anc_all.size();
        anc_all.insert(anc_all.end(), sg.anc.begin(), sg.anc.end());
        sg.side_off = side_total;
        side_total += (long long)sg.anc.size() * NB * (long long)sg.anc.size() * NB;
      }
      map_all.assign(segs.size() * (size_t)P, -1);
      for (size_t g = 0; g < segs.size(); ++g)
        for (size_t q = 0; q < segs[g].anc.size(); ++q) map_all[g * (size_t)P + (size_t)segs[g].anc[q]] = (int)q;
      side.assign(list.size(), -1);
      for (int K = 0; K < P; ++K) {
        const int g = seg_of[(size_t)K];
        for (int q = 0; q < cnt[K]; ++q) {
          const int I = list[(size_t)off[K] + q];
          if (I >= segs[(size_t)g].end) side[(size_t)off[K] + q] = map_all[(size_t)g * P + (size_t)I];  // -1 here would break the etree property
        }
      }
      critical = crit;
      if (segs.size() == 3 && segs[0].parent == 2 && segs[1].parent == 2) { split_a = segs[1].begin; split_t0 = segs[2].begin; }
    }
  }
  blk_off.assign((size_t)panels + 1, 0);
  for (int I = 0; I < panels; ++I) {
    const long long rows = std::min(NB, n - I * NB), width = (long long)NB * (row_off[I + 1] - row_off[I]);
    blk_off[I + 1] = blk_off[I] + rows * width;
  }
  return true;
}

int CholPlan::build(int n_cams, const uint8_t* pattern, int P, cudaStream_t s) {
  if (!compute(n_cams, pattern, P) || !use) return 0;
  // slot of every structural block (I, K), I in R_K: its position in row I's list (the triangular solve's forward sweep)
  col_slot.assign(list.size(), 0);
  for (int K = 0; K < panels; ++K)
    for (int q = 0; q < cnt[K]; ++q) {
      const int I = list[(size_t)off[K] + q];
      const int* rb = row_list.data() + row_off[(size_t)I];
      const int* re = row_list.data() + row_off[(size_t)I + 1];
      col_slot[(size_t)off[K] + q] = (int)(std::lower_bound(rb, re, K) - row_list.data());
    }
  // sweep order of the triangular solves: columns by height in the block elimination tree (leaves first), ties by index
  {
    std::vector<int> height((size_t)panels, 0);
    for (int K = 0; K < panels; ++K)
      if (cnt[K] > 0) { const int par = list[(size_t)off[K]]; height[(size_t)par] = std::max(height[(size_t)par], height[(size_t)K] + 1); }
    sweep.resize((size_t)panels);
    for (int K = 0; K < panels; ++K) sweep[(size_t)K] = K;
    std::stable_sort(sweep.begin(), sweep.end(), [&](int a, int b) { return height[(size_t)a] < height[(size_t)b]; });
  }
  const size_t need = 3 * list.size() + row_off.size() + row_list.size() + 3 * (size_t)panels + anc_all.size() + map_all.size() + 1;
  if (need > d_cap) {
    if (d_ints) cudaFree(d_ints);
    d_ints = nullptr; d_cap = 0;
    HSFM_CUDA(cudaMalloc(&d_ints, sizeof(int) * need));
    d_cap = need;
  }
  d_list = d_ints;
  d_row_off = d_list + list.size();
  d_row_list = d_row_off + row_off.size();
  d_col_off = d_row_list + row_list.size();
  d_col_cnt = d_col_off + panels;
  d_side = d_col_cnt + panels;
  d_anc = d_side + list.size();
  d_map = d_anc + anc_all.size();
  d_col_slot = d_map + map_all.size();
  d_sweep = d_col_slot + list.size();
  if ((size_t)panels + 1 > d_blk_cap) {
    if (d_blk_off) cudaFree(d_blk_off);
    d_blk_off = nullptr; d_blk_cap = 0;
    HSFM_CUDA(cudaMalloc(&d_blk_off, sizeof(long long) * ((size_t)panels + 1)));
    d_blk_cap = (size_t)panels + 1;
  }
  HSFM_CUDA(cudaMemcpyAsync(d_blk_off, blk_off.data(), sizeof(long long) * blk_off.size(), cudaMemcpyHostToDevice, s));
  if (!list.empty()) HSFM_CUDA(cudaMemcpyAsync(d_list, list.data(), sizeof(int) * list.size(), cudaMemcpyHostToDevice, s));
  HSFM_CUDA(cudaMemcpyAsync(d_row_off, row_off.data(), sizeof(int) * row_off.size(), cudaMemcpyHostToDevice, s));
  HSFM_CUDA(cudaMemcpyAsync(d_row_list, row_list.data(), sizeof(int) * row_list.size(), cudaMemcpyHostToDevice, s));
  HSFM_CUDA(cudaMemcpyAsync(d_col_off, off.data(), sizeof(int) * (size_t)panels, cudaMemcpyHostToDevice, s));
  HSFM_CUDA(cudaMemcpyAsync(d_col_cnt, cnt.data(), sizeof(int) * (size_t)panels, cudaMemcpyHostToDevice, s));
  if (!side.empty()) HSFM_CUDA(cudaMemcpyAsync(d_side, side.data(), sizeof(int) * side.size(), cudaMemcpyHostToDevice, s));
  if (!anc_all.empty()) HSFM_CUDA(cudaMemcpyAsync(d_anc, anc_all.data(), sizeof(int) * anc_all.size(), cudaMemcpyHostToDevice, s));
  if (!map_all.empty()) HSFM_CUDA(cudaMemcpyAsync(d_map, map_all.data(), sizeof(int) * map_all.size(), cudaMemcpyHostToDevice, s));
  if (!col_slot.empty()) HSFM_CUDA(cudaMemcpyAsync(d_col_slot, col_slot.data(), sizeof(int) * col_slot.size(), cudaMemcpyHostToDevice, s));
  HSFM_CUDA(cudaMemcpyAsync(d_sweep, sweep.data(), sizeof(int) * sweep.size(), cudaMemcpyHostToDevice, s));
  HSFM_CUDA(cudaStreamSynchronize(s));  // the host vectors may be rebuilt by the next problem
  return 0;
}
// One CTA per matrix row: the row's structural blocks <-> their place in the packed buffer (16-byte accesses), or
// (mode 2) zeroed in place.
__global__ void __launch_bounds__(256) structure_rows_kernel(double* __restrict__ S, int n, const int* __restrict__ row_off,
                                                             const int* __restrict__ row_list, const long long* __restrict__ blk_off,
                                                             double* __restrict__ packed, int mode) {
  const int r = blockIdx.x, I = r / NB;
  const int q0 = __ldg(row_off + I), nb = __ldg(row_off + I + 1) - q0;
  double2* pk = mode == 2 ? nullptr : reinterpret_cast<double2*>(packed + __ldg(blk_off + I) + (long long)(r - I * NB) * (NB * nb));
  for (int e = threadIdx.x; e < nb * (NB / 2); e += 256) {
    const int q = e / (NB / 2), c2 = (e - q * (NB / 2)) * 2;
    const int col = __ldg(row_list + q0 + q) * NB + c2;
    if (col >= n) continue;  // the last block's columns past n do not exist (n is even: 6 per camera)
    double2* cell = reinterpret_cast<double2*>(S + (size_t)r * n + col);
    if (mode == 2) *cell = make_double2(0.0, 0.0);
    else if (mode == 1) pk[e] = *cell;
    else *cell = pk[e];
  }
}
void launch_envelope_pack(const CholPlan& pl, double* S, double* packed, bool to_packed, cudaStream_t s) {
  structure_rows_kernel<<<pl.n, 256, 0, s>>>(S, pl.n, pl.d_row_off, pl.d_row_list, pl.d_blk_off, packed, to_packed ? 1 : 0);
}
void launch_structure_zero(const CholPlan& pl, double* S, cudaStream_t s) {
  structure_rows_kernel<<<pl.n, 256, 0, s>>>(S, pl.n, pl.d_row_off, pl.d_row_list, nullptr, nullptr, 2);
}

void CholPlan::destroy() {
  if (d_blk_off) cudaFree(d_blk_off);
  d_blk_off = nullptr; d_blk_cap = 0;
  if (d_ints) cudaFree(d_ints);
  d_ints = nullptr; d_list = d_row_off = d_row_list = d_col_off = d_col_cnt = nullptr; d_cap = 0; use = false; key.clear(); version = 0;
}

// Structured factorisation: per panel K only the active row blocks R_K are solved and updated.  The critical path is
// kept as short as the data dependence allows:
//   chain stream:  diag(K) -> tile (K+1,K+1) straight from the unsolved block A(K+1,K) -> diag(K+1)
//   bulk stream:   after diag(K) and once the chain has read A(K+1,K): panel solve of all active row blocks, then every
//                  active pair but (K+1,K+1).  All of it has one whole diagonal-block factorisation to finish before
//                  the chain needs it (tile (K+2,K+1) at the next step).
// run_chain() enqueues the panels [K_begin, K_end) of ONE chain on a (chain, bulk) stream pair.  own_end < panels:
// trailing updates whose COLUMN block is >= own_end (R_K is ascending, so the pairs among its last entries) do not go
// into S but into the side matrix Ctail (order (panels - own_end) * 128, zero at the start): see launch_cholesky_struct.
struct ChainStreams {
  cudaStream_t chain, bulk;
  cudaEvent_t ev_panel, ev_next, ev_rest;
  bool la;  // chain and bulk are different streams
  // updates of the side matrix (tail x tail pairs) need only the panel solve and are wanted only at the join: their own stream
  cudaStream_t tail = nullptr;
  cudaEvent_t ev_trsm = nullptr;
};
static void run_chain(double* S, int n, double* ws, int* fail_flag, const ChainStreams& cs, int64_t* launches, const CholPlan& pl, int n_sm,
                      int K_begin, int K_end, int own_end, double* Cside, int ld_side, cudaEvent_t start_ev) {
  const bool la = cs.la;
  cudaStream_t s = cs.bulk, s2 = cs.chain;
  // Grid caps of the panel solves / trailing updates: experiment knobs.  (Measured with two chains in flight, c4: capping
  // them to 48 CTAs so that the other chain's kernels find empty SMs makes the factorisation SLOWER, 2.27 -> 3.0 ms: the
  // bulk step, ~180 tiles of 8 us of DMMA issue each, is itself close to the panel period.)
  static const int cap_syrk_env = getenv("HSFM_BULK_SYRK_CTAS") ? atoi(getenv("HSFM_BULK_SYRK_CTAS")) : 0;
  static const int cap_trsm_env = getenv("HSFM_BULK_TRSM_CTAS") ? atoi(getenv("HSFM_BULK_TRSM_CTAS")) : 0;
  const int cap_syrk = cap_syrk_env > 0 ? cap_syrk_env : 2 * n_sm;
  const int cap_trsm = cap_trsm_env > 0 ? cap_trsm_env : n_sm;
  auto trsm = [&](cudaStream_t st, int k0, double* Linv, const int* list, int cnt) {
    int rows_per = 64;  // strips of 8..64 rows: as many CTAs as fit in ONE wave (the kernel runs one CTA per SM)
    while (rows_per > 8 && cnt * (NB / (rows_per >> 1)) <= cap_trsm) rows_per >>= 1;
    const int spb = NB / rows_per;
    chol_trsm_strip_kernel<<<cnt * spb, 256, rows_per * TS_LD * (int)sizeof(double), st>>>(S, n, k0, Linv, rows_per, fail_flag, list, spb);
    ++*launches;
  };
  auto syrk = [&](cudaStream_t st, int k0, int col_first, int tiles, const int* list, int cnt, double* C, int ldc, const int* side_list) {
    if (tiles <= 0) return;
    static const bool big_tiles = getenv("HSFM_SYRK_BIG_TILES") != nullptr;  // experiment knob: 128 x 64 tiles, two CTAs per SM
    if (big_tiles) chol_syrk_kernel<NB><<<min(tiles, cap_syrk), 2 * NB, kGemmSmem, st>>>(S, n, k0, 0, col_first, tiles, 0, fail_flag, list, cnt, C, ldc, side_list);
    else chol_syrk_kernel<64><<<min(2 * tiles, (3 * cap_syrk) / 2), 128, kGemmSmemHalf, st>>>(S, n, k0, 0, col_first, 2 * tiles, 0, fail_flag, list, cnt, C, ldc, side_list);
    ++*launches;
  };
  // pairs of the panel's rows beyond its own segment -> side matrix, right after the panel solve; off the bulk stream when the
  // chain has a tail stream (they are wanted only when the parent segment starts)
  auto tail_update = [&](int K, int k0, const int* list, int n_own, int ct) {
    if (ct <= 0) return;
    cudaStream_t st = s;
    if (la && cs.tail) {
      cudaEventRecord(cs.ev_trsm, s);
      cudaStreamWaitEvent(cs.tail, cs.ev_trsm, 0);
      st = cs.tail;
    }
    syrk(st, k0, -1, ct * (ct + 1), list + n_own, ct, Cside, ld_side, pl.d_side + pl.off[K] + n_own);
  };
  // chain-stream launches carry the programmatic-serialisation attribute (see pdl_enter)
  static const bool no_pdl = getenv("HSFM_NO_PDL") != nullptr;  // experiment knob
  cudaLaunchAttribute pdl_attr[1];
  pdl_attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  pdl_attr[0].val.programmaticStreamSerializationAllowed = 1;
  auto chain_cfg = [&](int grid, int block, size_t smem) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = s2;
    cfg.attrs = pdl_attr; cfg.numAttrs = (la && !no_pdl) ? 1 : 0;
    return cfg;
  };
  bool rest_pending = false;
  if (la && start_ev) cudaStreamWaitEvent(s2, start_ev, 0);
  for (int K = K_begin, k0 = K_begin * NB; K < K_end; ++K, k0 += NB) {
    const int kb = min(NB, n - k0);
    double* Linv = ws + (size_t)K * NB * NB;
    {
      cudaLaunchConfig_t cfg = chain_cfg(1, 256, kDiagSmem);
      cudaLaunchKernelEx(&cfg, chol_diag_kernel, S, n, k0, kb, Linv, fail_flag);
    }
    ++*launches;
    const int cnt = pl.cnt[K];
    // everything the bulk stream did one step ago must be complete before the chain touches column K+1 / block K+1
    if (la && rest_pending) { cudaStreamWaitEvent(s2, cs.ev_rest, 0); rest_pending = false; }
    if (cnt == 0) continue;
    const int* hl = pl.list.data() + pl.off[K];
    const int* list = pl.d_list + pl.off[K];
    int n_own = cnt;  // active rows whose block column is updated in S here
    if (own_end < pl.panels) { n_own = 0; while (n_own < cnt && hl[n_own] < own_end) ++n_own; }
    const int ct = cnt - n_own;
    const bool has_col = hl[0] == K + 1 && n_own > 0;
    if (la) cudaEventRecord(cs.ev_panel, s2);  // diag(K) done
    const int pairs_own = n_own * cnt - n_own * (n_own - 1) / 2;  // pairs (a, b), b < n_own, a = b .. cnt-1
    // experiment knob HSFM_CHAIN_SPLIT=1: panel solve of row block K+1 + 32x32 tile update as two light kernels instead of the
    // one-kernel form (c4, two chains with the side-matrix updates on their own streams: 2.01 vs 1.88 ms; one chain: 3.41 vs 3.16)
    static const int split_env = getenv("HSFM_CHAIN_SPLIT") ? atoi(getenv("HSFM_CHAIN_SPLIT")) : -1;
    const bool split_chain = split_env > 0;
    if (has_col && !split_chain) {
      // tile (K+1, K+1) straight from the unsolved block A(K+1, K): one kernel between two diagonal blocks
      cudaLaunchConfig_t cfg = chain_cfg(2 * (NB / 8), 256, kChainTileSmem);
      cudaLaunchKernelEx(&cfg, chol_chain_tile_kernel, S, n, k0, (const double*)Linv, list, (const int*)fail_flag);
      ++*launches;
      if (la) cudaEventRecord(cs.ev_next, s2);  // A(K+1, K) has been read: the bulk stream may overwrite it with L(K+1, K)
      if (la) cudaStreamWaitEvent(s, cs.ev_panel, 0);
      if (la) cudaStreamWaitEvent(s, cs.ev_next, 0);
      trsm(s, k0, Linv, list, cnt);
      tail_update(K, k0, list, n_own, ct);
      syrk(s, k0, -2, 2 * (pairs_own - 1), list, cnt, nullptr, 0, nullptr);
      if (la) { cudaEventRecord(cs.ev_rest, s); rest_pending = true; }
      continue;
    }
    if (has_col) {
      {  // L(K+1, K): 16 strips of 8 rows
        cudaLaunchConfig_t cfg = chain_cfg(NB / 8, 256, 8 * TS_LD * sizeof(double));
        cudaLaunchKernelEx(&cfg, chol_trsm_strip_kernel, S, n, k0, (const double*)Linv, 8, (const int*)fail_flag, list, NB / 8);
        ++*launches;
      }
      if (la && cnt > 1) cudaEventRecord(cs.ev_next, s2);
      {  // tile (K+1, K+1): 10 active CTAs of 32 x 32
        cudaLaunchConfig_t cfg = chain_cfg(16, 128, kSyrkLatSmem / 2);
        cudaLaunchKernelEx(&cfg, chol_syrk_col_kernel<32>, S, n, k0, list, (const int*)fail_flag, 0);
        ++*launches;
      }
    }
    const int rest_cnt = has_col ? cnt - 1 : cnt;
    if (rest_cnt > 0) {
      const int* rl = has_col ? list + 1 : list;
      if (la) cudaStreamWaitEvent(s, cs.ev_panel, 0);
      trsm(s, k0, Linv, rl, rest_cnt);
      tail_update(K, k0, list, n_own, ct);
      if (has_col) {
        // one update launch for column K+1 below its diagonal tile and every other active pair
        if (la) cudaStreamWaitEvent(s, cs.ev_next, 0);
        syrk(s, k0, -2, 2 * (pairs_own - 1), list, cnt, nullptr, 0, nullptr);
      } else {
        syrk(s, k0, -1, 2 * pairs_own, list, cnt, nullptr, 0, nullptr);
      }
      if (la) { cudaEventRecord(cs.ev_rest, s); rest_pending = true; }
    }
  }
  // leave with the chain stream caught up with its bulk stream
  if (la && rest_pending) cudaStreamWaitEvent(s2, cs.ev_rest, 0);
}

// Join of the segment tree: a segment that starts folds the side matrix of one child (order na_c * 128, over the child's
// ancestor panels anc_c) into S — the blocks of its OWN columns [own_begin, own_end) — and into its own side matrix — the
// blocks of ITS ancestors' columns (map_p: panel -> position in the parent's side matrix).  One CTA per row of the child's
// side matrix, lower block triangle, whole diagonal blocks.  Children are folded one after the other on one stream: the
// summation order is fixed.
__global__ void __launch_bounds__(256) side_join_kernel(double* __restrict__ S, int n, const double* __restrict__ Tc, int na_c,
                                                        const int* __restrict__ anc_c, int own_begin, int own_end, double* __restrict__ Tp,
                                                        int ld_p, const int* __restrict__ map_p) {
  const int lr = blockIdx.x, lb = lr / NB, rr = lr - lb * NB;
  const int I = __ldg(anc_c + lb), r = I * NB + rr;
  if (r >= n) return;
  const int ld_c = na_c * NB;
  const double2* src = reinterpret_cast<const double2*>(Tc + (size_t)lr * ld_c);
  for (int e = threadIdx.x; e < (lb + 1) * (NB / 2); e += 256) {
    const int q = e / (NB / 2), c2 = (e - q * (NB / 2)) * 2;
    const int J = __ldg(anc_c + q);
    if (J * NB + c2 >= n) continue;
    double2* dst;
    if (J >= own_begin && J < own_end) dst = reinterpret_cast<double2*>(S + (size_t)r * n + (size_t)J * NB + c2);
    else dst = reinterpret_cast<double2*>(Tp + (size_t)(__ldg(map_p + I) * NB + rr) * ld_p + (size_t)__ldg(map_p + J) * NB + c2);
    double2 v = *dst;
    const double2 a = src[e];
    v.x += a.x; v.y += a.y;
    *dst = v;
  }
}

// One chain over all panels, or — when the plan carries a segment tree (CholPlan::segs: the two-ended / nested camera
// orders of ba_order.cpp) — every segment as a chain of its own on one of the stream sets of CholAux, started as soon as
// its children are done.  Updates a segment makes to its ancestors' blocks go to its side matrix (after the L11^-1 blocks
// in the workspace) and are folded in by side_join_kernel when the parent starts: concurrent segments never write the
// same memory and the summation order is fixed.
static void launch_cholesky_struct(double* S, int n, double* ws, int* fail_flag, cudaStream_t s, CholAux* aux, int64_t* launches,
                                   const CholPlan& pl, int n_sm) {
  static const bool no_la = getenv("HSFM_NO_LOOKAHEAD") != nullptr;
  const bool la = aux && aux->stream && !no_la;
  if (!la || pl.segs.size() < 2 || aux->sets.empty() || aux->ev_seg.size() < pl.segs.size()) {
    ChainStreams A{la ? aux->stream : s, s, la ? aux->ev_panel : nullptr, la ? aux->ev_next : nullptr, la ? aux->ev_rest : nullptr, la};
    if (la) {
      cudaEventRecord(aux->ev_fork, s);
      cudaStreamWaitEvent(A.chain, aux->ev_fork, 0);
    }
    run_chain(S, n, ws, fail_flag, A, launches, pl, n_sm, 0, pl.panels, pl.panels, nullptr, 0, nullptr);
    if (la) {
      cudaEventRecord(aux->ev_join, A.chain);
      cudaStreamWaitEvent(s, aux->ev_join, 0);
    }
    return;
  }
  double* side0 = ws + cholesky_workspace_doubles(n);
  const int n_sets = (int)aux->sets.size();
  std::vector<char> forked((size_t)n_sets, 0);
  cudaEventRecord(aux->ev_fork, s);
  // the side matrices start at zero: cleared on the first set's tail stream while the leaf chains are already running
  {
    CholChainSet& z = aux->sets[0];
    cudaStreamWaitEvent(z.tail, aux->ev_fork, 0);
    cudaMemsetAsync(side0, 0, sizeof(double) * (size_t)pl.side_total, z.tail);
    cudaEventRecord(aux->ev_zero, z.tail);
  }
  for (size_t g = 0; g < pl.segs.size(); ++g) {
    const CholPlan::Seg& sg = pl.segs[g];
    CholChainSet& st = aux->sets[(size_t)(sg.slot % n_sets)];
    if (!forked[(size_t)(sg.slot % n_sets)]) {
      forked[(size_t)(sg.slot % n_sets)] = 1;
      cudaStreamWaitEvent(st.chain, aux->ev_fork, 0);
      cudaStreamWaitEvent(st.bulk, aux->ev_fork, 0);
      cudaStreamWaitEvent(st.tail, aux->ev_zero, 0);
    }
    const int na = (int)sg.anc.size();
    double* side = side0 + sg.side_off;
    cudaEvent_t start_ev = nullptr;
    if (!sg.children.empty()) {
      cudaStreamWaitEvent(st.bulk, aux->ev_zero, 0);
      for (int c : sg.children) cudaStreamWaitEvent(st.bulk, aux->ev_seg[(size_t)c], 0);
      for (int c : sg.children) {
        const CholPlan::Seg& ch = pl.segs[(size_t)c];
        side_join_kernel<<<(int)ch.anc.size() * NB, 256, 0, st.bulk>>>(S, n, side0 + ch.side_off, (int)ch.anc.size(), pl.d_anc + ch.anc_off, sg.begin,
                                                                       sg.end, side, na * NB, pl.d_map + g * (size_t)pl.panels);
        ++*launches;
      }
      cudaEventRecord(st.ev_joined, st.bulk);
      cudaStreamWaitEvent(st.tail, st.ev_joined, 0);  // the segment's own side-matrix updates come after the folded-in ones
      start_ev = st.ev_joined;
    }
    ChainStreams cs{st.chain, st.bulk, st.ev_panel, st.ev_next, st.ev_rest, true};
    cs.tail = st.tail; cs.ev_trsm = st.ev_trsm;
    run_chain(S, n, ws, fail_flag, cs, launches, pl, n_sm, sg.begin, sg.end, sg.parent >= 0 ? sg.end : pl.panels, side, na * NB, start_ev);
    // done = chain, bulk and tail streams of the segment
    cudaEventRecord(st.ev_panel, st.chain);
    cudaStreamWaitEvent(st.bulk, st.ev_panel, 0);
    cudaEventRecord(st.ev_tail, st.tail);
    cudaStreamWaitEvent(st.bulk, st.ev_tail, 0);
    cudaEventRecord(aux->ev_seg[g], st.bulk);
  }
  // every segment is a descendant of a root, but every stream that was forked has to join the caller's stream (graph capture)
  for (size_t g = 0; g < pl.segs.size(); ++g)
    if (pl.segs[g].parent < 0) cudaStreamWaitEvent(s, aux->ev_seg[g], 0);
  cudaStreamWaitEvent(s, aux->ev_zero, 0);
}

void launch_cholesky(double* S, int n, double* ws, int* fail_flag, cudaStream_t s, CholAux* aux, int64_t* launches, const CholPlan* plan) {
  const int n_sm = (aux && aux->n_sm > 0) ? aux->n_sm : 148;
  static int bulk_ctas = 0;
  if (!bulk_ctas) {
    const char* e = getenv("HSFM_BULK_CTAS");  // experiment knob: <0 means n_sm + value
    bulk_ctas = e ? atoi(e) : 2 * n_sm;
    if (bulk_ctas == 0) bulk_ctas = 1 << 30;  // one tile per CTA: slots turn over every tile, chain kernels slip in
    else if (bulk_ctas < 0) bulk_ctas = n_sm + bulk_ctas;
  }
  if (plan && plan->use && plan->n == n) {
    // The launch sequence (4-5 kernels and 3 event operations per panel, two streams) is static per problem:
    // capture it once and replay it as a CUDA graph — no host launch cost, tighter kernel-to-kernel hand-over.
    static const bool no_graph = getenv("HSFM_NO_GRAPH") != nullptr;  // experiment knob
    if (aux && aux->stream && !no_graph) {
      const bool valid = aux->graph && aux->g_S == S && aux->g_ws == ws && aux->g_fail == fail_flag && aux->g_stream == (const void*)s &&
                         aux->g_n == n && aux->g_version == plan->version;
      if (!valid) {
        if (aux->graph) { cudaGraphExecDestroy(aux->graph); aux->graph = nullptr; }
        cudaGraph_t g = nullptr;
        int64_t counted = 0;
        if (cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
          launch_cholesky_struct(S, n, ws, fail_flag, s, aux, &counted, *plan, n_sm);
          if (cudaStreamEndCapture(s, &g) == cudaSuccess && g && cudaGraphInstantiate(&aux->graph, g, 0) == cudaSuccess) {
            aux->g_S = S; aux->g_ws = ws; aux->g_fail = fail_flag; aux->g_stream = (const void*)s; aux->g_n = n;
            aux->g_version = plan->version; aux->g_launches = counted;
          } else {
            aux->graph = nullptr;
          }
          if (g) cudaGraphDestroy(g);
        }
        cudaGetLastError();
      }
      if (aux->graph && cudaGraphLaunch(aux->graph, s) == cudaSuccess) {
        *launches += aux->g_launches;
        return;
      }
      cudaGetLastError();
    }
    launch_cholesky_struct(S, n, ws, fail_flag, s, aux, launches, *plan, n_sm);
    return;
  }
  const int panels = (n + NB - 1) / NB;
  static const bool no_la = getenv("HSFM_NO_LOOKAHEAD") != nullptr;  // experiment knob
  const bool la = aux && aux->stream && panels > 2 && !no_la;
  cudaStream_t s2 = la ? aux->stream : s;
  if (la) {
    cudaEventRecord(aux->ev_fork, s);
    cudaStreamWaitEvent(s2, aux->ev_fork, 0);
  }
  for (int panel = 0, k0 = 0; k0 < n; k0 += NB, ++panel) {
    const int kb = min(NB, n - k0);
    double* Linv = ws + (size_t)panel * NB * NB;
    chol_diag_kernel<<<1, 256, kDiagSmem, s2>>>(S, n, k0, kb, Linv, fail_flag);
    ++*launches;
    const int m = n - (k0 + kb);
    if (m <= 0) break;
    const int rb = (m + NB - 1) / NB;
    int rows_per = ((m + n_sm - 1) / n_sm + 7) & ~7;
    rows_per = max(8, min(TS_MAXROWS, rows_per));
    chol_trsm_strip_kernel<<<(m + rows_per - 1) / rows_per, 256, rows_per * TS_LD * (int)sizeof(double), s2>>>(S, n, k0, Linv, rows_per,
                                                                                                                   fail_flag, nullptr, 1);
    ++*launches;
    if (la) {
      cudaEventRecord(aux->ev_panel, s2);
      if (panel > 0) cudaStreamWaitEvent(s2, aux->ev_rest, 0);  // "ahead k-1": cb 2,3 of the previous step = cb 0,1 of this one
    }
    chol_syrk_kernel<NB><<<min(2 * rb, 2 * n_sm), 2 * NB, kGemmSmem, s2>>>(S, n, k0, 0, 0, 2 * rb, 0, fail_flag, nullptr, 0, nullptr, 0, nullptr);
    ++*launches;
    if (rb > 1) {
      if (la) cudaStreamWaitEvent(s, aux->ev_panel, 0);
      chol_syrk_kernel<NB><<<min(2 * (rb - 1), 2 * n_sm), 2 * NB, kGemmSmem, s>>>(S, n, k0, 1, 2, 2 * (rb - 1), 0, fail_flag, nullptr, 0, nullptr, 0, nullptr);
      ++*launches;
    }
    if (la) cudaEventRecord(aux->ev_rest, s);
    if (rb > 2) {
      chol_syrk_kernel<NB><<<min((rb - 2) * (rb - 1), bulk_ctas), 2 * NB, kGemmSmem, s>>>(S, n, k0, 0, -4, (rb - 2) * (rb - 1), 0, fail_flag, nullptr, 0, nullptr, 0, nullptr);
      ++*launches;
    }
  }
  if (la) {
    cudaEventRecord(aux->ev_join, s2);
    cudaStreamWaitEvent(s, aux->ev_join, 0);
  }
}

// Solves S y = b with the factor from launch_cholesky.  b holds 3n doubles: [0,n) right-hand side
// in / solution out, [n,2n) the forward vector, [2n,3n) the backward vector (scratch).
void launch_chol_solve(const double* L, int n, double* ws, double* b, cudaStream_t s, int64_t* launches, const CholAux* aux, const CholPlan* plan) {
  double* w = b + n;
  const int panels = (n + NB - 1) / NB;
  const int n_sm = (aux && aux->n_sm > 0) ? aux->n_sm : 148;
  const int coop = aux ? aux->coop : 0;
  if (coop) {
    double* x = b + 2 * (size_t)n;  // backward result; copied to b afterwards
    const bool st = plan && plan->use && plan->n == n;
    double* slots = ws + (((size_t)panels * NB * NB + panels + 1) & ~(size_t)1);
    const size_t n_slots = st ? plan->row_list.size() : (size_t)panels * (panels + 1) / 2;
    // w, x and the slots: the sentinel the consumers poll against
    cudaMemsetAsync(w, 0xff, sizeof(double) * 2 * (size_t)n, s);
    cudaMemsetAsync(slots, 0xff, sizeof(double) * n_slots * NB, s);
    const double* LinvAll = ws;
    int nn = n;
    const int* row_off = st ? plan->d_row_off : nullptr;
    const int* col_off = st ? plan->d_col_off : nullptr;
    const int* col_cnt = st ? plan->d_col_cnt : nullptr;
    const int* col_list = st ? plan->d_list : nullptr;
    const int* col_slot = st ? plan->d_col_slot : nullptr;
    const int* sweep = (st && panels > n_sm) ? plan->d_sweep : nullptr;  // every column has a CTA of its own otherwise
    void* args[] = {(void*)&L,       (void*)&nn,      (void*)&LinvAll, (void*)&b,        (void*)&w,        (void*)&x,
                    (void*)&slots,   (void*)&row_off, (void*)&col_off, (void*)&col_cnt,  (void*)&col_list, (void*)&col_slot, (void*)&sweep};
    const int grid = std::min(panels, n_sm);
    if (cudaLaunchCooperativeKernel((const void*)trsv_coop_kernel, dim3(grid), dim3(256), args, kTrsvSmem, s) == cudaSuccess) {
      ++*launches;
      cudaMemcpyAsync(b, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s);
      return;
    }
    cudaGetLastError();  // fall through to the per-panel launches
  }
  int panel = 0;
  for (int k0 = 0; k0 < n; k0 += NB, ++panel) {
    const int kb = min(NB, n - k0);
    const int m = n - (k0 + kb);
    trsv_fwd_kernel<<<1 + (m + NB - 1) / NB, 256, 0, s>>>(L, n, k0, kb, ws + (size_t)panel * NB * NB, b, w);
    ++*launches;
  }
  for (--panel; panel >= 0; --panel) {
    const int k0 = panel * NB;
    const int kb = min(NB, n - k0);
    trsv_bwd_kernel<<<1 + (k0 + NB - 1) / NB, 256, 0, s>>>(L, n, k0, kb, ws + (size_t)panel * NB * NB, w, b);
    ++*launches;
  }
}

}  // namespace hsfm
