// ba_kernels.cu — sm_100a kernels of the bundle-adjustment hot path.
//
// Data layout: observations are kept twice, both static after upload:
//   * point-major (sorted by point), cut into TILES of <= 256 observations that contain whole
//     points only.  One CTA = one tile, one thread = one observation.  All point-keyed sums
//     (Hpp, g_p, W^T y, back-substitution) are staged per observation in shared memory and added
//     per point in observation order — deterministic, no atomics.
//   * camera-major (sorted by camera), cut into warp CHUNKS of <= 256 observations of ONE
//     camera.  Camera-keyed sums that are needed once per linearisation (Hcc, g_c, the
//     block-Jacobi diagonal) are recomputed here from 20 B/observation + a point gather
//     instead of being scattered with 27 atomics per observation.
// The linearisation (r, Jc, Jp) is materialised point-major as SoA planes of double2 = (du,dv)
// so every load/store of the streaming passes is a fully coalesced 16-byte access.
#include <cooperative_groups.h>

#include "ba_internal.h"

namespace hsfm {

namespace {

constexpr unsigned FULL = 0xffffffffu;

// D(8x8) += A(8x4) * B(4x8), FP64 tensor core.  Lane l holds A[l>>2][l&3], B[l&3][l>>2], D[l>>2][2(l&3) .. +1].
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

__device__ __forceinline__ CamRow load_row(const CamRow* rows, int c) {
  CamRow C;
  const double2* src = reinterpret_cast<const double2*>(rows + c);
  double2* dst = reinterpret_cast<double2*>(&C);
#pragma unroll
  for (int q = 0; q < 12; ++q) dst[q] = __ldg(src + q);
  return C;
}

// Everything a CTA needs to know about its tile, from one 32-byte record.
struct TileHdr { int b, e, p_first, n_p, ncam; };
__device__ __forceinline__ TileHdr tile_hdr(const Dev& d) {
  const int4 h0 = __ldg(d.tile_hdr + 2 * (size_t)blockIdx.x), h1 = __ldg(d.tile_hdr + 2 * (size_t)blockIdx.x + 1);
  TileHdr t;
  t.b = h0.x; t.e = h0.y; t.p_first = h0.z; t.n_p = h0.w - h0.z; t.ncam = h1.x;
  return t;
}

// Shared-memory staging of a tile's camera rows and points.  The index holds, per tile, the ascending list of its
// distinct cameras (~40 at Venice shape, against ~18 distinct rows touched by EVERY one of the twelve 16-byte
// gathers a warp would otherwise issue) and, per observation, the slot of its camera in that list; a tile's
// points are the contiguous range [p_first, p_first + n_p).  All 256 threads copy rows and points with coalesced
// loads that depend on nothing but the tile header, so the whole input of the tile is two memory round trips
// away; the observations then read their row IN PLACE from shared memory.  Row stride 208 B: eight
// consecutive slots fall in eight different 16-byte bank groups.
constexpr int ROW_LD = 13;  // double2 per staged row (12 + 1 pad)
__device__ __forceinline__ void stage_rows(const Dev& d, const TileHdr& t, const CamRow* __restrict__ rows, double2* sR) {
  const int* list = d.tile_cams + (size_t)blockIdx.x * TILE;
  const int m = t.ncam * 12;
  for (int q = threadIdx.x; q < m; q += TILE) {
    const int sl = q / 12, piece = q - sl * 12;
    sR[sl * ROW_LD + piece] = __ldg(reinterpret_cast<const double2*>(rows + __ldg(list + sl)) + piece);
  }
}
__device__ __forceinline__ void stage_points(const TileHdr& t, const double* __restrict__ src, double* sX) {
  const double* g = src + 3 * (size_t)t.p_first;
  for (int q = threadIdx.x; q < 3 * t.n_p; q += TILE) sX[q] = __ldg(g + q);
}
__device__ __forceinline__ const CamRow& staged_row(const double2* sR, int slot) {
  return *reinterpret_cast<const CamRow*>(sR + slot * ROW_LD);
}
// dynamic shared memory of the kernels that stage rows (doubles): rows, then X, then whatever the kernel adds
__host__ __device__ constexpr int rows_doubles(int row_cap) { return 2 * ROW_LD * row_cap; }

__device__ __forceinline__ constexpr int tri6(int a, int b) { return a * 6 - (a * (a - 1)) / 2 + (b - a); }  // a <= b

// Sum over the CTA; result valid on thread 0.  sm must hold 8 doubles... (256 threads).
__device__ __forceinline__ double block_sum_256(double v, double* sm) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(FULL, v, d);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sm[w];
  }
  __syncthreads();
  return t;
}

// Per-thread view of a point-major tile.
struct TileThread {
  int i;      // observation index (point-major), valid iff i < tile_e
  int tile_b, tile_e;
  int p;      // point id
  int seg_b, seg_e;  // observation range of the point
  bool valid;
};

__device__ __forceinline__ TileThread tile_thread(const Dev& d, int tile) {
  TileThread t;
  t.tile_b = __ldg(d.tile_start + tile);
  t.tile_e = __ldg(d.tile_start + tile + 1);
  t.i = t.tile_b + (int)threadIdx.x;
  t.valid = t.i < t.tile_e;
  t.p = t.valid ? __ldg(d.pt + t.i) : 0;
  t.seg_b = t.valid ? __ldg(d.pt_start + t.p) : 0;
  t.seg_e = t.valid ? __ldg(d.pt_start + t.p + 1) : 0;
  return t;
}
__device__ __forceinline__ TileThread tile_thread(const Dev& d) { return tile_thread(d, (int)blockIdx.x); }

// ------------------------------------------------------------------------------------------
__global__ void cam_prep_kernel(int n_cams, const double* __restrict__ cam, const double* __restrict__ K9,
                                const double* __restrict__ dist, const double* __restrict__ scale_c, CamRow* __restrict__ rows) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cams) return;
  CamRow C;
  cam_prepare(cam + 6 * c, K9 + 9 * c, dist ? dist + 2 * c : nullptr, scale_c ? scale_c + 6 * c : nullptr, C);
  rows[c] = C;
}

// Point-keyed sums of a tile through shared memory.  Every valid thread has stored its K values at
// sv[k * SV_LD + threadIdx.x]; after the barrier one thread per (point of the tile, k) adds the L values
// of that point in observation order (deterministic) and hands the total to out(point, k, total).
// SV_LD = 257: the K lanes of one point fall in K distinct 8-byte banks.  The CSR bounds of the first
// two task rounds are fetched by prefetch_point_bounds at kernel start, off the critical path.
constexpr int SV_LD = TILE + 1;
template <int K>
__device__ __forceinline__ void prefetch_point_bounds(const Dev& d, const TileHdr& t, int (&bnd)[4]) {
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int task = threadIdx.x + r * TILE;
    bnd[2 * r] = bnd[2 * r + 1] = 0;
    if (task < t.n_p * K) {
      const int p = t.p_first + task / K;
      bnd[2 * r] = __ldg(d.pt_start + p) - t.b;
      bnd[2 * r + 1] = __ldg(d.pt_start + p + 1) - t.b;
    }
  }
}
template <int K, class Out>
__device__ __forceinline__ void tile_point_sums(const Dev& d, const TileHdr& t, const int (&bnd)[4], const double* sv, Out&& out) {
  const int n_task = t.n_p * K;
#pragma unroll 1
  for (int r = 0, task = threadIdx.x; task < n_task; task += TILE, ++r) {
    const int pl = task / K, k = task - pl * K;
    const int p = t.p_first + pl;
    int b, e;
    if (r == 0) { b = bnd[0]; e = bnd[1]; }
    else if (r == 1) { b = bnd[2]; e = bnd[3]; }
    else { b = __ldg(d.pt_start + p) - t.b; e = __ldg(d.pt_start + p + 1) - t.b; }
    const double* src = sv + k * SV_LD;
    double s0 = 0.0, s1 = 0.0;
    int l = b;
    for (; l + 1 < e; l += 2) { s0 += src[l]; s1 += src[l + 1]; }
    if (l < e) s0 += src[l];
    out(p, k, s0 + s1);
  }
}

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// Kernel RJ: residual + Jacobian per observation (a1), fused with the point-keyed part of the
// normal equations (a7: Hpp = sum Jp^T Jp, g_p = sum Jp^T r) and the cost partial (a6).
// Algorithmic bytes / observation: 24 in (obs 16, cam 4, pt 4) + 160 out (r 16, Jc 96, Jp 48).
//
// PERSISTENT and double-buffered: two CTAs per SM walk the tiles.  Everything a tile needs is brought into shared
// memory by cp.async one tile ahead — camera rows, points, scales, observations, point ids, slots, CSR bounds —
// and the header + camera list ("meta") two tiles ahead, because the row copies of tile t+1 take their source
// addresses from it.  The three dependent memory round trips of a tile (header -> lists -> rows) are never
// waited for after the first tile, and being asynchronous copies the prefetches cannot be scheduled late by
// the compiler the way register prefetches are.  Every Jacobian column is stored the moment it is formed; the
// camera row is read in place from shared memory.  One barrier per tile (+ one before the point sums).
struct RjStage {          // byte offsets inside one stage, all 16-byte aligned
  int rows, X, S, obs, pt, slotw, bnd, bytes;
};
__host__ __device__ constexpr RjStage rj_stage(int row_cap) {
  RjStage g{};
  g.rows = 0;
  g.X = g.rows + 16 * ROW_LD * row_cap;
  g.S = g.X + 24 * TILE;
  g.obs = g.S + 24 * TILE;
  g.pt = g.obs + 16 * TILE;
  g.slotw = g.pt + 4 * TILE;
  g.bnd = g.slotw + 4 * TILE;      // [2 rounds][TILE][2] ints
  g.bytes = g.bnd + 16 * TILE;
  return g;
}
constexpr int RJ_META_BYTES = 32 + 4 * TILE;  // header + camera list
__host__ constexpr size_t rj_smem_bytes(int row_cap, bool form) {
  return 2 * (size_t)rj_stage(row_cap).bytes + 3 * RJ_META_BYTES + (form ? sizeof(double) * 9 * SV_LD : 0);
}
template <bool WRITE_J, bool FORM>
__global__ void __launch_bounds__(TILE, 2) rj_kernel(Dev d, const CamRow* __restrict__ rows, const double* __restrict__ pts,
                                                      const double* __restrict__ scale_p, double* __restrict__ Hpp,
                                                      double* __restrict__ gp, double* __restrict__ cost_part, int row_cap) {
  extern __shared__ __align__(16) unsigned char dynb[];
  const RjStage L = rj_stage(row_cap);
  unsigned char* meta0 = dynb + 2 * L.bytes;
  double* sv = reinterpret_cast<double*>(meta0 + 3 * RJ_META_BYTES);  // [9][SV_LD] (FORM)
  __shared__ double sm[8];
  const int tid = threadIdx.x;
  const int rounds = (row_cap * 12 + TILE - 1) / TILE;

  // three meta slots: tile t computes from slot t % 3 while t+1's is read for its copies and t+2's is being written
  auto meta_hdr = [&](int seq) { return reinterpret_cast<const int*>(meta0 + (seq % 3) * RJ_META_BYTES); };
  auto meta_list = [&](int seq) { return reinterpret_cast<const int*>(meta0 + (seq % 3) * RJ_META_BYTES + 32); };
  // header (two 16-byte pieces) + the 64 16-byte pieces of the camera list of one tile
  auto issue_meta = [&](int seq, int tile) {
    unsigned char* m = meta0 + (seq % 3) * RJ_META_BYTES;
    if (tid < 2) cp_async16(m + 16 * tid, d.tile_hdr + 2 * (size_t)tile + tid);
    else if (tid < 2 + TILE / 4) cp_async16(m + 32 + 16 * (tid - 2), d.tile_cams + (size_t)tile * TILE + 4 * (tid - 2));
  };
  // all inputs of the tile whose meta is in slot seq -> stage seq & 1
  auto issue_stage = [&](int seq) {
    const int* h = meta_hdr(seq);
    const int* lst = meta_list(seq);
    const int tb = h[0], te = h[1], p_first = h[2], n_p = h[3] - h[2], ncam = h[4];
    unsigned char* st = dynb + (seq & 1) * L.bytes;
    double2* sR = reinterpret_cast<double2*>(st + L.rows);
    for (int r = 0; r < rounds; ++r) {
      const int q = tid + r * TILE;
      if (q < ncam * 12) {
        const int sl = q / 12, piece = q - sl * 12;
        cp_async16(sR + sl * ROW_LD + piece, reinterpret_cast<const double2*>(rows + lst[sl]) + piece);
      }
    }
    double* sX = reinterpret_cast<double*>(st + L.X);
    double* sS = reinterpret_cast<double*>(st + L.S);
    for (int q = tid; q < 3 * n_p; q += TILE) {
      cp_async8(sX + q, pts + 3 * (size_t)p_first + q);
      if (scale_p) cp_async8(sS + q, scale_p + 3 * (size_t)p_first + q);
    }
    const int i = tb + tid;
    if (i < te) {
      cp_async16(st + L.obs + 16 * tid, d.obs + i);
      cp_async4(st + L.pt + 4 * tid, d.pt + i);
      cp_async4(st + L.slotw + 4 * tid, d.slot + (i & ~3));  // the aligned word that holds this observation's slot byte
    }
    if (FORM) {
      int* sb = reinterpret_cast<int*>(st + L.bnd);
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int task = tid + r * TILE;
        if (task < n_p * 9) {  // {start, end} of the task's point
          cp_async4(sb + 2 * task, d.pt_start + p_first + task / 9);
          cp_async4(sb + 2 * task + 1, d.pt_start + p_first + task / 9 + 1);
        }
      }
    }
  };

  const int tile0 = blockIdx.x;
  if (tile0 >= d.n_tiles) return;
  // prologue: meta of the first two tiles, then the first stage
  issue_meta(0, tile0);
  if (tile0 + (int)gridDim.x < d.n_tiles) issue_meta(1, tile0 + gridDim.x);
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  issue_stage(0);
  cp_async_commit();
  double cost_acc = 0.0;  // this thread's share of the cost over all tiles of the CTA
  for (int seq = 0, tile = tile0; tile < d.n_tiles; ++seq, tile += gridDim.x) {
    cp_async_wait<0>();  // this tile's stage (and the next tile's meta) have landed ...
    __syncthreads();     // ... for everyone, and everyone has left the previous tile: its stage, meta slot and sv are free
    const int* h = meta_hdr(seq);
    TileHdr t;
    t.b = h[0]; t.e = h[1]; t.p_first = h[2]; t.n_p = h[3] - h[2]; t.ncam = h[4];
    const int nxt = tile + gridDim.x, nn = nxt + gridDim.x;
    if (nxt < d.n_tiles) issue_stage(seq + 1);
    if (nn < d.n_tiles) issue_meta(seq + 2, nn);
    cp_async_commit();

    const unsigned char* st = dynb + (seq & 1) * L.bytes;
    const double2* sR = reinterpret_cast<const double2*>(st + L.rows);
    const double* sX = reinterpret_cast<const double*>(st + L.X);
    const double* sS = reinterpret_cast<const double*>(st + L.S);
    const int i = t.b + tid;
    if (i < t.e) {
      const double2 o = *reinterpret_cast<const double2*>(st + L.obs + 16 * tid);
      const int slot = (int)((*reinterpret_cast<const unsigned*>(st + L.slotw + 4 * tid) >> (8 * (i & 3))) & 0xffu);
      const int pl = *reinterpret_cast<const int*>(st + L.pt + 4 * tid) - t.p_first;
      const CamRow& C = staged_row(sR, slot);
      const double X0 = sX[3 * pl], X1 = sX[3 * pl + 1], X2 = sX[3 * pl + 2];
      double sp[3] = {1.0, 1.0, 1.0};
      if (scale_p) { sp[0] = sS[3 * pl]; sp[1] = sS[3 * pl + 1]; sp[2] = sS[3 * pl + 2]; }
      const size_t n = (size_t)d.n_obs;
      double r0, r1, jp[6];
      reproject_lin(C, X0, X1, X2, o.x, o.y, sp, r0, r1, [&](int col, double du, double dv) {
        if (col < 6) {
          if (WRITE_J) d.Jc[col * n + i] = make_double2(du, dv);
        } else {
          if (WRITE_J) d.Jp[(col - 6) * n + i] = make_double2(du, dv);
          jp[col - 6] = du; jp[col - 3] = dv;
        }
      });
      if (WRITE_J) d.r[i] = make_double2(r0, r1);
      cost_acc += r0 * r0 + r1 * r1;
      if (FORM) {
        double* v = sv + tid;
        v[0 * SV_LD] = jp[0] * jp[0] + jp[3] * jp[3];
        v[1 * SV_LD] = jp[0] * jp[1] + jp[3] * jp[4];
        v[2 * SV_LD] = jp[0] * jp[2] + jp[3] * jp[5];
        v[3 * SV_LD] = jp[1] * jp[1] + jp[4] * jp[4];
        v[4 * SV_LD] = jp[1] * jp[2] + jp[4] * jp[5];
        v[5 * SV_LD] = jp[2] * jp[2] + jp[5] * jp[5];
        v[6 * SV_LD] = jp[0] * r0 + jp[3] * r1;
        v[7 * SV_LD] = jp[1] * r0 + jp[4] * r1;
        v[8 * SV_LD] = jp[2] * r0 + jp[5] * r1;
      }
    }
    if (FORM) {
      __syncthreads();
      const int* sb = reinterpret_cast<const int*>(st + L.bnd);
      int bnd[4] = {0, 0, 0, 0};
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int task = tid + r * TILE;
        if (task < t.n_p * 9) { bnd[2 * r] = sb[2 * task] - t.b; bnd[2 * r + 1] = sb[2 * task + 1] - t.b; }
      }
      tile_point_sums<9>(d, t, bnd, sv, [&](int p, int k, double tot) {
        if (k < 6) Hpp[6 * (size_t)p + k] = tot;
        else gp[3 * (size_t)p + (k - 6)] = tot;
      });
    }
    if (tid == 0 && tile != tile0) cost_part[tile] = 0.0;  // the CTA's cost goes to the slot of its first tile
  }
  cp_async_wait<0>();
  const double cs = block_sum_256(cost_acc, sm);
  if (tid == 0) cost_part[tile0] = cs;
}

// Candidate cost: residual only (a9), 24 B/observation.
__global__ void __launch_bounds__(TILE) cost_kernel(Dev d, const CamRow* __restrict__ rows, const double* __restrict__ pts,
                                                     double* __restrict__ cost_part, int row_cap) {
  extern __shared__ __align__(16) double dyn[];
  double2* sR = reinterpret_cast<double2*>(dyn);
  double* sX = dyn + rows_doubles(row_cap);
  __shared__ double sm[8];
  const TileHdr t = tile_hdr(d);
  const int i = t.b + (int)threadIdx.x;
  const bool valid = i < t.e;
  double2 o = make_double2(0.0, 0.0);
  int pl = 0, slot = 0;
  if (valid) {
    o = __ldg(d.obs + i);
    slot = __ldg(d.slot + i);
    pl = __ldg(d.pt + i) - t.p_first;
  }
  stage_points(t, pts, sX);
  stage_rows(d, t, rows, sR);
  __syncthreads();
  double e = 0.0;
  if (valid) {
    double r0, r1;
    reproject_r(staged_row(sR, slot), sX[3 * pl], sX[3 * pl + 1], sX[3 * pl + 2], o.x, o.y, r0, r1);
    e = r0 * r0 + r1 * r1;
  }
  const double cs = block_sum_256(e, sm);
  if (threadIdx.x == 0) cost_part[blockIdx.x] = cs;
}
__host__ constexpr size_t cost_smem_bytes(int row_cap) { return sizeof(double) * (size_t)(rows_doubles(row_cap) + 3 * TILE); }

// Camera-major pass: one warp per chunk of <= 256 observations of one camera; recomputes the
// (scaled) Jacobian from 20 B/observation + a point gather and reduces camera-keyed sums.
//   WHAT == 0: 21 upper-triangle entries of Jc^T Jc and the 6 of Jc^T r          (a7: Hcc, g_c)
//   WHAT == 1: 21 entries of Jc^T (Jp Hpp^-1 Jp^T) Jc  (diagonal blocks of W Hpp^-1 W^T)
template <int WHAT>
__global__ void __launch_bounds__(256) cam_major_kernel(Dev d, const CamRow* __restrict__ rows, const double* __restrict__ pts,
                                                        const double* __restrict__ scale_p, const double* __restrict__ Hpp_inv,
                                                        double* __restrict__ chunk_part) {
  constexpr int NV = (WHAT == 0) ? 27 : 21;
  const int lane = threadIdx.x & 31;
  const int wc = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (wc >= d.n_chunks) return;
  const int c = __ldg(d.chunk_cam + wc);
  const int b = __ldg(d.chunk_b + wc), e = __ldg(d.chunk_e + wc);
  const CamRow C = load_row(rows, c);
  double acc[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) acc[k] = 0.0;
  for (int j = b + lane; j < e; j += 32) {
    const double2 o = __ldg(d.obs_cs + j);
    const int p = __ldg(d.pt_cs + j);
    const double X0 = __ldg(pts + 3 * (size_t)p), X1 = __ldg(pts + 3 * (size_t)p + 1), X2 = __ldg(pts + 3 * (size_t)p + 2);
    double sp[3] = {1.0, 1.0, 1.0};
    if (scale_p) {
      sp[0] = __ldg(scale_p + 3 * (size_t)p); sp[1] = __ldg(scale_p + 3 * (size_t)p + 1); sp[2] = __ldg(scale_p + 3 * (size_t)p + 2);
    }
    double r0, r1, jc[12], jp[6];
    reproject_rj(C, X0, X1, X2, o.x, o.y, sp, r0, r1, jc, jp);
    if (WHAT == 0) {
#pragma unroll
      for (int a = 0; a < 6; ++a) {
#pragma unroll
        for (int bb = a; bb < 6; ++bb) acc[tri6(a, bb)] += jc[a] * jc[bb] + jc[6 + a] * jc[6 + bb];
        acc[21 + a] += jc[a] * r0 + jc[6 + a] * r1;
      }
    } else {
      double Hi[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) Hi[k] = __ldg(Hpp_inv + 6 * (size_t)p + k);
      // M = Jp Hi Jp^T (2x2 symmetric)
      double a0, a1, a2, b0, b1, b2;
      sym3_mul(Hi, jp[0], jp[1], jp[2], a0, a1, a2);
      sym3_mul(Hi, jp[3], jp[4], jp[5], b0, b1, b2);
      const double m00 = jp[0] * a0 + jp[1] * a1 + jp[2] * a2;
      const double m01 = jp[0] * b0 + jp[1] * b1 + jp[2] * b2;
      const double m11 = jp[3] * b0 + jp[4] * b1 + jp[5] * b2;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
        const double ta = m00 * jc[a] + m01 * jc[6 + a];  // (M Jc)[0][a]
        const double tb = m01 * jc[a] + m11 * jc[6 + a];  // (M Jc)[1][a]
#pragma unroll
        for (int bb = a; bb < 6; ++bb) acc[tri6(a, bb)] += jc[bb] * ta + jc[6 + bb] * tb;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double v = acc[k];
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) v += __shfl_xor_sync(FULL, v, dd);
    acc[k] = v;
  }
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) chunk_part[(size_t)wc * 27 + k] = acc[k];
  }
}

// Per-keyframe reprojection statistics after the adjustment (the block the reference keeps
// commented out at CSfM.cpp:350-382 and asks for in its TODOs, CSfM.cpp:24-27): for every camera the sum and the
// maximum of the squared reprojection distance over its observations.  Camera-major, one warp per chunk.
__global__ void __launch_bounds__(256) frame_stats_kernel(Dev d, const CamRow* __restrict__ rows, const double* __restrict__ pts,
                                                          double* __restrict__ chunk_part) {
  const int lane = threadIdx.x & 31;
  const int wc = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (wc >= d.n_chunks) return;
  const int c = __ldg(d.chunk_cam + wc);
  const int b = __ldg(d.chunk_b + wc), e = __ldg(d.chunk_e + wc);
  const CamRow C = load_row(rows, c);
  const double sp[3] = {1.0, 1.0, 1.0};
  double sum = 0.0, mx = 0.0;
  for (int j = b + lane; j < e; j += 32) {
    const double2 o = __ldg(d.obs_cs + j);
    const int p = __ldg(d.pt_cs + j);
    double r0, r1, jc[12], jp[6];
    reproject_rj(C, __ldg(pts + 3 * (size_t)p), __ldg(pts + 3 * (size_t)p + 1), __ldg(pts + 3 * (size_t)p + 2), o.x, o.y, sp, r0, r1, jc, jp);
    const double d2 = r0 * r0 + r1 * r1;
    sum += d2;
    mx = fmax(mx, d2);
  }
#pragma unroll
  for (int dd = 16; dd > 0; dd >>= 1) {
    sum += __shfl_xor_sync(FULL, sum, dd);
    mx = fmax(mx, __shfl_xor_sync(FULL, mx, dd));
  }
  if (lane == 0) {
    chunk_part[(size_t)wc * 27] = sum;
    chunk_part[(size_t)wc * 27 + 1] = mx;
    chunk_part[(size_t)wc * 27 + 2] = (double)(e - b);
  }
}
// out[c] = sum, out[n_cams + c] = observation count, out[2 n_cams + c] = max of the squared distances of camera c
__global__ void frame_stats_reduce_kernel(int n_cams, const int* __restrict__ cam_chunk_start, const double* __restrict__ chunk_part,
                                          double* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cams) return;
  double sum = 0.0, mx = 0.0, cnt = 0.0;
  for (int w = cam_chunk_start[c]; w < cam_chunk_start[c + 1]; ++w) {
    sum += chunk_part[(size_t)w * 27];
    mx = fmax(mx, chunk_part[(size_t)w * 27 + 1]);
    cnt += chunk_part[(size_t)w * 27 + 2];
  }
  out[c] = sum; out[n_cams + c] = cnt; out[2 * n_cams + c] = mx;
}

// ---- camera-keyed sums through (tile, slot) entries ---------------------------------------------------------
// Every observation has staged K values at sq[k * SV_LD + srank] (srank = its position when the tile is ordered by
// camera slot).  One thread per (slot, k) adds the slot's contiguous run in a fixed order and writes the entry's
// partial: part[(ent_off[tile] + slot) * part_ld + k0 + k].  Call between two barriers.
// s_sb: the tile's slot_begin table in shared memory (stage_slot_begin: ncam entries + the observation count).
__device__ __forceinline__ void stage_slot_begin(const Dev& d, int tile, int ncam, int cnt, unsigned short* s_sb) {
  const unsigned short* sb = d.slot_begin + (size_t)tile * TILE;
  for (int q = threadIdx.x; q <= ncam; q += TILE) s_sb[q] = q < ncam ? __ldg(sb + q) : (unsigned short)cnt;
}
template <int K>
__device__ __forceinline__ void slot_reduce(int e0, int ncam, const unsigned short* s_sb, const double* sq, double* __restrict__ part, int part_ld, int k0) {
  for (int task = threadIdx.x; task < ncam * K; task += TILE) {
    const int sl = task / K, k = task - sl * K;
    const int b = s_sb[sl], e = s_sb[sl + 1];
    const double* src = sq + k * SV_LD;
    double a0 = 0.0, a1 = 0.0;
    int l = b;
    for (; l + 1 < e; l += 2) { a0 += src[l]; a1 += src[l + 1]; }
    if (l < e) a0 += src[l];
    part[(size_t)(e0 + sl) * part_ld + k0 + k] = a0 + a1;
  }
}

// A camera's sum = the sum of its entries' partials, in tile order: one warp per camera, lanes stride the entry list,
// fixed shuffle tree.  out[c][k] = (base ? base[c][k] : 0) + sign * sum (reduced rhs, implicit mat-vec).
template <int K>
__global__ void __launch_bounds__(256) slot_gather_kernel(Dev d, const double* __restrict__ part, const double* __restrict__ base, double sign,
                                                          double* __restrict__ out, const int* __restrict__ skip_flag) {
  // two cameras per CTA, four warps each (a camera has a few hundred entries: one warp alone is latency-bound)
  __shared__ double sp[2][4][K];
  if (skip_flag && *skip_flag) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = warp & 3, half = warp >> 2;
  const int c = blockIdx.x * 2 + half;
  double acc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) acc[k] = 0.0;
  if (c < d.n_cams) {
    const int b = __ldg(d.cs_start + c), e = __ldg(d.cs_start + c + 1);
    for (int q = b + sub * 32 + lane; q < e; q += 128) {
      const double* src = part + (size_t)__ldg(d.cs_ent + q) * K;
#pragma unroll
      for (int k = 0; k < K; ++k) acc[k] += __ldg(src + k);
    }
  }
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double v = acc[k];
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) v += __shfl_xor_sync(FULL, v, dd);
    if (lane == 0) sp[half][sub][k] = v;
  }
  __syncthreads();
  if (sub == 0 && lane < K && c < d.n_cams) {
    const int k = lane;
    const double v = ((sp[half][0][k] + sp[half][1][k]) + sp[half][2][k]) + sp[half][3][k];  // fixed order
    out[(size_t)c * K + k] = (base ? base[(size_t)c * K + k] : 0.0) + sign * v;
  }
}

// Fold the chunk partials of each camera in chunk order (deterministic).
__global__ void cam_reduce_kernel(int n_cams, int nv, const int* __restrict__ cam_chunk_start, const double* __restrict__ chunk_part,
                                  double* __restrict__ out21, double* __restrict__ out6) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_cams * nv) return;
  const int c = idx / nv, k = idx % nv;
  double s = 0.0;
  for (int w = cam_chunk_start[c]; w < cam_chunk_start[c + 1]; ++w) s += chunk_part[(size_t)w * 27 + k];
  if (k < 21) out21[(size_t)c * 21 + k] = s;
  else out6[(size_t)c * 6 + (k - 21)] = s;
}

// A CONSTANT parameter block (hsfm_ba_problem_set_constant; what ceres::Problem::SetParameterBlockConstant does) gets
// column scale 0: its Jacobian columns, gradient, normal-equation blocks and step vanish, the candidate keeps the
// block's value bit for bit, and "scale == 0" is how every other kernel recognises such a block (a free block's
// scale 1/(1+|col|) is never 0).
__global__ void jacobi_scale_kernel(int n_cams, int n_pts, const double* __restrict__ Hcc, const double* __restrict__ Hpp,
                                    double* __restrict__ scale_c, double* __restrict__ scale_p, int enabled,
                                    const unsigned char* __restrict__ cam_fixed, const unsigned char* __restrict__ pt_fixed) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int nc6 = 6 * n_cams;
  if (idx < nc6) {
    const int c = idx / 6, a = idx % 6;
    double v = enabled ? 1.0 / (1.0 + sqrt(Hcc[(size_t)c * 21 + tri6(a, a)])) : 1.0;
    if (cam_fixed && cam_fixed[c]) v = 0.0;
    scale_c[idx] = v;
  } else if (idx < nc6 + 3 * n_pts) {
    const int j = idx - nc6;
    const int p = j / 3, a = j % 3;
    const int diag = (a == 0) ? 0 : (a == 1 ? 3 : 5);
    double v = enabled ? 1.0 / (1.0 + sqrt(Hpp[(size_t)p * 6 + diag])) : 1.0;
    if (pt_fixed && pt_fixed[p]) v = 0.0;
    scale_p[j] = v;
  }
}

// 1/2 sum |r|^2 over the observations whose parameter blocks are ALL constant (by mask or by mode): such residuals
// are not part of the reduced program (ceres Summary::fixed_cost) and never change during a solve.  Once per solve.
__global__ void __launch_bounds__(TILE) fixed_cost_kernel(Dev d, const CamRow* __restrict__ rows, const double* __restrict__ pts,
                                                           const unsigned char* __restrict__ cam_fixed, const unsigned char* __restrict__ pt_fixed,
                                                           int cams_free, int pts_free, double* __restrict__ part) {
  __shared__ double sm[8];
  const int i = blockIdx.x * TILE + threadIdx.x;
  double e = 0.0;
  if (i < d.n_obs) {
    const int c = __ldg(d.cam + i), p = __ldg(d.pt + i);
    if ((!cams_free || cam_fixed[c]) && (!pts_free || pt_fixed[p])) {
      double r0, r1;
      const double2 o = __ldg(d.obs + i);
      reproject_r(load_row(rows, c), pts[3 * (size_t)p], pts[3 * (size_t)p + 1], pts[3 * (size_t)p + 2], o.x, o.y, r0, r1);
      e = r0 * r0 + r1 * r1;
    }
  }
  const double cs = block_sum_256(e, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = cs;
}

// Per attempt: Hpp_inv = (Hpp + D_p^2)^-1 with D = sqrt(clamp(diag)/radius) (a7), z_p = Hpp_inv g_p.
__global__ void point_invert_kernel(int n_pts, const double* __restrict__ Hpp, const double* __restrict__ gp, double radius,
                                    double min_diag, double max_diag, double* __restrict__ Hpp_inv, double* __restrict__ zp,
                                    int* __restrict__ flags) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pts) return;
  double H[6], Hi[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) H[k] = Hpp[6 * (size_t)p + k];
  const int dg[3] = {0, 3, 5};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const double D = sqrt(fmin(fmax(H[dg[a]], min_diag), max_diag) / radius);
    H[dg[a]] += D * D;
  }
  if (!inv3_spd(H, Hi)) {
    atomicOr(flags + F_INVALID, 1);
#pragma unroll
    for (int k = 0; k < 6; ++k) Hi[k] = 0.0;
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) Hpp_inv[6 * (size_t)p + k] = Hi[k];
  double z0, z1, z2;
  sym3_mul(Hi, gp[3 * (size_t)p], gp[3 * (size_t)p + 1], gp[3 * (size_t)p + 2], z0, z1, z2);
  zp[3 * (size_t)p] = z0; zp[3 * (size_t)p + 1] = z1; zp[3 * (size_t)p + 2] = z2;
}

__global__ void cam_diag_kernel(int n_cams, const double* __restrict__ Hcc, double radius, double min_diag, double max_diag,
                                double* __restrict__ Dc2) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 6 * n_cams) return;
  const int c = idx / 6, a = idx % 6;
  const double D = sqrt(fmin(fmax(Hcc[(size_t)c * 21 + tri6(a, a)], min_diag), max_diag) / radius);
  Dc2[idx] = D * D;
}

// Minv[c] = (Hcc + Dc2 - E)^-1 : block-Jacobi preconditioner (diagonal blocks of S), or with
// with_E == 0 the pose-only normal matrix inverse.
__global__ void precond_invert_kernel(int n_cams, const double* __restrict__ Hcc, const double* __restrict__ Dc2,
                                      const double* __restrict__ E, int with_E, double* __restrict__ Minv, int* __restrict__ flags) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cams) return;
  double A[36];
#pragma unroll
  for (int a = 0; a < 6; ++a)
#pragma unroll
    for (int b = a; b < 6; ++b) {
      double v = Hcc[(size_t)c * 21 + tri6(a, b)];
      if (with_E) v -= E[(size_t)c * 21 + tri6(a, b)];
      if (a == b) v += Dc2[6 * c + a];
      A[a * 6 + b] = v;
      A[b * 6 + a] = v;
    }
  if (!inv6_spd(A)) {
    atomicOr(flags + F_INVALID, 2);
#pragma unroll
    for (int k = 0; k < 36; ++k) A[k] = 0.0;
  }
#pragma unroll
  for (int k = 0; k < 36; ++k) Minv[(size_t)c * 36 + k] = A[k];
}

// Streaming pass over the stored linearisation, point-major tiles.
//   MODE 0 (RHS)    : rhs[c] += Jc^T (-Jp z_p)                      (a7: g_c - W Hpp^-1 g_p; rhs preloaded with g_c)
//   MODE 1 (MATVEC) : q[c]   += Jc^T (I - Jp Hpp^-1 Jp^T) Jc p      (implicit S p without the D_c^2 p term)
//   MODE 2 (BACKSUB): y_p = z_p - Hpp^-1 W^T y_c ; candidate point ; model cost change ; norms   (a8, a9)
// Algorithmic bytes / observation: Jc 96 + Jp 48 + cam 4 + pt 4 (+ r 16 in MODE 2).
#ifndef SCHUR_MATVEC_CTAS
#define SCHUR_MATVEC_CTAS 4
#endif
template <int MODE>
__global__ void __launch_bounds__(TILE, MODE == 0 ? 8 : (MODE == 1 ? SCHUR_MATVEC_CTAS : 4)) schur_pass_kernel(Dev d, const double* __restrict__ Hpp_inv, const double* __restrict__ zp,
                                                          const double* __restrict__ vec_in /* p or y_c */, double* __restrict__ vec_out,
                                                          const double* __restrict__ pts_x, const double* __restrict__ scale_p,
                                                          double* __restrict__ yp, double* __restrict__ pts_c,
                                                          double* __restrict__ part, int part_stride, const int* __restrict__ flags,
                                                          int cams_free, int pts_free) {
  // MODE 0/1: sq = the observations' camera-space 6-vectors, in slot order.  In MODE 1 it shares its storage with sP (the
  // staged vec_in blocks are dead once u = Jc p has been formed, two barriers before sq is written).
  __shared__ __align__(16) double sq_sp[MODE == 2 ? 6 * TILE : 6 * SV_LD];
  double* const sq = sq_sp;
  double2* const sP = reinterpret_cast<double2*>(sq_sp);  // MODE 1/2: vec_in blocks of the tile's cameras, by slot [3 * TILE]
  __shared__ unsigned short s_sb[MODE == 2 ? 2 : TILE + 1];
  __shared__ double sv[MODE == 0 ? 1 : 3 * SV_LD];  // per-observation W^T(.) pieces
  __shared__ double sw[MODE == 0 ? 1 : 3 * TILE];   // per-point result handed back to the observations
  __shared__ double sm[8];
  if (MODE == 1) {
    if (flags[F_PCG_DONE]) return;
  }
  const TileHdr t = tile_hdr(d);
  const int i = t.b + (int)threadIdx.x;
  const bool valid = i < t.e;
  const size_t n = (size_t)d.n_obs;
  const bool use_vec = MODE == 1 || (MODE == 2 && cams_free);
  int e0 = 0;
  if (MODE != 2) {
    e0 = __ldg(d.ent_off + blockIdx.x);
    stage_slot_begin(d, (int)blockIdx.x, t.ncam, t.e - t.b, s_sb);
  }
  if (MODE != 0 && use_vec) {
    const int* list = d.tile_cams + (size_t)blockIdx.x * TILE;
    for (int q = threadIdx.x; q < 3 * t.ncam; q += TILE) {
      const int sl = q / 3, piece = q - sl * 3;
      sP[q] = __ldg(reinterpret_cast<const double2*>(vec_in + 6 * (size_t)__ldg(list + sl)) + piece);
    }
  }
  double jc[12], jp[6];
  int slot = 0, pl = 0, rk = 0;
  if (valid) {
    if (MODE != 0) slot = __ldg(d.slot + i);
    if (MODE != 2) rk = __ldg(d.srank + i);
    pl = __ldg(d.pt + i) - t.p_first;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const double2 v = d.Jc[k * n + i];
      jc[k] = v.x; jc[6 + k] = v.y;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double2 v = d.Jp[k * n + i];
      jp[k] = v.x; jp[3 + k] = v.y;
    }
  }
  if (MODE == 0) {
    // t_i = Jc_i^T (Jp_i z_p): summed per camera slot of the tile in shared memory, one partial per (tile, slot) entry goes
    // out; slot_gather_kernel<6> adds a camera's entries — no atomics, fixed order, no per-observation round trip
    if (valid) {
      const size_t p = (size_t)(t.p_first + pl);
      const double z0 = __ldg(zp + 3 * p), z1 = __ldg(zp + 3 * p + 1), z2 = __ldg(zp + 3 * p + 2);
      const double y0 = jp[0] * z0 + jp[1] * z1 + jp[2] * z2;
      const double y1 = jp[3] * z0 + jp[4] * z1 + jp[5] * z2;
#pragma unroll
      for (int k = 0; k < 6; ++k) sq[k * SV_LD + rk] = jc[k] * y0 + jc[6 + k] * y1;
    }
    __syncthreads();
    slot_reduce<6>(e0, t.ncam, s_sb, sq, vec_out, 6, 0);
    return;
  }
  // the per-point thread's inputs depend on the tile header only: fetch them before the barrier
  const bool own_pt = (int)threadIdx.x < t.n_p;
  const int pp = t.p_first + (int)threadIdx.x;
  int sb = 0, se = 0;
  double Hi[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  double z0 = 0.0, z1 = 0.0, z2 = 0.0;
  if (own_pt) {
    sb = __ldg(d.pt_start + pp) - t.b; se = __ldg(d.pt_start + pp + 1) - t.b;
#pragma unroll
    for (int k = 0; k < 6; ++k) Hi[k] = __ldg(Hpp_inv + 6 * (size_t)pp + k);
    if (MODE == 2 && pts_free) { z0 = __ldg(zp + 3 * (size_t)pp); z1 = __ldg(zp + 3 * (size_t)pp + 1); z2 = __ldg(zp + 3 * (size_t)pp + 2); }
  }
  __syncthreads();
  // u = Jc * vec_in[c];  this observation's piece of W^T vec_in goes to shared memory
  double u0 = 0.0, u1 = 0.0;
  if (valid) {
    if (use_vec) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double2 pk = sP[3 * slot + k];
        u0 += jc[2 * k] * pk.x + jc[2 * k + 1] * pk.y;
        u1 += jc[6 + 2 * k] * pk.x + jc[7 + 2 * k] * pk.y;
      }
    }
    sv[threadIdx.x] = jp[0] * u0 + jp[3] * u1;
    sv[SV_LD + threadIdx.x] = jp[1] * u0 + jp[4] * u1;
    sv[2 * SV_LD + threadIdx.x] = jp[2] * u0 + jp[5] * u1;
  }
  __syncthreads();
  // one thread per point of the tile: total over the point's observations (in observation order), then the
  // 3x3 point-block work — Hpp^-1 W^T(.) for the mat-vec; y_p, the candidate point and the norms for MODE 2
  double dn = 0.0, xn = 0.0;
  if (own_pt) {
    double v0 = 0.0, v1 = 0.0, v2 = 0.0;
    for (int l = sb; l < se; ++l) { v0 += sv[l]; v1 += sv[SV_LD + l]; v2 += sv[2 * SV_LD + l]; }
    double w0, w1, w2;
    sym3_mul(Hi, v0, v1, v2, w0, w1, w2);  // Hpp^-1 W^T (.)
    if (MODE == 2) {
      const size_t q = 3 * (size_t)pp;
      double y0 = 0.0, y1 = 0.0, y2 = 0.0;  // y_p
      if (pts_free && se > sb) {
        y0 = z0 - w0; y1 = z1 - w1; y2 = z2 - w2;
        yp[q] = y0; yp[q + 1] = y1; yp[q + 2] = y2;
        const double x0 = pts_x[q], x1 = pts_x[q + 1], x2 = pts_x[q + 2];
        const double c0 = x0 + (-y0) * scale_p[q], c1 = x1 + (-y1) * scale_p[q + 1], c2 = x2 + (-y2) * scale_p[q + 2];
        pts_c[q] = c0; pts_c[q + 1] = c1; pts_c[q + 2] = c2;
        const double e0 = x0 - c0, e1 = x1 - c1, e2 = x2 - c2;
        dn = e0 * e0 + e1 * e1 + e2 * e2;
        if (scale_p[q] != 0.0) xn = x0 * x0 + x1 * x1 + x2 * x2;  // |x| runs over the free blocks only
      }
      w0 = y0; w1 = y1; w2 = y2;
    }
    sw[3 * threadIdx.x] = w0; sw[3 * threadIdx.x + 1] = w1; sw[3 * threadIdx.x + 2] = w2;
  }
  __syncthreads();
  double w0 = 0.0, w1 = 0.0, w2 = 0.0;
  if (valid) { w0 = sw[3 * pl]; w1 = sw[3 * pl + 1]; w2 = sw[3 * pl + 2]; }
  if (MODE == 1) {
    // t_i = Jc_i^T (I - Jp Hpp^-1 Jp^T)(Jc p) per observation, summed per camera slot of the tile (see MODE 0)
    if (valid) {
      const double y0 = u0 - (jp[0] * w0 + jp[1] * w1 + jp[2] * w2);
      const double y1 = u1 - (jp[3] * w0 + jp[4] * w1 + jp[5] * w2);
#pragma unroll
      for (int k = 0; k < 6; ++k) sq[k * SV_LD + rk] = jc[k] * y0 + jc[6 + k] * y1;
    }
    __syncthreads();
    slot_reduce<6>(e0, t.ncam, s_sb, sq, vec_out, 6, 0);
    return;
  }
  // MODE 2: model cost change of step = -y;  m = J step
  double mcc = 0.0;
  if (valid) {
    const double m0 = -(u0 + jp[0] * w0 + jp[1] * w1 + jp[2] * w2);
    const double m1 = -(u1 + jp[3] * w0 + jp[4] * w1 + jp[5] * w2);
    const double2 r = d.r[i];
    mcc = -(m0 * (r.x + m0 / 2.0) + m1 * (r.y + m1 / 2.0));
  }
  const double s0 = block_sum_256(mcc, sm);
  const double s1 = block_sum_256(dn, sm);
  const double s2 = block_sum_256(xn, sm);
  if (threadIdx.x == 0) {
    part[P_MCC * (size_t)part_stride + blockIdx.x] = s0;
    part[P_DN_P * (size_t)part_stride + blockIdx.x] = s1;
    part[P_XN_P * (size_t)part_stride + blockIdx.x] = s2;
  }
}

// Candidate cameras: cand = x + (-y_c) * scale ; |delta|^2, |x|^2 partials over cameras that have observations.
__global__ void __launch_bounds__(256) cam_step_kernel(int n6, const double* __restrict__ x, const double* __restrict__ yc,
                                                       const double* __restrict__ scale_c, const int* __restrict__ cam_nobs,
                                                       double* __restrict__ cand, double* __restrict__ part, int part_stride,
                                                       int cams_free, int* __restrict__ flags) {
  __shared__ double sm[8];
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  double dn = 0.0, xn = 0.0;
  if (k < n6) {
    const double xv = x[k];
    double cv = xv;
    if (cams_free && cam_nobs[k / 6] > 0) {
      const double y = yc[k];
      if (!isfinite(y)) atomicOr(flags + F_INVALID, 4);
      cv = xv + (-y) * scale_c[k];
      const double e = xv - cv;
      dn = e * e;
      if (scale_c[k] != 0.0) xn = xv * xv;  // |x| runs over the free blocks only
    }
    cand[k] = cv;
  }
  const double s0 = block_sum_256(dn, sm);
  const double s1 = block_sum_256(xn, sm);
  if (threadIdx.x == 0) {
    part[P_DN_C * (size_t)part_stride + blockIdx.x] = s0;
    part[P_XN_C * (size_t)part_stride + blockIdx.x] = s1;
  }
}

// |x|^2 of the point block over points that have observations (used at linearisation).
__global__ void __launch_bounds__(256) block_solve_cams_kernel(int n_cams, const double* __restrict__ Minv,
                                                               const double* __restrict__ gc, double* __restrict__ yc) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= 6 * n_cams) return;
  const int c = k / 6, a = k % 6;
  double s = 0.0;
#pragma unroll
  for (int b = 0; b < 6; ++b) s += Minv[(size_t)c * 36 + a * 6 + b] * gc[6 * c + b];
  yc[k] = s;
}

// max_k |g_k / scale_k| partials (gradient of the unscaled problem, a6/a9).
__global__ void __launch_bounds__(256) grad_norm_kernel(int n, const double* __restrict__ g, const double* __restrict__ scale,
                                                        double* __restrict__ part_out) {
  __shared__ double sm[8];
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  double v = 0.0;
  if (k < n && scale[k] != 0.0) v = fabs(g[k] / scale[k]);  // constant blocks (scale 0) have no gradient
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v = fmax(v, __shfl_down_sync(FULL, v, d));
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double m = 0.0;
    for (int w = 0; w < 8; ++w) m = fmax(m, sm[w]);
    part_out[blockIdx.x] = m;
  }
}

struct RedTasks {
  const double* in[8];
  double* out[8];
  int n[8];
  int op[8];  // 0 sum, 1 max, 2 sum then *0.5
};
__global__ void __launch_bounds__(1024) reduce_tasks_kernel(RedTasks T) {
  __shared__ double sm[32];
  const int task = blockIdx.x;
  const double* in = T.in[task];
  const int n = T.n[task], op = T.op[task];
  double v = 0.0;
  if (op == 1) {
    for (int k = threadIdx.x; k < n; k += 1024) v = fmax(v, in[k]);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = fmax(v, __shfl_down_sync(FULL, v, d));
  } else {
    for (int k = threadIdx.x; k < n; k += 1024) v += in[k];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(FULL, v, d);
  }
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 32; ++w) s = (op == 1) ? fmax(s, sm[w]) : s + sm[w];
    if (op == 2) s *= 0.5;
    *T.out[task] = s;
  }
}

// ---- block-Jacobi PCG on the reduced camera system, vectors of 6 n_cams, one CTA -------------
__device__ __forceinline__ double block_sum_1024(double v, double* sm) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(FULL, v, d);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x < 32) {
    s = sm[threadIdx.x];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(FULL, s, d);
  }
  if (threadIdx.x == 0) sm[32] = s;
  __syncthreads();
  s = sm[32];
  __syncthreads();
  return s;
}

__global__ void __launch_bounds__(1024) pcg_init_kernel(int n, const double* __restrict__ rhs, const double* __restrict__ Minv,
                                                        double* x, double* r, double* z, double* p, double* q, double* scal, int* flags) {
  __shared__ double sm[33];
  double loc = 0.0;
  for (int k = threadIdx.x; k < n; k += 1024) {
    const int c = k / 6, a = k % 6;
    double s = 0.0;
#pragma unroll
    for (int b = 0; b < 6; ++b) s += Minv[(size_t)c * 36 + a * 6 + b] * rhs[6 * c + b];
    x[k] = 0.0; r[k] = rhs[k]; z[k] = s; p[k] = s; q[k] = 0.0;
    loc += rhs[k] * s;
  }
  const double rho = block_sum_1024(loc, sm);
  if (threadIdx.x == 0) {
    scal[S_PCG_RHO0] = rho;
    scal[S_PCG_RHO] = rho;
    flags[F_PCG_ITERS] = 0;
    flags[F_PCG_DONE] = (rho > 0.0 && isfinite(rho)) ? 0 : 1;
  }
}

__global__ void __launch_bounds__(1024) pcg_step_kernel(int n, const double* __restrict__ Minv, const double* __restrict__ Dc2,
                                                        double* x, double* r, double* z, double* p, double* q, double* scal, int* flags,
                                                        double tol2, int max_it) {
  __shared__ double sm[33];
  if (flags[F_PCG_DONE]) return;
  const double rho = scal[S_PCG_RHO], rho0 = scal[S_PCG_RHO0];
  double loc = 0.0;
  for (int k = threadIdx.x; k < n; k += 1024) {
    const double pk = p[k];
    const double qk = q[k] + Dc2[k] * pk;
    q[k] = qk;
    loc += pk * qk;
  }
  const double pq = block_sum_1024(loc, sm);
  if (!(pq > 0.0) || !isfinite(pq)) {
    if (threadIdx.x == 0) flags[F_PCG_DONE] = 2;
    return;
  }
  const double alpha = rho / pq;
  for (int k = threadIdx.x; k < n; k += 1024) {
    x[k] += alpha * p[k];
    r[k] -= alpha * q[k];
  }
  __syncthreads();
  loc = 0.0;
  for (int k = threadIdx.x; k < n; k += 1024) {
    const int c = k / 6, a = k % 6;
    double s = 0.0;
#pragma unroll
    for (int b = 0; b < 6; ++b) s += Minv[(size_t)c * 36 + a * 6 + b] * r[6 * c + b];
    z[k] = s;
    loc += r[k] * s;
  }
  const double rho_new = block_sum_1024(loc, sm);
  const int it = flags[F_PCG_ITERS] + 1;
  __syncthreads();
  const bool done = (rho_new <= tol2 * rho0) || (it >= max_it) || !isfinite(rho_new);
  const double beta = rho_new / rho;
  for (int k = threadIdx.x; k < n; k += 1024) {
    p[k] = z[k] + beta * p[k];
    q[k] = 0.0;
  }
  if (threadIdx.x == 0) {
    flags[F_PCG_ITERS] = it;
    scal[S_PCG_RHO] = rho_new;
    if (done) flags[F_PCG_DONE] = 1;
  }
}

// ---- explicit dense S (DENSE_SCHUR, a7/a8) ---------------------------------------------------
__global__ void dense_init_kernel(int n_cams, const double* __restrict__ Hcc, const double* __restrict__ Dc2, double* __restrict__ S) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_cams * 36) return;
  const int c = idx / 36, a = (idx % 36) / 6, b = idx % 6;
  const int lo = min(a, b), hi = max(a, b);
  double v = Hcc[(size_t)c * 21 + tri6(lo, hi)];
  if (a == b) v += Dc2[6 * c + a];
  const size_t n = 6 * (size_t)n_cams;
  S[(6 * (size_t)c + a) * n + 6 * (size_t)c + b] = v;
}

// S[ci,cj] -= W_i Hpp^-1 W_j^T for every pair (i,j) of observations of a point with cam(i) >= cam(j):
// lower block triangle (a7).  Phase 1: one thread per observation stages W_i = Jc_i^T Jp_i (6x3) and
// A_i = W_i Hpp^-1 in shared memory.  Phase 2: the tile's pairs are numbered by a prefix sum and dealt to
// the warps eight at a time; consecutive lanes take consecutive block elements, so every RED instruction carries whole
// 6-double rows of a 6x6 block (48 contiguous bytes) instead of 32 unrelated addresses — measured 2.2-2.8x
// the FP64 atomic rate of the one-thread-per-block scatter (scripts/micro/fp64_peaks.cu).  Observations of a
// point are stored with ascending camera (ba_index.cu), so pair (i,j), j <= i, lands in the lower triangle.
constexpr int kDenseSchurSmem = TILE * 36 * (int)sizeof(double) + (3 * TILE + 8) * (int)sizeof(int) + 8 * 8 * 36 * (int)sizeof(double) + 8 * 8 * (int)sizeof(long long);
// The sorted path runs 16 warps on a tile (512 threads; the first 256 stage the tile's observations), each with four
// staging slots: twice the warps per SM at the same shared-memory footprint — the pair loop is latency-bound.
// tile0: first tile of the launch.
template <bool SORTED>
__device__ __forceinline__ void dense_schur_tile(const Dev& d, const double* __restrict__ Hpp_inv, double* __restrict__ S, const int tile) {
  extern __shared__ __align__(16) unsigned char ds_raw[];
  double (*sA)[18] = reinterpret_cast<double (*)[18]>(ds_raw);
  double (*sW)[18] = reinterpret_cast<double (*)[18]>(ds_raw + TILE * 18 * sizeof(double));
  int* sCam = reinterpret_cast<int*>(ds_raw + TILE * 36 * sizeof(double));
  int* sOff = sCam + TILE;      // exclusive pair offsets per observation
  int* sSeg = sOff + TILE;      // tile-local index of the first observation of the point
  int* sWarpTot = sSeg + TILE;  // [8]
  const TileThread t = tile_thread(d, tile);
  const size_t n = (size_t)d.n_obs;
  if (t.valid) {
    double jc[12], jp[6], W[18], Hi[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const double2 v = d.Jc[k * n + t.i];
      jc[k] = v.x; jc[6 + k] = v.y;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double2 v = d.Jp[k * n + t.i];
      jp[k] = v.x; jp[3 + k] = v.y;
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) Hi[k] = __ldg(Hpp_inv + 6 * (size_t)t.p + k);
#pragma unroll
    for (int a = 0; a < 6; ++a) {
#pragma unroll
      for (int b = 0; b < 3; ++b) W[a * 3 + b] = jc[a] * jp[b] + jc[6 + a] * jp[3 + b];
      double a0, a1, a2;
      sym3_mul(Hi, W[a * 3], W[a * 3 + 1], W[a * 3 + 2], a0, a1, a2);
      if (SORTED) { a0 = -a0; a1 = -a1; a2 = -a2; }  // the sorted path accumulates -A W^T: what is added to S
      sA[threadIdx.x][a * 3] = a0; sA[threadIdx.x][a * 3 + 1] = a1; sA[threadIdx.x][a * 3 + 2] = a2;
      sW[threadIdx.x][a * 3] = W[a * 3]; sW[threadIdx.x][a * 3 + 1] = W[a * 3 + 1]; sW[threadIdx.x][a * 3 + 2] = W[a * 3 + 2];
    }
    sCam[threadIdx.x] = __ldg(d.cam + t.i);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int total_pairs = 0;
  if (!SORTED) {
    // pairs owned by observation i: (i, j) for j = segment start .. i  →  exclusive scan over the tile
    const int cnt = t.valid ? (t.i - t.seg_b + 1) : 0;
    int incl = cnt;
#pragma unroll
    for (int dd = 1; dd < 32; dd <<= 1) {
      const int up = __shfl_up_sync(FULL, incl, dd);
      if (lane >= dd) incl += up;
    }
    if (lane == 31) sWarpTot[warp] = incl;
    sSeg[threadIdx.x] = t.valid ? (t.seg_b - t.tile_b) : 0;
    __syncthreads();
    int warp_off = 0;
#pragma unroll
    for (int w = 0; w < TILE / 32; ++w) {
      const int v = sWarpTot[w];
      if (w < warp) warp_off += v;
      total_pairs += v;
    }
    sOff[threadIdx.x] = warp_off + incl - cnt;
  }
  __syncthreads();
  const size_t ld = 6 * (size_t)d.n_cams;
  // Each 6x6 block A_i W_j^T is ONE FP64 tensor-core instruction (mma m8n8k4: rows = the six camera parameters
  // of i, columns = those of j, k = the three point coordinates; two LDS.64 per lane instead of six per
  // element).  The accumulator fragments go through a warp-private 8 x 36 staging buffer so that the reduction
  // into S keeps its efficient shape: nine full-warp passes in which consecutive lanes take consecutive block
  // elements (whole 48-byte rows per RED instruction).
  constexpr int SLOTS = SORTED ? 4 : 8, NW = SORTED ? 2 * TILE / 32 : TILE / 32, NPASS = (SLOTS * 36 + 31) / 32;
  double* sC = reinterpret_cast<double*>(sWarpTot + 8) + warp * (SLOTS * 36);                // [SLOTS][36] per warp
  long long* sBase = reinterpret_cast<long long*>(reinterpret_cast<double*>(sWarpTot + 8) + 8 * 8 * 36) + warp * SLOTS;  // [SLOTS] per warp
  const int fr = lane >> 2, fk = lane & 3;
  const bool frag_ok = fr < 6 && fk < 3;
  // which pair and which element of its block this lane reduces in each of the nine passes: lane constants
  int pass_pl[NPASS];
  size_t pass_off[NPASS], pass_offT[NPASS];
#pragma unroll
  for (int pass = 0; pass < NPASS; ++pass) {
    const int g = pass * 32 + lane, pl = g / 36, e = g - pl * 36, a = e / 6, b = e - a * 6;
    pass_pl[pass] = pl;
    pass_off[pass] = (size_t)a * ld + b;
    pass_offT[pass] = (size_t)b * ld + a;
  }
  // element (a, b) of the staged block goes to S[6 ci + a][6 cj + b] (ci < cj cannot happen with cameras ascending
  // inside a point, but then the transposed element is addressed)
  auto block_base = [&](int i, int j) -> long long {
    const int ci = sCam[i], cj = sCam[j];
    return (ci >= cj) ? (long long)((6 * (size_t)ci) * ld + 6 * (size_t)cj) : -1 - (long long)((6 * (size_t)cj) * ld + 6 * (size_t)ci);
  };
  auto reduce_staged = [&](int cnt) {
    __syncwarp();
#pragma unroll
    for (int pass = 0; pass < NPASS; ++pass) {
      if (pass_pl[pass] < cnt) {
        const long long bs = sBase[pass_pl[pass]];
        double* dst = (bs >= 0) ? S + bs + pass_off[pass] : S + (-1 - bs) + pass_offT[pass];
        atomicAdd(dst, -sC[pass * 32 + lane]);
      }
    }
    __syncwarp();
  };
  if constexpr (SORTED) {
    // The tile's pairs come from the static list SORTED BY DESTINATION BLOCK (ba_index / tile_pairs_kernel): a
    // warp walks a contiguous piece of the list and keeps accumulating into the same tensor-core fragment while
    // the destination stays the same, so the ~2.1 pairs per distinct block of a tile cost one reduction, not 2.1.
    const int p_off = __ldg(d.pair_off + tile);
    const int n_pairs = __ldg(d.pair_off + tile + 1) - p_off;
    const unsigned* __restrict__ list = d.pair_list + p_off;
    const int chunk = (n_pairs + NW - 1) / NW;
    const int beg = warp * chunk, end = min(n_pairs, beg + chunk);
    int nbuf = 0, li = 0, lj = 0;
    unsigned cur = 0x10000u;  // no destination yet
    bool diag_dest = false;
    double c0 = 0.0, c1 = 0.0;
    // The pair loop is instruction-issue-bound (ncu: 64 % issue-active, no dominant stall), so the hand-over of a
    // finished block is kept short: the base offset is one 32x32->64-bit multiply-add computed by every lane (no
    // divergent single-lane section), offsets inside a block are 32-bit lane constants, and the staged reduction
    // assumes ci >= cj — what the index guarantees (cameras ascend inside a point); a block with ci < cj would go
    // out transposed, element by element, straight from the fragment registers.
    const unsigned ld32 = (unsigned)ld;
    unsigned off32[NPASS];
#pragma unroll
    for (int pass = 0; pass < NPASS; ++pass) off32[pass] = (unsigned)pass_off[pass];
    unsigned long long* sBaseU = reinterpret_cast<unsigned long long*>(sBase);
    auto reduce_sorted = [&](int cnt) {
      __syncwarp();
#pragma unroll
      for (int pass = 0; pass < NPASS; ++pass) {
        if (pass_pl[pass] < cnt) atomicAdd(S + (sBaseU[pass_pl[pass]] + off32[pass]), sC[pass * 32 + lane]);
      }
      __syncwarp();
    };
    auto flush = [&]() {
      const unsigned ci = (unsigned)sCam[li], cj = (unsigned)sCam[lj];
      if (ci >= cj) {  // warp-uniform
        if (frag_ok) *reinterpret_cast<double2*>(sC + nbuf * 36 + fr * 6 + 2 * fk) = make_double2(c0, c1);
        const unsigned long long bs = (unsigned long long)(6u * ci) * ld32 + 6u * cj;
        if (lane == 0) sBaseU[nbuf] = bs;
        if (++nbuf == SLOTS) { reduce_sorted(SLOTS); nbuf = 0; }
      } else if (frag_ok) {
        double* dst = S + (size_t)(6u * cj + 2 * fk) * ld + 6u * ci + fr;
        atomicAdd(dst, c0);
        atomicAdd(dst + ld, c1);
      }
    };
    for (int base = beg; base < end; base += 32) {
      const unsigned key = (base + lane < end) ? __ldg(list + base + lane) : 0xffffffffu;
      const int cnt = min(32, end - base);
#pragma unroll 4
      for (int pl = 0; pl < cnt; ++pl) {
        const unsigned k = __shfl_sync(FULL, key, pl);
        const unsigned dest = k >> 16;
        const int i = (k >> 8) & 0xff, j = k & 0xff;
        const double a_frag = frag_ok ? sA[i][fr * 3 + fk] : 0.0;
        const double b_frag = frag_ok ? sW[j][fr * 3 + fk] : 0.0;
        if (dest != cur) {  // warp-uniform
          if (cur != 0x10000u) flush();
          cur = dest; c0 = 0.0; c1 = 0.0;
          diag_dest = (dest >> 8) == (dest & 0xff);
        }
        li = i; lj = j;
        dmma884(c0, c1, a_frag, b_frag);
        if (diag_dest && i != j) {  // a camera listed twice in a track: diagonal block, M + M^T
          const double at = frag_ok ? sW[j][fr * 3 + fk] : 0.0, bt = frag_ok ? sA[i][fr * 3 + fk] : 0.0;
          dmma884(c0, c1, at, bt);
        }
      }
    }
    if (cur != 0x10000u) flush();
    if (nbuf) reduce_sorted(nbuf);
  } else {
  // Unsorted fallback (no pair list): every warp takes 8 pairs (288 block elements) at a time, round-robin over the
  // tile, so the work of a long track is spread over all warps.
  for (int base = warp * 8; base < total_pairs; base += (TILE / 32) * 8) {
    int packed = -1;
    {
      const int q = base + (lane & 7);
      if (q < total_pairs) {
        int lo = 0, hi = TILE - 1;  // largest i with sOff[i] <= q (ties: the last such i owns pairs)
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int mid = (lo + hi + 1) >> 1;
          if (sOff[mid] <= q) lo = mid; else hi = mid - 1;
        }
        const int i = lo, j = sSeg[i] + (q - sOff[i]);
        packed = (i << 16) | j;
        if (lane < 8) sBase[lane] = block_base(i, j);
      }
    }
    const int cnt = min(8, total_pairs - base);
    // all eight pairs in flight: operand loads first, then the tensor-core products, then the staging stores
    double af[8], bf[8];
    bool dup[8];
#pragma unroll
    for (int pl = 0; pl < 8; ++pl) {
      const int ij = __shfl_sync(FULL, packed, pl);
      const int i = (ij >> 16) & 0xff, j = ij & 0xff;  // packed = -1 past the end: any valid row, result unused
      af[pl] = frag_ok ? sA[i][fr * 3 + fk] : 0.0;
      bf[pl] = frag_ok ? sW[j][fr * 3 + fk] : 0.0;
      dup[pl] = ij >= 0 && i != j && sCam[i] == sCam[j];
    }
#pragma unroll
    for (int pl = 0; pl < 8; ++pl) {
      double c0 = 0.0, c1 = 0.0;
      dmma884(c0, c1, af[pl], bf[pl]);  // (A_i W_j^T)[fr][2 fk], [2 fk + 1]
      if (dup[pl]) {                     // a camera listed twice in a track: the block is diagonal, it needs M + M^T
        const int ij = __shfl_sync(FULL, packed, pl);
        const int i = ij >> 16, j = ij & 0xffff;
        double at = 0.0, bt = 0.0;
        if (frag_ok) { at = sW[j][fr * 3 + fk]; bt = sA[i][fr * 3 + fk]; }
        dmma884(c0, c1, at, bt);
      }
      if (frag_ok) *reinterpret_cast<double2*>(sC + pl * 36 + fr * 6 + 2 * fk) = make_double2(c0, c1);
    }
    reduce_staged(cnt);
  }
  }
}

template <bool SORTED>
__global__ void __launch_bounds__(SORTED ? 2 * TILE : TILE, 2) dense_schur_kernel(Dev d, const double* __restrict__ Hpp_inv, double* __restrict__ S, int tile0) {
  dense_schur_tile<SORTED>(d, Hpp_inv, S, (int)blockIdx.x + tile0);
}

// Static pair lists of dense_schur (one CTA per tile, once per problem): every (i, j) pair of observations of one
// point, j <= i, keyed by its destination block (slot of camera i, slot of camera j) and sorted by that key with
// a bitonic network in shared memory.  Tiles with more pairs than the network holds keep generation order (the
// consumer only loses the merging of equal destinations there).
constexpr int PAIR_SORT_CAP = 4096;
// The pairs are generated in ascending (i, j) order, so sorting them by destination = a STABLE sort on the two slot
// bytes: two passes of a counting sort (slot of j, then slot of i).  Every warp ranks a contiguous piece of the list,
// 32 consecutive pairs per step (__match_any_sync: position among the lanes with the same digit + a warp-private
// running count per digit), the per-warp counts are prefix-summed over digits and warps, and the pairs are scattered.
// (The bitonic network this replaces needed 66 compare-exchange passes over up to 4096 keys: 1.17 ms at Venice size.)
__global__ void __launch_bounds__(TILE) tile_pairs_kernel(Dev d, unsigned* __restrict__ list) {
  __shared__ unsigned sk[2][PAIR_SORT_CAP];
  __shared__ unsigned short sRank[PAIR_SORT_CAP];
  __shared__ unsigned short sHist[TILE / 32][256];
  __shared__ int sOff[TILE], sSeg[TILE], sWarpTot[TILE / 32];
  __shared__ unsigned char sSlot[TILE];
  const TileThread t = tile_thread(d);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cnt = t.valid ? (t.i - t.seg_b + 1) : 0;
  int incl = cnt;
#pragma unroll
  for (int dd = 1; dd < 32; dd <<= 1) {
    const int up = __shfl_up_sync(FULL, incl, dd);
    if (lane >= dd) incl += up;
  }
  if (lane == 31) sWarpTot[warp] = incl;
  sSeg[threadIdx.x] = t.valid ? (t.seg_b - t.tile_b) : 0;
  sSlot[threadIdx.x] = t.valid ? d.slot[t.i] : 0;
  __syncthreads();
  int warp_off = 0;
#pragma unroll
  for (int w = 0; w < TILE / 32; ++w)
    if (w < warp) warp_off += sWarpTot[w];
  sOff[threadIdx.x] = warp_off + incl - cnt;
  __syncthreads();
  const int p_off = d.pair_off[blockIdx.x];
  const int total = d.pair_off[blockIdx.x + 1] - p_off;
  const bool sorted = total <= PAIR_SORT_CAP;
  for (int q = threadIdx.x; q < total; q += TILE) {
    int lo = 0, hi = TILE - 1;  // largest i with sOff[i] <= q
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int mid = (lo + hi + 1) >> 1;
      if (sOff[mid] <= q) lo = mid; else hi = mid - 1;
    }
    const int i = lo, j = sSeg[i] + (q - sOff[i]);
    const unsigned key = ((unsigned)sSlot[i] << 24) | ((unsigned)sSlot[j] << 16) | ((unsigned)i << 8) | (unsigned)j;
    if (sorted) sk[0][q] = key; else list[p_off + q] = key;
  }
  if (!sorted) return;
  const int seg = ((total + TILE - 1) / TILE) * 32;  // pairs per warp: a multiple of 32, the warps' pieces in list order
  const int w0 = min(total, warp * seg), w1 = min(total, w0 + seg);
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
    const int shift = pass ? 24 : 16;
    const unsigned* src = sk[pass];
    unsigned* dst = sk[pass ^ 1];
    for (int e = threadIdx.x; e < (TILE / 32) * 256; e += TILE) (&sHist[0][0])[e] = 0;
    __syncthreads();  // also: the keys of this pass are in place
    for (int base = w0; base < w1; base += 32) {
      const int idx = base + lane;
      const bool valid = idx < w1;
      const int dg = valid ? (int)((src[idx] >> shift) & 0xffu) : 0x100 + lane;  // idle lanes: a digit of their own
      const unsigned peers = __match_any_sync(FULL, dg);
      const int leader = __ffs(peers) - 1;
      unsigned old = 0;
      if (valid && lane == leader) {
        old = sHist[warp][dg];
        sHist[warp][dg] = (unsigned short)(old + __popc(peers));
      }
      old = __shfl_sync(FULL, old, leader);
      if (valid) sRank[idx] = (unsigned short)(old + __popc(peers & ((1u << lane) - 1u)));
      __syncwarp();
    }
    __syncthreads();
    {
      // thread = digit: where the pairs of (warp, digit) start = all smaller digits + the same digit in earlier warps
      int tot = 0;
      unsigned short pre[TILE / 32];
#pragma unroll
      for (int w = 0; w < TILE / 32; ++w) { pre[w] = (unsigned short)tot; tot += sHist[w][threadIdx.x]; }
      int inc = tot;
#pragma unroll
      for (int dd = 1; dd < 32; dd <<= 1) {
        const int up = __shfl_up_sync(FULL, inc, dd);
        if (lane >= dd) inc += up;
      }
      if (lane == 31) sWarpTot[warp] = inc;
      __syncthreads();
      int before = inc - tot;
#pragma unroll
      for (int w = 0; w < TILE / 32; ++w)
        if (w < warp) before += sWarpTot[w];
#pragma unroll
      for (int w = 0; w < TILE / 32; ++w) sHist[w][threadIdx.x] = (unsigned short)(before + pre[w]);
    }
    __syncthreads();
    for (int idx = w0 + lane; idx < w1; idx += 32) {
      const unsigned k = src[idx];
      dst[sHist[warp][(k >> shift) & 0xffu] + sRank[idx]] = k;
    }
    __syncthreads();
  }
  for (int q = threadIdx.x; q < total; q += TILE) list[p_off + q] = sk[0][q];
}

__global__ void mirror_lower_kernel(int n, double* __restrict__ S) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (j < n && j > i) S[(size_t)i * n + j] = S[(size_t)j * n + i];
}

// ---- parity taps -----------------------------------------------------------------------------
__global__ void scatter_eval_kernel(Dev d, const int* __restrict__ order_pm, double* __restrict__ r_out, double* __restrict__ jac_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= d.n_obs) return;
  const size_t n = (size_t)d.n_obs;
  const size_t o = (size_t)order_pm[i];
  if (r_out) {
    const double2 r = d.r[i];
    r_out[2 * o] = r.x; r_out[2 * o + 1] = r.y;
  }
  if (jac_out) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const double2 v = d.Jc[k * n + i];
      jac_out[18 * o + k] = v.x; jac_out[18 * o + 9 + k] = v.y;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double2 v = d.Jp[k * n + i];
      jac_out[18 * o + 6 + k] = v.x; jac_out[18 * o + 15 + k] = v.y;
    }
  }
}

__global__ void export_blocks_kernel(int n_cams, int n_pts, const double* __restrict__ Hcc, const double* __restrict__ Dc2,
                                     const double* __restrict__ Hpp, double radius, double min_diag, double max_diag,
                                     double* __restrict__ Hcc36, double* __restrict__ Hpp9) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (Hcc36 && idx < n_cams * 36) {
    const int c = idx / 36, a = (idx % 36) / 6, b = idx % 6;
    double v = Hcc[(size_t)c * 21 + tri6(min(a, b), max(a, b))];
    if (a == b) v += Dc2[6 * c + a];
    Hcc36[idx] = v;
  }
  if (Hpp9 && idx < n_pts * 9) {
    const int p = idx / 9, a = (idx % 9) / 3, b = idx % 3;
    const int lo = min(a, b), hi = max(a, b);
    const int t3 = lo * 3 - (lo * (lo - 1)) / 2 + (hi - lo);
    double v = Hpp[(size_t)p * 6 + t3];
    if (a == b) {
      const double D = sqrt(fmin(fmax(v, min_diag), max_diag) / radius);
      v += D * D;
    }
    Hpp9[idx] = v;
  }
}

inline int cdiv(int a, int b) { return (a + b - 1) / b; }

}  // namespace

// ==========================================================================================
// launchers
// ==========================================================================================
struct PendingTask { const double* in; int n; int scal_slot; int op; };
static thread_local std::vector<PendingTask> g_pending;  // per calling thread; flushed by launch_finalize_scalars

static void add_task(Problem& P, int part_slot, int n, int scal_slot, int op) {
  g_pending.push_back({P.part + (size_t)part_slot * P.part_stride, n, scal_slot, op});
}

void launch_cam_prep(Problem& P, const double* cam, const double* scale_c, CamRow* rows, cudaStream_t s) {
  cam_prep_kernel<<<cdiv(P.n_cams, 128), 128, 0, s>>>(P.n_cams, cam, P.K9, P.dist, scale_c, rows);
  P.launches++;
}

// Largest dynamic shared memory any staging kernel can ask for (256 distinct cameras in a tile).
constexpr int kStageSmemMax = 2 * rj_stage(TILE).bytes + 3 * RJ_META_BYTES + (int)sizeof(double) * 9 * SV_LD;
template <class Kern>
static void allow_stage_smem(Kern k) { cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, kStageSmemMax); }

// Dynamic shared-memory opt-ins of this translation unit.  Function attributes are PER DEVICE: called by
// hsfm_ba_create for the context's device (after cudaSetDevice), not once per process.
int init_kernel_attributes() {
  allow_stage_smem(rj_kernel<true, true>); allow_stage_smem(rj_kernel<true, false>);
  allow_stage_smem(rj_kernel<false, true>); allow_stage_smem(rj_kernel<false, false>);
  allow_stage_smem(cost_kernel);
  cudaFuncSetAttribute(dense_schur_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDenseSchurSmem);
  cudaFuncSetAttribute(dense_schur_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDenseSchurSmem);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : 1000 + (int)e;
}

void launch_rj(Problem& P, const CamRow* rows, const double* pts, const double* scale_p, bool write_j, bool form, cudaStream_t s) {
  if (P.n_tiles == 0) return;
  double* cp = P.part + (size_t)P_COST * P.part_stride;
  const int cap = P.max_tile_cams;
  const size_t sh = rj_smem_bytes(cap, form);
  const int grid = std::min(P.n_tiles, 2 * P.n_sm);  // persistent: two CTAs per SM walk the tiles
  if (write_j && form) rj_kernel<true, true><<<grid, TILE, sh, s>>>(P.d, rows, pts, scale_p, P.Hpp, P.gp, cp, cap);
  else if (write_j) rj_kernel<true, false><<<grid, TILE, sh, s>>>(P.d, rows, pts, scale_p, P.Hpp, P.gp, cp, cap);
  else if (form) rj_kernel<false, true><<<grid, TILE, sh, s>>>(P.d, rows, pts, scale_p, P.Hpp, P.gp, cp, cap);
  else rj_kernel<false, false><<<grid, TILE, sh, s>>>(P.d, rows, pts, scale_p, P.Hpp, P.gp, cp, cap);
  P.launches++;
  add_task(P, P_COST, P.n_tiles, S_COST, 2);
}

void launch_cost(Problem& P, const CamRow* rows, const double* pts, int part_slot, cudaStream_t s) {
  if (P.n_tiles == 0) return;
  cost_kernel<<<P.n_tiles, TILE, cost_smem_bytes(P.max_tile_cams), s>>>(P.d, rows, pts, P.part + (size_t)part_slot * P.part_stride,
                                                                         P.max_tile_cams);
  P.launches++;
  add_task(P, part_slot, P.n_tiles, S_CAND_COST, 2);
}

// Camera-major passes: what 0 -> Hcc, g_c ; what 1 -> E (diagonal blocks of W Hpp^-1 W^T, needs Hpp_inv).  Both recompute the
// Jacobian from 20 B/observation + a point gather.  (Round 2 measured the alternative for what 0 — a streaming pass over the
// stored Jc, r with per-(tile, slot) shared-memory sums of the 27 values and a per-camera gather: 0.33 ms against 0.22 ms
// for this recompute at Venice shape; the 27-value slot reduction costs more than re-deriving the Jacobian.  Not adopted.)
void launch_cam_major(Problem& P, const CamRow* rows, const double* pts, const double* scale_p, int what, cudaStream_t s) {
  if (P.n_chunks > 0) {
    if (what == 0) cam_major_kernel<0><<<cdiv(P.n_chunks, 8), 256, 0, s>>>(P.d, rows, pts, scale_p, P.Hpp_inv, P.chunk_part);
    else cam_major_kernel<1><<<cdiv(P.n_chunks, 8), 256, 0, s>>>(P.d, rows, pts, scale_p, P.Hpp_inv, P.chunk_part);
    P.launches++;
  }
  const int nv = what == 0 ? 27 : 21;
  cam_reduce_kernel<<<cdiv(P.n_cams * nv, 256), 256, 0, s>>>(P.n_cams, nv, P.d.cam_chunk_start, P.chunk_part, what == 0 ? P.Hcc : P.E, P.gc);
  P.launches++;
}

void launch_jacobi_scale(Problem& P, bool enabled, cudaStream_t s) {
  const int n = 6 * P.n_cams + 3 * P.n_pts;
  jacobi_scale_kernel<<<cdiv(n, 256), 256, 0, s>>>(P.n_cams, P.n_pts, P.Hcc, P.Hpp, P.scale_c, P.scale_p, enabled ? 1 : 0,
                                                   P.has_fixed ? P.cam_fixed : nullptr, P.has_fixed ? P.pt_fixed : nullptr);
  P.launches++;
}

void launch_fixed_cost(Problem& P, const CamRow* rows, const double* pts, cudaStream_t s) {
  if (P.n_obs == 0) return;
  const int g = cdiv(P.n_obs, TILE);  // <= n_tiles <= part_stride (tiles hold at most TILE observations)
  fixed_cost_kernel<<<g, TILE, 0, s>>>(P.d, rows, pts, P.cam_fixed, P.pt_fixed, P.mode != 0 ? 1 : 0, P.mode != 1 ? 1 : 0,
                                       P.part + (size_t)P_GMAX_C * P.part_stride);
  P.launches++;
  add_task(P, P_GMAX_C, g, S_FIXED_COST, 2);
}

void launch_point_invert(Problem& P, double radius, double min_diag, double max_diag, cudaStream_t s) {
  if (P.n_pts == 0) return;
  point_invert_kernel<<<cdiv(P.n_pts, 256), 256, 0, s>>>(P.n_pts, P.Hpp, P.gp, radius, min_diag, max_diag, P.Hpp_inv, P.zp, P.flags);
  P.launches++;
}

void launch_cam_diag(Problem& P, double radius, double min_diag, double max_diag, cudaStream_t s) {
  cam_diag_kernel<<<cdiv(6 * P.n_cams, 256), 256, 0, s>>>(P.n_cams, P.Hcc, radius, min_diag, max_diag, P.Dc2);
  P.launches++;
}

void launch_precond_invert(Problem& P, bool with_E, cudaStream_t s) {
  precond_invert_kernel<<<cdiv(P.n_cams, 64), 64, 0, s>>>(P.n_cams, P.Hcc, P.Dc2, P.E, with_E ? 1 : 0, P.Minv, P.flags);
  P.launches++;
}

void launch_schur_rhs(Problem& P, bool preload_gc, cudaStream_t s) {
  if (P.n_tiles > 0) {
    schur_pass_kernel<0><<<P.n_tiles, TILE, 0, s>>>(P.d, P.Hpp_inv, P.zp, nullptr, P.ent_part, nullptr, nullptr, nullptr, nullptr, P.part,
                                                    P.part_stride, P.flags, 1, 1);
    P.launches++;
  }
  if (P.n_cams > 0) {
    slot_gather_kernel<6><<<cdiv(P.n_cams, 2), 256, 0, s>>>(P.d, P.ent_part, preload_gc ? P.gc : nullptr, -1.0, P.rhs, nullptr);
    P.launches++;
  }
}

void launch_schur_matvec(Problem& P, const double* pvec, double* q, cudaStream_t s) {
  // q = (Jc^T (I - Jp Hpp^-1 Jp^T) Jc) p  (overwrites q; honours F_PCG_DONE)
  if (P.n_tiles > 0) {
    schur_pass_kernel<1><<<P.n_tiles, TILE, 0, s>>>(P.d, P.Hpp_inv, P.zp, pvec, P.ent_part, nullptr, nullptr, nullptr, nullptr, P.part,
                                                    P.part_stride, P.flags, 1, 1);
    P.launches++;
  }
  if (P.n_cams > 0) {
    slot_gather_kernel<6><<<cdiv(P.n_cams, 2), 256, 0, s>>>(P.d, P.ent_part, nullptr, 1.0, q, P.flags + F_PCG_DONE);
    P.launches++;
  }
}

void launch_backsub(Problem& P, cudaStream_t s) {
  if (P.n_tiles == 0) return;
  const int cams_free = P.mode != 0, pts_free = P.mode != 1;
  schur_pass_kernel<2><<<P.n_tiles, TILE, 0, s>>>(P.d, P.Hpp_inv, P.zp, P.yc, nullptr, P.pts_x, P.scale_p, P.yp, P.pts_c, P.part,
                                                  P.part_stride, P.flags, cams_free, pts_free);
  P.launches++;
  add_task(P, P_MCC, P.n_tiles, S_MCC, 0);
  add_task(P, P_DN_P, P.n_tiles, S_DN_P, 0);
  add_task(P, P_XN_P, P.n_tiles, S_XN_P, 0);
}

void launch_cam_step(Problem& P, cudaStream_t s) {
  const int n6 = 6 * P.n_cams, g = cdiv(n6, 256);
  cam_step_kernel<<<g, 256, 0, s>>>(n6, P.cam_x, P.yc, P.scale_c, P.cam_nobs, P.cam_c, P.part, P.part_stride, P.mode != 0 ? 1 : 0, P.flags);
  P.launches++;
  add_task(P, P_DN_C, g, S_DN_C, 0);
  add_task(P, P_XN_C, g, S_XN_C, 0);
}

void launch_block_solve_cams(Problem& P, cudaStream_t s) {
  block_solve_cams_kernel<<<cdiv(6 * P.n_cams, 256), 256, 0, s>>>(P.n_cams, P.Minv, P.gc, P.yc);
  P.launches++;
}

void launch_grad_norm(Problem& P, cudaStream_t s) {
  if (P.mode != 0) {
    const int n = 6 * P.n_cams, g = cdiv(n, 256);
    grad_norm_kernel<<<g, 256, 0, s>>>(n, P.gc, P.scale_c, P.part + (size_t)P_GMAX_C * P.part_stride);
    P.launches++;
    add_task(P, P_GMAX_C, g, S_GMAX_C, 1);
  }
  if (P.mode != 1 && P.n_pts > 0) {
    const int n = 3 * P.n_pts, g = cdiv(n, 256);
    grad_norm_kernel<<<g, 256, 0, s>>>(n, P.gp, P.scale_p, P.part + (size_t)P_GMAX_P * P.part_stride);
    P.launches++;
    add_task(P, P_GMAX_P, g, S_GMAX_P, 1);
  }
}

void launch_finalize_scalars(Problem& P, cudaStream_t s) {
  size_t k = 0;
  while (k < g_pending.size()) {
    RedTasks T;
    int cnt = 0;
    for (; cnt < 8 && k < g_pending.size(); ++cnt, ++k) {
      T.in[cnt] = g_pending[k].in;
      T.n[cnt] = g_pending[k].n;
      T.out[cnt] = P.scal + g_pending[k].scal_slot;
      T.op[cnt] = g_pending[k].op;
    }
    reduce_tasks_kernel<<<cnt, 1024, 0, s>>>(T);
    P.launches++;
  }
  g_pending.clear();
}

void launch_pcg_init(Problem& P, cudaStream_t s) {
  pcg_init_kernel<<<1, 1024, 0, s>>>(6 * P.n_cams, P.rhs, P.Minv, P.pcg_x, P.pcg_r, P.pcg_z, P.pcg_p, P.pcg_q, P.scal, P.flags);
  P.launches++;
}

void launch_pcg_step(Problem& P, double tol, int max_it, cudaStream_t s) {
  pcg_step_kernel<<<1, 1024, 0, s>>>(6 * P.n_cams, P.Minv, P.Dc2, P.pcg_x, P.pcg_r, P.pcg_z, P.pcg_p, P.pcg_q, P.scal, P.flags, tol * tol, max_it);
  P.launches++;
}

void launch_dense_init(Problem& P, bool with_diag, cudaStream_t s, const CholPlan* plan) {
  const size_t n = 6 * (size_t)P.n_cams;
  if (plan && plan->use && plan->n == (int)n) {
    // nothing outside the block structure is ever read by the structured factorisation / solves
    launch_structure_zero(*plan, P.S, s);
    P.launches++;
  } else {
    cudaMemsetAsync(P.S, 0, sizeof(double) * n * n, s);
  }
  if (!with_diag) return;
  dense_init_kernel<<<cdiv(P.n_cams * 36, 256), 256, 0, s>>>(P.n_cams, P.Hcc, P.Dc2, P.S);
  P.launches++;
}

void launch_frame_stats(Problem& P, const CamRow* rows, const double* pts, double* out3, cudaStream_t s) {
  if (P.n_cams == 0) return;
  if (P.n_chunks > 0) {
    frame_stats_kernel<<<cdiv(P.n_chunks, 8), 256, 0, s>>>(P.d, rows, pts, P.chunk_part);
    P.launches++;
  }
  frame_stats_reduce_kernel<<<cdiv(P.n_cams, 256), 256, 0, s>>>(P.n_cams, P.d.cam_chunk_start, P.chunk_part, out3);
  P.launches++;
}

void launch_dense_schur(Problem& P, cudaStream_t s, bool sorted_pairs) {
  if (P.n_tiles == 0) return;
  static const bool no_sort = getenv("HSFM_SCHUR_UNSORTED") != nullptr;  // experiment knob
  if (!P.pairs_ready && P.d.pair_off && !no_sort && sorted_pairs) {
    // first dense solve of this problem: build the destination-sorted pair lists (static, like the rest of the index)
    const size_t need = (size_t)std::max<long long>(P.total_pairs, 1);
    if (P.pair_capacity < need) {
      if (P.pair_list) cudaFree(P.pair_list);
      P.pair_list = nullptr; P.pair_capacity = 0;
      if (cudaMalloc(&P.pair_list, sizeof(unsigned) * need) == cudaSuccess) P.pair_capacity = need;
      else cudaGetLastError();
    }
    if (P.pair_list) {
      tile_pairs_kernel<<<P.n_tiles, TILE, 0, s>>>(P.d, P.pair_list);
      P.launches++;
      P.d.pair_list = P.pair_list;
      P.pairs_ready = true;
    }
  }
  if (P.pairs_ready && sorted_pairs) dense_schur_kernel<true><<<P.n_tiles, 2 * TILE, kDenseSchurSmem, s>>>(P.d, P.Hpp_inv, P.S, 0);
  else dense_schur_kernel<false><<<P.n_tiles, TILE, kDenseSchurSmem, s>>>(P.d, P.Hpp_inv, P.S, 0);
  P.launches++;
}

void launch_mirror_lower(Problem& P, cudaStream_t s) {
  const int n = 6 * P.n_cams;
  if (n == 0) return;
  dim3 g(cdiv(n, 256), n);
  mirror_lower_kernel<<<g, 256, 0, s>>>(n, P.S);
  P.launches++;
}

void launch_scatter_eval(Problem& P, double* r_out, double* jac_out, cudaStream_t s) {
  if (P.n_obs == 0) return;
  scatter_eval_kernel<<<cdiv(P.n_obs, 256), 256, 0, s>>>(P.d, P.order_pm, r_out, jac_out);
  P.launches++;
}

void launch_export_blocks(Problem& P, double* Hcc36, double* Hpp9, cudaStream_t s) {
  // uses the radius baked into Dc2 for cameras; points recompute from the diagonal — caller passes radius via P.scal_host[31]
  const int n = max(P.n_cams * 36, P.n_pts * 9);
  if (n == 0) return;
  export_blocks_kernel<<<cdiv(n, 256), 256, 0, s>>>(P.n_cams, P.n_pts, P.Hcc, P.Dc2, P.Hpp, P.scal_host[31], P.scal_host[30], P.scal_host[29],
                                                    Hcc36, Hpp9);
  P.launches++;
}

}  // namespace hsfm
