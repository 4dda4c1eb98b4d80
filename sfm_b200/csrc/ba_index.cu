// ba_index.cu — upload + one-time index build of a resident problem (the packed
// arrays CSfM::bundleAdjustment assembles at CSfM.cpp:311-341 arrive camera-major; the solver
// wants them point-major in tiles plus a camera-major copy).  Sorting is a plain library
// operation done once per problem: CUB radix sort (stable, so equal keys keep input order and
// every later floating-point sum has a fixed order).
#include <cub/cub.cuh>

#include <algorithm>
#include <chrono>

#include "ba_internal.h"

namespace hsfm {

void Problem::release() {
  // keep what outlives a problem: the arena, the dense matrix, the pinned scalar mirrors
  char* a = arena; const size_t cap = arena_cap;
  double* s_keep = S; const size_t s_cap = S_capacity;
  unsigned* pl_keep = pair_list; const size_t pl_cap = pair_capacity;
  double* ep_keep = ent_part; const size_t ep_cap = ent_part_cap;
  double* sh = scal_host; int* fh = flags_host;
  int* ih = idx_host; const size_t ih_cap = idx_host_cap;
  const int sm = n_sm;
  OrderCache oc;
  std::swap(oc, order_cache);
  cudaStream_t cs = copy_stream; cudaEvent_t e0 = ev_copy_go, e1 = ev_copy_done;
  const HostStager st = stager;
  *this = Problem();
  std::swap(oc, order_cache);
  copy_stream = cs; ev_copy_go = e0; ev_copy_done = e1;
  stager = st;
  n_sm = sm;
  arena = a; arena_cap = cap; arena_off = 0;
  S = s_keep; S_capacity = s_cap;
  pair_list = pl_keep; pair_capacity = pl_cap;
  ent_part = ep_keep; ent_part_cap = ep_cap;
  scal_host = sh; flags_host = fh;
  idx_host = ih; idx_host_cap = ih_cap;
}

void Problem::destroy() {
  if (arena) cudaFree(arena);
  if (S) cudaFree(S);
  if (pair_list) cudaFree(pair_list);
  if (ent_part) cudaFree(ent_part);
  if (scal_host) cudaFreeHost(scal_host);
  if (flags_host) cudaFreeHost(flags_host);
  if (idx_host) cudaFreeHost(idx_host);
  stager.destroy();
  if (copy_stream) cudaStreamDestroy(copy_stream);
  if (ev_copy_go) cudaEventDestroy(ev_copy_go);
  if (ev_copy_done) cudaEventDestroy(ev_copy_done);
  *this = Problem();
}

int Problem::arena_reserve(size_t bytes) {
  arena_off = 0;
  if (bytes <= arena_cap) return 0;
  if (arena) cudaFree(arena);
  arena = nullptr; arena_cap = 0;
  const size_t want = bytes + bytes / 4;  // head-room: the scan grows by a keyframe per call
  if (cudaMalloc(&arena, want) != cudaSuccess) {
    cudaGetLastError();
    if (cudaMalloc(&arena, bytes) != cudaSuccess) { cudaGetLastError(); arena = nullptr; return HSFM_BA_ERR_NOMEM; }
    arena_cap = bytes;
    return 0;
  }
  arena_cap = want;
  return 0;
}

void* Problem::arena_take(size_t bytes) {
  const size_t need = (bytes + 255) & ~(size_t)255;
  if (arena_off + need > arena_cap) return nullptr;
  void* p = arena + arena_off;
  arena_off += need;
  return p;
}

namespace {

template <typename T>
int dev_alloc(Problem& P, T** out, size_t count) {
  if (count == 0) count = 1;
  void* p = P.arena_take(count * sizeof(T));
  if (!p) return HSFM_BA_ERR_NOMEM;
  *out = reinterpret_cast<T*>(p);
  return 0;
}
#define ALLOC(ptr, count)                         \
  do {                                            \
    int rc__ = dev_alloc(P, &(ptr), (count));     \
    if (rc__) return rc__;                        \
  } while (0)

__global__ void iota_kernel(int n, int* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = i;
}
// Point-major / camera-major gathers, integer keys and pixel coordinates apart: the index build only needs the keys, so
// the observations' coordinates (the bulk of the upload) can still be on their way while it runs.
__global__ void gather_key_kernel(int n, const int* __restrict__ order, const int* __restrict__ key_in, int* __restrict__ key) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) key[i] = key_in[order[i]];
}
__global__ void gather_xy_kernel(int n, const int* __restrict__ order_pm, const int* __restrict__ pos_cs, const double2* __restrict__ obs_in,
                                 double2* __restrict__ obs, double2* __restrict__ obs_cs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  obs[i] = obs_in[order_pm[i]];
  obs_cs[i] = obs_in[order_pm[pos_cs[i]]];
}
// start[k] = first index i with keys[i] >= k   (keys sorted ascending), k = 0..n_keys
__global__ void lower_bound_kernel(int n, const int* __restrict__ keys, int n_keys, int* __restrict__ start) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n_keys) return;
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (keys[mid] < k) lo = mid + 1; else hi = mid;
  }
  start[k] = lo;
}
// pairs[p] = L (L + 1) / 2 observation pairs of point p (L = its track length), pairs[n_pts] = 0; *max_track = longest track
__global__ void point_pairs_kernel(int n_pts, const int* __restrict__ pt_start, long long* __restrict__ pairs, int* __restrict__ max_track) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  int L = 0;
  if (p < n_pts) L = pt_start[p + 1] - pt_start[p];
  if (p <= n_pts) pairs[p] = (long long)L * (L + 1) / 2;
  const int m = __reduce_max_sync(0xffffffffu, L);
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(max_track, m);
}
// pair_off[k] = pairs before the first point of tile k (k = n_tiles: before the end point of the last tile = all pairs)
__global__ void tile_pair_off_kernel(int n_tiles, const int* __restrict__ tile_hdr, const long long* __restrict__ pairs_before, int* __restrict__ pair_off) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n_tiles) return;
  const int p = k < n_tiles ? tile_hdr[8 * (size_t)k + 2] : tile_hdr[8 * (size_t)(n_tiles - 1) + 3];
  pair_off[k] = (int)pairs_before[p];
}
// ---- point renumbering: internal id = rank of (smallest observing camera, caller id) ----
__global__ void fill_int_kernel(int n, int v, int* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = v;
}
__global__ void min_cam_kernel(int n, const int* __restrict__ pt, const int* __restrict__ cam, int* __restrict__ min_cam) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) atomicMin(min_cam + pt[i], cam[i]);
}
// the same on indices that have not been validated yet (out-of-range entries are skipped; the upload then fails)
__global__ void min_cam_checked_kernel(int n, const int* __restrict__ pt, const int* __restrict__ cam, int n_cams, int n_pts,
                                       int* __restrict__ min_cam) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && (unsigned)cam[i] < (unsigned)n_cams && (unsigned)pt[i] < (unsigned)n_pts) atomicMin(min_cam + pt[i], cam[i]);
}
// first[c] = smallest camera that shares a point with camera c (first[] preset to c): the envelope of the reduced
// camera matrix in the CALLER's camera order.  Integer atomicMin: order-independent, once per upload.
__global__ void cam_first_raw_kernel(int n, const int* __restrict__ pt, const int* __restrict__ cam, int n_cams, int n_pts,
                                     const int* __restrict__ min_cam, int* __restrict__ first) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool ok = i < n && (unsigned)cam[i] < (unsigned)n_cams && (unsigned)pt[i] < (unsigned)n_pts;
  const int c = ok ? cam[i] : -1;
  int v = ok ? min_cam[pt[i]] : 0x7fffffff;
  // observations arrive grouped by camera (CSfM.cpp:321-341): the lanes of a warp mostly name the same camera, and a few
  // thousand atomics on one address serialise (1.46 ms at Venice shape).  Lanes with the same camera combine first.
  const unsigned peers = __match_any_sync(0xffffffffu, c);
  v = __reduce_min_sync(peers, v);
  if (ok && (int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicMin(first + c, v);
}
// One edge of the camera graph per observation: (its camera, the smallest camera observing its point) — a star per point,
// a subgraph of the co-visibility graph with the same connected components and distances within a factor of two, n_obs
// edges instead of sum of track^2.  Key = (larger << 32 | smaller); an observation of the hub itself gives no edge.
__global__ void star_edges_kernel(int n, const int* __restrict__ pt, const int* __restrict__ cam, const int* __restrict__ min_cam,
                                  unsigned long long* __restrict__ keys) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = cam[i], m = min_cam[pt[i]];
  keys[i] = (c == m) ? ~0ull : (((unsigned long long)(unsigned)max(c, m) << 32) | (unsigned)min(c, m));
}
__global__ void invert_perm_kernel(int n, const int* __restrict__ old_of_new, int* __restrict__ new_of_old) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) new_of_old[old_of_new[k]] = k;
}
__global__ void gather_int_kernel(int n, const int* __restrict__ idx, const int* __restrict__ table, int* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = table[idx[i]];
}
// dst[k][0..w) = src[row_of[k]][0..w)
__global__ void gather_rows_kernel(int n, int w, const int* __restrict__ row_of, const double* __restrict__ src, double* __restrict__ dst) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (size_t)n * w) return;
  const int k = (int)(e / w), c = (int)(e % w);
  dst[e] = src[(size_t)row_of[k] * w + c];
}
__global__ void diff_kernel(int n, const int* __restrict__ start, int* __restrict__ cnt) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) cnt[k] = start[k + 1] - start[k];
}

inline int cdiv(int a, int b) { return (a + b - 1) / b; }
inline int cdivz(size_t a, int b) { return (int)((a + b - 1) / b); }
double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

void launch_rows_to_internal(Problem& P, int w, const double* caller_order, double* internal_order, cudaStream_t s) {
  if (P.n_pts == 0) return;
  gather_rows_kernel<<<cdivz((size_t)P.n_pts * w, 256), 256, 0, s>>>(P.n_pts, w, P.pt_old_of_new, caller_order, internal_order);
  P.launches++;
}
void launch_rows_to_caller(Problem& P, int w, const double* internal_order, double* caller_order, cudaStream_t s) {
  if (P.n_pts == 0) return;
  gather_rows_kernel<<<cdivz((size_t)P.n_pts * w, 256), 256, 0, s>>>(P.n_pts, w, P.pt_new_of_old, internal_order, caller_order);
  P.launches++;
}

__global__ void gather_bytes_kernel(int n, const int* __restrict__ row_of, const unsigned char* __restrict__ src, unsigned char* __restrict__ dst) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) dst[k] = src[row_of ? row_of[k] : k] ? 1 : 0;
}
// Constant-block masks from the caller's camera / point order (device staging buffers) into the internal order.
void launch_set_fixed(Problem& P, const unsigned char* cam_caller, const unsigned char* pt_caller, cudaStream_t s) {
  if (P.n_cams > 0) {
    if (cam_caller) gather_bytes_kernel<<<cdiv(P.n_cams, 256), 256, 0, s>>>(P.n_cams, P.cam_old_of_new, cam_caller, P.cam_fixed);
    else cudaMemsetAsync(P.cam_fixed, 0, (size_t)P.n_cams, s);
  }
  if (P.n_pts > 0) {
    if (pt_caller) gather_bytes_kernel<<<cdiv(P.n_pts, 256), 256, 0, s>>>(P.n_pts, P.pt_old_of_new, pt_caller, P.pt_fixed);
    else cudaMemsetAsync(P.pt_fixed, 0, (size_t)P.n_pts, s);
  }
  P.launches += 2;
  P.has_fixed = cam_caller || pt_caller;
}

void launch_cam_rows_to_caller(Problem& P, int w, const double* internal_order, double* caller_order, cudaStream_t s) {
  if (P.n_cams == 0) return;
  gather_rows_kernel<<<cdivz((size_t)P.n_cams * w, 256), 256, 0, s>>>(P.n_cams, w, P.cam_new_of_old, internal_order, caller_order);
  P.launches++;
}

__global__ void validate_idx_kernel(int n, const int* __restrict__ cam, const int* __restrict__ pt, int n_cams, int n_pts,
                                    int* __restrict__ bad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && ((unsigned)cam[i] >= (unsigned)n_cams || (unsigned)pt[i] >= (unsigned)n_pts)) *bad = 1;
}

// One CTA per tile: the tile's distinct cameras in ascending order and the slot of every observation's
// camera in that list.  O(TILE^2) comparisons through shared-memory broadcasts; runs once per upload.
__global__ void __launch_bounds__(TILE) tile_cams_kernel(const int* __restrict__ tile_start, const int* __restrict__ cam,
                                                          unsigned char* __restrict__ slot, int* __restrict__ tile_cams,
                                                          int* __restrict__ tile_hdr, int* __restrict__ max_ncam,
                                                          unsigned char* __restrict__ srank, unsigned char* __restrict__ sorder,
                                                          unsigned short* __restrict__ slot_begin, int* __restrict__ tile_ncam) {
  __shared__ __align__(16) int sc[TILE];
  __shared__ __align__(16) int sfirst[TILE];
  const int b = tile_start[blockIdx.x], e = tile_start[blockIdx.x + 1];
  const int tid = threadIdx.x, i = b + tid;
  const int c = i < e ? cam[i] : 0x7fffffff;  // past the tile: larger than every camera, never "smaller than"
  sc[tid] = c;
  __syncthreads();
  // position of this observation when the tile is ordered by camera (ties keep tile order) — camera-keyed sums of a tile
  // are then sums over contiguous runs of a shared-memory array, in a fixed order — in ONE pass over the tile's cameras
  // (four per shared-memory broadcast); an observation is the first of its camera when none of the same camera precedes it
  int before = 0, same_before = 0;
#pragma unroll 4
  for (int j = 0; j < TILE; j += 4) {
    const int4 v = *reinterpret_cast<const int4*>(sc + j);
    before += (v.x < c) + (v.y < c) + (v.z < c) + (v.w < c);
    same_before += ((v.x == c) & (j < tid)) + ((v.y == c) & (j + 1 < tid)) + ((v.z == c) & (j + 2 < tid)) + ((v.w == c) & (j + 3 < tid));
  }
  const int first = (i < e) && same_before == 0;
  sfirst[tid] = first ? c : 0x7fffffff;  // the camera where an observation is the first of it, else "none"
  const int ncam = __syncthreads_count(first);
  // slot = number of distinct cameras below this one
  int rank = 0;
#pragma unroll 4
  for (int j = 0; j < TILE; j += 4) {
    const int4 v = *reinterpret_cast<const int4*>(sfirst + j);
    rank += (v.x < c) + (v.y < c) + (v.z < c) + (v.w < c);
  }
  if (i < e) {
    slot[i] = (unsigned char)rank;
    srank[i] = (unsigned char)(before + same_before);
    sorder[b + before + same_before] = (unsigned char)tid;  // inverse: which observation of the tile sits at that position
    if (first) {
      tile_cams[(size_t)blockIdx.x * TILE + rank] = c;
      slot_begin[(size_t)blockIdx.x * TILE + rank] = (unsigned short)before;
    }
  }
  if (tid == 0) {
    tile_hdr[8 * (size_t)blockIdx.x + 4] = ncam;
    tile_ncam[blockIdx.x] = ncam;
    atomicMax(max_ncam, ncam);
  }
}
// entry e = ent_off[tile] + slot of the (tile, slot) list: its camera (sort key) and itself (value)
__global__ void __launch_bounds__(TILE) tile_entries_kernel(const int* __restrict__ ent_off, const int* __restrict__ tile_cams, int* __restrict__ key,
                                                             int* __restrict__ val) {
  const int o = ent_off[blockIdx.x], n = ent_off[blockIdx.x + 1] - o;
  if ((int)threadIdx.x < n) {
    key[o + threadIdx.x] = tile_cams[(size_t)blockIdx.x * TILE + threadIdx.x];
    val[o + threadIdx.x] = o + threadIdx.x;
  }
}

// ---- pointer identity -> point ids (hsfm_ba_solve_ptr) -------------------------------------------------------
namespace {
__global__ void ptr_heads_kernel(int n, const unsigned long long* __restrict__ sorted, int* __restrict__ head) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) head[i] = (i == 0 || sorted[i] != sorted[i - 1]) ? 1 : 0;
}
// id of sorted element i = (inclusive scan of heads) - 1; scattered to the observation it came from; distinct pointers by id
__global__ void ptr_ids_kernel(int n, const unsigned long long* __restrict__ sorted, const int* __restrict__ src, const int* __restrict__ head,
                               const int* __restrict__ incl, int* __restrict__ pt_idx, unsigned long long* __restrict__ uniq) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int id = incl[i] - 1;
  pt_idx[src[i]] = id;
  if (head[i]) uniq[id] = sorted[i];
}
}  // namespace

void PtrScratch::destroy() {
  if (buf) cudaFree(buf);
  if (uniq_host) cudaFreeHost(uniq_host);
  if (pts_host) cudaFreeHost(pts_host);
  pts_host = nullptr; pts_cap = 0;
  *this = PtrScratch();
}

int dedupe_pointers(PtrScratch& W, int64_t n64, double* const* host_ptrs, int** d_pt_idx_out, double* const** uniq_out, int64_t* n_uniq_out,
                    cudaStream_t s, HostStager* stager) {
  *uniq_out = nullptr;
  *n_uniq_out = 0;
  *d_pt_idx_out = nullptr;
  if (n64 >= (int64_t)1 << 31) return HSFM_BA_ERR_UNSUPPORTED;
  const int n = (int)n64;
  if (n == 0) return 0;
  static_assert(sizeof(double*) == sizeof(unsigned long long), "64-bit pointers");
  size_t t_sort = 0, t_scan = 0;
  unsigned long long* nullk = nullptr;
  int* nulli = nullptr;
  cub::DeviceRadixSort::SortPairs(nullptr, t_sort, nullk, nullk, nulli, nulli, n, 0, 48, s);  // user-space addresses: 48 significant bits
  cub::DeviceScan::InclusiveSum(nullptr, t_scan, nulli, nulli, n, s);
  const size_t tmp = (std::max(t_sort, t_scan) + 255) & ~(size_t)255;
  const size_t n8 = ((size_t)n * 8 + 255) & ~(size_t)255, n4 = ((size_t)n * 4 + 255) & ~(size_t)255;
  const size_t need = 3 * n8 + 5 * n4 + tmp;
  if (W.cap < need) {
    if (W.buf) cudaFree(W.buf);
    W.buf = nullptr; W.cap = 0;
    if (cudaMalloc(&W.buf, need + need / 4) != cudaSuccess) { cudaGetLastError(); return HSFM_BA_ERR_NOMEM; }
    W.cap = need + need / 4;
  }
  char* b = W.buf;
  auto take = [&](size_t bytes) { char* p = b; b += bytes; return p; };
  unsigned long long* keys = (unsigned long long*)take(n8);
  unsigned long long* sorted = (unsigned long long*)take(n8);
  unsigned long long* uniq = (unsigned long long*)take(n8);
  int* iota = (int*)take(n4);
  int* src = (int*)take(n4);
  int* head = (int*)take(n4);
  int* incl = (int*)take(n4);
  int* pt_idx = (int*)take(n4);
  void* d_tmp = take(tmp);
  if (stager && sizeof(unsigned long long) * (size_t)n >= HostStager::MIN_BYTES && HostStager::pageable(host_ptrs)) {
    int device = 0;
    HSFM_CUDA(cudaGetDevice(&device));
    const int rc = stager->copy({{keys, host_ptrs, sizeof(unsigned long long) * (size_t)n}}, s, device);
    if (rc) return rc;
  } else {
    HSFM_CUDA(cudaMemcpyAsync(keys, host_ptrs, sizeof(unsigned long long) * (size_t)n, cudaMemcpyHostToDevice, s));
  }
  iota_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, iota);
  size_t tb = tmp;
  cub::DeviceRadixSort::SortPairs(d_tmp, tb, keys, sorted, iota, src, n, 0, 48, s);
  ptr_heads_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, sorted, head);
  tb = tmp;
  cub::DeviceScan::InclusiveSum(d_tmp, tb, head, incl, n, s);
  ptr_ids_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, sorted, src, head, incl, pt_idx, uniq);
  int n_uniq = 0;
  unsigned long long first = 0;
  HSFM_CUDA(cudaMemcpyAsync(&n_uniq, incl + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, s));
  HSFM_CUDA(cudaMemcpyAsync(&first, sorted, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
  HSFM_CUDA(cudaStreamSynchronize(s));
  if (first == 0ull) return HSFM_BA_ERR_ARG;  // a NULL point pointer (sorted first)
  if (W.uniq_cap < (size_t)n_uniq) {
    if (W.uniq_host) cudaFreeHost(W.uniq_host);
    W.uniq_host = nullptr; W.uniq_cap = 0;
    const size_t want = (size_t)n_uniq + (size_t)n_uniq / 4 + 1024;
    HSFM_CUDA(cudaMallocHost(&W.uniq_host, sizeof(unsigned long long) * want));
    W.uniq_cap = want;
  }
  HSFM_CUDA(cudaMemcpyAsync(W.uniq_host, uniq, sizeof(unsigned long long) * (size_t)n_uniq, cudaMemcpyDeviceToHost, s));
  HSFM_CUDA(cudaStreamSynchronize(s));
  *uniq_out = reinterpret_cast<double* const*>(W.uniq_host);  // the pinned mirror itself: valid until the next call
  *n_uniq_out = n_uniq;
  // pinned staging of the distinct points for the packed call (grow-only)
  if (W.pts_cap < 3 * (size_t)n_uniq) {
    if (W.pts_host) cudaFreeHost(W.pts_host);
    W.pts_host = nullptr; W.pts_cap = 0;
    const size_t want = 3 * ((size_t)n_uniq + (size_t)n_uniq / 4 + 1024);
    HSFM_CUDA(cudaMallocHost(&W.pts_host, sizeof(double) * want));
    W.pts_cap = want;
  }
  *d_pt_idx_out = pt_idx;
  return 0;
}

// Camera graph -> reverse Cuthill-McKee positions (host), for a problem whose cameras do not arrive in a banded order.
// Transient device scratch (this is the slow path of an upload); indices are validated already.
static int rcm_positions(int n, const int* pt_in, const int* cam_in, const int* min_cam, int n_cams, int cbits, std::vector<int>& pos, cudaStream_t s) {
  unsigned long long *keys = nullptr, *sorted = nullptr, *uniq = nullptr;
  int* d_num = nullptr;
  void* tmp = nullptr;
  size_t t_sort = 0, t_uniq = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, t_sort, keys, sorted, n, 0, 64, s);
  cub::DeviceSelect::Unique(nullptr, t_uniq, sorted, uniq, d_num, n, s);
  const size_t n8 = ((size_t)n * 8 + 255) & ~(size_t)255, tb = std::max(t_sort, t_uniq);
  char* buf = nullptr;
  if (cudaMalloc(&buf, 3 * n8 + 256 + tb) != cudaSuccess) { cudaGetLastError(); return HSFM_BA_ERR_NOMEM; }
  keys = (unsigned long long*)buf; sorted = (unsigned long long*)(buf + n8); uniq = (unsigned long long*)(buf + 2 * n8);
  d_num = (int*)(buf + 3 * n8); tmp = buf + 3 * n8 + 256;
  int rc = 0, h_num = 0;
  std::vector<unsigned long long> h_edges;
  auto run = [&]() -> int {
    star_edges_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, pt_in, cam_in, min_cam, keys);
    size_t t1 = tb, t2 = tb;
    cub::DeviceRadixSort::SortKeys(tmp, t1, keys, sorted, n, 0, 32 + cbits, s);  // the sentinel's high bits are not sorted on: it may land anywhere ...
    cub::DeviceSelect::Unique(tmp, t2, sorted, uniq, d_num, n, s);
    HSFM_CUDA(cudaMemcpyAsync(&h_num, d_num, sizeof(int), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    h_edges.resize((size_t)h_num);
    HSFM_CUDA(cudaMemcpyAsync(h_edges.data(), uniq, sizeof(unsigned long long) * (size_t)h_num, cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    return 0;
  };
  rc = run();
  cudaFree(buf);
  if (rc) return rc;
  std::vector<std::pair<int, int>> edges;
  edges.reserve(h_edges.size());
  for (unsigned long long k : h_edges) {
    if (k == ~0ull) continue;  // ... and is dropped here
    edges.emplace_back((int)(unsigned)(k & 0xffffffffu), (int)(unsigned)(k >> 32));
  }
  rcm_order(n_cams, edges, pos);
  return 0;
}

int build_problem(Problem& P, int mode, int64_t n_obs64, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                  int32_t n_cams, const double* K9, const double* dist, const double* cam_rt, int64_t n_pts64, const double* pts,
                  Comm& comm, cudaStream_t s, double* setup_ms, double* h2d_ms, bool natural_order) {
  if (P.copy_stream) cudaStreamSynchronize(P.copy_stream);  // a previous upload that returned early may still be copying into the arena
  P.release();
  if (n_obs64 >= (int64_t)1 << 31 || n_pts64 >= (int64_t)1 << 30) return HSFM_BA_ERR_UNSUPPORTED;
  const int n = (int)n_obs64, n_pts = (int)n_pts64;
  P.mode = mode; P.n_obs = n; P.n_cams = n_cams; P.n_pts = n_pts;

  // indices are validated on the device right after the upload (the reference indexes unchecked:
  // CTracker.cpp:676-690), before any kernel uses them as addresses

  {
    // upper bound of everything carved below (per observation / point / camera / tile / chunk) + sort scratch
    const size_t no = (size_t)n, np = (size_t)n_pts, nc = (size_t)n_cams;
    size_t sort_tmp = 0, t2 = 0;
    int* nullk = nullptr;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp, nullk, nullk, nullk, nullk, std::max({n, n_pts, 1}), 0, 32, s);
    t2 = sort_tmp;
    const size_t need = 376 * no + 360 * np + 2600 * nc + 1400 * (no / 128 + nc + np / 256 + 8) + 2 * t2 + ((size_t)1 << 20);
    int rc = P.arena_reserve(need);
    if (rc) return rc;
  }
  const double t_h2d0 = now_ms();
  double copy_wait_ms = 0.0;
  double2* obs_in; int *cam_in, *pt_in;
  ALLOC(obs_in, n); ALLOC(cam_in, n); ALLOC(pt_in, n);
  ALLOC(P.K9, 9 * (size_t)n_cams); ALLOC(P.dist, 2 * (size_t)n_cams);
  ALLOC(P.cam_x, 6 * (size_t)n_cams); ALLOC(P.pts_x, 3 * (size_t)n_pts);
  ALLOC(P.cam_x0, 6 * (size_t)n_cams); ALLOC(P.pts_x0, 3 * (size_t)n_pts); ALLOC(P.pts_io, 3 * (size_t)n_pts);
  // cudaMemcpyDefault: the sources are host buffers (one-shot calls) or the device-resident scan store (ba_scan.cu).
  // The index arrays go first on the solver's stream; pixel coordinates and points (two thirds of the bytes) follow on a
  // copy stream of their own and are only waited for at the end of the index build, which needs none of them.
  if (!P.copy_stream) {
    HSFM_CUDA(cudaStreamCreateWithFlags(&P.copy_stream, cudaStreamNonBlocking));
    HSFM_CUDA(cudaEventCreateWithFlags(&P.ev_copy_go, cudaEventDisableTiming));
    HSFM_CUDA(cudaEventCreateWithFlags(&P.ev_copy_done, cudaEventDisableTiming));
  }
  int device = 0;
  HSFM_CUDA(cudaGetDevice(&device));
  // pageable sources (the reference's std::vectors) are staged through pinned rings by a few threads (HostStager)
  auto staged = [&](const void* p, size_t bytes) { return bytes >= HostStager::MIN_BYTES && HostStager::pageable(p); };
  {
    std::vector<HostStager::Job> jobs;
    if (staged(cam_idx, sizeof(int) * (size_t)n)) jobs.push_back({cam_in, cam_idx, sizeof(int) * (size_t)n});
    else HSFM_CUDA(cudaMemcpyAsync(cam_in, cam_idx, sizeof(int) * (size_t)n, cudaMemcpyDefault, s));
    if (staged(pt_idx, sizeof(int) * (size_t)n)) jobs.push_back({pt_in, pt_idx, sizeof(int) * (size_t)n});
    else HSFM_CUDA(cudaMemcpyAsync(pt_in, pt_idx, sizeof(int) * (size_t)n, cudaMemcpyDefault, s));
    if (!jobs.empty()) {
      const int rc = P.stager.copy(jobs, s, device);
      if (rc) return rc;
    }
  }
  HSFM_CUDA(cudaMemcpyAsync(P.K9, K9, sizeof(double) * 9 * n_cams, cudaMemcpyDefault, s));
  if (dist) HSFM_CUDA(cudaMemcpyAsync(P.dist, dist, sizeof(double) * 2 * n_cams, cudaMemcpyDefault, s));
  else HSFM_CUDA(cudaMemsetAsync(P.dist, 0, sizeof(double) * 2 * std::max(1, n_cams), s));
  HSFM_CUDA(cudaMemcpyAsync(P.cam_x0, cam_rt, sizeof(double) * 6 * n_cams, cudaMemcpyDefault, s));
  // after the index arrays (they are needed first and share the link), and after everything that used the arena before
  HSFM_CUDA(cudaEventRecord(P.ev_copy_go, s));
  HSFM_CUDA(cudaStreamWaitEvent(P.copy_stream, P.ev_copy_go, 0));
  struct Joiner {  // background staging threads reference this frame: no return without them
    std::vector<std::thread> th;
    void join() { for (auto& t : th) if (t.joinable()) t.join(); th.clear(); }
    ~Joiner() { join(); }
  } bg;
  int bg_rc = 0;
  {
    std::vector<HostStager::Job> jobs;
    if (staged(obs_xy, sizeof(double) * 2 * (size_t)n)) jobs.push_back({obs_in, obs_xy, sizeof(double) * 2 * (size_t)n});
    else HSFM_CUDA(cudaMemcpyAsync(obs_in, obs_xy, sizeof(double) * 2 * n, cudaMemcpyDefault, P.copy_stream));
    if (staged(pts, sizeof(double) * 3 * (size_t)n_pts)) jobs.push_back({P.pts_io, pts, sizeof(double) * 3 * (size_t)n_pts});
    else HSFM_CUDA(cudaMemcpyAsync(P.pts_io, pts, sizeof(double) * 3 * (size_t)n_pts, cudaMemcpyDefault, P.copy_stream));
    if (!jobs.empty()) P.stager.start(jobs, P.copy_stream, device, bg.th, &bg_rc);
  }
  // Envelope of the reduced camera matrix in the caller's camera order (over all ranks): what the camera ordering and
  // the block structure of the Cholesky are derived from.  Only where a structured factorisation can pay (>= 4 panels).
  int *min_cam, *min_cam_sorted;
  ALLOC(min_cam, n_pts); ALLOC(min_cam_sorted, n_pts);
  const bool want_structure = mode == 2 && n > 0 && n_pts > 0 && 6 * (int64_t)n_cams >= 4 * 128;
  int* d_first = nullptr;
  if (want_structure) {
    ALLOC(d_first, n_cams);
    fill_int_kernel<<<cdiv(n_pts, 256), 256, 0, s>>>(n_pts, n_cams, min_cam);
    min_cam_checked_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, pt_in, cam_in, n_cams, n_pts, min_cam);
    iota_kernel<<<cdiv(n_cams, 256), 256, 0, s>>>(n_cams, d_first);
    cam_first_raw_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, pt_in, cam_in, n_cams, n_pts, min_cam, d_first);
    if (comm.n_ranks > 1) {
      int rc = comm.allreduce_min_i32(d_first, (size_t)n_cams, s);
      if (rc) return rc;
    }
    P.cam_first.assign((size_t)n_cams, 0);
    HSFM_CUDA(cudaMemcpyAsync(P.cam_first.data(), d_first, sizeof(int) * (size_t)n_cams, cudaMemcpyDeviceToHost, s));
  }
  {
    int* d_bad;
    ALLOC(d_bad, 1);
    int h_bad = 0;
    HSFM_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), s));
    if (n > 0) validate_idx_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, cam_in, pt_in, n_cams, n_pts, d_bad);
    if (comm.n_ranks > 1) {  // a bad index in one shard: every rank refuses (nobody is left waiting in a later collective)
      const int rc = comm.allreduce_sum_i32(d_bad, 1, s);
      if (rc) return rc;
    }
    HSFM_CUDA(cudaMemcpyAsync(&h_bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    if (h_bad) return HSFM_BA_ERR_ARG;
  }
  const double t_setup0 = now_ms();
  if (h2d_ms) *h2d_ms = t_setup0 - t_h2d0;
  static const bool stage_timing = getenv("HSFM_SETUP_TIMING") != nullptr;  // experiment knob: where the index build spends its time
  double t_stage = t_setup0;
  auto stamp = [&](const char* what) {
    if (!stage_timing) return;
    cudaStreamSynchronize(s);
    const double t = now_ms();
    std::fprintf(stderr, "[setup] %-28s %7.3f ms\n", what, t - t_stage);
    t_stage = t;
  };

  // ---- cameras that do not arrive in a banded order: reverse Cuthill-McKee first (one GPU; the ranks of a multi-GPU
  //      solve would each see their shard's graph only) ----
  // "Base order" below = the caller's order, or the RCM order when that is adopted; the envelope (P.cam_first) and the
  // structure code work in base order, P.h_cam_new_of_old maps the CALLER's cameras to internal ones.
  std::vector<int> base_of_caller;  // empty: base order = caller order
  static const bool no_rcm = getenv("HSFM_NO_RCM") != nullptr;  // experiment knob
  if (want_structure && !natural_order && !no_rcm && comm.n_ranks == 1 && n_cams >= 256) {
    auto env_fill = [&](const std::vector<int>& first) {
      double env = 0.0;
      for (int c = 0; c < n_cams; ++c) env += c - std::min(first[(size_t)c], c) + 1;
      return env / (0.5 * (double)n_cams * (n_cams + 1));
    };
    const double fill_caller = env_fill(P.cam_first);
    if (fill_caller > 0.35) {
      Problem::OrderCache& oc = P.order_cache;
      if (!(oc.rcm_valid && oc.first_caller == P.cam_first)) {
        int cb = 1; while ((1ll << cb) < std::max(2, (int)n_cams) + 1) ++cb;
        const int rc = rcm_positions(n, pt_in, cam_in, min_cam, n_cams, cb, oc.base_of_caller, s);
        if (rc) return rc;
        oc.first_caller = P.cam_first;
        oc.rcm_valid = true;
      }
      // the envelope in RCM order, measured on the data (a cached permutation stays valid for any graph: it is a permutation)
      int *d_base, *cam_rcm;
      ALLOC(d_base, n_cams); ALLOC(cam_rcm, n);
      HSFM_CUDA(cudaMemcpyAsync(d_base, oc.base_of_caller.data(), sizeof(int) * (size_t)n_cams, cudaMemcpyHostToDevice, s));
      gather_int_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, cam_in, d_base, cam_rcm);
      fill_int_kernel<<<cdiv(n_pts, 256), 256, 0, s>>>(n_pts, n_cams, min_cam);
      min_cam_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, pt_in, cam_rcm, min_cam);
      iota_kernel<<<cdiv(n_cams, 256), 256, 0, s>>>(n_cams, d_first);
      cam_first_raw_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, pt_in, cam_rcm, n_cams, n_pts, min_cam, d_first);
      std::vector<int> first_rcm((size_t)n_cams);
      HSFM_CUDA(cudaMemcpyAsync(first_rcm.data(), d_first, sizeof(int) * (size_t)n_cams, cudaMemcpyDeviceToHost, s));
      HSFM_CUDA(cudaStreamSynchronize(s));
      if (env_fill(first_rcm) < 0.6 * fill_caller) {
        base_of_caller = oc.base_of_caller;
        P.cam_first.swap(first_rcm);
      }
    }
  }
  stamp("camera graph / RCM");

  // ---- internal camera order (ba_order.cpp): a band is eliminated from both ends at once ----
  bool reorder = false;
  P.h_struct_new_of_base.clear();
  if (want_structure && !natural_order) {
    Problem::OrderCache& oc = P.order_cache;
    if (!(oc.valid && oc.first == P.cam_first)) {
      oc.reordered = choose_camera_order(n_cams, P.cam_first.data(), oc.new_of_old, &oc.split_cams, &oc.tail_cams);
      oc.first = P.cam_first;
      oc.valid = true;
    }
    if (oc.reordered) { P.h_struct_new_of_base = oc.new_of_old; P.split_cams = oc.split_cams; P.tail_cams = oc.tail_cams; }
    // caller -> internal: through the base order
    if (!base_of_caller.empty()) {
      P.h_cam_new_of_old.resize((size_t)n_cams);
      for (int c = 0; c < n_cams; ++c)
        P.h_cam_new_of_old[(size_t)c] = oc.reordered ? oc.new_of_old[(size_t)base_of_caller[(size_t)c]] : base_of_caller[(size_t)c];
      reorder = true;
    } else if (oc.reordered) {
      P.h_cam_new_of_old = oc.new_of_old;
      reorder = true;
    }
  }
  if (reorder) {
    int* cam_tmp;
    ALLOC(P.cam_new_of_old, n_cams); ALLOC(P.cam_old_of_new, n_cams); ALLOC(cam_tmp, n); ALLOC(P.cam_io, 9 * (size_t)n_cams);
    HSFM_CUDA(cudaMemcpyAsync(P.cam_new_of_old, P.h_cam_new_of_old.data(), sizeof(int) * (size_t)n_cams, cudaMemcpyHostToDevice, s));
    invert_perm_kernel<<<cdiv(n_cams, 256), 256, 0, s>>>(n_cams, P.cam_new_of_old, P.cam_old_of_new);
    gather_int_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, cam_in, P.cam_new_of_old, cam_tmp);
    cam_in = cam_tmp;
    // camera tables: row k of the internal order = row old_of_new[k] of the caller's
    struct { double* a; int w; } tabs[3] = {{P.K9, 9}, {P.dist, 2}, {P.cam_x0, 6}};
    for (auto& t : tabs) {
      gather_rows_kernel<<<cdivz((size_t)n_cams * t.w, 256), 256, 0, s>>>(n_cams, t.w, P.cam_old_of_new, t.a, P.cam_io);
      HSFM_CUDA(cudaMemcpyAsync(t.a, P.cam_io, sizeof(double) * (size_t)t.w * n_cams, cudaMemcpyDeviceToDevice, s));
    }
  } else {
    P.h_cam_new_of_old.clear();
    P.split_cams = -1;
    P.tail_cams = -1;
  }
  HSFM_CUDA(cudaMemcpyAsync(P.cam_x, P.cam_x0, sizeof(double) * 6 * n_cams, cudaMemcpyDeviceToDevice, s));
  stamp("camera order");

  // ---- static arrays ----
  double2 *obs, *obs_cs; int *cam, *pt, *pt_cs, *pt_start, *cam_start, *iota, *pos_cs, *cam_sorted, *order1, *key2;
  ALLOC(obs, n); ALLOC(cam, n); ALLOC(pt, n); ALLOC(obs_cs, n); ALLOC(pt_cs, n);
  ALLOC(pt_start, (size_t)n_pts + 1); ALLOC(cam_start, (size_t)n_cams + 1);
  const int n_iota = std::max(n, n_pts);
  ALLOC(iota, n_iota); ALLOC(P.order_pm, n); ALLOC(pos_cs, n); ALLOC(cam_sorted, n); ALLOC(order1, n); ALLOC(key2, n);
  ALLOC(P.pt_old_of_new, n_pts); ALLOC(P.pt_new_of_old, n_pts);
  if (n_iota > 0) iota_kernel<<<cdiv(n_iota, 256), 256, 0, s>>>(n_iota, iota);
  int pbits = 1; while ((1ll << pbits) < std::max(2, n_pts)) ++pbits;
  int cbits = 1; while ((1ll << cbits) < std::max(2, (int)n_cams) + 1) ++cbits;
  size_t tmp_bytes = 0, tmp2 = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, pt_in, pt, iota, P.order_pm, std::max(n, 1), 0, pbits, s);
  cub::DeviceRadixSort::SortPairs(nullptr, tmp2, cam, cam_sorted, iota, pos_cs, std::max({n, n_pts, 1}), 0, cbits, s);
  tmp_bytes = std::max(tmp_bytes, tmp2);
  cub::DeviceScan::InclusiveSum(nullptr, tmp2, iota, iota, std::max({n, n_pts, 1}), s);  // prefix sum over the tiles' camera counts
  tmp_bytes = std::max(tmp_bytes, tmp2);
  long long* pt_pairs;   // [n_pts + 1] observation pairs per point, then (exclusive scan) before every point
  int* d_max_track;
  ALLOC(pt_pairs, (size_t)n_pts + 1); ALLOC(d_max_track, 1);
  cub::DeviceScan::ExclusiveSum(nullptr, tmp2, pt_pairs, pt_pairs, n_pts + 1, s);
  tmp_bytes = std::max(tmp_bytes, tmp2);
  void* tmp = P.arena_take(tmp_bytes);
  if (!tmp) return HSFM_BA_ERR_NOMEM;
  // Internal point ids: points ordered by their smallest observing camera (ties: caller id).  Points that
  // share cameras become neighbours, so a tile's camera rows, the camera-vector atomics of the mat-vec
  // and above all the 6x6 blocks of S that dense_schur updates stay within a small, L2-resident window.
  if (n_pts > 0) {
    fill_int_kernel<<<cdiv(n_pts, 256), 256, 0, s>>>(n_pts, n_cams, min_cam);  // no observation: sorted last
    if (n > 0) min_cam_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, pt_in, cam_in, min_cam);
    cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, min_cam, min_cam_sorted, iota, P.pt_old_of_new, n_pts, 0, cbits, s);
    invert_perm_kernel<<<cdiv(n_pts, 256), 256, 0, s>>>(n_pts, P.pt_old_of_new, P.pt_new_of_old);
  }
  if (n > 0) {
    // point-major with cameras ascending inside a point: stable sort by camera, then stable sort by
    // internal point id (ties keep input order, so every later floating-point sum has a fixed order)
    cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, cam_in, cam_sorted, iota, order1, n, 0, cbits, s);
    gather_int_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, order1, pt_in, key2);
    gather_int_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, key2, P.pt_new_of_old, pos_cs /*scratch*/);
    cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, pos_cs, pt, order1, P.order_pm, n, 0, pbits, s);
    gather_key_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, P.order_pm, cam_in, cam);
    // camera-major: stable sort of point-major positions by camera
    cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, cam, cam_sorted, iota, pos_cs, n, 0, cbits, s);
    gather_key_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, pos_cs, pt, pt_cs);
    lower_bound_kernel<<<cdiv(n_pts + 1, 256), 256, 0, s>>>(n, pt, n_pts, pt_start);
    lower_bound_kernel<<<cdiv(n_cams + 1, 256), 256, 0, s>>>(n, cam_sorted, n_cams, cam_start);
    HSFM_CUDA(cudaMemsetAsync(d_max_track, 0, sizeof(int), s));
    point_pairs_kernel<<<cdiv(n_pts + 1, 256), 256, 0, s>>>(n_pts, pt_start, pt_pairs, d_max_track);
    size_t tb = tmp_bytes;
    cub::DeviceScan::ExclusiveSum(tmp, tb, pt_pairs, pt_pairs, n_pts + 1, s);
  } else {
    HSFM_CUDA(cudaMemsetAsync(pt_start, 0, sizeof(int) * ((size_t)n_pts + 1), s));
    HSFM_CUDA(cudaMemsetAsync(cam_start, 0, sizeof(int) * ((size_t)n_cams + 1), s));
  }
  HSFM_CUDA(cudaStreamSynchronize(s));
  stamp("sorts (point / camera major)");

  // ---- tiles: greedy packing of whole points into <= TILE observations ----
  // The host walks the point offsets tile by tile (one binary search per tile, ~n_obs / 250 of them: a loop over all
  // points cost 2.8 ms at Venice size); track lengths and pair counts are reduced on the device.
  // the point offsets come back through a pinned, grow-only staging buffer (a pageable 4 MB copy per call costs ~1 ms)
  if (P.idx_host_cap < (size_t)n_pts + 8) {
    if (P.idx_host) cudaFreeHost(P.idx_host);
    P.idx_host = nullptr; P.idx_host_cap = 0;
    const size_t want = ((size_t)n_pts + 8) * 5 / 4 + 1024;
    HSFM_CUDA(cudaMallocHost(&P.idx_host, sizeof(int) * want));
    P.idx_host_cap = want;
  }
  int* const h_pt_start = P.idx_host;
  int* const h_small = P.idx_host + (size_t)n_pts + 2;  // [0] longest track
  std::vector<int> h_cam_start((size_t)n_cams + 1);
  HSFM_CUDA(cudaMemcpyAsync(h_pt_start, pt_start, sizeof(int) * ((size_t)n_pts + 1), cudaMemcpyDeviceToHost, s));
  HSFM_CUDA(cudaMemcpyAsync(h_cam_start.data(), cam_start, sizeof(int) * ((size_t)n_cams + 1), cudaMemcpyDeviceToHost, s));
  long long h_total_pairs = 0;
  h_small[0] = 0;
  if (n > 0) {
    HSFM_CUDA(cudaMemcpyAsync(h_small, d_max_track, sizeof(int), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaMemcpyAsync(&h_total_pairs, pt_pairs + n_pts, sizeof(long long), cudaMemcpyDeviceToHost, s));
  }
  HSFM_CUDA(cudaStreamSynchronize(s));
  const int max_track = h_small[0];
  P.max_track = max_track;
  {
    // all ranks leave together (the next collective is the observation count per camera): a long track in ONE shard
    int too_long = max_track > MAX_TRACK ? 1 : 0;
    if (comm.n_ranks > 1) {
      HSFM_CUDA(cudaMemcpyAsync(d_max_track, &too_long, sizeof(int), cudaMemcpyHostToDevice, s));
      const int rc = comm.allreduce_sum_i32(d_max_track, 1, s);
      if (rc) return rc;
      HSFM_CUDA(cudaMemcpyAsync(&too_long, d_max_track, sizeof(int), cudaMemcpyDeviceToHost, s));
      HSFM_CUDA(cudaStreamSynchronize(s));
    }
    if (too_long) return HSFM_BA_ERR_UNSUPPORTED;
  }
  std::vector<int> tiles, tile_p;  // first observation / first point of every tile (+ end sentinels)
  tiles.reserve((size_t)n / (TILE - 32) + 2);
  tile_p.reserve((size_t)n / (TILE - 32) + 2);
  if (n > 0) {
    // points without observations are sorted last: p_end = one past the last point that has some
    const int p_end = (int)(std::lower_bound(h_pt_start, h_pt_start + n_pts + 1, n) - h_pt_start);
    int p = 0;
    tiles.push_back(0); tile_p.push_back(0);
    while (p < p_end) {
      // the tile takes the points p .. q-1 with h_pt_start[q] - h_pt_start[p] <= TILE (at least one: max_track <= TILE)
      const int q = (int)(std::upper_bound(h_pt_start + p + 1, h_pt_start + p_end + 1, h_pt_start[p] + TILE) - h_pt_start) - 1;
      p = std::max(q, p + 1);
      tiles.push_back(h_pt_start[p]); tile_p.push_back(p);
    }
  }
  stamp("tile packing (host)");
  P.n_tiles = tiles.empty() ? 0 : (int)tiles.size() - 1;
  int* tile_start;
  ALLOC(tile_start, tiles.size() + 1);
  if (!tiles.empty()) HSFM_CUDA(cudaMemcpyAsync(tile_start, tiles.data(), sizeof(int) * tiles.size(), cudaMemcpyHostToDevice, s));
  // per-tile offsets into the (lazily built) sorted pair lists of dense_schur: sum over the tile's points of L(L+1)/2
  // (filled on the device once the tile headers are there)
  int* pair_off = nullptr;
  P.total_pairs = h_total_pairs;
  P.pairs_ready = false;
  if (h_total_pairs < (1ll << 31) && P.n_tiles > 0) ALLOC(pair_off, (size_t)P.n_tiles + 1);
  // distinct cameras of every tile + each observation's slot in that list: the streaming kernels stage the
  // tile's camera rows / camera vectors in shared memory once instead of gathering them per observation
  unsigned char *slot, *srank, *sorder; unsigned short* slot_begin; int *tile_cams, *tile_hdr, *d_max_ncam, *ent_off;
  ALLOC(slot, n); ALLOC(tile_cams, (size_t)P.n_tiles * TILE); ALLOC(tile_hdr, 8 * (size_t)std::max(1, P.n_tiles)); ALLOC(d_max_ncam, 1);
  ALLOC(srank, n); ALLOC(sorder, n); ALLOC(slot_begin, (size_t)P.n_tiles * TILE); ALLOC(ent_off, (size_t)P.n_tiles + 2);
  int *cs_start = nullptr, *cs_ent = nullptr;
  ALLOC(cs_start, (size_t)n_cams + 1);
  P.n_entries = 0;
  P.max_tile_cams = 0;
  if (P.n_tiles) {
    std::vector<int> hdr(8 * (size_t)P.n_tiles, 0);
    for (int k = 0; k < P.n_tiles; ++k) {
      hdr[8 * (size_t)k] = tiles[k]; hdr[8 * (size_t)k + 1] = tiles[k + 1];
      hdr[8 * (size_t)k + 2] = tile_p[k]; hdr[8 * (size_t)k + 3] = tile_p[k + 1];
    }
    HSFM_CUDA(cudaMemcpyAsync(tile_hdr, hdr.data(), sizeof(int) * hdr.size(), cudaMemcpyHostToDevice, s));
    if (pair_off) tile_pair_off_kernel<<<cdiv(P.n_tiles + 1, 256), 256, 0, s>>>(P.n_tiles, tile_hdr, pt_pairs, pair_off);
    HSFM_CUDA(cudaMemsetAsync(d_max_ncam, 0, sizeof(int), s));
    HSFM_CUDA(cudaMemsetAsync(ent_off, 0, sizeof(int), s));
    tile_cams_kernel<<<P.n_tiles, TILE, 0, s>>>(tile_start, cam, slot, tile_cams, tile_hdr, d_max_ncam, srank, sorder, slot_begin, ent_off + 1);
    {  // (tile, slot) entries: ent_off = prefix sum of the tiles' camera counts
      size_t tb = tmp_bytes;
      size_t need_scan = 0;
      cub::DeviceScan::InclusiveSum(nullptr, need_scan, ent_off + 1, ent_off + 1, P.n_tiles, s);
      if (need_scan > tmp_bytes) return HSFM_BA_ERR_NOMEM;
      cub::DeviceScan::InclusiveSum(tmp, tb, ent_off + 1, ent_off + 1, P.n_tiles, s);
      HSFM_CUDA(cudaMemcpyAsync(&P.n_entries, ent_off + P.n_tiles, sizeof(int), cudaMemcpyDeviceToHost, s));
    }
    HSFM_CUDA(cudaMemcpyAsync(&P.max_tile_cams, d_max_ncam, sizeof(int), cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));  // hdr is a local vector; max_tile_cams is needed by the launches
    // every camera's (tile, slot) entries, in tile order: what the camera-keyed gathers walk (stable sort by camera)
    int *ek, *ev, *ek2;
    ALLOC(ek, P.n_entries); ALLOC(ev, P.n_entries); ALLOC(ek2, P.n_entries); ALLOC(cs_ent, P.n_entries);
    tile_entries_kernel<<<P.n_tiles, TILE, 0, s>>>(ent_off, tile_cams, ek, ev);
    {
      size_t need_sort = 0, tb = tmp_bytes;
      cub::DeviceRadixSort::SortPairs(nullptr, need_sort, ek, ek2, ev, cs_ent, P.n_entries, 0, cbits, s);
      if (need_sort > tmp_bytes) return HSFM_BA_ERR_NOMEM;
      cub::DeviceRadixSort::SortPairs(tmp, tb, ek, ek2, ev, cs_ent, P.n_entries, 0, cbits, s);
    }
    lower_bound_kernel<<<cdiv(n_cams + 1, 256), 256, 0, s>>>(P.n_entries, ek2, n_cams, cs_start);
  } else {
    HSFM_CUDA(cudaMemsetAsync(cs_start, 0, sizeof(int) * ((size_t)n_cams + 1), s));
    HSFM_CUDA(cudaMemsetAsync(ent_off, 0, sizeof(int) * 2, s));
  }

  stamp("tile cameras / entries");
  // ---- camera-major warp chunks ----
  std::vector<int> ch_cam, ch_b, ch_e, cam_ch_start((size_t)n_cams + 1, 0);
  for (int c = 0; c < n_cams; ++c) {
    cam_ch_start[c] = (int)ch_cam.size();
    for (int b = h_cam_start[c]; b < h_cam_start[c + 1]; b += CHUNK) {
      ch_cam.push_back(c); ch_b.push_back(b); ch_e.push_back(std::min(b + CHUNK, h_cam_start[c + 1]));
    }
  }
  cam_ch_start[n_cams] = (int)ch_cam.size();
  P.n_chunks = (int)ch_cam.size();
  int *d_ch_cam, *d_ch_b, *d_ch_e, *d_cam_ch_start;
  ALLOC(d_ch_cam, ch_cam.size()); ALLOC(d_ch_b, ch_cam.size()); ALLOC(d_ch_e, ch_cam.size()); ALLOC(d_cam_ch_start, (size_t)n_cams + 1);
  if (P.n_chunks) {
    HSFM_CUDA(cudaMemcpyAsync(d_ch_cam, ch_cam.data(), sizeof(int) * ch_cam.size(), cudaMemcpyHostToDevice, s));
    HSFM_CUDA(cudaMemcpyAsync(d_ch_b, ch_b.data(), sizeof(int) * ch_cam.size(), cudaMemcpyHostToDevice, s));
    HSFM_CUDA(cudaMemcpyAsync(d_ch_e, ch_e.data(), sizeof(int) * ch_cam.size(), cudaMemcpyHostToDevice, s));
  }
  HSFM_CUDA(cudaMemcpyAsync(d_cam_ch_start, cam_ch_start.data(), sizeof(int) * ((size_t)n_cams + 1), cudaMemcpyHostToDevice, s));

  // ---- observation counts per camera over ALL ranks (cameras without observations are not
  //      parameter blocks of the reference problem: CTracker.cpp:676-696 adds blocks per residual) ----
  ALLOC(P.cam_nobs, (size_t)n_cams);
  ALLOC(P.cam_fixed, (size_t)n_cams); ALLOC(P.pt_fixed, (size_t)n_pts);
  P.has_fixed = false;
  diff_kernel<<<cdiv(std::max(1, (int)n_cams), 256), 256, 0, s>>>(n_cams, cam_start, P.cam_nobs);
  if (comm.n_ranks > 1) {
    int rc = comm.allreduce_sum_i32(P.cam_nobs, (size_t)n_cams, s);
    if (rc) return rc;
  }

  // ---- linearisation + solver state ----
  double2 *r, *Jc, *Jp;
  ALLOC(r, n); ALLOC(Jc, 6 * (size_t)n); ALLOC(Jp, 3 * (size_t)n);
  ALLOC(P.cam_c, 6 * (size_t)n_cams); ALLOC(P.pts_c, 3 * (size_t)n_pts);
  ALLOC(P.rows_x, n_cams); ALLOC(P.rows_c, n_cams);
  ALLOC(P.scale_c, 6 * (size_t)n_cams); ALLOC(P.scale_p, 3 * (size_t)n_pts);
  ALLOC(P.Hpp, 6 * (size_t)n_pts); ALLOC(P.gp, 3 * (size_t)n_pts); ALLOC(P.Hpp_inv, 6 * (size_t)n_pts); ALLOC(P.zp, 3 * (size_t)n_pts);
  ALLOC(P.Hcc, 21 * (size_t)n_cams); ALLOC(P.gc, 6 * (size_t)n_cams); ALLOC(P.E, 21 * (size_t)n_cams);
  ALLOC(P.chunk_part, 27 * (size_t)std::max(1, P.n_chunks));
  P.tobs = nullptr;  // (round 1: per-observation camera-space vectors; replaced by per-(tile, slot) entry partials, Problem::ent_part)
  ALLOC(P.Dc2, 6 * (size_t)n_cams); ALLOC(P.Minv, 36 * (size_t)n_cams);
  ALLOC(P.rhs, 6 * (size_t)n_cams); ALLOC(P.yc, 18 * (size_t)n_cams);  /* solution | forward vector | backward vector of the triangular solves */ ALLOC(P.yp, 3 * (size_t)n_pts);
  ALLOC(P.pcg_x, 6 * (size_t)n_cams); ALLOC(P.pcg_r, 6 * (size_t)n_cams); ALLOC(P.pcg_z, 6 * (size_t)n_cams);
  ALLOC(P.pcg_p, 6 * (size_t)n_cams); ALLOC(P.pcg_q, 6 * (size_t)n_cams);
  P.part_stride = std::max({P.n_tiles, cdiv(6 * n_cams, 256), cdiv(3 * n_pts, 256), 1});
  ALLOC(P.part, (size_t)PART_SLOTS * P.part_stride);
  ALLOC(P.scal, SCAL_SLOTS); ALLOC(P.flags, FLAG_SLOTS);
  if (!P.scal_host) HSFM_CUDA(cudaMallocHost(&P.scal_host, sizeof(double) * SCAL_SLOTS));
  if (!P.flags_host) HSFM_CUDA(cudaMallocHost(&P.flags_host, sizeof(int) * FLAG_SLOTS));
  HSFM_CUDA(cudaMemsetAsync(P.scal, 0, sizeof(double) * SCAL_SLOTS, s));
  HSFM_CUDA(cudaMemsetAsync(P.flags, 0, sizeof(int) * FLAG_SLOTS, s));
  HSFM_CUDA(cudaMemsetAsync(P.part, 0, sizeof(double) * (size_t)PART_SLOTS * P.part_stride, s));
  // blocks of points / cameras without observations are never written by the kernels: keep them zero
  HSFM_CUDA(cudaMemsetAsync(P.Hpp, 0, sizeof(double) * 6 * std::max<size_t>(1, n_pts), s));
  HSFM_CUDA(cudaMemsetAsync(P.gp, 0, sizeof(double) * 3 * std::max<size_t>(1, n_pts), s));
  HSFM_CUDA(cudaMemsetAsync(P.yp, 0, sizeof(double) * 3 * std::max<size_t>(1, n_pts), s));
  HSFM_CUDA(cudaMemsetAsync(P.yc, 0, sizeof(double) * 18 * std::max<size_t>(1, n_cams), s));
  // ---- the bulk of the upload has had the whole index build to arrive: pixel coordinates and points into their orders ----
  {
    const double t_w0 = now_ms();
    if (stage_timing) cudaStreamSynchronize(s);
    bg.join();
    if (bg_rc) return bg_rc;
    HSFM_CUDA(cudaEventRecord(P.ev_copy_done, P.copy_stream));
    HSFM_CUDA(cudaEventSynchronize(P.ev_copy_done));
    copy_wait_ms = now_ms() - t_w0;
    if (stage_timing) std::fprintf(stderr, "[setup] index arrays + envelope: %.3f ms; wait for the rest of the upload: %.3f ms\n", t_setup0 - t_h2d0, copy_wait_ms);
  }
  if (n > 0) gather_xy_kernel<<<cdiv(n, 256), 256, 0, s>>>(n, P.order_pm, pos_cs, obs_in, obs, obs_cs);
  if (n_pts > 0) gather_rows_kernel<<<cdivz((size_t)n_pts * 3, 256), 256, 0, s>>>(n_pts, 3, P.pt_old_of_new, P.pts_io, P.pts_x0);
  HSFM_CUDA(cudaMemcpyAsync(P.pts_x, P.pts_x0, sizeof(double) * 3 * (size_t)n_pts, cudaMemcpyDeviceToDevice, s));
  HSFM_CUDA(cudaMemcpyAsync(P.pts_c, P.pts_x, sizeof(double) * 3 * (size_t)n_pts, cudaMemcpyDeviceToDevice, s));

  {
    const size_t need_ep = 6 * (size_t)std::max(1, P.n_entries);
    if (P.ent_part_cap < need_ep) {
      if (P.ent_part) cudaFree(P.ent_part);
      P.ent_part = nullptr; P.ent_part_cap = 0;
      if (cudaMalloc(&P.ent_part, sizeof(double) * (need_ep + need_ep / 4)) != cudaSuccess) { cudaGetLastError(); return HSFM_BA_ERR_NOMEM; }
      P.ent_part_cap = need_ep + need_ep / 4;
    }
  }
  Dev& d = P.d;
  d.n_obs = n; d.n_cams = n_cams; d.n_pts = n_pts; d.n_tiles = P.n_tiles; d.n_chunks = P.n_chunks;
  d.obs = obs; d.cam = cam; d.pt = pt; d.pt_start = pt_start; d.tile_start = tile_start;
  d.tile_cams = tile_cams; d.tile_hdr = reinterpret_cast<const int4*>(tile_hdr); d.slot = slot;
  d.pair_off = pair_off; d.pair_list = nullptr;
  d.srank = srank; d.sorder = sorder; d.slot_begin = slot_begin; d.ent_off = ent_off; d.cs_start = cs_start; d.cs_ent = cs_ent; d.n_entries = P.n_entries;
  d.obs_cs = obs_cs; d.pt_cs = pt_cs; d.pos_cs = pos_cs; d.chunk_cam = d_ch_cam; d.chunk_b = d_ch_b; d.chunk_e = d_ch_e; d.cam_chunk_start = d_cam_ch_start;
  d.r = r; d.Jc = Jc; d.Jp = Jp;
  HSFM_CUDA(cudaStreamSynchronize(s));
  stamp("chunks, state, clears");
  // the wait for the tail of the upload is transfer time, not index-build time
  if (setup_ms) *setup_ms = now_ms() - t_setup0 - copy_wait_ms;
  if (h2d_ms) *h2d_ms += copy_wait_ms;
  return 0;
}

}  // namespace hsfm
