// ba_order.cpp — host side of the reduced-camera-system structure: camera ordering, 128-block pattern and symbolic
// block factorisation for the Cholesky (the reference hands S to a dense LLT whatever its
// structure, CTracker.cpp:574 — same factor, the zeros are simply not computed).
//
// Plain C++ (no CUDA): compiled into libhsfm_ba.so and, for the CPU test-suite, reachable through the
// hsfm_ba_debug_structure tap of ba_api.cu.
//
// The reduced camera matrix of a scan is a band: camera c shares points only with cameras first[c] .. c (keyframes
// see the points of nearby keyframes), plus loop closures.  A right-looking blocked factorisation of a band is a
// serial chain of n/128 panel steps — latency-bound (one diagonal block at a time), the bulk of the GPU idle.  A
// band can be eliminated FROM BOTH ENDS AT ONCE without creating any fill ("twisted" / burn-at-both-ends order):
//
//     order = [ cameras 0 .. m-1 ascending | cameras n-1 .. m descending, those coupled to the first part LAST ]
//
// In this order S is two independent bands (two concurrent chains of half the length) that meet in a small tail
// block; cameras of the second part that share points with the first part (the band neighbours of the split and
// any loop closure across it) are ordered last, so a closure costs a few extra rows in the chains' panels instead of
// serialising them.  The structure handed to the factorisation is a general block pattern (not an envelope), closed
// under elimination by a symbolic block factorisation.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <vector>

#include "ba_internal.h"

namespace hsfm {

namespace {
constexpr int NB = 128;
constexpr int CAM_ALIGN = 64;  // 64 cameras = 384 rows = 3 panels: chain boundaries must be panel boundaries
}  // namespace

// ---- recursive two-ended dissection of a band ----
// A set of cameras (ascending caller ids) that form a band among themselves is eliminated as
//   depth 0:  [ first part ascending | second part descending | T = cameras of the second part coupled to the first ]
//   depth d:  [ order(lower half, d-1) | order(upper half minus M, d-1) | M = cameras of the upper half coupled to the lower ]
// Every piece but the very last one of the whole order holds a multiple of CAM_ALIGN cameras (chain boundaries are
// panel boundaries): a separator's surplus cameras move UP into the separator of the enclosing level (always valid:
// that one is eliminated later still), returned in `up`.
namespace {
bool order_band(const std::vector<int>& cams, const int* first, int depth, bool is_root, std::vector<int>& out, std::vector<int>& up,
                int* two_ended_sizes = nullptr /*[2]: cameras of the first and of the second chain (root, depth 0)*/) {
  const int k = (int)cams.size();
  if (k < 4 * CAM_ALIGN) return false;
  std::vector<int> sep;
  if (depth == 0) {
    std::vector<int> w((size_t)k);
    for (int q = 0; q < k; ++q) w[q] = cams[q] - std::min(first[cams[q]], cams[q]);
    std::nth_element(w.begin(), w.begin() + (size_t)(0.9 * (k - 1)), w.end());
    const int bw = std::min(w[(size_t)(0.9 * (k - 1))], k / 2);  // typical half-bandwidth in cameras (closure rows are outliers)
    const int mpos = ((k - bw) / 2) / CAM_ALIGN * CAM_ALIGN;
    if (mpos < CAM_ALIGN) return false;
    const int m = cams[(size_t)mpos];
    std::vector<int> second;
    for (int q = mpos; q < k; ++q) (first[cams[q]] < m ? sep : second).push_back(cams[q]);
    // the second chain must hold whole panels as well: its cameras nearest to the split join the separator
    const size_t ex = second.size() % CAM_ALIGN;
    sep.insert(sep.end(), second.begin(), second.begin() + ex);
    second.erase(second.begin(), second.begin() + ex);
    if ((int)second.size() < CAM_ALIGN) return false;
    out.insert(out.end(), cams.begin(), cams.begin() + mpos);
    out.insert(out.end(), second.rbegin(), second.rend());
    if (two_ended_sizes) { two_ended_sizes[0] = mpos; two_ended_sizes[1] = (int)second.size(); }
  } else {
    const int hpos = (k / 2) / CAM_ALIGN * CAM_ALIGN;
    const int h = cams[(size_t)hpos];
    std::vector<int> lower(cams.begin(), cams.begin() + hpos), upper, moved;
    for (int q = hpos; q < k; ++q) (first[cams[q]] < h ? sep : upper).push_back(cams[q]);
    if (!order_band(lower, first, depth - 1, false, out, moved)) return false;
    if (!order_band(upper, first, depth - 1, false, out, moved)) return false;
    sep.insert(sep.end(), moved.begin(), moved.end());
  }
  std::sort(sep.begin(), sep.end());
  if (!is_root) {
    const size_t ex = sep.size() % CAM_ALIGN;
    up.insert(up.end(), sep.end() - ex, sep.end());
    sep.resize(sep.size() - ex);
  }
  out.insert(out.end(), sep.begin(), sep.end());
  return true;
}
}  // namespace

// Column structure of L from a pattern closed under elimination: R_K ascending per panel.
void structure_lists(int P, const std::vector<uint8_t>& st, std::vector<int>& off, std::vector<int>& cnt, std::vector<int>& list) {
  off.assign((size_t)P, 0); cnt.assign((size_t)P, 0); list.clear();
  for (int K = 0; K < P; ++K) {
    off[K] = (int)list.size();
    for (int I = K + 1; I < P; ++I)
      if (st[(size_t)I * P + K]) list.push_back(I);
    cnt[K] = (int)list.size() - off[K];
  }
}

// Panel chains of the factorisation = paths of the block elimination tree (parent(K) = min R_K): a new segment starts
// at every panel that is not the only child of its predecessor.  Segments are numbered in panel order, so children
// precede parents.
void panel_segments(int P, const std::vector<int>& off, const std::vector<int>& cnt, const std::vector<int>& list, std::vector<PanelSeg>& segs,
                    std::vector<int>& seg_of) {
  segs.clear();
  seg_of.assign((size_t)P, 0);
  std::vector<int> parent((size_t)P, -1), nchild((size_t)P, 0);
  for (int K = 0; K < P; ++K)
    if (cnt[K] > 0) { parent[K] = list[(size_t)off[K]]; ++nchild[(size_t)parent[K]]; }
  for (int K = 0; K < P; ++K) {
    if (K == 0 || parent[K - 1] != K || nchild[K] != 1) segs.push_back(PanelSeg{K, K, -1});
    seg_of[K] = (int)segs.size() - 1;
    segs.back().end = K + 1;
  }
  for (auto& g : segs) g.parent = parent[g.end - 1] >= 0 ? seg_of[(size_t)parent[g.end - 1]] : -1;
  // A chain of one or two panels that its parent continues directly is not worth a start of its own (a join and a
  // stream hand-over): it becomes the head of the parent, which then waits for the children of both.
  for (bool merged = true; merged;) {
    merged = false;
    for (size_t g = 0; g + 1 < segs.size(); ++g) {
      if (segs[g].parent != (int)g + 1 || segs[g].end - segs[g].begin > 2) continue;
      segs[g + 1].begin = segs[g].begin;
      segs.erase(segs.begin() + (long)g);
      for (auto& t : segs) {
        if (t.parent == (int)g) t.parent = (int)g;            // its children now hang off the merged segment (which took index g)
        else if (t.parent > (int)g) --t.parent;
      }
      merged = true;
      break;
    }
  }
  for (size_t g = 0; g < segs.size(); ++g)
    for (int K = segs[g].begin; K < segs[g].end; ++K) seg_of[(size_t)K] = (int)g;
}

// Longest root-to-leaf path of the segment tree, in panels.
int critical_panels(const std::vector<PanelSeg>& segs) {
  std::vector<int> depth(segs.size(), 0);
  int best = 0;
  for (int g = (int)segs.size() - 1; g >= 0; --g) {  // parents have larger indices: their depth is final first
    depth[g] = (segs[g].end - segs[g].begin) + (segs[g].parent >= 0 ? depth[(size_t)segs[g].parent] : 0);
    best = std::max(best, depth[g]);
  }
  return best;
}

namespace {
// Estimated time of the structured factorisation + triangular solves in a given camera order.  Segments at the same depth below the root run
// concurrently; a level costs the larger of its longest chain (38 us per panel step, 20 % more per concurrent chain
// beyond two: the chains' kernels share SMs and L2 with each other's bulk updates) and the tensor-pipe time of its
// structural flops (18 TFLOP/s: what the 128-deep trailing updates sustain with several chains in flight).  Constants
// fitted to c3 / c4 at order depths 0-2 on a B200; deep trees are over-estimated (levels overlap).
double order_cost(int n_cams, const int* first, const int* new_of_old, int* crit_out) {
  std::vector<uint8_t> pat;
  int P = 0;
  block_pattern_from_envelope(n_cams, first, new_of_old, pat, &P);
  symbolic_block_factor(P, pat);
  std::vector<int> off, cnt, list, seg_of;
  structure_lists(P, pat, off, cnt, list);
  std::vector<PanelSeg> segs;
  panel_segments(P, off, cnt, list, segs, seg_of);
  if (segs.size() > (size_t)MAX_CHOL_SEGMENTS) { segs.assign(1, PanelSeg{0, P, -1}); }
  if (crit_out) *crit_out = critical_panels(segs);
  int levels = 0;
  std::vector<int> lev(segs.size(), 0);
  for (int g = (int)segs.size() - 1; g >= 0; --g) {
    lev[(size_t)g] = segs[(size_t)g].parent >= 0 ? lev[(size_t)segs[(size_t)g].parent] + 1 : 0;
    levels = std::max(levels, lev[(size_t)g] + 1);
  }
  double total = 0.0;
  for (int l = 0; l < levels; ++l) {
    int chains = 0, longest = 0;
    double tiles = 0.0;  // 128^3 products: diagonal block, panel solve, pairs
    for (size_t g = 0; g < segs.size(); ++g) {
      if (lev[g] != l) continue;
      ++chains;
      longest = std::max(longest, segs[g].end - segs[g].begin);
      for (int K = segs[g].begin; K < segs[g].end; ++K) tiles += 0.5 * cnt[K] * (cnt[K] + 1.0) + 0.5 * cnt[K] + 1.0 / 6;
    }
    const double t_chain = longest * 38e-6 * (1.0 + 0.2 * std::max(0, chains - 2)), t_flops = tiles * 2.0 * NB * NB * NB / 18e12;
    total += std::max(t_chain, t_flops);
  }
  // the two triangular sweeps follow the same chains: two hand-overs of ~4.2 us per critical panel
  return total + 8.4e-6 * critical_panels(segs);
}
}  // namespace

// Decide the internal camera order from the caller-order envelope first[c] (smallest camera sharing a point with c).
// Returns true and fills new_of_old when a reordering pays; *split_cams / *tail_cams: first camera position of the second
// chain / of the tail when the order is the plain two-ended one (depth 0), else -1.  force_depth >= 0 (tests): that
// dissection depth, whatever the cost model says.
bool choose_camera_order(int n_cams, const int* first, std::vector<int>& new_of_old, int* split_cams, int* tail_cams, int force_depth) {
  *split_cams = -1;
  if (tail_cams) *tail_cams = -1;
  new_of_old.clear();
  static const bool off = getenv("HSFM_NO_TWIST") != nullptr;  // experiment knob
  static const int max_depth = getenv("HSFM_ORDER_DEPTH") ? atoi(getenv("HSFM_ORDER_DEPTH")) : 2;
  static const int force_env = getenv("HSFM_ORDER_FORCE") ? atoi(getenv("HSFM_ORDER_FORCE")) : -1;  // experiment knob: that depth, no cost model
  if (force_depth < 0) force_depth = force_env;
  if (off || n_cams < 4 * CAM_ALIGN) return false;
  // how banded is the natural order?
  double env = 0.0;
  for (int c = 0; c < n_cams; ++c) env += c - std::min(first[c], c) + 1;
  const double fill = env / (0.5 * (double)n_cams * (n_cams + 1));
  if (fill > 0.35) return false;
  std::vector<int> cams((size_t)n_cams), best;
  std::iota(cams.begin(), cams.end(), 0);
  int crit = 0;
  double best_cost = order_cost(n_cams, first, nullptr, &crit);
  int best_depth = -1, best_sizes[2] = {-1, -1};
  for (int depth = force_depth >= 0 ? force_depth : 0; depth <= (force_depth >= 0 ? force_depth : max_depth); ++depth) {
    std::vector<int> order, up, cand((size_t)n_cams);
    order.reserve((size_t)n_cams);
    int sizes[2] = {-1, -1};
    if (!order_band(cams, first, depth, true, order, up, sizes) || (int)order.size() != n_cams) break;
    for (int pos = 0; pos < n_cams; ++pos) cand[(size_t)order[(size_t)pos]] = pos;
    const double cost = order_cost(n_cams, first, cand.data(), &crit);
    static const bool verbose = getenv("HSFM_ORDER_VERBOSE") != nullptr;
    if (verbose) std::fprintf(stderr, "[order] depth %d: critical panels %d, estimated %.3f ms (best so far %.3f)\n", depth, crit, 1e3 * cost, 1e3 * best_cost);
    if (cost < (depth == 0 ? 0.9 : 0.98) * best_cost || force_depth >= 0) { best_cost = cost; best.swap(cand); best_depth = depth; best_sizes[0] = sizes[0]; best_sizes[1] = sizes[1]; }
    else if (depth > 0) break;  // a deeper dissection that does not pay ends the search
  }
  if (best_depth < 0) return false;
  new_of_old.swap(best);
  if (best_depth == 0) { *split_cams = best_sizes[0]; if (tail_cams) *tail_cams = best_sizes[0] + best_sizes[1]; }
  return true;
}

// 128-block pattern of S (lower triangle, row-major P x P bytes) in the INTERNAL camera order, from the caller-order
// envelope: cameras c and d (first[c] <= d <= c) may share points.  A superset of the true pattern.
void block_pattern_from_envelope(int n_cams, const int* first, const int* new_of_old /*null = identity*/, std::vector<uint8_t>& pattern, int* panels) {
  const int n = 6 * n_cams, P = (n + NB - 1) / NB;
  *panels = P;
  pattern.assign((size_t)P * P, 0);
  // first and last 128-row block of every camera, by CALLER index (a camera's six rows straddle a block boundary at most once)
  std::vector<int> pos((size_t)n_cams), b0((size_t)n_cams), b1((size_t)n_cams);
  for (int c = 0; c < n_cams; ++c) {
    const int ic = new_of_old ? new_of_old[c] : c;
    pos[c] = ic; b0[c] = (6 * ic) / NB; b1[c] = (6 * ic + 5) / NB;
  }
  uint8_t* pat = pattern.data();
  for (int c = 0; c < n_cams; ++c) {
    const int ic = pos[c], cb0 = b0[c], cb1 = b1[c];
    for (int d = std::min(first[c], c); d <= c; ++d) {
      // block rows of the later camera (in the internal order) x block columns of the earlier one, lower triangle
      const bool c_later = ic >= pos[d];
      const int r0 = c_later ? cb0 : b0[d], r1 = c_later ? cb1 : b1[d], q0 = c_later ? b0[d] : cb0, q1 = c_later ? b1[d] : cb1;
      pat[(size_t)r0 * P + q0] = 1;
      if (r1 != r0 || q1 != q0) {
        for (int bi = r0; bi <= r1; ++bi)
          for (int bj = q0; bj <= std::min(q1, bi); ++bj) pat[(size_t)bi * P + bj] = 1;
      }
    }
  }
  for (int I = 0; I < P; ++I) pattern[(size_t)I * P + I] = 1;
}

// Symbolic block factorisation: closes the pattern under elimination in place (pattern becomes the structure of L).
// Column K's structure R_K = { I > K : L(I,K) != 0 }; eliminating K couples every pair of R_K.  With parent(K) =
// min R_K it is enough to merge R_K \ {parent} into the parent's column (elimination-tree property).
void symbolic_block_factor(int P, std::vector<uint8_t>& pattern) {
  std::vector<int> rk;
  for (int K = 0; K < P; ++K) {
    rk.clear();
    for (int I = K + 1; I < P; ++I)
      if (pattern[(size_t)I * P + K]) rk.push_back(I);
    if (rk.size() < 2) continue;
    const int par = rk[0];
    for (size_t q = 1; q < rk.size(); ++q) pattern[(size_t)rk[q] * P + par] = 1;
  }
}

}  // namespace hsfm

namespace hsfm {

// Reverse Cuthill-McKee on the camera graph (undirected edges a < b, each once; any order): new position of every
// camera.  For a camera graph that is a band in SOME order (a scan whose keyframes arrive shuffled) this finds such an
// order again; the envelope of the result is then measured, not assumed.  Components are ordered one after the other,
// each from a pseudo-peripheral start node (two rounds of breadth-first search from a minimum-degree node).
void rcm_order(int n, const std::vector<std::pair<int, int>>& edges, std::vector<int>& pos_of_cam) {
  std::vector<int> deg((size_t)n, 0), off((size_t)n + 1, 0);
  for (const auto& e : edges) { ++deg[(size_t)e.first]; ++deg[(size_t)e.second]; }
  for (int c = 0; c < n; ++c) off[(size_t)c + 1] = off[(size_t)c] + deg[(size_t)c];
  std::vector<int> adj((size_t)off[(size_t)n]), fillp(off.begin(), off.end() - 1);
  for (const auto& e : edges) { adj[(size_t)fillp[(size_t)e.first]++] = e.second; adj[(size_t)fillp[(size_t)e.second]++] = e.first; }
  for (int c = 0; c < n; ++c)  // neighbours by ascending degree (ties: index): the Cuthill-McKee visiting order
    std::sort(adj.begin() + off[(size_t)c], adj.begin() + off[(size_t)c + 1],
              [&](int a, int b) { return deg[(size_t)a] != deg[(size_t)b] ? deg[(size_t)a] < deg[(size_t)b] : a < b; });
  std::vector<int> order, level((size_t)n, -1), queue;
  order.reserve((size_t)n);
  std::vector<char> done((size_t)n, 0);
  // breadth-first search over the not-yet-ordered part from `start`; returns the last node of the deepest level with the smallest degree
  auto bfs = [&](int start, std::vector<int>& out) {
    out.clear();
    out.push_back(start);
    level[(size_t)start] = 0;
    for (size_t h = 0; h < out.size(); ++h) {
      const int u = out[h];
      for (int q = off[(size_t)u]; q < off[(size_t)u + 1]; ++q) {
        const int v = adj[(size_t)q];
        if (done[(size_t)v] || level[(size_t)v] >= 0) continue;
        level[(size_t)v] = level[(size_t)u] + 1;
        out.push_back(v);
      }
    }
    const int deepest = level[(size_t)out.back()];
    int best = out.back();
    for (size_t h = out.size(); h-- > 0 && level[(size_t)out[h]] == deepest;)
      if (deg[(size_t)out[h]] < deg[(size_t)best]) best = out[h];
    for (int u : out) level[(size_t)u] = -1;
    return best;
  };
  std::vector<int> by_degree((size_t)n);
  std::iota(by_degree.begin(), by_degree.end(), 0);
  std::stable_sort(by_degree.begin(), by_degree.end(), [&](int a, int b) { return deg[(size_t)a] < deg[(size_t)b]; });
  for (int seed : by_degree) {
    if (done[(size_t)seed]) continue;
    int start = seed;
    for (int round = 0; round < 2; ++round) start = bfs(start, queue);  // towards one end of the component
    bfs(start, queue);                                                   // the Cuthill-McKee order of the component
    for (int u : queue) { done[(size_t)u] = 1; order.push_back(u); }
  }
  pos_of_cam.assign((size_t)n, 0);
  for (int k = 0; k < n; ++k) pos_of_cam[(size_t)order[(size_t)(n - 1 - k)]] = k;  // reversed
}

}  // namespace hsfm
