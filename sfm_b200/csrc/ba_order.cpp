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

// Decide the internal camera order from the caller-order envelope first[c] (smallest camera sharing a point with c).
// Returns true and fills new_of_old when a reordering pays; *split_cams receives m (the first camera position of the
// second chain) or -1, *tail_cams (optional) the first camera position of the tail.
bool choose_camera_order(int n_cams, const int* first, std::vector<int>& new_of_old, int* split_cams, int* tail_cams) {
  *split_cams = -1;
  if (tail_cams) *tail_cams = -1;
  new_of_old.clear();
  static const bool off = getenv("HSFM_NO_TWIST") != nullptr;  // experiment knob
  if (off || n_cams < 4 * CAM_ALIGN) return false;
  // how banded is the natural order?
  double env = 0.0;
  std::vector<int> w((size_t)n_cams);
  for (int c = 0; c < n_cams; ++c) { w[c] = c - std::min(first[c], c); env += w[c] + 1; }
  const double fill = env / (0.5 * (double)n_cams * (n_cams + 1));
  if (fill > 0.35) return false;
  std::vector<int> ws(w);
  std::nth_element(ws.begin(), ws.begin() + (size_t)(0.9 * (n_cams - 1)), ws.end());
  const int bw = ws[(size_t)(0.9 * (n_cams - 1))];  // typical half-bandwidth in cameras (closure rows are outliers)
  int m = ((n_cams - bw) / 2) / CAM_ALIGN * CAM_ALIGN;
  if (m < CAM_ALIGN) return false;
  // cameras of the second part coupled to the first part: ordered last
  std::vector<char> in_tail((size_t)n_cams, 0);
  int n_tail = 0;
  for (int c = m; c < n_cams; ++c)
    if (first[c] < m) { in_tail[c] = 1; ++n_tail; }
  // the tail must start on a panel boundary as well: move the nearest cameras of the second chain into it
  int rest = n_cams - m - n_tail;
  for (int c = m; c < n_cams && rest % CAM_ALIGN != 0; ++c)
    if (!in_tail[c]) { in_tail[c] = 1; ++n_tail; --rest; }
  if (rest < CAM_ALIGN) return false;
  // critical path (in cameras) of the two-chain schedule against the single chain
  if ((double)(std::max(m, rest) + n_tail) > 0.8 * n_cams) return false;
  new_of_old.resize((size_t)n_cams);
  int pos = 0;
  for (int c = 0; c < m; ++c) new_of_old[c] = pos++;
  for (int c = n_cams - 1; c >= m; --c)
    if (!in_tail[c]) new_of_old[c] = pos++;
  for (int c = n_cams - 1; c >= m; --c)
    if (in_tail[c]) new_of_old[c] = pos++;
  *split_cams = m;
  if (tail_cams) *tail_cams = n_cams - n_tail;
  return true;
}

// 128-block pattern of S (lower triangle, row-major P x P bytes) in the INTERNAL camera order, from the caller-order
// envelope: cameras c and d (first[c] <= d <= c) may share points.  A superset of the true pattern.
void block_pattern_from_envelope(int n_cams, const int* first, const int* new_of_old /*null = identity*/, std::vector<uint8_t>& pattern, int* panels) {
  const int n = 6 * n_cams, P = (n + NB - 1) / NB;
  *panels = P;
  pattern.assign((size_t)P * P, 0);
  for (int c = 0; c < n_cams; ++c) {
    const int ic = new_of_old ? new_of_old[c] : c;
    for (int d = std::min(first[c], c); d <= c; ++d) {
      const int id = new_of_old ? new_of_old[d] : d;
      const int hi = std::max(ic, id), lo = std::min(ic, id);
      const int bi0 = (6 * hi) / NB, bi1 = (6 * hi + 5) / NB, bj0 = (6 * lo) / NB, bj1 = (6 * lo + 5) / NB;
      for (int bi = bi0; bi <= bi1; ++bi)
        for (int bj = bj0; bj <= std::min(bj1, bi); ++bj) pattern[(size_t)bi * P + bj] = 1;
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
