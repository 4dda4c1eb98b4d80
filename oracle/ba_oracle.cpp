// ba_oracle.cpp — CPU ORACLE (test infrastructure; see ba_oracle.h).  PARITY UNPINNED by the
// reference's own tests (it has none); cross-checked in tests/ against cv2 / mpmath / numpy
// golden vectors under tests/golden/.
//
// Build: g++ -O2 -ffp-contract=off -fopenmp (no -ffast-math, no FMA contraction: the dual-number
// evaluation must round like a plain IEEE build of the reference would).
#include "ba_oracle.h"
#include "ba_oracle_jet.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using namespace ba_oracle;

// The reference functor body, CTracker.cpp:585-604 (identical in :615-636 and :647-668 apart
// from which arguments are constants).  k = Matx33d::val row-major; k[0],k[1],k[2],k[4],k[5] used.
// dist = {k1,k2} radial extension or NULL; with NULL the code path is the
// reference's statement for statement.
template <typename T>
inline void reprojection(const T R[3], const T t[3], const T point[3], const double* k, const double* dist,
                         double obs_x, double obs_y, T residuals[2]) {
  T p[3];
  AngleAxisRotatePoint(R, point, p);       // :588
  p[0] += t[0]; p[1] += t[1]; p[2] += t[2];  // :590
  T xp = p[0] / p[2];                      // :593
  T yp = p[1] / p[2];                      // :594
  if (dist) {
    const T r2 = xp * xp + yp * yp;
    const T f = T(1.0) + T(dist[0]) * r2 + T(dist[1]) * (r2 * r2);
    xp = xp * f;
    yp = yp * f;
  }
  T predicted_x = T(k[0]) * xp + T(k[1]) * yp + T(k[2]);  // :597
  T predicted_y = T(k[4]) * yp + T(k[5]);                 // :598
  residuals[0] = predicted_x - T(obs_x);                  // :601
  residuals[1] = predicted_y - T(obs_y);                  // :602
}

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ------------------------------------------------------------------------------------------
struct Problem {
  int mode = 2;
  int64_t n_obs = 0, n_pts = 0;
  int32_t n_cams = 0;
  const double* obs = nullptr;
  const int32_t* cam_idx = nullptr;
  const int32_t* pt_idx = nullptr;
  const double* K9 = nullptr;
  const double* dist = nullptr;
  int threads = 1;
  // constant parameter blocks: what ceres::Problem::SetParameterBlockConstant does to a camera
  // (both its R and t blocks) or a point — the block keeps its value and is not a column of the Jacobian
  const uint8_t* cam_fixed = nullptr;  // [n_cams] or null
  const uint8_t* pt_fixed = nullptr;   // [n_pts] or null
  // residual + Jacobian evaluation through a caller-supplied callback instead of the built-in restatement of the
  // functor: oracle/_ref plugs the reference's OWN functor text in here (oracle/make_ref.py)
  ba_oracle_eval_fn eval_fn = nullptr;
  void* eval_user = nullptr;
  bool cam_is_fixed(int64_t c) const { return cam_fixed && cam_fixed[c]; }
  bool pt_is_fixed(int64_t p) const { return pt_fixed && pt_fixed[p]; }
  // CSR by point
  std::vector<int64_t> pt_start;
  std::vector<int64_t> pt_obs;
  std::vector<uint8_t> cam_used;  // parameter blocks exist only for cameras / points some residual references (CTracker.cpp:676-696)
  void build_index() {
    cam_used.assign((size_t)n_cams, 0);
    for (int64_t i = 0; i < n_obs; ++i) cam_used[cam_idx[i]] = 1;
    pt_start.assign(n_pts + 1, 0);
    for (int64_t i = 0; i < n_obs; ++i) pt_start[pt_idx[i] + 1]++;
    for (int64_t p = 0; p < n_pts; ++p) pt_start[p + 1] += pt_start[p];
    pt_obs.resize(n_obs);
    std::vector<int64_t> cur(pt_start.begin(), pt_start.end() - 1);
    for (int64_t i = 0; i < n_obs; ++i) pt_obs[cur[pt_idx[i]]++] = i;
  }
  bool cams_free() const { return mode != 0; }
  bool pts_free() const { return mode != 1; }
};

int validate(int64_t n_obs, const double* obs, const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
             const double* K9, const double* cam_rt, int64_t n_pts, const double* pts, bool external_eval = false) {
  if (n_obs < 0 || n_cams < 0 || n_pts < 0) return -1;
  if (n_obs > 0 && (!cam_idx || !pt_idx || !cam_rt || !pts)) return -1;
  if (n_obs > 0 && !external_eval && (!obs || !K9)) return -1;
  for (int64_t i = 0; i < n_obs; ++i)
    if (cam_idx[i] < 0 || cam_idx[i] >= n_cams || pt_idx[i] < 0 || pt_idx[i] >= n_pts) return -1;
  return 0;
}

// r, J (unscaled) — Jc [n_obs][12] = rows (du/d[R t], dv/d[R t]); Jp [n_obs][6] = rows (du/dX, dv/dX).
// A residual whose parameter blocks are ALL constant is not part of the reduced program (Ceres removes it and
// carries its cost as Summary::fixed_cost): it adds nothing to the returned cost; *fixed_cost (optional) receives
// 1/2 sum of those.  Jacobian columns of constant blocks are zeroed (they do not exist in the reduced program).
double evaluate(const Problem& P, const double* cam, const double* pts, double* r, double* Jc, double* Jp, double* fixed_cost = nullptr) {
  double cost = 0.0, fcost = 0.0;
  const bool want_j = (Jc != nullptr);
  if (P.eval_fn) {
    std::vector<double> rtmp;
    if (!r) { rtmp.resize((size_t)P.n_obs * 2); r = rtmp.data(); }
    if (P.eval_fn(P.eval_user, cam, pts, r, Jc, Jp)) return std::numeric_limits<double>::quiet_NaN();
    for (int64_t i = 0; i < P.n_obs; ++i) {
      const bool cf = !P.cams_free() || P.cam_is_fixed(P.cam_idx[i]), pf = !P.pts_free() || P.pt_is_fixed(P.pt_idx[i]);
      const double e = r[2 * i] * r[2 * i] + r[2 * i + 1] * r[2 * i + 1];
      if (cf && pf) fcost += e; else cost += e;
      if (want_j && P.cam_is_fixed(P.cam_idx[i])) for (int j = 0; j < 12; ++j) Jc[12 * i + j] = 0.0;
      if (want_j && P.pt_is_fixed(P.pt_idx[i])) for (int j = 0; j < 6; ++j) Jp[6 * i + j] = 0.0;
    }
    if (fixed_cost) *fixed_cost = 0.5 * fcost;
    return 0.5 * cost;
  }
#pragma omp parallel for schedule(static) reduction(+ : cost, fcost) num_threads(P.threads)
  for (int64_t i = 0; i < P.n_obs; ++i) {
    const int c = P.cam_idx[i];
    const int64_t p = P.pt_idx[i];
    const bool both_fixed = (!P.cams_free() || P.cam_is_fixed(c)) && (!P.pts_free() || P.pt_is_fixed(p)) && (P.cam_fixed || P.pt_fixed);
    const double* k = P.K9 + 9 * (int64_t)c;
    const double* d = P.dist ? P.dist + 2 * (int64_t)c : nullptr;
    if (want_j) {
      typedef Jet<9> J9;
      J9 R[3], t[3], X[3], res[2];
      for (int j = 0; j < 3; ++j) {
        R[j] = J9(cam[6 * c + j], j);
        t[j] = J9(cam[6 * c + 3 + j], 3 + j);
        X[j] = J9(pts[3 * p + j], 6 + j);
      }
      reprojection(R, t, X, k, d, P.obs[2 * i], P.obs[2 * i + 1], res);
      r[2 * i] = res[0].a;
      r[2 * i + 1] = res[1].a;
      const bool cfx = P.cam_is_fixed(c), pfx = P.pt_is_fixed(p);
      for (int row = 0; row < 2; ++row) {
        for (int j = 0; j < 6; ++j) Jc[12 * i + 6 * row + j] = cfx ? 0.0 : res[row].v[j];
        for (int j = 0; j < 3; ++j) Jp[6 * i + 3 * row + j] = pfx ? 0.0 : res[row].v[6 + j];
      }
      const double e = res[0].a * res[0].a + res[1].a * res[1].a;
      if (both_fixed) fcost += e; else cost += e;
    } else {
      double res[2];
      reprojection<double>(cam + 6 * c, cam + 6 * c + 3, pts + 3 * p, k, d, P.obs[2 * i], P.obs[2 * i + 1], res);
      if (r) { r[2 * i] = res[0]; r[2 * i + 1] = res[1]; }
      const double e = res[0] * res[0] + res[1] * res[1];
      if (both_fixed) fcost += e; else cost += e;
    }
  }
  if (fixed_cost) *fixed_cost = 0.5 * fcost;
  return 0.5 * cost;
}

// 3x3 SPD inverse through Cholesky (Ceres' InvertPSDMatrix uses LLT).  H row-major full.  false if not PD.
bool inv3_spd(const double* H, double* Hi) {
  const double l00s = H[0]; if (!(l00s > 0.0)) return false;
  const double l00 = std::sqrt(l00s);
  const double l10 = H[3] / l00, l20 = H[6] / l00;
  const double l11s = H[4] - l10 * l10; if (!(l11s > 0.0)) return false;
  const double l11 = std::sqrt(l11s);
  const double l21 = (H[7] - l20 * l10) / l11;
  const double l22s = H[8] - l20 * l20 - l21 * l21; if (!(l22s > 0.0)) return false;
  const double l22 = std::sqrt(l22s);
  // inverse of L (lower)
  const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
  const double i10 = -l10 * i00 * i11;
  const double i21 = -l21 * i11 * i22;
  const double i20 = -(l20 * i00 + l21 * i10) * i22;
  // Hi = Li^T Li
  Hi[0] = i00 * i00 + i10 * i10 + i20 * i20;
  Hi[1] = Hi[3] = i10 * i11 + i20 * i21;
  Hi[2] = Hi[6] = i20 * i22;
  Hi[4] = i11 * i11 + i21 * i21;
  Hi[5] = Hi[7] = i21 * i22;
  Hi[8] = i22 * i22;
  return true;
}

}  // namespace

// Blocked Cholesky lives in ba_oracle_chol.cpp (compiled with vector flags; rounding there does
// not have to mimic the reference functor).
extern "C" int ba_oracle_cholesky(double* A, int64_t n, int num_threads);

namespace {

// n×n SPD solve, in place on a copy-free lower factor: A row-major lower = L; b <- solution.
void chol_solve_inplace(const double* L, int64_t n, double* b) {
  for (int64_t i = 0; i < n; ++i) {
    double s = b[i];
    const double* Li = L + i * n;
    for (int64_t j = 0; j < i; ++j) s -= Li[j] * b[j];
    b[i] = s / Li[i];
  }
  // back substitution with L^T: column access; do it row-oriented (axpy form) to stay contiguous
  for (int64_t i = n - 1; i >= 0; --i) {
    const double* Li = L + i * n;
    const double xi = b[i] / Li[i];
    b[i] = xi;
    for (int64_t j = 0; j < i; ++j) b[j] -= Li[j] * xi;
  }
}

// small dense SPD solve used for 6x6 (mode 1) — unblocked, returns false if not PD
bool small_spd_solve(double* A, int n, double* b) {
  for (int j = 0; j < n; ++j) {
    double d = A[j * n + j];
    for (int k = 0; k < j; ++k) d -= A[j * n + k] * A[j * n + k];
    if (!(d > 0.0)) return false;
    d = std::sqrt(d);
    A[j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = A[i * n + j];
      for (int k = 0; k < j; ++k) s -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = s / d;
    }
  }
  for (int i = 0; i < n; ++i) {
    double s = b[i];
    for (int j = 0; j < i; ++j) s -= A[i * n + j] * b[j];
    b[i] = s / A[i * n + i];
  }
  for (int i = n - 1; i >= 0; --i) {
    double s = b[i];
    for (int j = i + 1; j < n; ++j) s -= A[j * n + i] * b[j];
    b[i] = s / A[i * n + i];
  }
  return true;
}

bool small_spd_inverse(const double* A, int n, double* Ainv) {
  std::vector<double> L(A, A + n * n), e(n);
  // factor once
  for (int j = 0; j < n; ++j) {
    double d = L[j * n + j];
    for (int k = 0; k < j; ++k) d -= L[j * n + k] * L[j * n + k];
    if (!(d > 0.0)) return false;
    d = std::sqrt(d);
    L[j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = L[i * n + j];
      for (int k = 0; k < j; ++k) s -= L[i * n + k] * L[j * n + k];
      L[i * n + j] = s / d;
    }
  }
  for (int c = 0; c < n; ++c) {
    for (int i = 0; i < n; ++i) e[i] = (i == c) ? 1.0 : 0.0;
    for (int i = 0; i < n; ++i) {
      double s = e[i];
      for (int j = 0; j < i; ++j) s -= L[i * n + j] * e[j];
      e[i] = s / L[i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
      double s = e[i];
      for (int j = i + 1; j < n; ++j) s -= L[j * n + i] * e[j];
      e[i] = s / L[i * n + i];
    }
    for (int i = 0; i < n; ++i) Ainv[i * n + c] = e[i];
  }
  return true;
}

// Normal-equation blocks from (scaled) J and r.  Hcc [Nc][36] full symmetric, gc [Nc][6], Hpp [Np][9], gp [Np][3].
// No LM diagonal added here.
void form_blocks(const Problem& P, const double* r, const double* Jc, const double* Jp, std::vector<double>& Hcc,
                 std::vector<double>& gc, std::vector<double>& Hpp, std::vector<double>& gp) {
  Hcc.assign((size_t)P.n_cams * 36, 0.0);
  gc.assign((size_t)P.n_cams * 6, 0.0);
  Hpp.assign((size_t)P.n_pts * 9, 0.0);
  gp.assign((size_t)P.n_pts * 3, 0.0);
  for (int64_t i = 0; i < P.n_obs; ++i) {
    const int c = P.cam_idx[i];
    const int64_t p = P.pt_idx[i];
    const double* jc = Jc + 12 * i;
    const double* jp = Jp + 6 * i;
    const double r0 = r[2 * i], r1 = r[2 * i + 1];
    if (P.cams_free()) {
      double* H = &Hcc[(size_t)c * 36];
      double* g = &gc[(size_t)c * 6];
      for (int a = 0; a < 6; ++a) {
        for (int b = 0; b < 6; ++b) H[a * 6 + b] += jc[a] * jc[b] + jc[6 + a] * jc[6 + b];
        g[a] += jc[a] * r0 + jc[6 + a] * r1;
      }
    }
    if (P.pts_free()) {
      double* H = &Hpp[(size_t)p * 9];
      double* g = &gp[(size_t)p * 3];
      for (int a = 0; a < 3; ++a) {
        for (int b = 0; b < 3; ++b) H[a * 3 + b] += jp[a] * jp[b] + jp[3 + a] * jp[3 + b];
        g[a] += jp[a] * r0 + jp[3 + a] * r1;
      }
    }
  }
}

struct LinearResult {
  bool ok = false;
  int iterations = 0;
};

// Solve (J^T J + D^2) y = J^T r by eliminating the points.
// Dc2 [6Nc], Dp2 [3Np] = squared LM diagonal.  Outputs yc [6Nc], yp [3Np].
// S_out/rhs_out (optional): dense S (both triangles) and reduced rhs, for the KAT tap.
LinearResult solve_schur(const Problem& P, const ba_oracle_options& o, const double* Jc, const double* Jp,
                         const std::vector<double>& Hcc, const std::vector<double>& gc, const std::vector<double>& Hpp,
                         const std::vector<double>& gp, const double* Dc2, const double* Dp2, double* yc, double* yp,
                         double* S_out, double* rhs_out, bool factor_and_solve) {
  LinearResult res;
  const int64_t n = 6 * (int64_t)P.n_cams;
  const int64_t Np = P.n_pts;
  std::vector<double> Hinv((size_t)Np * 9);
  std::vector<double> zp((size_t)Np * 3);  // Hpp^-1 g_p
  bool pd = true;
#pragma omp parallel for schedule(static) num_threads(P.threads)
  for (int64_t p = 0; p < Np; ++p) {
    double H[9];
    for (int k = 0; k < 9; ++k) H[k] = Hpp[9 * p + k];
    H[0] += Dp2[3 * p]; H[4] += Dp2[3 * p + 1]; H[8] += Dp2[3 * p + 2];
    if (!inv3_spd(H, &Hinv[9 * p])) {
#pragma omp atomic write
      pd = false;
      continue;
    }
    const double* Hi = &Hinv[9 * p];
    for (int a = 0; a < 3; ++a) zp[3 * p + a] = Hi[3 * a] * gp[3 * p] + Hi[3 * a + 1] * gp[3 * p + 1] + Hi[3 * a + 2] * gp[3 * p + 2];
  }
  if (!pd) return res;

  // W_i = Jc_i^T Jp_i (6x3)
  auto W_of = [&](int64_t i, double* W) {
    const double* jc = Jc + 12 * i;
    const double* jp = Jp + 6 * i;
    for (int a = 0; a < 6; ++a)
      for (int b = 0; b < 3; ++b) W[a * 3 + b] = jc[a] * jp[b] + jc[6 + a] * jp[3 + b];
  };

  // reduced rhs
  std::vector<double> rhs(n);
  for (int64_t k = 0; k < n; ++k) rhs[k] = gc[k];
  for (int64_t i = 0; i < P.n_obs; ++i) {
    double W[18];
    W_of(i, W);
    const int c = P.cam_idx[i];
    const double* z = &zp[3 * (int64_t)P.pt_idx[i]];
    for (int a = 0; a < 6; ++a) rhs[6 * c + a] -= W[a * 3] * z[0] + W[a * 3 + 1] * z[1] + W[a * 3 + 2] * z[2];
  }
  if (rhs_out) std::memcpy(rhs_out, rhs.data(), sizeof(double) * n);

  const bool use_pcg = (o.linear_solver == 2) && !S_out;
  if (!use_pcg) {
    // ---- explicit dense S (lower triangle accumulated, mirrored at the end) ----
    std::vector<double> S;
    try { S.assign((size_t)n * n, 0.0); } catch (...) { return res; }
    for (int c = 0; c < P.n_cams; ++c)
      for (int a = 0; a < 6; ++a)
        for (int b = 0; b < 6; ++b) S[(size_t)(6 * c + a) * n + 6 * c + b] = Hcc[(size_t)c * 36 + a * 6 + b] + (a == b ? Dc2[6 * c + a] : 0.0);
#pragma omp parallel for schedule(dynamic, 64) num_threads(P.threads)
    for (int64_t p = 0; p < Np; ++p) {
      const int64_t b0 = P.pt_start[p], b1 = P.pt_start[p + 1];
      const double* Hi = &Hinv[9 * p];
      for (int64_t ii = b0; ii < b1; ++ii) {
        const int64_t i = P.pt_obs[ii];
        const int ci = P.cam_idx[i];
        double Wi[18], A[18];
        W_of(i, Wi);
        for (int a = 0; a < 6; ++a)
          for (int b = 0; b < 3; ++b) A[a * 3 + b] = Wi[a * 3] * Hi[b] + Wi[a * 3 + 1] * Hi[3 + b] + Wi[a * 3 + 2] * Hi[6 + b];
        for (int64_t jj = b0; jj < b1; ++jj) {
          const int64_t j = P.pt_obs[jj];
          const int cj = P.cam_idx[j];
          if (cj > ci) continue;  // lower triangle of blocks (ci >= cj)
          double Wj[18];
          W_of(j, Wj);
          for (int a = 0; a < 6; ++a)
            for (int b = 0; b < 6; ++b) {
              if (ci == cj && b > a) continue;
              const double v = A[a * 3] * Wj[b * 3] + A[a * 3 + 1] * Wj[b * 3 + 1] + A[a * 3 + 2] * Wj[b * 3 + 2];
              double* dst = &S[(size_t)(6 * ci + a) * n + 6 * cj + b];
              if (P.threads > 1) {
#pragma omp atomic
                *dst -= v;
              } else {
                *dst -= v;
              }
            }
        }
      }
    }
    if (S_out) {
      for (int64_t a = 0; a < n; ++a)
        for (int64_t b = 0; b <= a; ++b) { S_out[a * n + b] = S[a * n + b]; S_out[b * n + a] = S[a * n + b]; }
    }
    if (!factor_and_solve) { res.ok = true; return res; }
    if (ba_oracle_cholesky(S.data(), n, P.threads) != 0) return res;
    for (int64_t k = 0; k < n; ++k) yc[k] = rhs[k];
    chol_solve_inplace(S.data(), n, yc);
  } else {
    // ---- implicit Schur complement, block-Jacobi PCG (deviation used only where dense S is
    //      infeasible; equals Ceres ITERATIVE_SCHUR + JACOBI in exact arithmetic)
    const int Nc = P.n_cams;
    std::vector<double> Minv((size_t)Nc * 36);
    {
      std::vector<double> Sd((size_t)Nc * 36);
      for (int c = 0; c < Nc; ++c)
        for (int a = 0; a < 6; ++a)
          for (int b = 0; b < 6; ++b) Sd[(size_t)c * 36 + a * 6 + b] = Hcc[(size_t)c * 36 + a * 6 + b] + (a == b ? Dc2[6 * c + a] : 0.0);
      for (int64_t i = 0; i < P.n_obs; ++i) {
        double W[18], A[18];
        W_of(i, W);
        const double* Hi = &Hinv[9 * (int64_t)P.pt_idx[i]];
        for (int a = 0; a < 6; ++a)
          for (int b = 0; b < 3; ++b) A[a * 3 + b] = W[a * 3] * Hi[b] + W[a * 3 + 1] * Hi[3 + b] + W[a * 3 + 2] * Hi[6 + b];
        double* D = &Sd[(size_t)P.cam_idx[i] * 36];
        for (int a = 0; a < 6; ++a)
          for (int b = 0; b < 6; ++b) D[a * 6 + b] -= A[a * 3] * W[b * 3] + A[a * 3 + 1] * W[b * 3 + 1] + A[a * 3 + 2] * W[b * 3 + 2];
      }
      for (int c = 0; c < Nc; ++c)
        if (!small_spd_inverse(&Sd[(size_t)c * 36], 6, &Minv[(size_t)c * 36])) return res;
    }
    std::vector<double> x(n, 0.0), rr(rhs), z(n), pv(n), q(n), vp((size_t)Np * 3);
    auto apply_M = [&](const std::vector<double>& in, std::vector<double>& out) {
      for (int c = 0; c < Nc; ++c)
        for (int a = 0; a < 6; ++a) {
          double s = 0.0;
          for (int b = 0; b < 6; ++b) s += Minv[(size_t)c * 36 + a * 6 + b] * in[6 * c + b];
          out[6 * c + a] = s;
        }
    };
    auto apply_S = [&](const std::vector<double>& in, std::vector<double>& out) {
      // out = (Hcc + Dc2) in - Jc^T Jp Hpp^-1 Jp^T Jc in
      std::fill(vp.begin(), vp.end(), 0.0);
      std::vector<double> u((size_t)P.n_obs * 2);
      for (int64_t i = 0; i < P.n_obs; ++i) {
        const double* jc = Jc + 12 * i; const double* jp = Jp + 6 * i;
        const double* pc = &in[6 * (int64_t)P.cam_idx[i]];
        double u0 = 0, u1 = 0;
        for (int a = 0; a < 6; ++a) { u0 += jc[a] * pc[a]; u1 += jc[6 + a] * pc[a]; }
        u[2 * i] = u0; u[2 * i + 1] = u1;
        double* v = &vp[3 * (int64_t)P.pt_idx[i]];
        for (int b = 0; b < 3; ++b) v[b] += jp[b] * u0 + jp[3 + b] * u1;
      }
      for (int64_t k = 0; k < n; ++k) out[k] = Dc2[k] * in[k];
      for (int64_t i = 0; i < P.n_obs; ++i) {
        const double* jc = Jc + 12 * i; const double* jp = Jp + 6 * i;
        const int64_t p = P.pt_idx[i];
        const double* Hi = &Hinv[9 * p]; const double* v = &vp[3 * p];
        double zz[3];
        for (int a = 0; a < 3; ++a) zz[a] = Hi[3 * a] * v[0] + Hi[3 * a + 1] * v[1] + Hi[3 * a + 2] * v[2];
        const double y0 = u[2 * i] - (jp[0] * zz[0] + jp[1] * zz[1] + jp[2] * zz[2]);
        const double y1 = u[2 * i + 1] - (jp[3] * zz[0] + jp[4] * zz[1] + jp[5] * zz[2]);
        double* oc = &out[6 * (int64_t)P.cam_idx[i]];
        for (int a = 0; a < 6; ++a) oc[a] += jc[a] * y0 + jc[6 + a] * y1;
      }
    };
    apply_M(rr, z);
    pv = z;
    double rho = 0.0;
    for (int64_t k = 0; k < n; ++k) rho += rr[k] * z[k];
    const double rho0 = rho;
    int it = 0;
    if (rho0 > 0.0) {
      for (; it < o.pcg_max_iterations;) {
        apply_S(pv, q);
        double pq = 0.0;
        for (int64_t k = 0; k < n; ++k) pq += pv[k] * q[k];
        if (!(pq > 0.0)) break;
        const double alpha = rho / pq;
        for (int64_t k = 0; k < n; ++k) { x[k] += alpha * pv[k]; rr[k] -= alpha * q[k]; }
        apply_M(rr, z);
        double rho_new = 0.0;
        for (int64_t k = 0; k < n; ++k) rho_new += rr[k] * z[k];
        ++it;
        if (rho_new <= o.pcg_tolerance * o.pcg_tolerance * rho0) { rho = rho_new; break; }
        const double beta = rho_new / rho;
        rho = rho_new;
        for (int64_t k = 0; k < n; ++k) pv[k] = z[k] + beta * pv[k];
      }
    }
    res.iterations = it;
    for (int64_t k = 0; k < n; ++k) yc[k] = x[k];
  }
  // back-substitution y_p = Hpp^-1 (g_p - sum_i W_i^T y_c)
  std::vector<double> acc((size_t)Np * 3, 0.0);
  for (int64_t i = 0; i < P.n_obs; ++i) {
    double W[18];
    W_of(i, W);
    const double* y = yc + 6 * (int64_t)P.cam_idx[i];
    double* a3 = &acc[3 * (int64_t)P.pt_idx[i]];
    for (int b = 0; b < 3; ++b) {
      double s = 0.0;
      for (int a = 0; a < 6; ++a) s += W[a * 3 + b] * y[a];
      a3[b] += s;
    }
  }
  for (int64_t p = 0; p < Np; ++p) {
    const double* Hi = &Hinv[9 * p];
    double g[3] = {gp[3 * p] - acc[3 * p], gp[3 * p + 1] - acc[3 * p + 1], gp[3 * p + 2] - acc[3 * p + 2]};
    for (int a = 0; a < 3; ++a) yp[3 * p + a] = Hi[3 * a] * g[0] + Hi[3 * a + 1] * g[1] + Hi[3 * a + 2] * g[2];
  }
  res.ok = true;
  return res;
}

// Column scaling at x (Ceres: 1/(1+sqrt(sum_i J_ij^2)) from the unscaled J of iteration 0).
void jacobi_scale(const Problem& P, const double* Jc, const double* Jp, bool enabled, std::vector<double>& sc,
                  std::vector<double>& sp) {
  sc.assign((size_t)P.n_cams * 6, 0.0);
  sp.assign((size_t)P.n_pts * 3, 0.0);
  for (int64_t i = 0; i < P.n_obs; ++i) {
    const double* jc = Jc + 12 * i; const double* jp = Jp + 6 * i;
    double* a = &sc[6 * (int64_t)P.cam_idx[i]];
    double* b = &sp[3 * (int64_t)P.pt_idx[i]];
    for (int k = 0; k < 6; ++k) a[k] += jc[k] * jc[k] + jc[6 + k] * jc[6 + k];
    for (int k = 0; k < 3; ++k) b[k] += jp[k] * jp[k] + jp[3 + k] * jp[3 + k];
  }
  for (auto& v : sc) v = enabled ? 1.0 / (1.0 + std::sqrt(v)) : 1.0;
  for (auto& v : sp) v = enabled ? 1.0 / (1.0 + std::sqrt(v)) : 1.0;
}

void apply_scale(const Problem& P, double* Jc, double* Jp, const std::vector<double>& sc, const std::vector<double>& sp) {
#pragma omp parallel for schedule(static) num_threads(P.threads)
  for (int64_t i = 0; i < P.n_obs; ++i) {
    const double* a = &sc[6 * (int64_t)P.cam_idx[i]];
    const double* b = &sp[3 * (int64_t)P.pt_idx[i]];
    for (int k = 0; k < 6; ++k) { Jc[12 * i + k] *= a[k]; Jc[12 * i + 6 + k] *= a[k]; }
    for (int k = 0; k < 3; ++k) { Jp[6 * i + k] *= b[k]; Jp[6 * i + 3 + k] *= b[k]; }
  }
}

}  // namespace

extern "C" {

void ba_oracle_default_options(ba_oracle_options* o) {
  std::memset(o, 0, sizeof(*o));
  o->max_num_iterations = 50;
  o->max_consecutive_invalid_steps = 5;
  o->jacobi_scaling = 1;
  o->linear_solver = 1;  // DENSE_SCHUR, CTracker.cpp:574
  o->dense_max_cams = 512;
  o->pcg_max_iterations = 1000;
  o->verbose = 0;
  o->num_threads = 1;  // the reference never sets num_threads (CTracker.cpp:571-577)
  o->initial_trust_region_radius = 1e4;
  o->max_trust_region_radius = 1e16;
  o->min_trust_region_radius = 1e-32;
  o->min_relative_decrease = 1e-3;
  o->min_lm_diagonal = 1e-6;
  o->max_lm_diagonal = 1e32;
  o->function_tolerance = 1e-6;
  o->gradient_tolerance = 1e-10;
  o->parameter_tolerance = 1e-8;
  o->pcg_tolerance = 1e-10;
}

int ba_oracle_eval(int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
                   const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts,
                   double* residuals, double* jac, double* cost, int num_threads) {
  if (validate(n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, cam_rt, n_pts, pts)) return -1;
  Problem P;
  P.n_obs = n_obs; P.n_pts = n_pts; P.n_cams = n_cams; P.obs = obs_xy; P.cam_idx = cam_idx; P.pt_idx = pt_idx;
  P.K9 = K9; P.dist = dist; P.threads = std::max(1, num_threads);
  std::vector<double> r((size_t)n_obs * 2), Jc, Jp;
  if (jac) { Jc.resize((size_t)n_obs * 12); Jp.resize((size_t)n_obs * 6); }
  const double c = evaluate(P, cam_rt, pts, r.data(), jac ? Jc.data() : nullptr, jac ? Jp.data() : nullptr);
  if (cost) *cost = c;
  if (residuals) std::memcpy(residuals, r.data(), sizeof(double) * 2 * n_obs);
  if (jac)
    for (int64_t i = 0; i < n_obs; ++i)
      for (int row = 0; row < 2; ++row) {
        for (int j = 0; j < 6; ++j) jac[18 * i + 9 * row + j] = Jc[12 * i + 6 * row + j];
        for (int j = 0; j < 3; ++j) jac[18 * i + 9 * row + 6 + j] = Jp[6 * i + 3 * row + j];
      }
  return 0;
}

int ba_oracle_schur(int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx, int32_t n_cams,
                    const double* K9, const double* dist, const double* cam_rt, int64_t n_pts, const double* pts,
                    const ba_oracle_options* opts, double radius, double* S_dense, double* rhs, double* Hcc_out,
                    double* gc_out, double* Hpp_out, double* gp_out, double* scale_out) {
  if (validate(n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, cam_rt, n_pts, pts) || !opts) return -1;
  Problem P;
  P.n_obs = n_obs; P.n_pts = n_pts; P.n_cams = n_cams; P.obs = obs_xy; P.cam_idx = cam_idx; P.pt_idx = pt_idx;
  P.K9 = K9; P.dist = dist; P.threads = 1;
  P.build_index();
  std::vector<double> r((size_t)n_obs * 2), Jc((size_t)n_obs * 12), Jp((size_t)n_obs * 6);
  evaluate(P, cam_rt, pts, r.data(), Jc.data(), Jp.data());
  std::vector<double> sc, sp;
  jacobi_scale(P, Jc.data(), Jp.data(), opts->jacobi_scaling != 0, sc, sp);
  apply_scale(P, Jc.data(), Jp.data(), sc, sp);
  std::vector<double> Hcc, gc, Hpp, gp;
  form_blocks(P, r.data(), Jc.data(), Jp.data(), Hcc, gc, Hpp, gp);
  std::vector<double> Dc2((size_t)n_cams * 6), Dp2((size_t)n_pts * 3);
  for (int c = 0; c < n_cams; ++c)
    for (int a = 0; a < 6; ++a)
      Dc2[6 * c + a] = std::min(std::max(Hcc[(size_t)c * 36 + a * 7], opts->min_lm_diagonal), opts->max_lm_diagonal) / radius;
  for (int64_t p = 0; p < n_pts; ++p)
    for (int a = 0; a < 3; ++a)
      Dp2[3 * p + a] = std::min(std::max(Hpp[(size_t)p * 9 + a * 4], opts->min_lm_diagonal), opts->max_lm_diagonal) / radius;
  std::vector<double> yc((size_t)n_cams * 6), yp((size_t)n_pts * 3);
  ba_oracle_options o = *opts;
  o.linear_solver = 1;
  std::vector<double> S_tmp;
  double* S_ptr = S_dense;
  if (!S_ptr) { S_tmp.resize((size_t)36 * n_cams * n_cams); S_ptr = S_tmp.data(); }
  LinearResult lr = solve_schur(P, o, Jc.data(), Jp.data(), Hcc, gc, Hpp, gp, Dc2.data(), Dp2.data(), yc.data(), yp.data(),
                                S_ptr, rhs, /*factor_and_solve=*/false);
  if (!lr.ok) return -2;
  if (Hcc_out)
    for (int c = 0; c < n_cams; ++c)
      for (int a = 0; a < 6; ++a)
        for (int b = 0; b < 6; ++b) Hcc_out[(size_t)c * 36 + a * 6 + b] = Hcc[(size_t)c * 36 + a * 6 + b] + (a == b ? Dc2[6 * c + a] : 0.0);
  if (gc_out) std::memcpy(gc_out, gc.data(), sizeof(double) * gc.size());
  if (Hpp_out)
    for (int64_t p = 0; p < n_pts; ++p)
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) Hpp_out[9 * p + a * 3 + b] = Hpp[9 * p + a * 3 + b] + (a == b ? Dp2[3 * p + a] : 0.0);
  if (gp_out) std::memcpy(gp_out, gp.data(), sizeof(double) * gp.size());
  if (scale_out) {
    std::memcpy(scale_out, sc.data(), sizeof(double) * sc.size());
    std::memcpy(scale_out + sc.size(), sp.data(), sizeof(double) * sp.size());
  }
  return 0;
}

int ba_oracle_solve(int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                    int32_t n_cams, const double* K9, const double* dist, double* cam_rt, int64_t n_pts, double* pts,
                    const ba_oracle_options* opts, ba_oracle_summary* summary) {
  return ba_oracle_solve_ex(mode, n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, dist, cam_rt, n_pts, pts, nullptr, nullptr, opts, summary,
                            nullptr, nullptr);
}

int ba_oracle_solve_ex(int mode, int64_t n_obs, const double* obs_xy, const int32_t* cam_idx, const int32_t* pt_idx,
                       int32_t n_cams, const double* K9, const double* dist, double* cam_rt, int64_t n_pts, double* pts,
                       const uint8_t* cam_fixed, const uint8_t* pt_fixed, const ba_oracle_options* opts,
                       ba_oracle_summary* summary, ba_oracle_eval_fn eval_fn, void* eval_user) {
  if (!opts) return -1;
  if (mode < 0 || mode > 2) return 0;  // CTracker.cpp:692-693: nothing is added, nothing happens
  if (validate(n_obs, obs_xy, cam_idx, pt_idx, n_cams, K9, cam_rt, n_pts, pts, eval_fn != nullptr)) return -1;
  const ba_oracle_options& o = *opts;
  const double t_begin = now_s();
  Problem P;
  P.mode = mode; P.n_obs = n_obs; P.n_pts = n_pts; P.n_cams = n_cams; P.obs = obs_xy; P.cam_idx = cam_idx;
  P.pt_idx = pt_idx; P.K9 = K9; P.dist = dist; P.threads = std::max(1, (int)o.num_threads);
  P.cam_fixed = cam_fixed; P.pt_fixed = pt_fixed; P.eval_fn = eval_fn; P.eval_user = eval_user;
  P.build_index();

  const int64_t nc6 = 6 * (int64_t)n_cams, np3 = 3 * n_pts;
  std::vector<double> xc(cam_rt, cam_rt + nc6), xp(pts, pts + np3);
  std::vector<double> cand_c(nc6), cand_p(np3);
  std::vector<double> r((size_t)n_obs * 2), Jc((size_t)n_obs * 12), Jp((size_t)n_obs * 6);
  std::vector<double> sc, sp, Hcc, gc, Hpp, gp;
  std::vector<double> diag_c(nc6), diag_p(np3), Dc2(nc6), Dp2(np3), yc(nc6, 0.0), yp(np3, 0.0);
  std::vector<double> step_c(nc6, 0.0), step_p(np3, 0.0);
  double eval_s = 0.0, lin_s = 0.0;

  ba_oracle_summary local;
  std::memset(&local, 0, sizeof(local));
  ba_oracle_summary& S = summary ? *summary : local;
  ba_oracle_iteration* trace = summary ? summary->trace : nullptr;
  const int trace_cap = summary ? summary->trace_capacity : 0;
  int trace_n = 0;
  auto push_trace = [&](const ba_oracle_iteration& it) { if (trace && trace_n < trace_cap) trace[trace_n] = it; ++trace_n; };

  bool have_scale = false;
  double x_cost = 0.0, gmax = 0.0, x_norm = 0.0, fixed_cost = 0.0;
  auto linearize = [&]() {
    const double t0 = now_s();
    x_cost = evaluate(P, xc.data(), xp.data(), r.data(), Jc.data(), Jp.data(), &fixed_cost);
    // gradient from the UNscaled Jacobian
    std::vector<double> Hc_, gc_, Hp_, gp_;
    if (!have_scale) {
      jacobi_scale(P, Jc.data(), Jp.data(), o.jacobi_scaling != 0, sc, sp);
      have_scale = true;
    }
    apply_scale(P, Jc.data(), Jp.data(), sc, sp);
    form_blocks(P, r.data(), Jc.data(), Jp.data(), Hcc, gc, Hpp, gp);
    gmax = 0.0;
    if (P.cams_free()) for (int64_t k = 0; k < nc6; ++k) gmax = std::max(gmax, std::fabs(gc[k] / sc[k]));
    if (P.pts_free()) for (int64_t k = 0; k < np3; ++k) gmax = std::max(gmax, std::fabs(gp[k] / sp[k]));
    double s = 0.0;
    // |x| over the parameter blocks of the reduced program: referenced by a residual and not constant
    if (P.cams_free()) for (int64_t k = 0; k < nc6; ++k) if (P.cam_used[k / 6] && !P.cam_is_fixed(k / 6)) s += xc[k] * xc[k];
    if (P.pts_free()) for (int64_t k = 0; k < np3; ++k) if (P.pt_start[k / 3 + 1] > P.pt_start[k / 3] && !P.pt_is_fixed(k / 3)) s += xp[k] * xp[k];
    x_norm = std::sqrt(s);
    eval_s += now_s() - t0;
  };

  linearize();
  const double setup_s = now_s() - t_begin;
  S.initial_cost = x_cost;
  double radius = o.initial_trust_region_radius;
  double decrease_factor = 2.0;
  bool reuse_diag = false;
  int invalid_run = 0;
  int num_success = 0, total_lin = 0;
  int term = 1;  // NO_CONVERGENCE
  {
    ba_oracle_iteration it0;
    std::memset(&it0, 0, sizeof(it0));
    it0.cost = x_cost; it0.gradient_max_norm = gmax; it0.trust_region_radius = radius; it0.step_is_valid = 1; it0.step_is_successful = 1;
    push_trace(it0);
  }
  int iter = 0;
  if (gmax <= o.gradient_tolerance) term = 0;
  else
    while (true) {
      if (iter >= o.max_num_iterations) { term = 1; break; }
      ++iter;
      ba_oracle_iteration its;
      std::memset(&its, 0, sizeof(its));
      const double tl0 = now_s();
      if (!reuse_diag) {
        for (int c = 0; c < n_cams; ++c)
          for (int a = 0; a < 6; ++a) diag_c[6 * c + a] = std::min(std::max(Hcc[(size_t)c * 36 + a * 7], o.min_lm_diagonal), o.max_lm_diagonal);
        for (int64_t p = 0; p < n_pts; ++p)
          for (int a = 0; a < 3; ++a) diag_p[3 * p + a] = std::min(std::max(Hpp[(size_t)p * 9 + a * 4], o.min_lm_diagonal), o.max_lm_diagonal);
      }
      // lm_diagonal = sqrt(diag/radius); the solver uses its square
      for (int64_t k = 0; k < nc6; ++k) { const double d = std::sqrt(diag_c[k] / radius); Dc2[k] = d * d; }
      for (int64_t k = 0; k < np3; ++k) { const double d = std::sqrt(diag_p[k] / radius); Dp2[k] = d * d; }
      reuse_diag = true;
      bool ok = true;
      int lin_its = 0;
      if (mode == 2) {
        LinearResult lr = solve_schur(P, o, Jc.data(), Jp.data(), Hcc, gc, Hpp, gp, Dc2.data(), Dp2.data(), yc.data(), yp.data(),
                                      nullptr, nullptr, true);
        ok = lr.ok; lin_its = lr.iterations;
      } else if (mode == 1) {
        for (int c = 0; c < n_cams && ok; ++c) {
          double A[36], b[6];
          for (int a = 0; a < 6; ++a) { for (int bb = 0; bb < 6; ++bb) A[a * 6 + bb] = Hcc[(size_t)c * 36 + a * 6 + bb]; A[a * 7] += Dc2[6 * c + a]; b[a] = gc[6 * c + a]; }
          bool any = false; for (int a = 0; a < 36; ++a) any |= (Hcc[(size_t)c * 36 + a] != 0.0);
          if (!any) { for (int a = 0; a < 6; ++a) yc[6 * c + a] = 0.0; continue; }
          ok = small_spd_solve(A, 6, b);
          for (int a = 0; a < 6; ++a) yc[6 * c + a] = b[a];
        }
      } else {
        for (int64_t p = 0; p < n_pts && ok; ++p) {
          double H[9], Hi[9];
          for (int k = 0; k < 9; ++k) H[k] = Hpp[9 * p + k];
          H[0] += Dp2[3 * p]; H[4] += Dp2[3 * p + 1]; H[8] += Dp2[3 * p + 2];
          if (P.pt_start[p] == P.pt_start[p + 1]) { yp[3 * p] = yp[3 * p + 1] = yp[3 * p + 2] = 0.0; continue; }
          ok = inv3_spd(H, Hi);
          for (int a = 0; a < 3; ++a) yp[3 * p + a] = Hi[3 * a] * gp[3 * p] + Hi[3 * a + 1] * gp[3 * p + 1] + Hi[3 * a + 2] * gp[3 * p + 2];
        }
      }
      total_lin += lin_its;
      its.linear_iterations = lin_its;
      double model_cost_change = 0.0;
      if (ok) {
        bool finite = true;
        if (P.cams_free()) for (int64_t k = 0; k < nc6; ++k) { step_c[k] = -yc[k]; finite &= std::isfinite(step_c[k]); } else std::fill(step_c.begin(), step_c.end(), 0.0);
        if (P.pts_free()) for (int64_t k = 0; k < np3; ++k) { step_p[k] = -yp[k]; finite &= std::isfinite(step_p[k]); } else std::fill(step_p.begin(), step_p.end(), 0.0);
        ok = finite;
      }
      if (ok) {
        double acc = 0.0;
        for (int64_t i = 0; i < n_obs; ++i) {
          const double* jc = &Jc[12 * i]; const double* jp = &Jp[6 * i];
          const double* dc = &step_c[6 * (int64_t)cam_idx[i]]; const double* dp = &step_p[3 * (int64_t)pt_idx[i]];
          double m0 = 0.0, m1 = 0.0;
          if (P.cams_free()) for (int a = 0; a < 6; ++a) { m0 += jc[a] * dc[a]; m1 += jc[6 + a] * dc[a]; }
          if (P.pts_free()) for (int a = 0; a < 3; ++a) { m0 += jp[a] * dp[a]; m1 += jp[3 + a] * dp[a]; }
          acc += m0 * (r[2 * i] + m0 / 2.0) + m1 * (r[2 * i + 1] + m1 / 2.0);
        }
        model_cost_change = -acc;
        if (!(model_cost_change > 0.0)) ok = false;
      }
      lin_s += now_s() - tl0;
      its.model_cost_change = model_cost_change;
      if (!ok) {
        // HandleInvalidStep / LevenbergMarquardtStrategy::StepIsInvalid
        ++invalid_run;
        its.step_is_valid = 0; its.cost = x_cost; its.gradient_max_norm = gmax;
        radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diag = true;
        its.trust_region_radius = radius;
        push_trace(its);
        if (invalid_run >= o.max_consecutive_invalid_steps) { term = 2; break; }
        if (radius < o.min_trust_region_radius) { term = 0; break; }
        continue;
      }
      invalid_run = 0;
      its.step_is_valid = 1;
      // candidate = x + step ∘ scale
      double dn = 0.0;
      cand_c = xc; cand_p = xp;
      if (P.cams_free()) for (int64_t k = 0; k < nc6; ++k) { const double d = step_c[k] * sc[k]; cand_c[k] = xc[k] + d; const double e = xc[k] - cand_c[k]; dn += e * e; }
      if (P.pts_free()) for (int64_t k = 0; k < np3; ++k) { const double d = step_p[k] * sp[k]; cand_p[k] = xp[k] + d; const double e = xp[k] - cand_p[k]; dn += e * e; }
      const double te0 = now_s();
      double cand_cost = evaluate(P, cand_c.data(), cand_p.data(), nullptr, nullptr, nullptr);
      eval_s += now_s() - te0;
      if (!std::isfinite(cand_cost)) cand_cost = std::numeric_limits<double>::max();
      its.step_norm = std::sqrt(dn);
      its.cost_change = x_cost - cand_cost;
      its.cost = x_cost; its.gradient_max_norm = gmax; its.trust_region_radius = radius;
      if (its.step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) { term = 0; push_trace(its); break; }
      if (std::fabs(its.cost_change) <= o.function_tolerance * x_cost) { term = 0; push_trace(its); break; }
      const double rho = its.cost_change / model_cost_change;
      its.relative_decrease = rho;
      if (rho > o.min_relative_decrease) {
        xc = cand_c; xp = cand_p;
        linearize();
        ++num_success;
        its.step_is_successful = 1;
        radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * rho - 1.0, 3));
        radius = std::min(o.max_trust_region_radius, radius);
        decrease_factor = 2.0; reuse_diag = false;
        its.cost = x_cost; its.gradient_max_norm = gmax; its.trust_region_radius = radius;
        push_trace(its);
        if (gmax <= o.gradient_tolerance) { term = 0; break; }
      } else {
        radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diag = true;
        its.trust_region_radius = radius;
        push_trace(its);
        if (radius < o.min_trust_region_radius) { term = 0; break; }
      }
      if (o.verbose)
        std::fprintf(stderr, "[oracle] it %3d cost %.9e dcost %.3e |g| %.3e |step| %.3e rho %.3e radius %.3e lin %d\n", iter, its.cost,
                     its.cost_change, its.gradient_max_norm, its.step_norm, its.relative_decrease, its.trust_region_radius, its.linear_iterations);
    }

  std::memcpy(cam_rt, xc.data(), sizeof(double) * nc6);
  std::memcpy(pts, xp.data(), sizeof(double) * np3);
  // ceres::Solver::Summary: initial/final cost = the reduced program's cost + fixed_cost (residuals whose blocks are all constant)
  S.initial_cost += fixed_cost;
  S.final_cost = x_cost + fixed_cost;
  S.fixed_cost = fixed_cost;
  S.setup_seconds = setup_s;
  S.rms_px = n_obs > 0 ? std::sqrt(2.0 * S.final_cost / (double)n_obs) : 0.0;
  S.num_iterations = iter;
  S.num_successful_steps = num_success;
  S.termination_type = term;
  S.linear_solver_used = (mode == 2 && o.linear_solver == 2) ? 2 : 1;
  S.total_linear_iterations = total_lin;
  S.n_threads = P.threads;
  S.eval_seconds = eval_s;
  S.linear_seconds = lin_s;
  S.total_seconds = now_s() - t_begin;
  S.trace_count = std::min(trace_n, trace_cap);
  return 0;
}

}  // extern "C"
