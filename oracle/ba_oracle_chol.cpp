// ba_oracle_chol.cpp — blocked right-looking Cholesky for the CPU ORACLE / CPU baseline
// (test infrastructure; see ba_oracle.h).  Stands in for the Eigen LLT that Ceres'
// DenseSchurComplementSolver calls (dense_linear_algebra_library_type = EIGEN);
// no BLAS/LAPACK is installed in this image.  Compiled with vector flags (rounding of the
// factorisation is not part of the functor parity).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {
constexpr int NB = 96;  // panel width

// unblocked lower Cholesky of the nb×nb diagonal block at A (leading dimension lda)
int potf2(double* A, int64_t lda, int nb) {
  for (int j = 0; j < nb; ++j) {
    double d = A[j * lda + j];
    for (int k = 0; k < j; ++k) d -= A[j * lda + k] * A[j * lda + k];
    if (!(d > 0.0) || !std::isfinite(d)) return j + 1;
    d = std::sqrt(d);
    A[j * lda + j] = d;
    const double inv = 1.0 / d;
    for (int i = j + 1; i < nb; ++i) {
      double s = A[i * lda + j];
      for (int k = 0; k < j; ++k) s -= A[i * lda + k] * A[j * lda + k];
      A[i * lda + j] = s * inv;
    }
  }
  return 0;
}
}  // namespace

extern "C" int ba_oracle_cholesky(double* A, int64_t n, int num_threads) {
  const int threads = std::max(1, num_threads);
  std::vector<double> PT;  // packed transposed panel: PT[k][row]  (kb × rows below)
  for (int64_t k0 = 0; k0 < n; k0 += NB) {
    const int kb = (int)std::min<int64_t>(NB, n - k0);
    double* Akk = A + k0 * n + k0;
    const int bad = potf2(Akk, n, kb);
    if (bad) return (int)(k0 + bad);
    const int64_t m = n - (k0 + kb);  // rows below the panel
    if (m <= 0) break;
    // TRSM: rows i > panel: L[i, k0:k0+kb] = A[i, k0:k0+kb] * Lkk^-T
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = k0 + kb; i < n; ++i) {
      double* Ai = A + i * n + k0;
      for (int j = 0; j < kb; ++j) {
        double s = Ai[j];
        const double* Lj = Akk + (int64_t)j * n;
        for (int k = 0; k < j; ++k) s -= Ai[k] * Lj[k];
        Ai[j] = s / Lj[j];
      }
    }
    // pack the panel transposed so the update's inner loop is a contiguous axpy over columns
    PT.resize((size_t)kb * m);
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < m; ++i) {
      const double* Ai = A + (k0 + kb + i) * n + k0;
      for (int k = 0; k < kb; ++k) PT[(size_t)k * m + i] = Ai[k];
    }
    // SYRK: trailing lower triangle  A[i, j] -= sum_k L[i,k] L[j,k]   (j <= i)
    const int64_t RB = 8;  // row micro-block
    const int64_t nblk = (m + RB - 1) / RB;
#pragma omp parallel for schedule(dynamic, 8) num_threads(threads)
    for (int64_t b = 0; b < nblk; ++b) {
      const int64_t i0 = b * RB, i1 = std::min(m, i0 + RB);
      for (int64_t i = i0; i < i1; ++i) {
        double* Ci = A + (k0 + kb + i) * n + (k0 + kb);
        const double* Li = A + (k0 + kb + i) * n + k0;
        const int64_t jn = i + 1;
        for (int k = 0; k < kb; ++k) {
          const double lik = Li[k];
          const double* pk = &PT[(size_t)k * m];
#pragma omp simd
          for (int64_t j = 0; j < jn; ++j) Ci[j] -= lik * pk[j];
        }
      }
    }
  }
  return 0;
}
