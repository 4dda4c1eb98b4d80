// Stand-in for ceres/ceres.h (Ceres Solver 1.12, absent from /root/reference and from this image) —
// CPU ORACLE / test infrastructure, see oracle/ba_oracle.h.  It is the subset of the public Ceres API
// that the bundle-adjustment section of the reference (CTracker.h:99-151, CTracker.cpp:571-702) uses:
//   ceres::CostFunction, ceres::AutoDiffCostFunction<Functor, 2, N0[, N1[, N2]]>  (CTracker.h:108-110,124-126,141-143)
//   ceres::Problem::AddResidualBlock(cost, NULL, x0[, x1[, x2]])                    (CTracker.cpp:682,686,690)
//   ceres::Solver::Options {linear_solver_type, minimizer_progress_to_stdout, logging_type}, ceres::DENSE_SCHUR,
//   ceres::LoggingType::SILENT                                                      (CTracker.cpp:574-576)
//   ceres::Solver::Summary, ceres::Solve(options, &problem, &summary)               (CTracker.cpp:700-701)
// AutoDiffCostFunction differentiates the functor with the oracle's dual numbers (ba_oracle::Jet, Ceres' jet.h
// rules); Solve packs the parameter blocks BY POINTER IDENTITY (what ceres::Problem does) and runs the oracle's
// Ceres-semantics trust-region loop (ba_oracle_solve_ex) with the problem's own cost functions as evaluator.
// The reference text compiled against this header is extracted verbatim by oracle/make_ref.py.
#ifndef REF_STANDIN_CERES_CERES_H_
#define REF_STANDIN_CERES_CERES_H_
#include <cstddef>
#include <cstdint>
#include <map>
#include <vector>

#include "../../ba_oracle.h"
#include "../../ba_oracle_jet.h"

namespace ceres {

template <int N> using Jet = ba_oracle::Jet<N>;

class LossFunction;  // only ever passed as NULL (CTracker.cpp:682,686,690)

class CostFunction {
 public:
  virtual ~CostFunction() {}
  // Ceres' contract: parameters[b] = block b; jacobians == NULL or jacobians[b] (may be NULL) = row-major
  // [num_residuals x block size]
  virtual bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const = 0;
  virtual int num_blocks() const = 0;
  virtual int block_size(int b) const = 0;
};

template <typename Functor, int kNumResiduals, int N0, int N1 = 0, int N2 = 0>
class AutoDiffCostFunction : public CostFunction {
 public:
  explicit AutoDiffCostFunction(Functor* f) : functor_(f) {}
  ~AutoDiffCostFunction() override { delete functor_; }
  int num_blocks() const override { return 1 + (N1 > 0) + (N2 > 0); }
  int block_size(int b) const override { return b == 0 ? N0 : b == 1 ? N1 : N2; }
  bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const override {
    if (!jacobians) return call(parameters[0], N1 ? parameters[1] : nullptr, N2 ? parameters[2] : nullptr, residuals);
    constexpr int N = N0 + N1 + N2;
    typedef Jet<N> J;
    J x0[N0], x1[N1 ? N1 : 1], x2[N2 ? N2 : 1], res[kNumResiduals];
    for (int j = 0; j < N0; ++j) x0[j] = J(parameters[0][j], j);
    for (int j = 0; j < N1; ++j) x1[j] = J(parameters[1][j], N0 + j);
    for (int j = 0; j < N2; ++j) x2[j] = J(parameters[2][j], N0 + N1 + j);
    if (!call(x0, x1, x2, res)) return false;
    const int off[3] = {0, N0, N0 + N1}, sz[3] = {N0, N1, N2};
    for (int r = 0; r < kNumResiduals; ++r) residuals[r] = res[r].a;
    for (int b = 0; b < num_blocks(); ++b) {
      if (!jacobians[b]) continue;
      for (int r = 0; r < kNumResiduals; ++r)
        for (int j = 0; j < sz[b]; ++j) jacobians[b][r * sz[b] + j] = res[r].v[off[b] + j];
    }
    return true;
  }

 private:
  template <typename T>
  bool call(const T* x0, const T* x1, const T* x2, T* res) const {
    if constexpr (N2 > 0) return (*functor_)(x0, x1, x2, res);
    else if constexpr (N1 > 0) return (*functor_)(x0, x1, res);
    else return (*functor_)(x0, res);
  }
  Functor* functor_;
};

struct ResidualBlock {
  CostFunction* cost;
  double* block[3];
  int n_blocks;
};

class Problem {
 public:
  Problem() {}
  ~Problem() { for (auto& rb : residual_blocks_) delete rb.cost; }  // ceres::Problem owns its cost functions
  void AddResidualBlock(CostFunction* c, LossFunction*, double* x0) { residual_blocks_.push_back({c, {x0, nullptr, nullptr}, 1}); }
  void AddResidualBlock(CostFunction* c, LossFunction*, double* x0, double* x1) { residual_blocks_.push_back({c, {x0, x1, nullptr}, 2}); }
  void AddResidualBlock(CostFunction* c, LossFunction*, double* x0, double* x1, double* x2) { residual_blocks_.push_back({c, {x0, x1, x2}, 3}); }
  const std::vector<ResidualBlock>& residual_blocks() const { return residual_blocks_; }

 private:
  Problem(const Problem&);
  std::vector<ResidualBlock> residual_blocks_;
};

enum LinearSolverType { DENSE_NORMAL_CHOLESKY, DENSE_QR, SPARSE_NORMAL_CHOLESKY, DENSE_SCHUR, SPARSE_SCHUR, ITERATIVE_SCHUR, CGNR };
enum LoggingType { SILENT, PER_MINIMIZER_ITERATION };

class Solver {
 public:
  struct Options {
    LinearSolverType linear_solver_type = SPARSE_NORMAL_CHOLESKY;
    bool minimizer_progress_to_stdout = false;
    LoggingType logging_type = PER_MINIMIZER_ITERATION;
    int num_threads = 1;
  };
  struct Summary {
    double initial_cost = 0.0, final_cost = 0.0, fixed_cost = 0.0;
    int num_successful_steps = 0, num_iterations = 0, termination_type = 0;
    int num_parameter_blocks = 0, num_residual_blocks = 0;
    std::vector<ba_oracle_iteration> iterations;
  };
};

// Defined in oracle/ref_entry.inc (compiled into oracle/_ref/libba_ref.so).
void Solve(const Solver::Options& options, Problem* problem, Solver::Summary* summary);

}  // namespace ceres
#endif
