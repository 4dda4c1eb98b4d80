// ba_solver.cu — Levenberg–Marquardt trust-region driver with Ceres' semantics (SURVEY.md
// Appendix A.3; replaces ceres::Solve at CTracker.cpp:700-701 with the options of
// CTracker.cpp:571-577).  All arithmetic runs in the kernels of ba_kernels.cu / ba_chol.cu; the
// host only sequences launches and reads back a 256-byte scalar block once per trust-region
// attempt to take the accept/reject decision.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "ba_internal.h"

namespace hsfm {
namespace {

struct Driver {
  Ctx& ctx;
  Problem& P;
  cudaStream_t s;
  const hsfm_ba_options& o;
  bool cams_free, pts_free;
  bool dense;
  int rc = 0;

  Driver(Ctx& c, const hsfm_ba_options& opt)
      : ctx(c), P(c.prob), s(c.stream), o(opt), cams_free(c.prob.mode != 0), pts_free(c.prob.mode != 1), dense(false) {}

  // phase timing: mark(p) says "phase p starts here"; elapsed times between consecutive marks are
  // summed per phase after the solve.
  std::vector<int> marks;
  void mark(int phase) {
    if (marks.size() >= ctx.phase_events.size()) {
      cudaEvent_t e;
      if (cudaEventCreate(&e) != cudaSuccess) return;
      ctx.phase_events.push_back(e);
    }
    cudaEventRecord(ctx.phase_events[marks.size()], s);
    marks.push_back(phase);
  }
  void collect(double* phase_ms) {
    for (size_t k = 0; k + 1 < marks.size(); ++k) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ctx.phase_events[k], ctx.phase_events[k + 1]) == cudaSuccess) phase_ms[marks[k]] += ms;
    }
  }

  bool multi() const { return ctx.comm.n_ranks > 1; }

  void chk(int r) { if (r && !rc) rc = r; }

  int read_back() {
    mark(5);
    HSFM_CUDA(cudaMemcpyAsync(P.scal_host, P.scal, sizeof(double) * 16, cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaMemcpyAsync(P.flags_host, P.flags, sizeof(int) * FLAG_SLOTS, cudaMemcpyDeviceToHost, s));
    HSFM_CUDA(cudaStreamSynchronize(s));
    HSFM_CUDA(cudaGetLastError());  // launch-configuration errors are not sticky: without this a failed launch would go unnoticed
    return 0;
  }

  // Jacobi column scaling from the unscaled Jacobian at the initial point (A.3 "iter 0").
  void compute_scaling() {
    mark(0);
    launch_cam_prep(P, P.cam_x, nullptr, P.rows_x, s);
    launch_rj(P, P.rows_x, P.pts_x, nullptr, /*write_j=*/false, /*form=*/true, s);
    launch_cam_major(P, P.rows_x, P.pts_x, nullptr, 0, s);
    if (multi()) chk(ctx.comm.allreduce_sum(P.Hcc, 21 * (size_t)P.n_cams, s));
    launch_jacobi_scale(P, o.jacobi_scaling != 0, s);
  }

  // r, J (scaled), cost, point blocks, camera blocks, gradient norm at the current x.
  void linearize() {
    mark(0);
    launch_cam_prep(P, P.cam_x, P.scale_c, P.rows_x, s);
    launch_rj(P, P.rows_x, P.pts_x, P.scale_p, true, true, s);
    launch_cam_major(P, P.rows_x, P.pts_x, P.scale_p, 0, s);
    if (multi()) {
      chk(ctx.comm.allreduce_sum(P.Hcc, 21 * (size_t)P.n_cams, s));
      chk(ctx.comm.allreduce_sum(P.gc, 6 * (size_t)P.n_cams, s));
    }
    launch_grad_norm(P, s);
    launch_finalize_scalars(P, s);
    if (multi()) {
      chk(ctx.comm.allreduce_sum(P.scal + S_COST, 1, s));
      chk(ctx.comm.allreduce_max(P.scal + S_GMAX_P, 1, s));
    }
  }

  // One linear solve + candidate evaluation.  Leaves scalars/flags on the host mirrors.
  int attempt(double radius, int* linear_its) {
    *linear_its = 0;
    mark(1);
    HSFM_CUDA(cudaMemsetAsync(P.flags, 0, sizeof(int) * FLAG_SLOTS, s));
    if (pts_free) launch_point_invert(P, radius, o.min_lm_diagonal, o.max_lm_diagonal, s);
    if (cams_free) launch_cam_diag(P, radius, o.min_lm_diagonal, o.max_lm_diagonal, s);
    const int n6 = 6 * P.n_cams;
    if (P.mode == 2) {
      const bool lead = ctx.comm.rank == 0;
      if (dense) {
        launch_dense_init(P, lead, s, (o.dense_flags & 1) ? nullptr : &ctx.chol_plan);
        launch_dense_schur(P, s, !(o.dense_flags & 2));
        launch_schur_rhs(P, lead, s);
        if (multi()) {
          const CholPlan& pl = ctx.chol_plan;
          if (pl.use && pl.n == n6 && !(o.dense_flags & 1)) {
            // only the envelope of S can be non-zero on any rank: sum that (packed behind the factorisation's
            // workspace) instead of the whole n x n matrix
            double* packed = P.S + (size_t)n6 * n6 + cholesky_workspace_doubles(n6) + (size_t)pl.tail_doubles();
            launch_envelope_pack(pl, P.S, packed, true, s);
            chk(ctx.comm.allreduce_sum(packed, (size_t)pl.packed_doubles(), s));
            launch_envelope_pack(pl, P.S, packed, false, s);
            P.launches += 2;
          } else {
            chk(ctx.comm.allreduce_sum(P.S, (size_t)n6 * n6, s));
          }
          chk(ctx.comm.allreduce_sum(P.rhs, n6, s));
        }
        mark(2);
        launch_cholesky(P.S, n6, P.S + (size_t)n6 * n6, P.flags + F_CHOL_FAIL, s, &ctx.chol_aux, &P.launches, (o.dense_flags & 1) ? nullptr : &ctx.chol_plan);
        mark(3);
        HSFM_CUDA(cudaMemcpyAsync(P.yc, P.rhs, sizeof(double) * n6, cudaMemcpyDeviceToDevice, s));
        launch_chol_solve(P.S, n6, P.S + (size_t)n6 * n6, P.yc, s, &P.launches, &ctx.chol_aux, (o.dense_flags & 1) ? nullptr : &ctx.chol_plan);
      } else {
        launch_cam_major(P, P.rows_x, P.pts_x, P.scale_p, 1, s);
        if (multi()) chk(ctx.comm.allreduce_sum(P.E, 21 * (size_t)P.n_cams, s));
        launch_precond_invert(P, true, s);
        launch_schur_rhs(P, lead, s);
        if (multi()) chk(ctx.comm.allreduce_sum(P.rhs, n6, s));
        mark(3);
        launch_pcg_init(P, s);
        const int batch = 8;
        int issued = 0;
        while (issued < o.pcg_max_iterations) {
          for (int b = 0; b < batch && issued < o.pcg_max_iterations; ++b, ++issued) {
            launch_schur_matvec(P, P.pcg_p, P.pcg_q, s);
            if (multi()) chk(ctx.comm.allreduce_sum(P.pcg_q, n6, s));
            launch_pcg_step(P, o.pcg_tolerance, o.pcg_max_iterations, s);
          }
          HSFM_CUDA(cudaMemcpyAsync(P.flags_host, P.flags, sizeof(int) * FLAG_SLOTS, cudaMemcpyDeviceToHost, s));
          HSFM_CUDA(cudaStreamSynchronize(s));
          if (P.flags_host[F_PCG_DONE]) break;
        }
        *linear_its = P.flags_host[F_PCG_ITERS];
        HSFM_CUDA(cudaMemcpyAsync(P.yc, P.pcg_x, sizeof(double) * n6, cudaMemcpyDeviceToDevice, s));
      }
    } else if (P.mode == 1) {
      launch_precond_invert(P, false, s);
      launch_block_solve_cams(P, s);
    }
    mark(4);
    launch_cam_step(P, s);
    launch_cam_prep(P, P.cam_c, P.scale_c, P.rows_c, s);
    launch_backsub(P, s);
    launch_cost(P, P.rows_c, P.pts_c, P_COST, s);
    launch_finalize_scalars(P, s);
    if (multi()) chk(ctx.comm.allreduce_sum(P.scal + S_MCC, 4, s));
    if (rc) return rc;
    return read_back();
  }
};

}  // namespace

int lm_solve(Ctx& ctx, const hsfm_ba_options& o, hsfm_ba_summary* summary) {
  Problem& P = ctx.prob;
  cudaStream_t s = ctx.stream;
  hsfm_ba_summary local;
  std::memset(&local, 0, sizeof(local));
  hsfm_ba_iteration* trace = summary ? summary->trace : nullptr;
  const int trace_cap = summary ? summary->trace_capacity : 0;
  hsfm_ba_summary& S = summary ? *summary : local;
  const double keep_setup = S.setup_ms, keep_h2d = S.h2d_ms;
  std::memset(&S, 0, sizeof(S));
  S.trace = trace; S.trace_capacity = trace_cap; S.setup_ms = keep_setup; S.h2d_ms = keep_h2d;
  S.n_ranks = ctx.comm.n_ranks;
  int trace_n = 0;
  auto push = [&](const hsfm_ba_iteration& it) { if (trace && trace_n < trace_cap) trace[trace_n] = it; ++trace_n; };
  const int64_t launches0 = P.launches;

  Driver D(ctx, o);
  const int n6 = 6 * P.n_cams;
  if (P.mode == 2) {
    auto ensure_plan = [&]() -> int {  // static per problem: which 128-blocks of S (and of L) can be non-zero
      if (ctx.chol_plan_valid) return 0;
      int rc;
      if ((int)P.cam_first.size() == P.n_cams && P.n_cams > 0) {
        // the same camera graph in the same internal order as the plan was built from: nothing to do
        if (ctx.chol_plan.version > 0 && ctx.chol_plan.n == n6 && ctx.plan_first == P.cam_first && ctx.plan_new_of_old == P.h_struct_new_of_base) {
          ctx.chol_plan_valid = true;
          return 0;
        }
        std::vector<uint8_t> pattern;
        int panels = 0;
        block_pattern_from_envelope(P.n_cams, P.cam_first.data(), P.h_struct_new_of_base.empty() ? nullptr : P.h_struct_new_of_base.data(), pattern, &panels);
        rc = ctx.chol_plan.build(P.n_cams, pattern.data(), panels, s);
        ctx.plan_first = P.cam_first; ctx.plan_new_of_old = P.h_struct_new_of_base;
      } else {
        ctx.plan_first.clear(); ctx.plan_new_of_old.clear();
        rc = ctx.chol_plan.build(P.n_cams, nullptr, 0, s);
      }
      if (rc) return rc;
      ctx.chol_plan_valid = true;
      return 0;
    };
    // Under a communicator every rank must take the same branch (the solve is a fixed sequence of collectives): decisions
    // that depend on rank-local state — free memory, an allocation that may fail — are taken on the minimum over ranks.
    auto agree = [&](bool local, bool* all) -> int {
      *all = local;
      if (ctx.comm.n_ranks <= 1) return 0;
      int v = local ? 1 : 0;
      HSFM_CUDA(cudaMemcpyAsync(P.flags + FLAG_SLOTS - 1, &v, sizeof(int), cudaMemcpyHostToDevice, s));
      const int rc = ctx.comm.allreduce_min_i32(P.flags + FLAG_SLOTS - 1, 1, s);
      if (rc) return rc;
      HSFM_CUDA(cudaMemcpyAsync(&v, P.flags + FLAG_SLOTS - 1, sizeof(int), cudaMemcpyDeviceToHost, s));
      HSFM_CUDA(cudaStreamSynchronize(s));
      *all = v != 0;
      return 0;
    };
    if (o.linear_solver == HSFM_BA_LINEAR_DENSE_SCHUR) D.dense = true;
    else if (o.linear_solver == HSFM_BA_LINEAR_PCG) D.dense = false;
    else {
      D.dense = P.n_cams <= o.dense_max_cams;
      if (!D.dense && P.n_cams <= 16 * (int64_t)o.dense_max_cams) {
        // Beyond dense_max_cams a dense factorisation is hopeless — unless S is a narrow envelope (a scan trajectory):
        // then the envelope factorisation costs a latency chain of n/128 panel steps, far less than the hundreds
        // of implicit-Schur mat-vecs PCG needs per LM iteration.  S is still stored as a full n x n array (only the
        // envelope is ever touched), so it must fit next to the problem.
        int rc = ensure_plan();
        if (rc) return rc;
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        const size_t need_b = sizeof(double) * ((size_t)n6 * n6 + cholesky_workspace_doubles(n6) + (size_t)ctx.chol_plan.packed_doubles() + (size_t)ctx.chol_plan.tail_doubles()) +
                              sizeof(unsigned) * (size_t)std::max<long long>(P.total_pairs, 0);
        const size_t have_b = free_b + sizeof(double) * P.S_capacity;
        // "narrow": the envelope in base order (what the camera graph is) or the block structure of the chosen order
        // (which carries the separator rows of a nested order on top) holds < 15 % of the triangle
        double env = 0.0;
        for (int c = 0; c < P.n_cams && c < (int)P.cam_first.size(); ++c) env += c - std::min(P.cam_first[(size_t)c], c) + 1;
        const double band_fill = P.cam_first.empty() ? 1.0 : env / (0.5 * (double)P.n_cams * (P.n_cams + 1));
        bool fits = false;
        rc = agree(have_b > need_b + ((size_t)4 << 30), &fits);  // free memory differs from rank to rank
        if (rc) return rc;
        if (ctx.chol_plan.use && std::min(ctx.chol_plan.fill, band_fill) < 0.15 && fits) D.dense = true;
      }
    }
    if (D.dense) {
      int rc = ensure_plan();
      if (rc) return rc;
      // workspace after S: L11^-1 blocks + flags, then (multi-GPU) the packed envelope for the all-reduce
      const size_t packed = (ctx.comm.n_ranks > 1 && ctx.chol_plan.use) ? (size_t)ctx.chol_plan.packed_doubles() : 0;
      const size_t need = (size_t)n6 * n6 + cholesky_workspace_doubles(n6) + (size_t)ctx.chol_plan.tail_doubles() + packed;
      bool have = true;
      if (P.S_capacity < need) {
        if (P.S) cudaFree(P.S);
        P.S = nullptr; P.S_capacity = 0;
        if (cudaMalloc(&P.S, sizeof(double) * std::max<size_t>(need, 1)) != cudaSuccess) {
          cudaGetLastError();
          have = false;
        } else {
          P.S_capacity = need;
        }
      }
      bool all_have = have;
      rc = agree(have, &all_have);  // one rank out of memory: every rank falls back (or fails) together
      if (rc) return rc;
      if (!all_have) {
        if (o.linear_solver == HSFM_BA_LINEAR_DENSE_SCHUR) return HSFM_BA_ERR_NOMEM;
        D.dense = false;
      }
    }
  }
  S.linear_solver_used = (P.mode == 2 && !D.dense) ? HSFM_BA_LINEAR_PCG : HSFM_BA_LINEAR_DENSE_SCHUR;
  S.cameras_reordered = P.cam_new_of_old ? 1 : 0;
  if (P.mode == 2 && D.dense) {
    const CholPlan& pl = ctx.chol_plan;
    S.chol_panels = (n6 + 127) / 128;
    S.chol_critical_panels = S.chol_panels;
    const double b3 = 128.0 * 128.0 * 128.0;
    S.chol_flops = (double)n6 * n6 * n6 / 3.0;
    if (!(o.dense_flags & 1) && pl.use) {
      S.chol_flops = 0.0;
      for (int K = 0; K < pl.panels; ++K) S.chol_flops += b3 / 3.0 + pl.cnt[K] * b3 + (double)pl.cnt[K] * (pl.cnt[K] + 1) * b3;
    }
    S.chol_chains = 1;
    if (!(o.dense_flags & 1) && pl.use) { S.chol_critical_panels = pl.critical; S.chol_chains = std::max<int>(1, (int)pl.segs.size()); }
  }

  struct EventPair {  // destroyed on every return path
    cudaEvent_t a = nullptr, b = nullptr;
    ~EventPair() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
  } evs;
  HSFM_CUDA(cudaEventCreate(&evs.a));
  HSFM_CUDA(cudaEventCreate(&evs.b));
  const cudaEvent_t ev0 = evs.a, ev1 = evs.b;
  HSFM_CUDA(cudaEventRecord(ev0, s));

  D.compute_scaling();
  if (P.has_fixed) {
    // residuals whose blocks are all constant: not part of the reduced program, constant over the solve
    launch_fixed_cost(P, P.rows_x, P.pts_x, s);
    launch_finalize_scalars(P, s);  // now: the partial-sum slot is reused by the gradient norm
    if (D.multi()) D.chk(ctx.comm.allreduce_sum(P.scal + S_FIXED_COST, 1, s));
  }
  D.linearize();
  if (D.rc) return D.rc;
  { int r = D.read_back(); if (r) return r; }
  auto gmax_of = [&]() {
    double g = 0.0;
    if (D.cams_free) g = std::max(g, P.scal_host[S_GMAX_C]);
    if (D.pts_free) g = std::max(g, P.scal_host[S_GMAX_P]);
    return g;
  };
  // the costs the trust-region loop sees are those of the reduced program (ceres: fixed_cost is added in the summary only)
  const double fixed_cost = P.has_fixed ? P.scal_host[S_FIXED_COST] : 0.0;
  double x_cost = P.scal_host[S_COST] - fixed_cost;
  double gmax = gmax_of();
  S.initial_cost = x_cost + fixed_cost;
  S.fixed_cost = fixed_cost;
  double radius = o.initial_trust_region_radius, decrease_factor = 2.0;
  int invalid_run = 0, num_success = 0, total_lin = 0, iter = 0;
  int term = HSFM_BA_NO_CONVERGENCE;
  {
    hsfm_ba_iteration it0;
    std::memset(&it0, 0, sizeof(it0));
    it0.cost = x_cost; it0.gradient_max_norm = gmax; it0.trust_region_radius = radius; it0.step_is_valid = 1; it0.step_is_successful = 1;
    push(it0);
  }
  if (!std::isfinite(x_cost)) {
    term = HSFM_BA_FAILURE;
  } else if (gmax <= o.gradient_tolerance || (!D.multi() && P.n_obs == 0)) {  // an empty SHARD is not an empty problem
    term = HSFM_BA_CONVERGENCE;
  } else {
    while (true) {
      if (iter >= o.max_num_iterations) { term = HSFM_BA_NO_CONVERGENCE; break; }
      ++iter;
      hsfm_ba_iteration its;
      std::memset(&its, 0, sizeof(its));
      int lin = 0;
      { int r = D.attempt(radius, &lin); if (r) return r; }
      total_lin += lin;
      its.linear_iterations = lin;
      const double mcc = P.scal_host[S_MCC];
      const double cand_cost_raw = P.scal_host[S_CAND_COST] - fixed_cost;
      const double dn2 = (D.cams_free ? P.scal_host[S_DN_C] : 0.0) + (D.pts_free ? P.scal_host[S_DN_P] : 0.0);
      const double xn2 = (D.cams_free ? P.scal_host[S_XN_C] : 0.0) + (D.pts_free ? P.scal_host[S_XN_P] : 0.0);
      const bool bad = P.flags_host[F_INVALID] != 0 || P.flags_host[F_CHOL_FAIL] != 0 || !std::isfinite(mcc) || !(mcc > 0.0) ||
                       !std::isfinite(dn2);
      its.model_cost_change = mcc;
      if (bad) {
        ++invalid_run;
        its.step_is_valid = 0; its.cost = x_cost; its.gradient_max_norm = gmax;
        radius /= decrease_factor; decrease_factor *= 2.0;
        its.trust_region_radius = radius;
        push(its);
        if (o.verbose) std::fprintf(stderr, "[hsfm_ba] it %3d INVALID step (flags %d chol %d mcc %.3e) radius -> %.3e\n", iter,
                                    P.flags_host[F_INVALID], P.flags_host[F_CHOL_FAIL], mcc, radius);
        if (invalid_run >= o.max_consecutive_invalid_steps) { term = HSFM_BA_FAILURE; break; }
        if (radius < o.min_trust_region_radius) { term = HSFM_BA_CONVERGENCE; break; }
        continue;
      }
      invalid_run = 0;
      its.step_is_valid = 1;
      double cand_cost = cand_cost_raw;
      if (!std::isfinite(cand_cost)) cand_cost = 1.7976931348623157e308;
      const double x_norm = std::sqrt(xn2);
      its.step_norm = std::sqrt(dn2);
      its.cost_change = x_cost - cand_cost;
      its.cost = x_cost; its.gradient_max_norm = gmax; its.trust_region_radius = radius;
      if (its.step_norm <= o.parameter_tolerance * (x_norm + o.parameter_tolerance)) { term = HSFM_BA_CONVERGENCE; push(its); break; }
      if (std::fabs(its.cost_change) <= o.function_tolerance * x_cost) { term = HSFM_BA_CONVERGENCE; push(its); break; }
      const double rho = its.cost_change / mcc;
      its.relative_decrease = rho;
      if (rho > o.min_relative_decrease) {
        std::swap(P.cam_x, P.cam_c);
        std::swap(P.pts_x, P.pts_c);
        D.linearize();
        if (D.rc) return D.rc;
        { int r = D.read_back(); if (r) return r; }
        x_cost = P.scal_host[S_COST] - fixed_cost;
        gmax = gmax_of();
        ++num_success;
        its.step_is_successful = 1;
        radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * rho - 1.0, 3));
        radius = std::min(o.max_trust_region_radius, radius);
        decrease_factor = 2.0;
        its.cost = x_cost; its.gradient_max_norm = gmax; its.trust_region_radius = radius;
        push(its);
        if (o.verbose)
          std::fprintf(stderr, "[hsfm_ba] it %3d cost %.9e dcost %.3e |g| %.3e |step| %.3e rho %.3e radius %.3e lin %d\n", iter, its.cost,
                       its.cost_change, its.gradient_max_norm, its.step_norm, rho, radius, lin);
        if (gmax <= o.gradient_tolerance) { term = HSFM_BA_CONVERGENCE; break; }
      } else {
        radius /= decrease_factor; decrease_factor *= 2.0;
        its.trust_region_radius = radius;
        push(its);
        if (o.verbose)
          std::fprintf(stderr, "[hsfm_ba] it %3d REJECTED cost %.9e cand %.9e rho %.3e radius -> %.3e lin %d\n", iter, x_cost, cand_cost, rho,
                       radius, lin);
        if (radius < o.min_trust_region_radius) { term = HSFM_BA_CONVERGENCE; break; }
      }
    }
  }
  D.mark(7);
  HSFM_CUDA(cudaEventRecord(ev1, s));
  HSFM_CUDA(cudaEventSynchronize(ev1));
  D.collect(S.phase_ms);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ev0, ev1);
  S.solve_ms = ms;
  S.final_cost = x_cost + fixed_cost;
  S.rms_px = 0.0;  // filled by the caller when the global observation count is known
  S.num_iterations = iter;
  S.num_successful_steps = num_success;
  S.termination_type = term;
  S.total_linear_iterations = total_lin;
  S.gpu_launches = P.launches - launches0;
  S.trace_count = std::min(trace_n, trace_cap);
  return 0;
}

}  // namespace hsfm
