// The ProxADMM iteration loop on the device.
// Reference: ProxADMMSolver::Solve (src/epsilon/algorithms/prox_admm.cc:131-169),
// ComputeResiduals (:178-217), GetSolution (:171-176).
#include <chrono>

#include "admm.h"

using namespace eps;

namespace {
double WallMs() {
  return std::chrono::duration<double, std::milli>(
             std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace

// for i: u += y_i; x_i = prox_i(u); y_i = A x_i; u -= y_i   (prox_admm.cc:141-147), stage by
// stage; yo holds the previous y_i, yn receives the new ones.
void eps_solver::RunStages(const double* yo, double* yn) {
  for (auto& stp : stages) {
    Stage& st = *stp;
    const int nseg = (int)st.members.size();
    switch (st.kind) {
      case ST_ELEM: {
        const ElemSeg& sg = ops[st.members[0]].seg;
        static const bool no_vec = getenv("EPS_PROX_SCALAR") != nullptr;
        const bool vec2 = nseg == 1 && !no_vec && ((sg.u_off | sg.y_off | sg.x_off) & 1) == 0 &&
                          (sg.g_off < 0 || (sg.g_off & 1) == 0) && sg.q_off < 0 && sg.d_off < 0;
        if (vec2) ElemProxStepSingle(ctx, sg, consts.p, u.p, yo, yn, x.p);
        else ElemProxStep(ctx, st.segs.p, st.prefix.p, nseg, st.total, st.type, consts.p, u.p, yo, yn, x.p);
        break;
      }
      case ST_NORM2:
        Norm2ProxStep(ctx, st.segs.p, st.n_segs, st.max_len, consts.p, u.p, yo, yn, x.p);
        break;
      case ST_TV:
        Tv1dProxStep(ctx, ops[st.members[0]].seg, consts.p, u.p, yo, yn, x.p, st.tv_work.p, st.tv_tag);
        break;
      case ST_VEC: {
        // general VectorProx reduction around a multi-argument operator (vector_prox.cc:131-137)
        ProxOp& op = ops[st.members[0]];
        double* us = u.p + op.span_lo;
        double* xi = x.p + op.x_off;
        VecAxpby(ctx, op.span_len, 1.0, us, 1.0, yo + op.y_off, us);      // w = u + y_i
        op.A.ApplyT(us, op.vtmp.p);                                       // A'w
        op.Hop.Apply(op.vtmp.p, op.vbuf.p);                               // H A'w
        const double* g = op.g_off >= 0 ? consts.p + op.g_off : nullptr;
        if (g) VecAxpby(ctx, op.rows_h, op.vec_bg, op.vbuf.p, 1.0, g, op.vbuf.p);
        else VecScale(ctx, op.rows_h, op.vec_bg, op.vbuf.p, op.vbuf.p);
        VecOpApply(ctx, op.vdesc, op.vbuf.p, &op.vwork);
        if (g) VecAxpby(ctx, op.rows_h, 1.0, op.vbuf.p, -1.0, g, op.vbuf.p);
        op.Hop.ApplyT(op.vbuf.p, xi);                                     // x = (1/beta) H'(x' - g)
        VecScale(ctx, op.n, op.vec_cinv, xi, xi);
        op.A.Apply(xi, yn + op.y_off);                                    // y_i = A x_i
        VecAxpby(ctx, op.span_len, 1.0, us, -1.0, yn + op.y_off, us);     // u = w - y_i
        break;
      }
      case ST_KKT: {
        ProxOp& op = ops[st.members[0]];
        double* us = u.p + op.span_lo;
        double* xi = x.p + op.x_off;
        if (op.simple_scalar && op.kkt->mode == EPS_KKT_DIAGONAL && !op.kkt->inner) {
          KktScalarDiagStep(ctx, op.n, op.a, op.kkt->c0.p, op.kkt->dinv.p, us, yo + op.y_off, yn + op.y_off, xi);
        } else if (op.simple_scalar) {
          KktScalarPre(ctx, op.n, op.a, op.kkt->c0.p, us, yo + op.y_off, rhs_n.p);
          op.kkt->Solve(rhs_n.p, xi);
          KktScalarPost(ctx, op.n, op.a, xi, us, yn + op.y_off);
        } else {
          if (op.span_len > 0 && op.a_sharded) {
            // rows of the sharded constraints: this rank's A_r'w, summed over the ranks; replicated rows as usual
            VecAxpby(ctx, op.span_len, 1.0, us, 1.0, yo + op.y_off, us);
            op.A_sh.ApplyT(us, tmp_n.p);
            AllReduceSum(ctx, tmp_n.p, op.n);
            if (!op.A_rep.blks.empty()) {
              op.A_rep.ApplyT(us, rhs_n.p);
              VecAxpby(ctx, op.n, 1.0, rhs_n.p, 1.0, tmp_n.p, rhs_n.p);
              VecAxpby(ctx, op.n, 1.0, rhs_n.p, 1.0, op.kkt->c0.p, rhs_n.p);
            } else {
              VecAxpby(ctx, op.n, 1.0, tmp_n.p, 1.0, op.kkt->c0.p, rhs_n.p);
            }
          } else if (op.span_len > 0) {
            VecAxpby(ctx, op.span_len, 1.0, us, 1.0, yo + op.y_off, us);
            op.A.ApplyT(us, rhs_n.p);
            VecAxpby(ctx, op.n, 1.0, rhs_n.p, 1.0, op.kkt->c0.p, rhs_n.p);
          } else {
            VecCopy(ctx, op.n, op.kkt->c0.p, rhs_n.p);
          }
          op.kkt->Solve(rhs_n.p, xi);
          if (op.span_len > 0) {
            op.A.Apply(xi, yn + op.y_off);
            VecAxpby(ctx, op.span_len, 1.0, us, -1.0, yn + op.y_off, us);
          }
        }
        break;
      }
    }
  }
}

// One iteration: prox_admm.cc:135-147.  ybuf[parity] is y (= y_prev during this
// iteration), ybuf[1-parity] receives the new y.
void eps_solver::Iteration() {
  if (two_block) { TwoBlockIteration(); return; }
  const double* yo = ybuf[parity].p;
  double* yn = ybuf[1 - parity].p;
  // u -= b; u -= sum_i y_i                                   (:137-139)
  if (has_b) VecAxpby(ctx, m_total, 1.0, u.p, -1.0, b.p, u.p);
  for (auto& stp : stages) {
    Stage& st = *stp;
    SegAxpy(ctx, st.pairs.p, st.prefix.p, (int)st.members.size(), st.total, u.p, yo, -1.0, nullptr, 0.0);
  }
  RunStages(yo, yn);
  parity = 1 - parity;
}

// ProxADMMTwoBlockSolver::Solve body (prox_admm_two_block.cc:99-113).  Every prox operator was
// initialised with A = I on its own variables, laid out exactly like the x arena, so
// "x_i = prox_i(z - u)" is RunStages on a scratch copy of z - u with zero previous y.
void eps_solver::TwoBlockIteration() {
  const double* z = zbuf[zpar].p;
  double* zn = zbuf[1 - zpar].p;   // becomes z; the old z becomes z_prev
  VecAxpby(ctx, n_all, 1.0, z, -1.0, ub.p, u.p);                 // zu = z - u          (:102)
  RunStages(zeros.p, ybuf[0].p);                                  // x = sum_i prox_i(zu) (:103-107)
  VecAxpby(ctx, n_all, 1.0, x.p, 1.0, ub.p, v_tmp.p);            // v = x + u
  const double* t = nullptr;
  if (m_c > 0) {
    // ZeroProx (prox/zero.cc:10-36) with A = I:  z = v - H'(HH')^-1 (H v + b)
    Hc.Apply(v_tmp.p, r_m.p);
    VecAxpby(ctx, m_c, 1.0, r_m.p, 1.0, hb.p, r_m.p);
    proj->Solve(r_m.p, y_m.p);
    Hc.ApplyT(y_m.p, diff_m.p);
    t = diff_m.p;
  }
  // z = v - t (restricted to the constraint variables), u += x - z (:111), residual sums
  TwoBlockDualStep(ctx, n_all, x.p, v_tmp.p, t, has_mask ? cmask.p : nullptr, z, zn, ub.p, resid.p);
  zpar = 1 - zpar;
}

void eps_solver::RunIteration() {
  const int p = parity;
  if (!graph_ok || eager_done[p] < 1) {
    Iteration();
    eager_done[p]++;
    return;
  }
  if (!graph_exec[p]) {
    cudaGraph_t graph = nullptr;
    ctx->capturing = true;
    ctx->capture_launches = 0;
    cudaError_t e = cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal);
    if (e == cudaSuccess) {
      try {
        Iteration();
      } catch (...) {
        cudaStreamEndCapture(ctx->stream, &graph);
        if (graph) cudaGraphDestroy(graph);
        ctx->capturing = false;
        throw;
      }
      e = cudaStreamEndCapture(ctx->stream, &graph);
    }
    ctx->capturing = false;
    parity = p;  // the captured call flipped it; the replay below flips it for real
    if (e != cudaSuccess || !graph) {
      cudaGetLastError();
      graph_ok = false;
      Iteration();
      return;
    }
    graph_nodes[p] = ctx->capture_launches;
    e = cudaGraphInstantiate(&graph_exec[p], graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) {
      cudaGetLastError();
      graph_exec[p] = nullptr;
      graph_ok = false;
      Iteration();
      return;
    }
  }
  CUDA_CHECK(cudaGraphLaunch(graph_exec[p], ctx->stream));
  ctx->launches += graph_nodes[p];
  parity = 1 - p;
}

// prox_admm.cc:178-217.  On entry ybuf[parity] = y, ybuf[1-parity] = y_prev.
// ComputeResiduals (prox_admm.cc:178-217) with row-sharded constraints: every sum over constraint rows splits
// into the rows this rank holds a shard of (summed over the ranks) and the replicated rows (counted once);
// products with A_i' of a replicated variable that enters sharded rows are all-reduced as n-vectors.  All scalar
// partial sums travel in ONE all-reduce.  Layout of resid_sh: [sharded | replicated] x [r^2, s_i^2[N], |A_i'u|^2[N], |y_i|^2[N]].
void eps_solver::ComputeResidualsSharded() {
  const double* yn = ybuf[parity].p;
  const double* yo = ybuf[1 - parity].p;
  const int K = 1 + 3 * N;
  double* SH = resid_sh.p;
  double* RP = resid_sh.p + K;
  resid_sh.zero();
  const int nc = (int)c_sharded.size();
  auto rows_dot = [&](const double* v, int64_t lo, int64_t len, int64_t base, int slot) {
    // sum over rows [lo, lo+len) of u-space (v indexed from `base`), split by constraint
    for (int c = 0; c < nc; c++) {
      const int64_t a = std::max(lo, constraint_off[c]), b = std::min(lo + len, constraint_off[c + 1]);
      if (b <= a) continue;
      VecDot(ctx, b - a, v + (a - base), v + (a - base), (c_sharded[c] ? SH : RP) + slot, true);
    }
  };
  // r = || b + sum_i A x_i ||
  VecCopy(ctx, m_total, b.p, tmp_m.p);
  for (auto& stp : stages) {
    Stage& st = *stp;
    SegAxpy(ctx, st.pairs.p, st.prefix.p, (int)st.members.size(), st.total, tmp_m.p, yn, 1.0, nullptr, 0.0);
  }
  rows_dot(tmp_m.p, 0, m_total, 0, 0);
  // || A x_i ||^2 = || y_i ||^2
  for (int i = 0; i < N; i++) {
    ProxOp& op = ops[i];
    if (op.span_len > 0) rows_dot(yn + op.y_off, op.span_lo, op.span_len, op.span_lo, 1 + 2 * N + i);
  }
  // A_i' d: scalar blocks give a^2 |d_rows|^2; a replicated variable in sharded rows needs the all-reduced n-vector
  auto at_norm = [&](ProxOp& op, const double* d, int slot) {
    if (op.span_len == 0) return;
    if (op.a_sharded) {
      op.A_sh.ApplyT(d + op.span_lo, tmp_n.p);
      AllReduceSum(ctx, tmp_n.p, op.n);
      if (!op.A_rep.blks.empty()) op.A_rep.ApplyT(d + op.span_lo, tmp_n.p, true);
      VecDot(ctx, op.n, tmp_n.p, tmp_n.p, RP + slot, true);
    } else if (op.simple_scalar) {
      rows_dot(d, op.span_lo, op.span_len, 0, slot);   // times a^2 on the host
    } else {
      op.A.ApplyT(d + op.span_lo, tmp_n.p);
      VecDot(ctx, op.n, tmp_n.p, tmp_n.p, (op.local ? SH : RP) + slot, true);
    }
  };
  // s^2 = sum_{i<N-1} || A_i' sum_{j>i} (y_j - y_prev_j) ||^2
  diff_m.zero();
  for (int i = N - 1; i >= 0; i--) {
    ProxOp& op = ops[i];
    if (i < N - 1) at_norm(op, diff_m.p, 1 + i);
    if (i > 0 && op.span_len > 0) {
      VecAxpby(ctx, op.span_len, 1.0, diff_m.p + op.span_lo, 1.0, yn + op.y_off, diff_m.p + op.span_lo);
      VecAxpby(ctx, op.span_len, 1.0, diff_m.p + op.span_lo, -1.0, yo + op.y_off, diff_m.p + op.span_lo);
    }
  }
  // || A' u ||^2 over all variables
  for (int i = 0; i < N; i++) at_norm(ops[i], u.p, 1 + N + i);
  AllReduceSum(ctx, SH, K);
  std::vector<double> h(2 * K);
  resid_sh.download(h.data(), 2 * K);
  auto tot = [&](int slot) { return h[slot] + h[K + slot]; };
  double s2 = 0, atu2 = 0, max_ax2 = 0;
  for (int i = 0; i < N; i++) {
    const double coef = ops[i].simple_scalar && !ops[i].a_sharded ? ops[i].a * ops[i].a : 1.0;
    s2 += coef * tot(1 + i);
    atu2 += coef * tot(1 + N + i);
    max_ax2 = std::fmax(max_ax2, tot(1 + 2 * N + i));
  }
  const double rho = params.rho;
  status.r_norm = std::sqrt(tot(0));
  status.s_norm = rho * std::sqrt(s2);
  status.epsilon_primal = params.abs_tol * std::sqrt((double)m_global) +
                          params.rel_tol * std::fmax(b_norm, std::sqrt(max_ax2));
  status.epsilon_dual = params.abs_tol * std::sqrt((double)n_total) + params.rel_tol * rho * std::sqrt(atu2);
  status.state = (status.r_norm <= status.epsilon_primal && status.s_norm <= status.epsilon_dual)
                     ? EPS_OPTIMAL : EPS_RUNNING;
}

void eps_solver::ComputeResiduals() {
  if (any_sharded) { ComputeResidualsSharded(); return; }
  if (two_block) {
    // prox_admm_two_block.cc:135-156; the five sums were left in `resid` by the last iteration
    CUDA_CHECK(cudaMemcpyAsync(h_resid, resid.p, 5 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    const double rho = params.rho;
    status.r_norm = std::sqrt(h_resid[0]);
    status.s_norm = rho * std::sqrt(h_resid[1]);
    status.epsilon_primal = params.abs_tol * std::sqrt((double)tb_n) +
                            params.rel_tol * std::fmax(std::sqrt(h_resid[2]), std::sqrt(h_resid[3]));
    status.epsilon_dual = params.abs_tol * std::sqrt((double)tb_n) + params.rel_tol * rho * std::sqrt(h_resid[4]);
    status.state = (status.r_norm <= status.epsilon_primal && status.s_norm <= status.epsilon_dual)
                       ? EPS_OPTIMAL : EPS_RUNNING;
    return;
  }
  const double* yn = ybuf[parity].p;
  const double* yo = ybuf[1 - parity].p;
  double* R = resid.p;  // [0] r^2  [1] s^2  [2] max_i ||A x_i||^2  [3] ||A'u||^2
  resid.zero();
  // r = || b + sum_i A x_i ||   (A x_i == y_i: the same products)
  VecCopy(ctx, m_total, b.p, tmp_m.p);
  for (auto& stp : stages) {
    Stage& st = *stp;
    const int nseg = (int)st.members.size();
    SegAxpy(ctx, st.pairs.p, st.prefix.p, nseg, st.total, tmp_m.p, yn, 1.0, nullptr, 0.0);
    SegSumSq(ctx, st.pairs.p, st.pairs_host.data(), nseg, nullptr, nullptr, yn, R + 2);
  }
  VecDot(ctx, m_total, tmp_m.p, tmp_m.p, R + 0, false);
  // s^2 = sum_{i<N-1} || AiT_i * sum_{j>i} (y_j - y_prev_j) ||^2, walking the stages backwards
  diff_m.zero();
  for (int s = (int)stages.size() - 1; s >= 0; s--) {
    Stage& st = *stages[s];
    const int nseg = (int)st.members.size();
    if (s < (int)stages.size() - 1) {
      if (st.scalar_a) {
        SegSumSq(ctx, st.pairs.p, st.pairs_host.data(), nseg, diff_m.p, R + 1, nullptr, nullptr);
      } else {
        for (int j : st.members) {
          ProxOp& op = ops[j];
          if (op.span_len == 0) continue;
          op.A.ApplyT(diff_m.p + op.span_lo, tmp_n.p);
          VecDot(ctx, op.n, tmp_n.p, tmp_n.p, R + 1, true);
        }
      }
    }
    if (s > 0) SegAxpy(ctx, st.pairs.p, st.prefix.p, nseg, st.total, diff_m.p, yn, 1.0, yo, -1.0);
  }
  // || A' u ||^2 over all variables
  for (auto& stp : stages) {
    Stage& st = *stp;
    const int nseg = (int)st.members.size();
    if (st.scalar_a) {
      SegSumSq(ctx, st.pairs.p, st.pairs_host.data(), nseg, u.p, R + 3, nullptr, nullptr);
    } else {
      for (int j : st.members) {
        ProxOp& op = ops[j];
        if (op.span_len == 0) continue;
        op.A.ApplyT(u.p + op.span_lo, tmp_n.p);
        VecDot(ctx, op.n, tmp_n.p, tmp_n.p, R + 3, true);
      }
    }
  }
  CUDA_CHECK(cudaMemcpyAsync(h_resid, R, 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  const double rho = params.rho;
  status.r_norm = std::sqrt(h_resid[0]);
  status.s_norm = rho * std::sqrt(h_resid[1]);
  const double max_ax = std::fmax(b_norm, std::sqrt(h_resid[2]));
  status.epsilon_primal = params.abs_tol * std::sqrt((double)m_total) + params.rel_tol * max_ax;
  status.epsilon_dual = params.abs_tol * std::sqrt((double)n_total) + params.rel_tol * rho * std::sqrt(h_resid[3]);
  status.state = (status.r_norm <= status.epsilon_primal && status.s_norm <= status.epsilon_dual)
                     ? EPS_OPTIMAL : EPS_RUNNING;
}

#define EPS_API_BEGIN try {
#define EPS_API_END                                  \
    return EPS_OK;                                   \
  } catch (const eps::Error& e) {                    \
    eps::SetLastError(e.what());                     \
    return e.code;                                   \
  } catch (const std::exception& e) {                \
    eps::SetLastError(e.what());                     \
    return EPS_ERR_INVALID;                          \
  }

extern "C" {

void eps_params_default(eps_params* p) {
  memset(p, 0, sizeof(*p));
  p->max_iterations = 10000;
  p->rho = 1.0;
  p->abs_tol = 1e-4;
  p->rel_tol = 1e-2;
  p->epoch_iterations = 10;
  p->warm_start = 0;
  p->kkt_mode = EPS_KKT_AUTO;
  p->cg_tol = 1e-12;
  p->cg_max_iterations = 500;
  p->use_graph = 1;
  p->profile = 0;
  p->gram_max_n = 60000;
  p->solver = EPS_SOLVER_PROX_ADMM;
}

int eps_solver_create(eps_problem* p, const eps_params* params, eps_solver** out) {
  EPS_API_BEGIN
  EPS_CHECK(p && out, "eps_solver_create arguments");
  CUDA_CHECK(cudaSetDevice(p->ctx->device));
  const double t0 = WallMs();
  auto s = std::make_unique<eps_solver>();
  if (params) s->params = *params; else eps_params_default(&s->params);
  EPS_CHECK(s->params.epoch_iterations > 0, "epoch_iterations = %d", s->params.epoch_iterations);
  s->Init(*p);
  CUDA_CHECK(cudaStreamSynchronize(p->ctx->stream));
  CUDA_CHECK(cudaGetLastError());
  s->status.init_ms = WallMs() - t0;
  *out = s.release();
  EPS_API_END
}

void eps_solver_destroy(eps_solver* s) {
  if (!s) return;
  if (eps::CtxIsLive(s->ctx)) {
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
  } else {
    cudaDeviceSynchronize();
  }
  delete s;
}

int eps_solver_solve(eps_solver* s, eps_status* out, eps_trace_fn trace, void* user) {
  EPS_API_BEGIN
  EPS_CHECK(s, "null solver");
  Ctx* ctx = s->ctx;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  ctx->profile = s->params.profile != 0;
  if (!s->params.warm_start || !s->started) s->ResetVariables();
  s->started = true;
  int64_t cg0 = 0, un0 = 0;
  for (auto& op : s->ops) if (op.kkt) { cg0 += op.kkt->cg_iters_total; un0 += op.kkt->cg_unconverged; }
  if (s->proj) { cg0 += s->proj->cg_iters_total; un0 += s->proj->cg_unconverged; }
  cudaEvent_t e0, e1;
  CUDA_CHECK(cudaEventCreate(&e0));
  CUDA_CHECK(cudaEventCreate(&e1));
  CUDA_CHECK(cudaEventRecord(e0, ctx->stream));
  const int max_it = s->params.max_iterations;
  int run = 0;
  bool stopped = false;
  s->status.state = EPS_RUNNING;
  for (s->iter = 0; s->iter < max_it; s->iter++) {
    s->RunIteration();
    run++;
    if (s->iter % s->params.epoch_iterations == 0) {
      s->ComputeResiduals();
      s->status.num_iterations = s->iter;
      s->status.iterations_run = run;
      if (trace && trace(&s->status, user)) { stopped = true; break; }
      if (s->status.state == EPS_OPTIMAL) break;
    }
  }
  if (!stopped && s->iter == max_it) {
    s->ComputeResiduals();
    s->status.num_iterations = s->iter;
    s->status.state = EPS_MAX_ITERATIONS_REACHED;
    s->status.iterations_run = run;
    if (trace) trace(&s->status, user);
  }
  CUDA_CHECK(cudaEventRecord(e1, ctx->stream));
  CUDA_CHECK(cudaEventSynchronize(e1));
  float ms = 0;
  CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  CUDA_CHECK(cudaGetLastError());
  s->status.loop_ms = ms;
  s->status.iterations_run = run;
  int64_t cg1 = 0, un1 = 0;
  for (auto& op : s->ops) if (op.kkt) { cg1 += op.kkt->cg_iters_total; un1 += op.kkt->cg_unconverged; }
  if (s->proj) { cg1 += s->proj->cg_iters_total; un1 += s->proj->cg_unconverged; }
  s->status.cg_iterations = cg1 - cg0;
  s->status.cg_unconverged = un1 - un0;
  ctx->profile = false;
  if (out) *out = s->status;
  EPS_API_END
}

int eps_solver_iterate(eps_solver* s, int iterations, float* elapsed_ms) {
  EPS_API_BEGIN
  EPS_CHECK(s && iterations >= 0, "eps_solver_iterate arguments");
  Ctx* ctx = s->ctx;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  ctx->profile = s->params.profile != 0;
  s->started = true;
  int64_t cg0 = 0, un0 = 0;
  for (auto& op : s->ops) if (op.kkt) { cg0 += op.kkt->cg_iters_total; un0 += op.kkt->cg_unconverged; }
  if (s->proj) { cg0 += s->proj->cg_iters_total; un0 += s->proj->cg_unconverged; }
  cudaEvent_t e0, e1;
  CUDA_CHECK(cudaEventCreate(&e0));
  CUDA_CHECK(cudaEventCreate(&e1));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  CUDA_CHECK(cudaEventRecord(e0, ctx->stream));
  for (int k = 0; k < iterations; k++) {
    s->RunIteration();
    s->iter++;
  }
  CUDA_CHECK(cudaEventRecord(e1, ctx->stream));
  CUDA_CHECK(cudaEventSynchronize(e1));
  float ms = 0;
  CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  CUDA_CHECK(cudaGetLastError());
  int64_t cg1 = 0, un1 = 0;
  for (auto& op : s->ops) if (op.kkt) { cg1 += op.kkt->cg_iters_total; un1 += op.kkt->cg_unconverged; }
  if (s->proj) { cg1 += s->proj->cg_iters_total; un1 += s->proj->cg_unconverged; }
  s->status.cg_iterations = cg1 - cg0;
  s->status.cg_unconverged = un1 - un0;
  s->status.iterations_run = iterations;
  s->status.loop_ms = ms;
  ctx->profile = false;
  if (elapsed_ms) *elapsed_ms = ms;
  EPS_API_END
}

int eps_solver_residuals(eps_solver* s, eps_status* out) {
  EPS_API_BEGIN
  EPS_CHECK(s && out, "eps_solver_residuals arguments");
  CUDA_CHECK(cudaSetDevice(s->ctx->device));
  s->ComputeResiduals();
  s->status.num_iterations = s->iter > 0 ? s->iter - 1 : 0;
  *out = s->status;
  EPS_API_END
}

int eps_solver_set_constant(eps_solver* s, int prox_index, int arg_or_constraint_index, const double* values) {
  EPS_API_BEGIN
  EPS_CHECK(s, "null solver");
  CUDA_CHECK(cudaSetDevice(s->ctx->device));
  s->SetConstant(prox_index, arg_or_constraint_index, values);
  EPS_API_END
}

int eps_solver_copy_state(eps_solver* dst, eps_solver* src) {
  EPS_API_BEGIN
  EPS_CHECK(dst && src && dst != src, "eps_solver_copy_state arguments");
  EPS_CHECK(dst->ctx == src->ctx, "solvers live on different contexts");
  EPS_CHECK(dst->two_block == src->two_block && dst->u.n == src->u.n && dst->x.n == src->x.n &&
                dst->ybuf[0].n == src->ybuf[0].n && dst->zbuf[0].n == src->zbuf[0].n && dst->ub.n == src->ub.n &&
                dst->y_m.n == src->y_m.n,
            "the two solvers do not share a variable / constraint layout");
  Ctx* ctx = dst->ctx;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  auto copy = [&](eps::DevBuf<double>& d, const eps::DevBuf<double>& s_) {
    if (d.n) CUDA_CHECK(cudaMemcpyAsync(d.p, s_.p, d.n * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
  };
  copy(dst->u, src->u);
  copy(dst->x, src->x);
  copy(dst->ybuf[0], src->ybuf[0]);
  copy(dst->ybuf[1], src->ybuf[1]);
  copy(dst->zbuf[0], src->zbuf[0]);
  copy(dst->zbuf[1], src->zbuf[1]);
  copy(dst->ub, src->ub);
  copy(dst->y_m, src->y_m);
  dst->parity = src->parity;
  dst->zpar = src->zpar;
  dst->started = src->started;
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  EPS_API_END
}

int eps_solver_get_variable(eps_solver* s, const char* var_id, double* out) {
  EPS_API_BEGIN
  EPS_CHECK(s && var_id && out, "eps_solver_get_variable arguments");
  CUDA_CHECK(cudaSetDevice(s->ctx->device));
  for (size_t v = 0; v < s->vars.size(); v++) {
    if (s->vars[v].id != var_id) continue;
    EPS_CHECK(s->var_x_off[v] >= 0, "variable %s is not owned by a prox function", var_id);
    CUDA_CHECK(cudaMemcpyAsync(out, s->x.p + s->var_x_off[v], s->vars[v].size * sizeof(double),
                               cudaMemcpyDeviceToHost, s->ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(s->ctx->stream));
    return EPS_OK;
  }
  throw Error(EPS_ERR_INVALID, Format("unknown variable %s", var_id));
  EPS_API_END
}

int eps_solver_profile(eps_solver* s, int max_classes, const char** names, int64_t* launches,
                       double* total_ms, double* bytes) {
  int n = 0;
  for (int k = 0; k < KC_COUNT && n < max_classes; k++) {
    const Ctx::Prof& p = s->ctx->prof[k];
    if (p.launches == 0) continue;
    names[n] = kKClassNames[k];
    launches[n] = p.launches;
    total_ms[n] = p.ms;
    bytes[n] = p.bytes;
    n++;
  }
  return n;
}

const char* eps_solver_describe(eps_solver* s) { return s->description.c_str(); }

// EvalProx: python/epopt/solvemodule.cc:189-242.
int eps_prox_eval(eps_problem* p, double lambda, const double* v_concat, double* x_concat) {
  EPS_API_BEGIN
  EPS_CHECK(p && v_concat && x_concat, "eps_prox_eval arguments");
  EPS_CHECK(p->prox.size() == 1 && p->constraints.empty(),
            "eps_prox_eval expects one prox function and no constraint");
  EPS_CHECK(lambda > 0, "lambda = %g", lambda);
  CUDA_CHECK(cudaSetDevice(p->ctx->device));
  // A(constraint:k, var_k) = (1/sqrt(lambda)) I   (solvemodule.cc:221-226)
  eps_problem q = *p;
  const double a = 1 / std::sqrt(lambda);
  std::vector<int64_t> off(q.vars.size() + 1, 0);
  for (size_t k = 0; k < q.vars.size(); k++) {
    AffineExpr e;
    e.rows = q.vars[k].size;
    Term t;
    auto d = std::make_shared<MapData>();
    d->ctx = p->ctx; d->kind = MK_SCALAR; d->m = d->n = e.rows; d->alpha = a;
    t.map = d;
    t.var = (int)k;
    e.terms.push_back(std::move(t));
    q.constraints.push_back(std::move(e));
    off[k + 1] = off[k] + q.vars[k].size;
  }
  eps_solver s;
  eps_params_default(&s.params);
  s.params.use_graph = 0;
  s.Init(q);
  // v = A * v_in  (solvemodule.cc:229), then one Apply
  std::vector<double> uv(off.back());
  for (int64_t i = 0; i < off.back(); i++) uv[i] = a * v_concat[i];
  s.u.upload(uv.data(), uv.size());
  s.Iteration();
  for (size_t k = 0; k < q.vars.size(); k++) {
    EPS_CHECK(s.var_x_off[k] >= 0, "variable %s unused", q.vars[k].id.c_str());
    CUDA_CHECK(cudaMemcpyAsync(x_concat + off[k], s.x.p + s.var_x_off[k], q.vars[k].size * sizeof(double),
                               cudaMemcpyDeviceToHost, p->ctx->stream));
  }
  CUDA_CHECK(cudaStreamSynchronize(p->ctx->stream));
  CUDA_CHECK(cudaGetLastError());
  EPS_API_END
}

}  // extern "C"
