// ProxADMMSolver::Init on the device side (reference: src/epsilon/algorithms/prox_admm.cc:25-129).
//   InitConstraints  (:25-43)  -> flat u-space, one contiguous row block per constraint
//   InitProxOperators(:45-94)  -> per f_i: H (args), A (its columns of A_), operator object
//   InitVariables    (:96-108) -> u = 0, x and y empty (= zero slabs)
#include <algorithm>
#include <chrono>
#include <numeric>

#include "admm.h"

namespace eps {

namespace {

std::shared_ptr<MapData> IdentityMap(Ctx* ctx, int n) {
  auto d = std::make_shared<MapData>();
  d->ctx = ctx; d->kind = MK_SCALAR; d->m = n; d->n = n; d->alpha = 1.0;
  return d;
}

Blk TermBlk(Ctx* ctx, const Term& t, int rows, int row_off, int col_off) {
  Blk b;
  b.row_off = row_off;
  b.col_off = col_off;
  b.map = t.map ? t.map : IdentityMap(ctx, rows);
  b.trans = t.map ? t.trans : false;
  b.scale = 1.0;
  return b;
}

// Sum of the constant terms of an affine expression (affine.cc:51-69), on the host.
bool EvalConstants(Ctx* ctx, const AffineExpr& e, std::vector<double>* out) {
  out->assign(e.rows, 0.0);
  bool any = false;
  for (const Term& t : e.terms) {
    if (t.var >= 0) continue;
    any = true;
    if (!t.map) {
      for (int i = 0; i < e.rows; i++) (*out)[i] += t.constant[i];
    } else if (t.map->kind == MK_SCALAR) {
      for (int i = 0; i < e.rows; i++) (*out)[i] += t.map->alpha * t.constant[i];
    } else {
      const int n_in = (int)t.constant.size();
      DevBuf<double> cin(ctx, n_in), cout(ctx, e.rows);
      cin.upload(t.constant.data(), n_in);
      MapApply(ctx, *t.map, t.trans, 1.0, cin.p, 0.0, cout.p);
      std::vector<double> tmp(e.rows);
      cout.download(tmp.data(), e.rows);
      for (int i = 0; i < e.rows; i++) (*out)[i] += tmp[i];
    }
  }
  return any;
}

bool IsKktType(int t) { return t == EPS_PROX_SUM_SQUARE || t == EPS_PROX_AFFINE || t == EPS_PROX_CONSTANT; }
bool IsElemType(int t) {
  return t == EPS_PROX_NORM_1 || t == EPS_PROX_SUM_HINGE || t == EPS_PROX_SUM_DEADZONE ||
         t == EPS_PROX_SUM_QUANTILE || t == EPS_PROX_NON_NEGATIVE || t == EPS_PROX_SUM_LOGISTIC ||
         t == EPS_PROX_SUM_EXP || t == EPS_PROX_SUM_INV_POS || t == EPS_PROX_SUM_NEG_ENTR ||
         t == EPS_PROX_SUM_NEG_LOG;
}

const char* TypeName(int t) {
  switch (t) {
    case EPS_PROX_AFFINE: return "AFFINE";
    case EPS_PROX_CONSTANT: return "CONSTANT";
    case EPS_PROX_ZERO: return "ZERO";
    case EPS_PROX_SUM_SQUARE: return "SUM_SQUARE";
    case EPS_PROX_NON_NEGATIVE: return "NON_NEGATIVE";
    case EPS_PROX_NORM_1: return "NORM_1";
    case EPS_PROX_SUM_DEADZONE: return "SUM_DEADZONE";
    case EPS_PROX_SUM_HINGE: return "SUM_HINGE";
    case EPS_PROX_SUM_LOGISTIC: return "SUM_LOGISTIC";
    case EPS_PROX_SUM_EXP: return "SUM_EXP";
    case EPS_PROX_SUM_INV_POS: return "SUM_INV_POS";
    case EPS_PROX_SUM_NEG_ENTR: return "SUM_NEG_ENTR";
    case EPS_PROX_SUM_NEG_LOG: return "SUM_NEG_LOG";
    case EPS_PROX_SUM_QUANTILE: return "SUM_QUANTILE";
    case EPS_PROX_NORM_2: return "NORM_2";
    case EPS_PROX_EXP: return "EXP";
    case EPS_PROX_MAX: return "MAX";
    case EPS_PROX_SUM_LARGEST: return "SUM_LARGEST";
    case EPS_PROX_LOG_SUM_EXP: return "LOG_SUM_EXP";
    case EPS_PROX_LAMBDA_MAX: return "LAMBDA_MAX";
    case EPS_PROX_NEG_LOG_DET: return "NEG_LOG_DET";
    case EPS_PROX_NORM_NUCLEAR: return "NORM_NUCLEAR";
    case EPS_PROX_SEMIDEFINITE: return "SEMIDEFINITE";
    case EPS_PROX_SECOND_ORDER_CONE: return "SECOND_ORDER_CONE";
    case EPS_PROX_SUM_KL_DIV: return "SUM_KL_DIV";
    case EPS_PROX_TOTAL_VARIATION_1D: return "TOTAL_VARIATION_1D";
    default: return "?";
  }
}

}  // namespace
}  // namespace eps

using namespace eps;

namespace {
// Consensus form of ProxADMMTwoBlockSolver (prox_admm_two_block.cc:66-71): every variable of
// f_i sees the constraint block A(var, var) = sqrt(rho) I.  The synthetic "constraints" are
// laid out in the order of the x arena (prox by prox, variable ids ascending inside one), so
// that each f_i touches one contiguous span with A = I.
eps_problem ConsensusForm(const eps_problem& p) {
  eps_problem q = p;
  q.constraints.clear();
  const int N = (int)p.prox.size();
  std::vector<int> owner(p.vars.size(), -1);
  for (int i = 0; i < N; i++)
    for (const AffineExpr& a : p.prox[i].args)
      for (const Term& t : a.terms)
        if (t.var >= 0 && owner[t.var] < 0) owner[t.var] = i;
  for (int i = 0; i < N; i++) {
    std::vector<int> vs;
    for (size_t v = 0; v < p.vars.size(); v++)
      if (owner[v] == i) vs.push_back((int)v);
    std::sort(vs.begin(), vs.end(), [&](int a, int b) { return p.vars[a].id < p.vars[b].id; });
    for (int v : vs) {
      AffineExpr e;
      e.rows = p.vars[v].size;
      Term t;
      t.var = v;
      e.terms.push_back(std::move(t));
      q.constraints.push_back(std::move(e));
    }
  }
  return q;
}
}  // namespace

void eps_solver::Init(const eps_problem& original) {
  two_block = params.solver == EPS_SOLVER_PROX_ADMM_TWO_BLOCK;
  EPS_CHECK(params.solver == EPS_SOLVER_PROX_ADMM || two_block, "unknown solver %d", params.solver);
  eps_problem consensus;
  if (two_block) consensus = ConsensusForm(original);
  const eps_problem& p = two_block ? consensus : original;
  ctx = p.ctx;
  vars = p.vars;
  N = (int)p.prox.size();
  EPS_CHECK(params.rho == 1.0, "rho = %g (the reference requires rho == 1, prox_admm.cc:50)", params.rho);
  EPS_CHECK(N > 0, "no prox functions");

  // ---- InitConstraints: row offsets, b ----
  std::vector<int64_t> coff(p.constraints.size() + 1, 0);
  for (size_t c = 0; c < p.constraints.size(); c++) coff[c + 1] = coff[c] + p.constraints[c].rows;
  m_total = coff.back();
  constraint_off = coff;
  b_host.assign(m_total, 0.0);
  std::vector<char> var_in_constraint(vars.size(), 0);
  for (size_t c = 0; c < p.constraints.size(); c++) {
    std::vector<double> bc;
    if (EvalConstants(ctx, p.constraints[c], &bc)) {
      has_b = true;
      std::copy(bc.begin(), bc.end(), b_host.begin() + coff[c]);
    }
    for (const Term& t : p.constraints[c].terms)
      if (t.var >= 0) var_in_constraint[t.var] = 1;
  }
  // ---- row-sharded constraints (eps_problem_constraint_sharded): which rows are this rank's shard, which
  // variables live on those rows, the global A_.m() / A_.n() of the tolerances (prox_admm.cc:41-42,206-207)
  c_sharded.assign(p.constraints.size(), 0);
  m_global = 0;
  std::vector<int64_t> var_global(vars.size(), -1);   // global size of a sharded variable
  for (size_t c = 0; c < p.constraints.size(); c++) {
    const bool sh = p.constraints[c].global_rows > 0 && ctx->n_ranks > 1;
    c_sharded[c] = sh;
    any_sharded = any_sharded || sh;
    m_global += sh ? p.constraints[c].global_rows : p.constraints[c].rows;
    if (!sh) continue;
    for (const Term& t : p.constraints[c].terms) {
      if (t.var < 0) continue;
      const bool diag = !t.map || t.map->kind == MK_SCALAR || t.map->kind == MK_DIAGONAL;
      if (diag) var_global[t.var] = p.constraints[c].global_rows;   // lives on the sharded rows
    }
  }
  if (any_sharded) {
    EPS_CHECK(!two_block, "row-sharded constraints are not supported by the two-block solver");
    EPS_CHECK(N <= 64, "row-sharded constraints: at most 64 prox functions (got %d)", N);
    for (size_t c = 0; c < p.constraints.size(); c++)
      for (const Term& t : p.constraints[c].terms) {
        if (t.var < 0 || var_global[t.var] < 0) continue;
        const bool diag = !t.map || t.map->kind == MK_SCALAR || t.map->kind == MK_DIAGONAL;
        EPS_CHECK(c_sharded[c] && diag, "variable %s lives on the rows of a sharded constraint but also enters "
                  "constraint %zu%s", vars[t.var].id.c_str(), c, diag ? "" : " through a general map");
      }
  }
  n_total = 0;
  for (size_t v = 0; v < vars.size(); v++)
    if (var_in_constraint[v]) n_total += var_global[v] >= 0 ? var_global[v] : vars[v].size;
  {
    double sq_sh = 0, sq_rep = 0;
    for (size_t c = 0; c < p.constraints.size(); c++)
      for (int64_t i = coff[c]; i < coff[c + 1]; i++) (c_sharded[c] ? sq_sh : sq_rep) += b_host[i] * b_host[i];
    if (any_sharded) {
      DevBuf<double> tmp(ctx, 1);
      tmp.upload(&sq_sh, 1);
      AllReduceSum(ctx, tmp.p, 1);
      tmp.download(&sq_sh, 1);
    }
    b_norm = std::sqrt(sq_sh + sq_rep);
  }

  // ---- variable ownership ----
  var_owner.assign(vars.size(), -1);
  for (int i = 0; i < N; i++)
    for (const AffineExpr& a : p.prox[i].args)
      for (const Term& t : a.terms)
        if (t.var >= 0) {
          EPS_CHECK(var_owner[t.var] < 0 || var_owner[t.var] == i,
                    "variable %s appears in prox %d and %d (problem is not separable)",
                    vars[t.var].id.c_str(), var_owner[t.var], i);
          var_owner[t.var] = i;
        }
  for (size_t v = 0; v < vars.size(); v++)
    EPS_CHECK(!var_in_constraint[v] || var_owner[v] >= 0,
              "variable %s appears in constraints but in no prox function", vars[v].id.c_str());

  // ---- per-operator layout ----
  ops.resize(N);
  var_x_off.assign(vars.size(), -1);
  int64_t x_total = 0, y_total = 0, max_n = 1;
  std::vector<std::vector<double>> g_hosts(N);
  std::vector<BlockOp> Hs(N);
  std::vector<int64_t> rows_h(N, 0);
  std::vector<char> h_sharded(N, 0);
  for (int i = 0; i < N; i++) {
    ProxOp& op = ops[i];
    const ProxFn& f = p.prox[i];
    op.type = f.type;
    EPS_CHECK(!f.epigraph || IsVecOpType(f.type, true),
              "prox %d: the epigraph form of type %d has no device operator", i, f.type);
    for (size_t v = 0; v < vars.size(); v++)
      if (var_owner[v] == i) op.vars.push_back((int)v);
    std::sort(op.vars.begin(), op.vars.end(),
              [&](int a, int b) { return vars[a].id < vars[b].id; });
    std::map<int, int64_t> voff;
    op.n = 0;
    for (int v : op.vars) {
      voff[v] = op.n;
      op.var_off.push_back(op.n);
      var_x_off[v] = x_total + op.n;
      op.n += vars[v].size;
    }
    op.x_off = x_total;
    x_total += op.n;
    max_n = std::max(max_n, op.n);

    // H, g
    BlockOp& H = Hs[i];
    H.ctx = ctx;
    std::vector<double>& g = g_hosts[i];
    bool any_g = false;
    int64_t hoff = 0;
    for (const AffineExpr& a : f.args) {
      op.arg_off.push_back(hoff);
      std::vector<double> ga;
      any_g = EvalConstants(ctx, a, &ga) || any_g;
      g.insert(g.end(), ga.begin(), ga.end());
      for (const Term& t : a.terms)
        if (t.var >= 0) H.blks.push_back(TermBlk(ctx, t, a.rows, (int)hoff, (int)voff[t.var]));
      if (a.global_rows > 0 && ctx->n_ranks > 1) h_sharded[i] = 1;
      hoff += a.rows;
    }
    op.arg_off.push_back(hoff);
    op.g_host = g;                 // zeros when no argument carries a constant
    if (!any_g) g.clear();
    rows_h[i] = hoff;
    H.rows = hoff;
    H.cols = op.n;

    // A restricted to this operator's variables, span
    int64_t lo = m_total, hi = 0;
    for (size_t c = 0; c < p.constraints.size(); c++)
      for (const Term& t : p.constraints[c].terms)
        if (t.var >= 0 && var_owner[t.var] == i) {
          lo = std::min(lo, coff[c]);
          hi = std::max(hi, coff[c + 1]);
        }
    if (hi <= lo) { lo = 0; hi = 0; }
    op.span_lo = lo;
    op.span_len = hi - lo;
    op.y_off = y_total;
    y_total += op.span_len;
    op.A.ctx = ctx;
    op.A.rows = op.span_len;
    op.A.cols = op.n;
    for (size_t c = 0; c < p.constraints.size(); c++)
      for (const Term& t : p.constraints[c].terms)
        if (t.var >= 0 && var_owner[t.var] == i)
          op.A.blks.push_back(TermBlk(ctx, t, p.constraints[c].rows, (int)(coff[c] - lo), (int)voff[t.var]));
    op.A.Compile();
    op.simple_scalar = op.A.simple_scalar && op.A.simple_row_off == 0 && op.span_len == op.n;
    op.a = op.simple_scalar ? op.A.simple_alpha : 0.0;
    if (any_sharded) {
      // local: every variable of f_i lives on sharded rows; a_sharded: replicated variables that enter sharded rows
      int n_local = 0;
      for (int v : op.vars) n_local += var_global[v] >= 0;
      EPS_CHECK(n_local == 0 || n_local == (int)op.vars.size(), "prox %d mixes sharded and replicated variables", i);
      op.local = n_local > 0;
      op.A_sh.ctx = op.A_rep.ctx = ctx;
      op.A_sh.rows = op.A_rep.rows = op.A.rows;
      op.A_sh.cols = op.A_rep.cols = op.A.cols;
      for (size_t c = 0; c < p.constraints.size(); c++)
        for (const Term& t : p.constraints[c].terms)
          if (t.var >= 0 && var_owner[t.var] == i)
            (c_sharded[c] ? op.A_sh : op.A_rep).blks.push_back(
                TermBlk(ctx, t, p.constraints[c].rows, (int)(coff[c] - lo), (int)voff[t.var]));
      op.a_sharded = !op.local && !op.A_sh.blks.empty();
      if (op.a_sharded) {
        EPS_CHECK(IsKktType(f.type), "prox %d (type %d): a replicated variable enters a row-sharded constraint; only "
                  "the linear-equality operators (SUM_SQUARE, AFFINE, CONSTANT) can own it", i, f.type);
        op.A_sh.Compile();
        op.A_rep.Compile();
      }
    }
  }

  // ---- operators ----
  for (int i = 0; i < N; i++) {
    ProxOp& op = ops[i];
    const ProxFn& f = p.prox[i];
    if (f.type == EPS_PROX_SECOND_ORDER_CONE) {
      // SecondOrderConeProx::Init (second_order_cone.cc:30-104): arguments (at t + gt, ax X + gx), both scalar
      // multiples of whole variables; A'A the same scalar on t and x.  In the stage's terms:
      // V = P (A'w / alphaA) + [gt/at ; gx/ax] with P the variable -> argument permutation,
      // x = P'(V' - shift).
      BlockOp& H = Hs[i];
      const int m = f.soc_m, n = f.soc_n;
      EPS_CHECK(f.args.size() == 2 && f.args[0].rows == m && f.args[1].rows == (int64_t)m * n && !f.epigraph,
                "prox %d (SECOND_ORDER_CONE): arguments (t: m, X: m x n) with m = %d, n = %d", i, m, n);
      EPS_CHECK(H.blks.size() == 2 && rows_h[i] == op.n, "prox %d (SECOND_ORDER_CONE): one variable per argument", i);
      double at = 0, ax = 0;
      op.Hop.ctx = ctx; op.Hop.rows = H.rows; op.Hop.cols = H.cols;
      for (auto& b : H.blks) {
        EPS_CHECK(b.map->kind == MK_SCALAR, "prox %d (SECOND_ORDER_CONE): arguments must be scalar multiples of variables", i);
        const double h = b.scale * b.map->alpha;
        if (b.row_off == 0) at = h; else ax = h;
        Blk pb;
        pb.row_off = b.row_off; pb.col_off = b.col_off;
        pb.map = IdentityMap(ctx, b.m());
        op.Hop.blks.push_back(pb);
      }
      EPS_CHECK(at > 0 && ax != 0, "prox %d (SECOND_ORDER_CONE): at = %g, ax = %g", i, at, ax);
      op.Hop.Compile();
      std::vector<double> da;
      EPS_CHECK(op.A.GramDiagonal(&da), "prox %d (SECOND_ORDER_CONE): A'A must be diagonal", i);
      for (int64_t j = 0; j < op.n; j++) EPS_CHECK(da[j] == da[0] && da[0] > 0, "A'A not scalar matrix");
      op.rows_h = rows_h[i];
      op.vec_bg = 1 / da[0];
      op.vec_cinv = 1.0;
      VecOpDesc& d = op.vdesc;
      d.type = f.type; d.epigraph = false;
      d.n0 = m; d.rows = rows_h[i]; d.soc_n = n; d.soc_a = at / std::fabs(ax);
      op.g_off = -1;
      if (!g_hosts[i].empty()) {
        op.g_off = (int64_t)consts_host.size();
        for (int64_t j = 0; j < op.rows_h; j++) consts_host.push_back(g_hosts[i][j] / (j < m ? at : ax));
      }
      op.g_div0 = at; op.g_div1 = ax; op.g_split = m;
      op.vbuf.alloc(ctx, op.rows_h);
      op.vtmp.alloc(ctx, op.n);
      VecOpInit(ctx, d, &op.vwork);
      description += Format("prox %d SECOND_ORDER_CONE m=%d n=%d a=%g span=[%lld,+%lld)\n", i, m, n, d.soc_a,
                            (long long)op.span_lo, (long long)op.span_len);
    } else if (IsVecOpType(f.type, f.epigraph)) {
      // VectorProx::InitScalar (vector_prox.cc:51-70) with several arguments: every argument is
      // a scalar multiple of one whole variable; H'H = beta I and H A'A H' = gamma I.
      BlockOp& H = Hs[i];
      EPS_CHECK(rows_h[i] == op.n, "prox %d (%s): arguments have %lld rows, variables %lld entries", i,
                TypeName(f.type), (long long)rows_h[i], (long long)op.n);
      std::vector<double> dh, da;
      H.Compile();
      EPS_CHECK(H.GramDiagonal(&dh) && op.A.GramDiagonal(&da),
                "prox %d (%s): H'H and A'A must be diagonal", i, TypeName(f.type));
      for (auto& b : H.blks)
        EPS_CHECK(b.map->kind == MK_SCALAR && b.m() == b.n(),
                  "prox %d (%s): every argument must be a scalar multiple of one variable", i, TypeName(f.type));
      // with scalar blocks H is a signed scaled permutation: H A'A H' = diag(dh .* da) permuted
      const double beta = dh[0], gamma = dh[0] * da[0];
      for (int64_t j = 0; j < op.n; j++)
        EPS_CHECK(dh[j] == beta && dh[j] * da[j] == gamma,
                  "prox %d (%s): affine transformation is not scalar (diagonal case is not supported)", i,
                  TypeName(f.type));
      EPS_CHECK(beta > 0 && gamma > 0, "prox %d: H'H = %g, H A'A H' = %g", i, beta, gamma);
      op.Hop.ctx = ctx; op.Hop.rows = H.rows; op.Hop.cols = H.cols; op.Hop.blks = H.blks;
      op.Hop.Compile();
      op.rows_h = rows_h[i];
      op.vec_bg = beta / gamma;
      op.vec_cinv = 1 / beta;
      VecOpDesc& d = op.vdesc;
      d.type = f.type;
      d.epigraph = f.epigraph;
      d.n0 = f.args[0].rows;
      d.rows = rows_h[i];
      d.lam = f.alpha * beta * beta / gamma;
      d.zone_m = f.type == EPS_PROX_SUM_DEADZONE ? f.zone_m : 0.0;
      d.zone_alpha = 1; d.zone_beta = f.type == EPS_PROX_SUM_HINGE ? 0 : 1;
      if (f.type == EPS_PROX_SUM_QUANTILE) {
        EPS_CHECK(f.q_alpha.size() == 1 && f.q_beta.size() == 1,
                  "SUM_QUANTILE epigraph: only constant alpha/beta (the reference indexes them by the "
                  "position in its permuted work array, scaled_zone.cc:238-252)");
        d.zone_alpha = f.q_alpha[0];
        d.zone_beta = f.q_beta[0];
      }
      d.mat_m = f.mat_m; d.mat_n = f.mat_n;
      if (f.type == EPS_PROX_EXP)
        EPS_CHECK(f.epigraph && f.args.size() == 2 && f.args[1].rows == f.args[0].rows,
                  "prox %d (EXP): elementwise epigraph form over (x, t) of equal size", i);
      else if (f.type == EPS_PROX_SUM_KL_DIV)
        EPS_CHECK(f.args.size() == (f.epigraph ? 3u : 2u) && f.args[0].rows == f.args[1].rows &&
                      (!f.epigraph || f.args[2].rows == 1),
                  "SUM_KL_DIV takes two equal-sized arguments (and t in the epigraph form)");
      else if (f.epigraph)
        EPS_CHECK(f.args.size() == 2 && f.args[1].rows == 1, "prox %d (%s): epigraph form takes (x, t)", i,
                  TypeName(f.type));
      if (f.type == EPS_PROX_LOG_SUM_EXP && !f.epigraph) EPS_CHECK(f.args.size() == 1, "LOG_SUM_EXP takes one argument");
      if (f.type == EPS_PROX_MAX && !f.epigraph) EPS_CHECK(f.args.size() == 1, "MAX takes one argument");
      if (f.type == EPS_PROX_SUM_LARGEST) {
        EPS_CHECK(f.args.size() == (f.epigraph ? 2u : 1u), "SUM_LARGEST takes one argument (and t in the epigraph form)");
        d.k = f.sum_largest_k;
      }
      op.g_off = -1;
      if (!g_hosts[i].empty()) {
        op.g_off = (int64_t)consts_host.size();
        consts_host.insert(consts_host.end(), g_hosts[i].begin(), g_hosts[i].end());
      }
      op.vbuf.alloc(ctx, op.rows_h);
      op.vtmp.alloc(ctx, op.n);
      VecOpInit(ctx, d, &op.vwork);
      description += Format("prox %d %s%s rows=%lld span=[%lld,+%lld) lambda=%.6g beta=%g gamma=%g\n", i,
                            TypeName(f.type), f.epigraph ? " (epigraph)" : "", (long long)op.rows_h,
                            (long long)op.span_lo, (long long)op.span_len, d.lam, beta, gamma);
    } else if (IsKktType(f.type)) {
      auto k = std::make_unique<KktSolver>();
      k->ctx = ctx;
      k->mode = params.kkt_mode;
      k->n = op.n;
      k->cg_tol = params.cg_tol;
      k->cg_max_iter = params.cg_max_iterations;
      k->gram_max_n = params.gram_max_n;
      k->rows_a = op.span_len;
      k->A.ctx = ctx; k->A.rows = op.A.rows; k->A.cols = op.A.cols; k->A.blks = op.A.blks;
      if (op.a_sharded) {
        k->a_sharded = true;
        k->A_sh.ctx = k->A_rep.ctx = ctx;
        k->A_sh.rows = k->A_rep.rows = op.A.rows;
        k->A_sh.cols = k->A_rep.cols = op.A.cols;
        k->A_sh.blks = op.A_sh.blks;
        k->A_rep.blks = op.A_rep.blks;
      }
      std::vector<double> c_host;
      if (f.type == EPS_PROX_SUM_SQUARE) {
        k->alpha2 = 2.0 * f.alpha;
        k->H.ctx = ctx; k->H.rows = Hs[i].rows; k->H.cols = Hs[i].cols; k->H.blks = Hs[i].blks;
        k->rows_h = rows_h[i];
        k->h_sharded = h_sharded[i] != 0;
        k->Init(g_hosts[i], c_host);
      } else {
        k->alpha2 = 0.0;
        k->H.ctx = ctx; k->H.rows = 0; k->H.cols = op.n;
        k->rows_h = 0;
        if (f.type == EPS_PROX_AFFINE) {
          // c = alpha * (1 x n linear map)'   (affine.cc:8-17,26-28)
          EPS_CHECK(rows_h[i] == 1, "AFFINE prox expects a 1 x n argument, got %lld rows", (long long)rows_h[i]);
          Hs[i].Compile();
          DevBuf<double> one(ctx, 1), cdev(ctx, op.n);
          VecFill(ctx, 1, f.alpha, one.p);
          Hs[i].ApplyT(one.p, cdev.p);
          c_host.resize(op.n);
          cdev.download(c_host.data(), op.n);
        }
        k->Init(std::vector<double>(), c_host);
      }
      description += Format("prox %d %s vars=%lld span=[%lld,+%lld) %s\n", i, TypeName(f.type),
                            (long long)op.n, (long long)op.span_lo, (long long)op.span_len,
                            k->description.c_str());
      op.kkt = std::move(k);
    } else if (IsElemType(f.type) || f.type == EPS_PROX_NORM_2 || f.type == EPS_PROX_TOTAL_VARIATION_1D) {
      // VectorProx::InitScalar (vector_prox.cc:51-70) for scalar H and A blocks.
      const BlockOp& H = Hs[i];
      EPS_CHECK(f.args.size() == 1 && op.vars.size() == 1,
                "prox %d (%s): one argument over one variable expected", i, TypeName(f.type));
      EPS_CHECK(H.blks.size() == 1 && H.blks[0].diag_like() && H.rows == op.n,
                "prox %d (%s): argument must be a scalar or diagonal multiple of its variable "
                "(a general H is a reference CHECK too, vector_prox.cc:118-120)", i, TypeName(f.type));
      EPS_CHECK(op.simple_scalar,
                "prox %d (%s): the variable must enter exactly one constraint through a scalar map", i,
                TypeName(f.type));
      const bool h_diagonal = H.blks[0].map->kind == MK_DIAGONAL;
      // diagonal H: the scalar reduction is tried first, as in VectorProx::Init (vector_prox.cc:118-119)
      bool h_const = true;
      if (h_diagonal)
        for (double d : H.blks[0].map->h_diag) h_const = h_const && d == H.blks[0].map->h_diag[0];
      const double h = H.blks[0].scale * (h_diagonal ? H.blks[0].map->h_diag[0] : H.blks[0].map->alpha);
      const double a = op.a;
      const bool diag_reduction = h_diagonal && (!h_const || h == 0.0);
      EPS_CHECK(!diag_reduction || IsElemType(f.type),
                "prox %d (%s): a diagonally weighted argument needs an elementwise lambda, which this operator "
                "does not take (VectorProxInput::lambda, vector_prox.cc:185-188)", i, TypeName(f.type));
      EPS_CHECK(f.axis < 0 || f.type == EPS_PROX_NORM_2,
                "prox %d (%s): has_axis is supported for NORM_2 only on the device path", i, TypeName(f.type));
      const double beta = diag_reduction ? 1.0 : h * h;
      const double gamma = diag_reduction ? 1.0 : ((h * a) * a) * h;
      EPS_CHECK(gamma > 0, "prox %d: H A'A H' = %g", i, gamma);
      ElemSeg& s = op.seg;
      s.u_off = op.span_lo;
      s.y_off = op.y_off;
      s.x_off = op.x_off;
      s.len = (int32_t)op.n;
      s.type = f.type;
      s.Bs = ((beta / gamma) * h) * a;
      s.Cs = (1 / beta) * h;
      s.As = a;
      s.lam = f.alpha * beta * beta / gamma;
      EPS_CHECK(s.lam >= 0, "lambda = %g < 0", s.lam);
      s.g_off = -1;
      if (!g_hosts[i].empty()) {
        s.g_off = (int64_t)consts_host.size();
        consts_host.insert(consts_host.end(), g_hosts[i].begin(), g_hosts[i].end());
      }
      s.q_off = -1;
      s.d_off = -1;
      s.stride = 1;
      if (diag_reduction) {
        // VectorProx::InitDiagonal (vector_prox.cc:72-116) with H = diag(h_i), A = a I:
        //   beta_i = h_i^2, gamma_i = h_i^2 a^2; where gamma_i == 0: lambda_i = 0, beta_i = gamma_i = delta_i = 1
        //   B = H diag(beta/gamma) A',  C = diag(1/beta) H',  D = (A'A)^-1 diag(delta) A',  lambda_i = alpha beta_i^2/gamma_i
        const std::vector<double>& hd = H.blks[0].map->h_diag;
        const double hs = H.blks[0].scale;
        s.d_off = (int64_t)consts_host.size();
        consts_host.resize(consts_host.size() + 4 * (size_t)op.n);
        double* Bv = consts_host.data() + s.d_off;
        double* Cv = Bv + op.n;
        double* Dv = Cv + op.n;
        double* Lv = Dv + op.n;
        for (int64_t j = 0; j < op.n; j++) {
          const double hj = hs * hd[j];
          double bj = hj * hj, gj = ((hj * a) * a) * hj, dj = 0.0, lj;
          if (gj != 0.0) {
            lj = f.alpha * bj * bj / gj;
          } else {
            lj = 0.0; bj = 1.0; gj = 1.0; dj = 1.0;
          }
          EPS_CHECK(lj >= 0, "lambda = %g < 0", lj);
          Bv[j] = (hj * (bj / gj)) * a;
          Cv[j] = (1 / bj) * hj;
          Dv[j] = (1 / (a * a)) * dj * a;
          Lv[j] = lj;
        }
      }
      s.zone_m = 0; s.zone_alpha = 1; s.zone_beta = 1;
      if (f.type == EPS_PROX_SUM_DEADZONE) s.zone_m = f.zone_m;
      if (f.type == EPS_PROX_SUM_HINGE) s.zone_beta = 0;
      if (f.type == EPS_PROX_SUM_QUANTILE) {
        EPS_CHECK(!f.q_alpha.empty(), "SUM_QUANTILE without alpha/beta");
        if (f.q_alpha.size() == 1) {
          s.zone_alpha = f.q_alpha[0];
          s.zone_beta = f.q_beta[0];
        } else {
          EPS_CHECK((int64_t)f.q_alpha.size() == op.n, "quantile vectors have %zu entries, need %lld",
                    f.q_alpha.size(), (long long)op.n);
          s.q_off = (int64_t)consts_host.size();
          consts_host.insert(consts_host.end(), f.q_alpha.begin(), f.q_alpha.end());
          consts_host.insert(consts_host.end(), f.q_beta.begin(), f.q_beta.end());
        }
      }
      if (f.axis >= 0) {
        // has_axis: one NORM_2 segment per column (axis 0) or row (axis 1) of mat(arg, m, n), column-major
        EPS_CHECK((int64_t)f.ax_m * f.ax_n == op.n, "prox %d: axis dims %d x %d vs %lld entries", i, f.ax_m, f.ax_n,
                  (long long)op.n);
        const int k = f.axis == 0 ? f.ax_n : f.ax_m;
        for (int q = 0; q < k; q++) {
          ElemSeg t = s;
          const int64_t first = f.axis == 0 ? (int64_t)q * f.ax_m : q;
          t.u_off += first; t.y_off += first; t.x_off += first;
          if (t.g_off >= 0) t.g_off += first;
          t.len = f.axis == 0 ? f.ax_m : f.ax_n;
          t.stride = f.axis == 0 ? 1 : f.ax_m;
          op.axis_segs.push_back(t);
        }
      }
      description += Format("prox %d %s n=%lld rows=[%lld,+%lld) lambda=%.6g B=%g C=%g A=%g%s%s\n", i,
                            TypeName(f.type), (long long)op.n, (long long)op.span_lo,
                            (long long)op.span_len, s.lam, s.Bs, s.Cs, s.As,
                            diag_reduction ? " (diagonal reduction: per-element B, C, D, lambda)" : "",
                            f.axis >= 0 ? (f.axis == 0 ? " axis=0 (columns)" : " axis=1 (rows)") : "");
    } else {
      throw Error(EPS_ERR_INVALID, Format("prox %d: type %d (%s) has no device operator", i, f.type, TypeName(f.type)));
    }
  }

  // ---- stages: consecutive same-type elementwise/block operators on disjoint rows ----
  for (int i = 0; i < N; i++) {
    const ProxOp& op = ops[i];
    int kind = ST_KKT;
    if (op.rows_h > 0) kind = ST_VEC;
    else if (IsElemType(op.type)) kind = ST_ELEM;
    else if (op.type == EPS_PROX_NORM_2) kind = ST_NORM2;
    else if (op.type == EPS_PROX_TOTAL_VARIATION_1D) kind = ST_TV;
    bool merged = false;
    if ((kind == ST_ELEM || kind == ST_NORM2) && !stages.empty()) {
      Stage& last = *stages.back();
      if (last.kind == kind && last.type == op.type) {
        bool disjoint = true;
        for (int j : last.members) {
          const ProxOp& o = ops[j];
          if (op.span_lo < o.span_lo + o.span_len && o.span_lo < op.span_lo + op.span_len) disjoint = false;
        }
        if (disjoint) { last.members.push_back(i); merged = true; }
      }
    }
    if (!merged) {
      auto st = std::make_unique<Stage>();
      st->kind = kind;
      st->type = op.type;
      st->members.push_back(i);
      stages.push_back(std::move(st));
    }
  }
  for (auto& stp : stages) {
    Stage& st = *stp;
    std::vector<ElemSeg> segs;
    std::vector<int64_t> prefix(1, 0);
    st.scalar_a = true;
    for (int j : st.members) {
      const ProxOp& op = ops[j];
      if (op.axis_segs.empty()) segs.push_back(op.seg);
      else segs.insert(segs.end(), op.axis_segs.begin(), op.axis_segs.end());
      SegPair pr;
      pr.a_off = op.span_lo;
      pr.b_off = op.y_off;
      pr.len = (int32_t)op.span_len;
      pr.coef = op.a * op.a;
      st.pairs_host.push_back(pr);
      prefix.push_back(prefix.back() + op.span_len);
      if (op.axis_segs.empty()) st.max_len = std::max<int>(st.max_len, (int)op.span_len);
      else st.max_len = std::max<int>(st.max_len, op.axis_segs[0].len);
      if (!op.simple_scalar) st.scalar_a = false;
    }
    st.n_segs = (int)segs.size();
    st.total = prefix.back();
    st.segs.alloc(ctx, segs.size());
    st.segs.upload(segs.data(), segs.size());
    st.pairs.alloc(ctx, st.pairs_host.size());
    st.pairs.upload(st.pairs_host.data(), st.pairs_host.size());
    st.prefix.alloc(ctx, prefix.size());
    st.prefix.upload(prefix.data(), prefix.size());
    if (st.kind == ST_TV) st.tv_work.alloc(ctx, Tv1dWorkDoubles(ops[st.members[0]].n));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }

  // ---- slabs ----
  u.alloc(ctx, std::max<int64_t>(m_total, 1));
  b.alloc(ctx, std::max<int64_t>(m_total, 1));
  b.upload(b_host.data(), m_total);
  x.alloc(ctx, std::max<int64_t>(x_total, 1));
  ybuf[0].alloc(ctx, std::max<int64_t>(y_total, 1));
  ybuf[1].alloc(ctx, std::max<int64_t>(y_total, 1));
  tmp_m.alloc(ctx, std::max<int64_t>(m_total, 1));
  diff_m.alloc(ctx, std::max<int64_t>(m_total, 1));
  tmp_n.alloc(ctx, max_n);
  rhs_n.alloc(ctx, max_n);
  resid.alloc(ctx, 8);
  consts.alloc(ctx, std::max<size_t>(consts_host.size(), 1));
  consts.upload(consts_host.data(), consts_host.size());
  CUDA_CHECK(cudaMallocHost(&h_resid, 8 * sizeof(double)));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  ResetVariables();

  if (two_block) InitTwoBlock(original);
  graph_ok = params.use_graph && !params.profile && !two_block;
  for (auto& op : ops)
    if (op.kkt && op.kkt->mode == EPS_KKT_CG) graph_ok = false;  // host-checked inner loop
  for (auto& stp : stages)
    if (stp->kind == ST_TV || stp->kind == ST_VEC) graph_ok = false;  // host-checked rounds / library sorts
  if (any_sharded) graph_ok = false;                                  // collectives inside the iteration
  memset(&status, 0, sizeof(status));
  status.state = EPS_NOT_STARTED;
  status.m = (int32_t)(any_sharded ? m_global : m_total);
  status.n = (int32_t)n_total;
  if (any_sharded) resid_sh.alloc(ctx, 2 * (1 + 3 * (size_t)N));
  if (two_block) { status.m = (int32_t)m_c; status.n = (int32_t)tb_n; }  // H.A.m(), H.A.n()
  status.N = N;
  description = Format("%s m=%lld n=%lld N=%d stages=%zu graph=%d\n", two_block ? "ProxADMMTwoBlock" : "ProxADMM",
                       (long long)status.m, (long long)status.n, N, stages.size(), graph_ok ? 1 : 0) + description;
}

// InitConstraints of the two-block solver (prox_admm_two_block.cc:21-49): H, b over the x/z
// layout, and the solve object of the ZeroProx projection.
void eps_solver::InitTwoBlock(const eps_problem& p) {
  n_all = 0;
  for (auto& op : ops) {
    EPS_CHECK(op.span_lo == op.x_off && op.span_len == op.n && (op.n == 0 || op.simple_scalar),
              "consensus layout: prox spans [%lld,+%lld) vs x offset %lld", (long long)op.span_lo,
              (long long)op.span_len, (long long)op.x_off);
    n_all += op.n;
  }
  EPS_CHECK(n_all == m_total, "consensus layout: %lld variables vs %lld rows", (long long)n_all, (long long)m_total);
  std::vector<int64_t> coff(p.constraints.size() + 1, 0);
  for (size_t c = 0; c < p.constraints.size(); c++) coff[c + 1] = coff[c] + p.constraints[c].rows;
  m_c = coff.back();
  Hc.ctx = ctx;
  Hc.rows = m_c;
  Hc.cols = n_all;
  std::vector<double> hb_host(m_c, 0.0), mask_host(n_all, 0.0);
  tb_n = 0;
  std::vector<char> seen(vars.size(), 0);
  for (size_t c = 0; c < p.constraints.size(); c++) {
    std::vector<double> bc;
    if (EvalConstants(ctx, p.constraints[c], &bc)) std::copy(bc.begin(), bc.end(), hb_host.begin() + coff[c]);
    for (const Term& t : p.constraints[c].terms) {
      if (t.var < 0) continue;
      EPS_CHECK(var_x_off[t.var] >= 0, "variable %s appears in constraints but in no prox function",
                vars[t.var].id.c_str());
      Hc.blks.push_back(TermBlk(ctx, t, p.constraints[c].rows, (int)coff[c], (int)var_x_off[t.var]));
      if (!seen[t.var]) {
        seen[t.var] = 1;
        tb_n += vars[t.var].size;
        std::fill(mask_host.begin() + var_x_off[t.var], mask_host.begin() + var_x_off[t.var] + vars[t.var].size, 1.0);
      }
    }
  }
  Hc.Compile();
  has_mask = tb_n != n_all;
  const int64_t na = std::max<int64_t>(n_all, 1), mc = std::max<int64_t>(m_c, 1);
  zbuf[0].alloc(ctx, na); zbuf[1].alloc(ctx, na); ub.alloc(ctx, na); zeros.alloc(ctx, na);
  v_tmp.alloc(ctx, na); hb.alloc(ctx, mc); r_m.alloc(ctx, mc); y_m.alloc(ctx, mc); cmask.alloc(ctx, na);
  zeros.zero();
  constraint_off = coff;       // eps_solver_set_constant addresses the user's constraints
  tb_hb_host = hb_host;
  hb.upload(hb_host.data(), m_c);
  cmask.upload(mask_host.data(), n_all);
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  if (m_c > 0) {
    // (H H') y = r  ==  KktSolver with "H" := H' (blocks transposed), 2a = 1, no A term
    auto k = std::make_unique<KktSolver>();
    k->ctx = ctx;
    k->mode = params.kkt_mode;
    k->n = m_c;
    k->rows_h = n_all;
    k->rows_a = 0;
    k->alpha2 = 1.0;
    k->cg_tol = params.cg_tol;
    k->cg_max_iter = params.cg_max_iterations;
    k->gram_max_n = params.gram_max_n;
    k->H.ctx = ctx; k->H.rows = n_all; k->H.cols = m_c;
    for (const Blk& b : Hc.blks) {
      Blk t = b;
      t.row_off = b.col_off;
      t.col_off = b.row_off;
      t.trans = !b.trans;
      k->H.blks.push_back(t);
    }
    k->A.ctx = ctx; k->A.rows = 0; k->A.cols = m_c;
    k->Init(std::vector<double>(), std::vector<double>());
    description += Format("two-block projection onto {Hz + b = 0}: H %lld x %lld, %s\n", (long long)m_c,
                          (long long)n_all, k->description.c_str());
    proj = std::move(k);
  }
  ResetVariables();
}

void eps_solver::ResetVariables() {
  if (two_block && ub.n) {
    zbuf[0].zero(); zbuf[1].zero(); ub.zero(); y_m.zero();
    zpar = 0;
  }
  // x is about to change under the CG solvers: the residual they carry between solves is void
  for (auto& op : ops)
    if (op.kkt) { op.kkt->carry_valid = false; if (op.kkt->inner) op.kkt->inner->carry_valid = false; }
  if (proj) proj->carry_valid = false;
  u.zero();
  x.zero();
  ybuf[0].zero();
  ybuf[1].zero();
  for (auto& stp : stages) { stp->tv_work.zero(); stp->tv_tag[0] = stp->tv_tag[1] = 0; }  // cold active set for the TV solver
  parity = 0;
  iter = 0;
  started = false;
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
}

void eps_solver::InvalidateGraphs() {
  for (int i = 0; i < 2; i++) {
    if (graph_exec[i]) cudaGraphExecDestroy(graph_exec[i]);
    graph_exec[i] = nullptr;
    eager_done[i] = 0;
  }
}

// Solver::SetParameterValue (solver.cc:109-116) for constant vectors.  The reference rewrites the Constant
// inside its Problem proto and re-runs InitConstraints / InitProxOperators on the next Solve()
// (prox_admm.cc:110-120), i.e. it re-reads b_ and every g_i and refactorises although no matrix changed;
// here only what depends on the constant is refreshed: b (and ||b||), the g of an elementwise / vector
// operator, or c0 = -2a H'g - c of a KKT operator.  Factors, SELL streams and the iterate stay.
void eps_solver::SetConstant(int prox, int index, const double* values) {
  EPS_CHECK(values, "null constant");
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  if (prox < 0) {
    EPS_CHECK(index >= 0 && index + 1 < (int)constraint_off.size(), "constraint %d of %d", index,
              (int)constraint_off.size() - 1);
    const int64_t lo = constraint_off[index], len = constraint_off[index + 1] - lo;
    if (two_block) {
      std::copy(values, values + len, tb_hb_host.begin() + lo);
      if (len) UploadBytes(ctx, hb.p + lo, tb_hb_host.data() + lo, len * sizeof(double));
    } else {
      std::copy(values, values + len, b_host.begin() + lo);
      if (len) UploadBytes(ctx, b.p + lo, b_host.data() + lo, len * sizeof(double));
      if (!has_b) { has_b = true; InvalidateGraphs(); }   // the captured iteration skipped "u -= b"
      b_norm = 0;
      for (double v : b_host) b_norm += v * v;
      b_norm = std::sqrt(b_norm);
    }
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    return;
  }
  EPS_CHECK(prox < N, "prox function %d of %d", prox, N);
  ProxOp& op = ops[prox];
  EPS_CHECK(index >= 0 && index + 1 < (int)op.arg_off.size(), "argument %d of prox function %d", index, prox);
  const int64_t lo = op.arg_off[index], len = op.arg_off[index + 1] - lo;
  std::copy(values, values + len, op.g_host.begin() + lo);
  if (op.kkt) {
    // AFFINE / CONSTANT: only the linear part of the argument enters (affine.cc:26-28)
    if (op.type == EPS_PROX_SUM_SQUARE) op.kkt->SetConstant(op.g_host);
    return;
  }
  const bool vec = op.rows_h > 0;   // ST_VEC keeps its offset in the operator, the others in the segment
  int64_t& goff = vec ? op.g_off : op.seg.g_off;
  const int64_t rows = (int64_t)op.g_host.size();
  if (goff < 0) {
    // the operator was built without a constant: give it a slot in the arena
    goff = (int64_t)consts_host.size();
    consts_host.resize(consts_host.size() + rows, 0.0);
    consts.alloc(ctx, consts_host.size());
    consts.upload(consts_host.data(), consts_host.size());
    InvalidateGraphs();
    for (auto& stp : stages) {
      Stage& st = *stp;
      for (size_t k = 0; k < st.members.size(); k++)
        if (st.members[k] == prox && st.segs.n > k)
          UploadBytes(ctx, st.segs.p + k, &op.seg, sizeof(ElemSeg));
    }
  }
  for (int64_t j = 0; j < rows; j++) {
    const double div = op.type == EPS_PROX_SECOND_ORDER_CONE ? (j < op.g_split ? op.g_div0 : op.g_div1) : 1.0;
    consts_host[goff + j] = op.g_host[j] / div;
  }
  UploadBytes(ctx, consts.p + goff, consts_host.data() + goff, rows * sizeof(double));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
}

eps_solver::~eps_solver() {
  for (int i = 0; i < 2; i++)
    if (graph_exec[i]) cudaGraphExecDestroy(graph_exec[i]);
  if (h_resid) cudaFreeHost(h_resid);
}
