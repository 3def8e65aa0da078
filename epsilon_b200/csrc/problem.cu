// C ABI: building the prox-affine problem (AffineOperator terms).
#include "problem.h"

using namespace eps;

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

int eps_problem_create(eps_ctx* ctx, eps_problem** out) {
  EPS_API_BEGIN
  EPS_CHECK(ctx && out, "eps_problem_create arguments");
  auto* p = new eps_problem();
  p->ctx = &ctx->c;
  *out = p;
  EPS_API_END
}

void eps_problem_destroy(eps_problem* p) { delete p; }

int eps_problem_add_variable(eps_problem* p, const char* var_id, int size, int* var_index) {
  EPS_API_BEGIN
  EPS_CHECK(p && var_id && size >= 0, "eps_problem_add_variable arguments");
  for (auto& v : p->vars) EPS_CHECK(v.id != var_id, "duplicate variable %s", var_id);
  p->vars.push_back({var_id, size});
  if (var_index) *var_index = (int)p->vars.size() - 1;
  EPS_API_END
}

int eps_problem_add_constraint(eps_problem* p, int rows, int* constraint_index) {
  EPS_API_BEGIN
  EPS_CHECK(p && rows >= 0, "eps_problem_add_constraint arguments");
  AffineExpr e;
  e.rows = rows;
  p->constraints.push_back(std::move(e));
  if (constraint_index) *constraint_index = (int)p->constraints.size() - 1;
  EPS_API_END
}

int eps_problem_add_prox(eps_problem* p, int prox_type, double alpha, int epigraph, int* prox_index) {
  EPS_API_BEGIN
  EPS_CHECK(p, "null problem");
  ProxFn f;
  f.type = prox_type;
  f.alpha = alpha;
  f.epigraph = epigraph != 0;
  p->prox.push_back(std::move(f));
  if (prox_index) *prox_index = (int)p->prox.size() - 1;
  EPS_API_END
}

int eps_problem_prox_zone(eps_problem* p, int prox_index, double zone_m, const double* q_alpha,
                          const double* q_beta, int n_q) {
  EPS_API_BEGIN
  EPS_CHECK(p && prox_index >= 0 && prox_index < (int)p->prox.size(), "prox index %d", prox_index);
  ProxFn& f = p->prox[prox_index];
  f.zone_m = zone_m;
  if (q_alpha && q_beta && n_q > 0) {
    f.q_alpha.assign(q_alpha, q_alpha + n_q);
    f.q_beta.assign(q_beta, q_beta + n_q);
  }
  EPS_API_END
}

int eps_problem_prox_sum_largest(eps_problem* p, int prox_index, int k) {
  EPS_API_BEGIN
  EPS_CHECK(p && prox_index >= 0 && prox_index < (int)p->prox.size() && k >= 0, "prox index %d, k %d", prox_index, k);
  p->prox[prox_index].sum_largest_k = k;
  EPS_API_END
}

int eps_problem_prox_soc_dims(eps_problem* p, int prox_index, int m, int n) {
  EPS_API_BEGIN
  EPS_CHECK(p && prox_index >= 0 && prox_index < (int)p->prox.size() && m > 0 && n > 0, "prox index %d, %d x %d",
            prox_index, m, n);
  p->prox[prox_index].soc_m = m;
  p->prox[prox_index].soc_n = n;
  EPS_API_END
}

int eps_problem_prox_mat_dims(eps_problem* p, int prox_index, int m, int n) {
  EPS_API_BEGIN
  EPS_CHECK(p && prox_index >= 0 && prox_index < (int)p->prox.size() && m > 0 && n > 0, "prox index %d, %d x %d",
            prox_index, m, n);
  p->prox[prox_index].mat_m = m;
  p->prox[prox_index].mat_n = n;
  EPS_API_END
}

int eps_problem_prox_axis(eps_problem* p, int prox_index, int axis, int m, int n) {
  EPS_API_BEGIN
  EPS_CHECK(p && prox_index >= 0 && prox_index < (int)p->prox.size() && (axis == 0 || axis == 1) && m > 0 && n > 0,
            "prox index %d, axis %d over %d x %d", prox_index, axis, m, n);
  p->prox[prox_index].axis = axis;
  p->prox[prox_index].ax_m = m;
  p->prox[prox_index].ax_n = n;
  EPS_API_END
}

int eps_problem_prox_add_arg(eps_problem* p, int prox_index, int rows, int* arg_index) {
  EPS_API_BEGIN
  EPS_CHECK(p && prox_index >= 0 && prox_index < (int)p->prox.size() && rows >= 0, "prox index %d", prox_index);
  AffineExpr e;
  e.rows = rows;
  p->prox[prox_index].args.push_back(std::move(e));
  if (arg_index) *arg_index = (int)p->prox[prox_index].args.size() - 1;
  EPS_API_END
}

int eps_problem_prox_arg_sharded(eps_problem* p, int prox_index, int arg_index, int64_t global_rows) {
  EPS_API_BEGIN
  EPS_CHECK(p && prox_index >= 0 && prox_index < (int)p->prox.size(), "prox index %d", prox_index);
  ProxFn& f = p->prox[prox_index];
  EPS_CHECK(arg_index >= 0 && arg_index < (int)f.args.size(), "arg index %d", arg_index);
  EPS_CHECK(global_rows >= f.args[arg_index].rows, "global rows %lld < local rows %d",
            (long long)global_rows, f.args[arg_index].rows);
  f.args[arg_index].global_rows = global_rows;
  EPS_API_END
}

int eps_problem_constraint_sharded(eps_problem* p, int constraint_index, int64_t global_rows) {
  EPS_API_BEGIN
  EPS_CHECK(p && constraint_index >= 0 && constraint_index < (int)p->constraints.size(), "constraint index %d",
            constraint_index);
  EPS_CHECK(global_rows >= p->constraints[constraint_index].rows, "global rows %lld < local rows %d",
            (long long)global_rows, p->constraints[constraint_index].rows);
  p->constraints[constraint_index].global_rows = global_rows;
  EPS_API_END
}

int eps_problem_add_term(eps_problem* p, int prox_index, int index, eps_map* map, int var_index,
                         const double* constant) {
  EPS_API_BEGIN
  EPS_CHECK(p, "null problem");
  AffineExpr* e = nullptr;
  if (prox_index < 0) {
    EPS_CHECK(index >= 0 && index < (int)p->constraints.size(), "constraint index %d", index);
    e = &p->constraints[index];
  } else {
    EPS_CHECK(prox_index < (int)p->prox.size(), "prox index %d", prox_index);
    EPS_CHECK(index >= 0 && index < (int)p->prox[prox_index].args.size(), "arg index %d", index);
    e = &p->prox[prox_index].args[index];
  }
  Term t;
  if (map) {
    EPS_CHECK(map->d->ctx == p->ctx, "map belongs to another context");
    EPS_CHECK(map->m() == e->rows, "map has %d rows, expression has %d", map->m(), e->rows);
    t.map = map->d;
    t.trans = map->trans;
  }
  const int width = map ? map->n() : e->rows;
  if (var_index >= 0) {
    EPS_CHECK(var_index < (int)p->vars.size(), "variable index %d", var_index);
    EPS_CHECK(p->vars[var_index].size == width, "variable %s has %d entries, map expects %d",
              p->vars[var_index].id.c_str(), p->vars[var_index].size, width);
    t.var = var_index;
  } else {
    EPS_CHECK(constant != nullptr, "constant term without data");
    t.constant.assign(constant, constant + width);
  }
  e->terms.push_back(std::move(t));
  EPS_API_END
}

}  // extern "C"
