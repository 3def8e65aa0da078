// See prox_admm_gpu.h.  Reference statements this follows, in order:
//   ProxADMMSolver::InitConstraints    src/epsilon/algorithms/prox_admm.cc:25-43
//   ProxADMMSolver::InitProxOperators  src/epsilon/algorithms/prox_admm.cc:45-94
//   ScaledZone GetParams (quantile)    src/epsilon/prox/scaled_zone.cc:62-71
//   ProxADMMSolver::Solve / LogStatus  src/epsilon/algorithms/prox_admm.cc:131-169,219-230
//   Solver::SetParameterValue          src/epsilon/algorithms/solver.cc:109-116 (re-read on every Solve)
#include "prox_admm_gpu.h"

#include <cstdlib>

#include <glog/logging.h>

#include "epsilon/expression/expression_util.h"
#include "epsilon/linear/dense_matrix_impl.h"
#include "epsilon/linear/diagonal_matrix_impl.h"
#include "epsilon/linear/kronecker_product_impl.h"
#include "epsilon/linear/scalar_matrix_impl.h"
#include "epsilon/linear/sparse_matrix_impl.h"
#include "epsilon/util/logging.h"
#include "epsilon/util/string.h"

// glog CHECK / LOG(FATAL) -> InstallFailureFunction(HandleFailure) -> longjmp -> _solve.error, exactly
// as for every other failure of the reference (python/epopt/solvemodule.cc:245-263).  The library never
// throws or exits; all device memory hangs off the context.
#define EPS_OK_OR_FATAL(call)                                  \
  do {                                                         \
    if ((call) != EPS_OK) LOG(FATAL) << #call << ": " << eps_last_error(); \
  } while (0)

namespace {

// One context per process and GPU (include/epsgpu.h); solvers come and go with every epopt.solve().
eps_ctx* Context() {
  static eps_ctx* ctx = nullptr;
  if (ctx == nullptr) {
    const char* dev = getenv("EPSILON_GPU_DEVICE");
    EPS_OK_OR_FATAL(eps_ctx_create(dev ? atoi(dev) : 0, nullptr, &ctx));
  }
  return ctx;
}

bool IsMatrixFunction(ProxFunction::Type t) { return t >= ProxFunction::LAMBDA_MAX; }

// LinearMapImpl::operator== is unimplemented for sparse matrices (sparse_matrix_impl.cc:80-90 returns
// false), so the compressed arrays are compared here; everything else defers to the reference.
bool SameMap(const linear_map::LinearMapImpl& a, const linear_map::LinearMapImpl& b) {
  if (a.type() != b.type() || a.m() != b.m() || a.n() != b.n()) return false;
  if (a.type() == linear_map::SPARSE_MATRIX) {
    const auto& A = static_cast<const linear_map::SparseMatrixImpl&>(a).sparse();
    const auto& B = static_cast<const linear_map::SparseMatrixImpl&>(b).sparse();
    if (A.nonZeros() != B.nonZeros()) return false;
    return memcmp(A.outerIndexPtr(), B.outerIndexPtr(), sizeof(int) * (A.cols() + 1)) == 0 &&
           memcmp(A.innerIndexPtr(), B.innerIndexPtr(), sizeof(int) * A.nonZeros()) == 0 &&
           memcmp(A.valuePtr(), B.valuePtr(), sizeof(double) * A.nonZeros()) == 0;
  }
  if (a.type() == linear_map::KRONECKER_PRODUCT) {
    const auto& A = static_cast<const linear_map::KroneckerProductImpl&>(a);
    const auto& B = static_cast<const linear_map::KroneckerProductImpl&>(b);
    return SameMap(A.A().impl(), B.A().impl()) && SameMap(A.B().impl(), B.B().impl());
  }
  return a == b;
}

}  // namespace

ProxADMMGpuSolver::ProxADMMGpuSolver(
    const Problem& problem, const DataMap& data_map, const SolverParams& params)
    : Solver(problem), data_map_(data_map), params_(params) {
  memset(&device_status_, 0, sizeof(device_status_));
}

ProxADMMGpuSolver::~ProxADMMGpuSolver() { Release(); }

void ProxADMMGpuSolver::Release() {
  if (solver_) eps_solver_destroy(solver_);
  if (prob_) eps_problem_destroy(prob_);
  solver_ = nullptr;
  prob_ = nullptr;
  var_index_.clear();
}

const char* ProxADMMGpuSolver::describe() const {
  return solver_ ? eps_solver_describe(solver_) : "";
}

// LinearMapImpl -> eps_map, by implementation type (src/epsilon/linear/linear_map.h:17-26).
eps_map* ProxADMMGpuSolver::Upload(const linear_map::LinearMap& L) {
  eps_ctx* ctx = Context();
  eps_map* m = nullptr;
  const linear_map::LinearMapImpl& impl = L.impl();
  switch (impl.type()) {
    case linear_map::DENSE_MATRIX: {
      // column-major storage shared between a matrix and its transpose (dense_matrix_impl.h:41-43)
      const auto& D = static_cast<const linear_map::DenseMatrixImpl&>(impl);
      const bool t = *D.trans() == 'T';
      EPS_OK_OR_FATAL(eps_map_dense(ctx, t ? D.n() : D.m(), t ? D.m() : D.n(), D.data(), &m));
      if (t) {
        eps_map* mt = nullptr;
        EPS_OK_OR_FATAL(eps_map_transpose(m, &mt));
        eps_map_destroy(m);
        m = mt;
      }
      break;
    }
    case linear_map::SPARSE_MATRIX: {
      // Eigen's compressed column storage IS the wire layout (vector_util.cc:261-281)
      const auto& S = static_cast<const linear_map::SparseMatrixImpl&>(impl).sparse();
      EPS_OK_OR_FATAL(eps_map_csc(ctx, S.rows(), S.cols(), S.nonZeros(), S.outerIndexPtr(),
                                  S.innerIndexPtr(), S.valuePtr(), &m));
      break;
    }
    case linear_map::DIAGONAL_MATRIX: {
      const auto& d = static_cast<const linear_map::DiagonalMatrixImpl&>(impl).diagonal().diagonal();
      EPS_OK_OR_FATAL(eps_map_diagonal(ctx, d.rows(), d.data(), &m));
      break;
    }
    case linear_map::SCALAR_MATRIX: {
      const auto& S = static_cast<const linear_map::ScalarMatrixImpl&>(impl);
      EPS_OK_OR_FATAL(eps_map_scalar(ctx, S.n(), S.alpha(), &m));
      break;
    }
    case linear_map::KRONECKER_PRODUCT: {
      const auto& K = static_cast<const linear_map::KroneckerProductImpl&>(impl);
      eps_map* a = Upload(K.A());
      eps_map* b = Upload(K.B());
      EPS_OK_OR_FATAL(eps_map_kronecker(ctx, a, b, &m));
      eps_map_destroy(a);   // the product holds its own references
      eps_map_destroy(b);
      break;
    }
    default:
      LOG(FATAL) << "linear map type " << impl.type() << " has no device counterpart";
  }
  return m;
}

// Every block A(row_key, var) becomes `map * variable`, b(row_key) the constant term.
void ProxADMMGpuSolver::AddTerms(const Affine& e) {
  for (const std::string& var_id : e.A.col_keys()) {
    if (!e.A.has_key(e.row_key, var_id)) continue;
    auto it = var_index_.find(var_id);
    CHECK(it != var_index_.end()) << "unknown variable " << var_id;
    eps_map* m = Upload(e.A(e.row_key, var_id));
    EPS_OK_OR_FATAL(eps_problem_add_term(prob_, e.prox, e.index, m, it->second, nullptr));
    eps_map_destroy(m);     // the problem keeps the device data alive
  }
  if (e.b.has_key(e.row_key)) {
    const BlockVector::DenseVector& c = e.b(e.row_key);
    CHECK_EQ(c.rows(), e.rows);
    EPS_OK_OR_FATAL(eps_problem_add_term(prob_, e.prox, e.index, nullptr, -1, c.data()));
  }
}

std::vector<ProxADMMGpuSolver::Affine> ProxADMMGpuSolver::Walk() {
  std::vector<Affine> out;
  for (int i = 0; i < problem().constraint_size(); i++) {          // prox_admm.cc:28-39
    const Expression& constr = problem().constraint(i);
    CHECK_EQ(Expression::INDICATOR, constr.expression_type());
    CHECK_EQ(Cone::ZERO, constr.cone().cone_type());
    CHECK_EQ(1, constr.arg_size());
    out.emplace_back();
    Affine& e = out.back();
    e.prox = -1; e.index = i; e.rows = GetDimension(constr.arg(0));
    e.row_key = affine::constraint_key(i);
    affine::BuildAffineOperator(constr.arg(0), data_map_, e.row_key, &e.A, &e.b);
  }
  CHECK_EQ(Expression::ADD, problem().objective().expression_type());   // prox_admm.cc:46
  for (int i = 0; i < problem().objective().arg_size(); i++) {
    const Expression& f_expr = problem().objective().arg(i);
    for (int j = 0; j < f_expr.arg_size(); j++) {                   // prox_admm.cc:60-67
      out.emplace_back();
      Affine& e = out.back();
      e.prox = i; e.index = j; e.rows = GetDimension(f_expr.arg(j));
      e.row_key = affine::arg_key(j);
      affine::BuildAffineOperator(f_expr.arg(j), data_map_, e.row_key, &e.A, &e.b);
    }
  }
  return out;
}

void ProxADMMGpuSolver::Build(const std::vector<Affine>& exprs) {
  BuildProblem(exprs);
}

void ProxADMMGpuSolver::BuildProblem(const std::vector<Affine>& exprs) {
  // a live solver whose matrices changed (SetParameterValue on a LinearMap constant): the reference keeps
  // x_, y_, u_ across the re-run of Init (prox_admm.cc:110-120); the old device solver hands its iterate on
  eps_solver* old_solver = solver_;
  eps_problem* old_prob = prob_;
  solver_ = nullptr;
  prob_ = nullptr;
  var_index_.clear();
  eps_ctx* ctx = Context();
  EPS_OK_OR_FATAL(eps_problem_create(ctx, &prob_));
  for (const Expression* var : GetVariables(problem())) {
    int idx = -1;
    const std::string& id = var->variable().variable_id();
    EPS_OK_OR_FATAL(eps_problem_add_variable(prob_, id.c_str(), GetDimension(*var), &idx));
    var_index_[id] = idx;
  }
  for (int i = 0; i < problem().constraint_size(); i++) {
    int ci = -1;
    EPS_OK_OR_FATAL(eps_problem_add_constraint(prob_, GetDimension(problem().constraint(i).arg(0)), &ci));
    CHECK_EQ(ci, i);
  }
  CHECK_EQ(1, params_.rho());                                       // prox_admm.cc:50
  for (int i = 0; i < problem().objective().arg_size(); i++) {
    const Expression& f_expr = problem().objective().arg(i);
    const ProxFunction& f = f_expr.prox_function();
    int fi = -1;
    EPS_OK_OR_FATAL(eps_problem_add_prox(prob_, f.prox_function_type(), f.alpha(), f.epigraph(), &fi));
    CHECK_EQ(fi, i);
    if (f.has_axis())                                               // vector_prox.cc:150-177
      EPS_OK_OR_FATAL(eps_problem_prox_axis(prob_, fi, f.axis(), f.arg_size(0).dim(0), f.arg_size(0).dim(1)));
    switch (f.prox_function_type()) {
      case ProxFunction::SUM_DEADZONE:
        EPS_OK_OR_FATAL(eps_problem_prox_zone(prob_, fi, f.scaled_zone_params().m(), nullptr, nullptr, 0));
        break;
      case ProxFunction::SUM_QUANTILE: {                            // scaled_zone.cc:62-71
        BlockVector tmp;
        affine::BuildAffineOperator(f.scaled_zone_params().alpha_expr(), data_map_, "alpha", nullptr, &tmp);
        affine::BuildAffineOperator(f.scaled_zone_params().beta_expr(), data_map_, "beta", nullptr, &tmp);
        CHECK_EQ(tmp("alpha").rows(), tmp("beta").rows());
        EPS_OK_OR_FATAL(eps_problem_prox_zone(prob_, fi, 0.0, tmp("alpha").data(), tmp("beta").data(),
                                              tmp("alpha").rows()));
        break;
      }
      case ProxFunction::SUM_LARGEST:
        EPS_OK_OR_FATAL(eps_problem_prox_sum_largest(prob_, fi, f.sum_largest_params().k()));
        break;
      case ProxFunction::SECOND_ORDER_CONE:                         // second_order_cone.cc:31-35
        EPS_OK_OR_FATAL(eps_problem_prox_soc_dims(prob_, fi, f.arg_size(1).dim(0), f.arg_size(1).dim(1)));
        break;
      default:
        if (IsMatrixFunction(f.prox_function_type()))               // ortho_invariant.cc:12-16
          EPS_OK_OR_FATAL(eps_problem_prox_mat_dims(prob_, fi, f.arg_size(0).dim(0), f.arg_size(0).dim(1)));
    }
    for (int j = 0; j < f_expr.arg_size(); j++) {
      int aj = -1;
      EPS_OK_OR_FATAL(eps_problem_prox_add_arg(prob_, fi, GetDimension(f_expr.arg(j)), &aj));
      CHECK_EQ(aj, j);
    }
  }
  for (const Affine& e : exprs) AddTerms(e);
  if (exprs_only_) {   // EvalProx: the problem is all that is needed
    if (old_solver) eps_solver_destroy(old_solver);
    if (old_prob) eps_problem_destroy(old_prob);
    return;
  }

  eps_params p;
  eps_params_default(&p);                                            // SolverParams -> eps_params
  p.max_iterations = params_.max_iterations();
  p.rho = params_.rho();
  p.abs_tol = params_.abs_tol();
  p.rel_tol = params_.rel_tol();
  p.epoch_iterations = params_.epoch_iterations();
  p.warm_start = params_.warm_start();
  p.solver = EPS_SOLVER_PROX_ADMM;
  if (const char* mode = getenv("EPSILON_GPU_KKT_MODE")) p.kkt_mode = atoi(mode);
  EPS_OK_OR_FATAL(eps_solver_create(prob_, &p, &solver_));
  if (old_solver && params_.warm_start()) EPS_OK_OR_FATAL(eps_solver_copy_state(solver_, old_solver));
  if (old_solver) eps_solver_destroy(old_solver);
  if (old_prob) eps_problem_destroy(old_prob);
  built_ = exprs;
}

// SetParameterValue rewrites Constant protos inside problem() (solver.cc:109-116); the reference then
// rebuilds everything on the next Solve().  Here: if every linear map is unchanged only the constant
// vectors are refreshed on the device (no upload of matrices, no refactorisation).
bool ProxADMMGpuSolver::SameMaps(const std::vector<Affine>& exprs) const {
  if (exprs.size() != built_.size()) return false;
  for (size_t k = 0; k < exprs.size(); k++) {
    const Affine& a = exprs[k];
    const Affine& b = built_[k];
    if (a.prox != b.prox || a.index != b.index || a.rows != b.rows) return false;
    if (a.A.col_keys() != b.A.col_keys()) return false;
    for (const std::string& var_id : a.A.col_keys()) {
      if (a.A.has_key(a.row_key, var_id) != b.A.has_key(b.row_key, var_id)) return false;
      if (a.A.has_key(a.row_key, var_id) &&
          !SameMap(a.A(a.row_key, var_id).impl(), b.A(b.row_key, var_id).impl()))
        return false;
    }
  }
  return true;
}

void ProxADMMGpuSolver::RefreshConstants(const std::vector<Affine>& exprs) {
  for (size_t k = 0; k < exprs.size(); k++) {
    const Affine& e = exprs[k];
    const bool now = e.b.has_key(e.row_key), before = built_[k].b.has_key(built_[k].row_key);
    if (!now && !before) continue;
    BlockVector::DenseVector c = e.b.Get(e.row_key, e.rows);
    if (now && before && c == built_[k].b(built_[k].row_key)) continue;
    EPS_OK_OR_FATAL(eps_solver_set_constant(solver_, e.prox, e.index, c.data()));
  }
  built_ = exprs;
}

int ProxADMMGpuSolver::OnEpoch(const eps_status* st, void* self_) {
  auto* self = static_cast<ProxADMMGpuSolver*>(self_);
  SolverStatus s;
  s.set_state(static_cast<SolverStatus::State>(st->state));          // same values, solver.proto:4-16
  s.set_num_iterations(st->num_iterations);
  SolverStatus::Residuals* r = s.mutable_residuals();
  r->set_r_norm(st->r_norm);
  r->set_s_norm(st->s_norm);
  r->set_epsilon_primal(st->epsilon_primal);
  r->set_epsilon_dual(st->epsilon_dual);
  self->epochs_.push_back(s);
  if (st->num_iterations % self->params_.log_iterations() == 0 && self->params_.verbose()) {
    LogVerbose(StringPrintf("iter=%d residuals primal=%.2e [%.2e] dual=%.2e [%.2e]", st->num_iterations,
                            st->r_norm, st->epsilon_primal, st->s_norm, st->epsilon_dual));
  }
  return self->HasExternalStop() ? 1 : 0;                             // stop_callback_, solver.h:60-63
}

BlockVector ProxADMMGpuSolver::EvalProx(double lambda, const BlockVector& v) {
  CHECK_EQ(0, problem().constraint_size());
  CHECK_EQ(1, problem().objective().arg_size());
  exprs_only_ = true;
  BuildProblem(Walk());
  exprs_only_ = false;
  // variables concatenated in the order they were registered (GetVariables: ordered by variable id)
  std::vector<double> vin, xout;
  for (const Expression* var : GetVariables(problem())) {
    const BlockVector::DenseVector vv = v.Get(var->variable().variable_id(), GetDimension(*var));
    vin.insert(vin.end(), vv.data(), vv.data() + vv.rows());
  }
  xout.resize(vin.size());
  EPS_OK_OR_FATAL(eps_prox_eval(prob_, lambda, vin.data(), xout.data()));
  BlockVector x;
  size_t off = 0;
  for (const Expression* var : GetVariables(problem())) {
    const int n = GetDimension(*var);
    x(var->variable().variable_id()) = Eigen::Map<const Eigen::VectorXd>(xout.data() + off, n);
    off += n;
  }
  return x;
}

BlockVector ProxADMMGpuSolver::Solve() {
  std::vector<Affine> exprs = Walk();
  if (solver_ != nullptr && params_.warm_start() && SameMaps(exprs)) RefreshConstants(exprs);
  else Build(exprs);

  epochs_.clear();
  EPS_OK_OR_FATAL(eps_solver_solve(solver_, &device_status_, &ProxADMMGpuSolver::OnEpoch, this));

  SolverStatus status;
  status.set_state(static_cast<SolverStatus::State>(device_status_.state));
  status.set_num_iterations(device_status_.num_iterations);
  SolverStatus::Residuals* r = status.mutable_residuals();
  r->set_r_norm(device_status_.r_norm);
  r->set_s_norm(device_status_.s_norm);
  r->set_epsilon_primal(device_status_.epsilon_primal);
  r->set_epsilon_dual(device_status_.epsilon_dual);
  if (params_.verbose()) {
    LogVerbose(StringPrintf("iter=%d residuals primal=%.2e [%.2e] dual=%.2e [%.2e]", status.num_iterations(),
                            r->r_norm(), r->epsilon_primal(), r->s_norm(), r->epsilon_dual()));
  }
  UpdateStatus(status);

  BlockVector x;                                                     // GetSolution, prox_admm.cc:171-176
  for (const Expression* var : GetVariables(problem())) {
    const std::string& id = var->variable().variable_id();
    BlockVector::DenseVector xv(GetDimension(*var));
    EPS_OK_OR_FATAL(eps_solver_get_variable(solver_, id.c_str(), xv.data()));
    x(id) = xv;
  }
  return x;
}
