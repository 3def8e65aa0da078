// epsilon_ref -- command-line driver around the UNMODIFIED reference solver.
//
// TEST INFRASTRUCTURE, not product code.  This file contains no solver math of
// its own: it parses a prox-affine `Problem` (protobuf text format), fills the
// reference's DataMap with raw bytes and calls the reference's own classes
// compiled from /root/reference by oracle/Makefile:
//
//   solve : ProxADMMSolver(problem, data, params).Solve()  -- the stock code
//           path of src/epsilon/algorithms/prox_admm.cc:131-169, untouched.
//   trace : the same objects stepped one iteration at a time so that every
//           epoch's residuals can be written at full precision.  Stepping uses
//           the `friend class ProxADMMSolverTest` declared by the reference
//           (prox_admm.h:66) and replays the statement sequence of
//           prox_admm.cc:134-154 on the reference's own members; the harness
//           cross-checks that the stepped run ends in the same status as
//           `solve` (tests/golden/make_golden.py: solve_case).
//   prox  : single operator, the statement sequence of EvalProx
//           (python/epopt/solvemodule.cc:189-242).
//   resolve     : Solve(), then Solver::SetParameterValue (solver.cc:109-116) for every
//                 <dir>/reparam/<parameter_id> (a text-format Constant) and Solve() again on the
//                 same solver object -- the warm-start path of solvemodule.cc:142-156.
//   gpu-solve, gpu-resolve : the same two flows through ProxADMMGpuSolver
//                 (adapter/prox_admm_gpu.cc: the reference-side binding of libepsgpu.so), only in
//                 the binary built with -DEPS_GPU_ADAPTER (oracle/_ref/epsilon_ref_gpu).  The DataMap
//                 strings it uploads from are ordinary pageable std::string buffers, as in
//                 solvemodule.cc:58-72.
//
// Directory protocol (all numbers little-endian, doubles as raw f64):
//   <dir>/problem.pbtxt   Problem            (solve/trace)
//   <dir>/params.pbtxt    SolverParams       (optional)
//   <dir>/f_expr.pbtxt    Expression         (prox)  + <dir>/lambda.txt
//   <dir>/data/<location> raw constant bytes, <location> with '/' -> '_'
//   <dir>/v/<var_id>      prox input per variable (prox)
// Outputs under <dir>/out_<mode>/ : status.txt, trace.txt, x/<var_id>.
#include <sys/stat.h>
#include <dirent.h>

#include <chrono>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include <glog/logging.h>
#include <google/protobuf/text_format.h>

#include "epsilon/affine/affine.h"
#include "epsilon/algorithms/prox_admm.h"
#include "epsilon/algorithms/prox_admm_two_block.h"
#include "epsilon/expression.pb.h"
#include "epsilon/expression/expression_util.h"
#include "epsilon/prox/prox.h"
#include "epsilon/solver.pb.h"
#include "epsilon/solver_params.pb.h"
#include "epsilon/vector/vector_util.h"

#ifdef EPS_GPU_ADAPTER
#include "prox_admm_gpu.h"
#endif

namespace {

double Now() {
  return std::chrono::duration<double>(
      std::chrono::steady_clock::now().time_since_epoch()).count();
}

std::string ReadFile(const std::string& path, bool required = true) {
  std::ifstream in(path, std::ios::binary);
  if (!in) {
    if (required) LOG(FATAL) << "cannot read " << path;
    return "";
  }
  std::ostringstream ss;
  ss << in.rdbuf();
  return ss.str();
}

void WriteFile(const std::string& path, const std::string& bytes) {
  std::ofstream out(path, std::ios::binary);
  CHECK(out.good()) << "cannot write " << path;
  out.write(bytes.data(), bytes.size());
}

void MakeDir(const std::string& path) { mkdir(path.c_str(), 0755); }

std::vector<std::string> ListDir(const std::string& path) {
  std::vector<std::string> names;
  DIR* d = opendir(path.c_str());
  if (d == nullptr) return names;
  while (dirent* e = readdir(d)) {
    std::string name = e->d_name;
    if (name != "." && name != "..") names.push_back(name);
  }
  closedir(d);
  return names;
}

DataMap LoadData(const std::string& dir) {
  DataMap data;
  for (const std::string& name : ListDir(dir + "/data"))
    data[name] = ReadFile(dir + "/data/" + name);
  return data;
}

void WriteVector(const std::string& path, const Eigen::VectorXd& x) {
  WriteFile(path, std::string(reinterpret_cast<const char*>(x.data()),
                              sizeof(double) * x.size()));
}

std::string StatusLine(const SolverStatus& s) {
  char buf[512];
  snprintf(buf, sizeof(buf), "%d %d %.17g %.17g %.17g %.17g",
           static_cast<int>(s.state()), s.num_iterations(),
           s.residuals().r_norm(), s.residuals().s_norm(),
           s.residuals().epsilon_primal(), s.residuals().epsilon_dual());
  return buf;
}

void WriteSolution(const std::string& out, const BlockVector& x) {
  MakeDir(out + "/x");
  for (const std::string& key : x.keys()) WriteVector(out + "/x/" + key, x(key));
}

}  // namespace

// Named as in the reference's friend declaration (prox_admm.h:66): gives the
// harness access to the solver's own Init()/ComputeResiduals()/members.
class ProxADMMSolverTest {
 public:
  static int Trace(ProxADMMSolver* s, const SolverParams& params,
                   const std::string& out, int dump_iter) {
    double t0 = Now();
    s->Init();
    double t_init = Now() - t0;
    std::ofstream trace(out + "/trace.txt");
    double t_loop0 = Now();
    double t_resid = 0;
    // Statement sequence of prox_admm.cc:134-154, on the reference's members.
    for (s->iter_ = 0; s->iter_ < params.max_iterations(); s->iter_++) {
      s->y_prev_ = s->y_;
      s->u_ -= s->b_;
      for (int i = 0; i < s->N_; i++) s->u_ -= s->y_[i];
      for (int i = 0; i < s->N_; i++) {
        s->u_ += s->y_[i];
        if (s->iter_ == dump_iter) {
          MakeDir(out + "/apply_in_" + std::to_string(i));
          for (const std::string& key : s->u_.keys())
            WriteVector(out + "/apply_in_" + std::to_string(i) + "/" + key,
                        s->u_(key));
        }
        s->x_[i] = s->prox_[i]->Apply(s->u_);
        if (s->iter_ == dump_iter) {
          MakeDir(out + "/apply_out_" + std::to_string(i));
          for (const std::string& key : s->x_[i].keys())
            WriteVector(out + "/apply_out_" + std::to_string(i) + "/" + key,
                        s->x_[i](key));
        }
        s->y_[i] = s->A_ * s->x_[i];
        s->u_ -= s->y_[i];
      }
      if (s->iter_ % params.epoch_iterations() == 0) {
        double tr = Now();
        s->ComputeResiduals();
        t_resid += Now() - tr;
        trace << StatusLine(s->status_) << "\n";
        if (s->status_.state() == SolverStatus::OPTIMAL) break;
      }
    }
    if (s->iter_ == params.max_iterations()) {
      s->ComputeResiduals();
      s->status_.set_state(SolverStatus::MAX_ITERATIONS_REACHED);
      trace << StatusLine(s->status_) << "\n";
    }
    double t_loop = Now() - t_loop0;
    int iters_run = s->iter_ < params.max_iterations() ? s->iter_ + 1 : s->iter_;
    char buf[512];
    snprintf(buf, sizeof(buf),
             "%s\ninit_s %.6f\nloop_s %.6f\nresid_s %.6f\niters_run %d\n"
             "m %d\nn %d\nN %d\n",
             StatusLine(s->status_).c_str(), t_init, t_loop, t_resid,
             iters_run, s->m_, s->n_, s->N_);
    WriteFile(out + "/status.txt", buf);
    WriteSolution(out, s->GetSolution());
    return 0;
  }
};

namespace {

int RunSolve(const std::string& dir, bool trace, int dump_iter) {
  Problem problem;
  CHECK(google::protobuf::TextFormat::ParseFromString(
      ReadFile(dir + "/problem.pbtxt"), &problem));
  SolverParams params;
  std::string params_txt = ReadFile(dir + "/params.pbtxt", false);
  if (!params_txt.empty())
    CHECK(google::protobuf::TextFormat::ParseFromString(params_txt, &params));
  DataMap data = LoadData(dir);

  std::string out = dir + (trace ? "/out_trace" : "/out_solve");
  MakeDir(out);
  // CreateSolver of python/epopt/solvemodule.cc:74-87
  std::unique_ptr<Solver> any;
  if (params.solver() == SolverParams::PROX_ADMM_TWO_BLOCK) {
    CHECK(!trace) << "stepping is only possible for ProxADMMSolver (friend class)";
    any.reset(new ProxADMMTwoBlockSolver(problem, data, params));
  } else {
    ProxADMMSolver* s = new ProxADMMSolver(problem, data, params);
    any.reset(s);
    if (trace) return ProxADMMSolverTest::Trace(s, params, out, dump_iter);
  }
  Solver& solver = *any;

  double t0 = Now();
  BlockVector x = solver.Solve();
  double t_total = Now() - t0;
  char buf[256];
  snprintf(buf, sizeof(buf), "%s\ntotal_s %.6f\n",
           StatusLine(solver.status()).c_str(), t_total);
  WriteFile(out + "/status.txt", buf);
  WriteSolution(out, x);
  return 0;
}

void WriteTrace(const std::string& path, const std::vector<SolverStatus>& epochs) {
  std::ofstream trace(path);
  for (const SolverStatus& e : epochs) trace << StatusLine(e) << "\n";
}

// solve / resolve through the Solver interface only (solver.h:42-102), CPU or device class.
int RunSolver(const std::string& dir, const std::string& mode) {
  const bool gpu = mode.compare(0, 4, "gpu-") == 0;
  const bool resolve = mode.find("resolve") != std::string::npos;
  Problem problem;
  CHECK(google::protobuf::TextFormat::ParseFromString(
      ReadFile(dir + "/problem.pbtxt"), &problem));
  SolverParams params;
  std::string params_txt = ReadFile(dir + "/params.pbtxt", false);
  if (!params_txt.empty())
    CHECK(google::protobuf::TextFormat::ParseFromString(params_txt, &params));
  if (resolve) params.set_warm_start(true);
  double t0 = Now();
  DataMap data = LoadData(dir);            // pageable std::string buffers (solvemodule.cc:58-72)
  double t_load = Now() - t0;
  std::string out = dir + "/out_" + mode;
  MakeDir(out);

  std::unique_ptr<Solver> solver;
#ifdef EPS_GPU_ADAPTER
  ProxADMMGpuSolver* dev = nullptr;
  if (gpu) {
    dev = new ProxADMMGpuSolver(problem, data, params);
    solver.reset(dev);
  }
#else
  CHECK(!gpu) << "this binary was built without the device adapter (make -C oracle gpu)";
#endif
  if (!solver) solver.reset(new ProxADMMSolver(problem, data, params));

  t0 = Now();
  BlockVector x = solver->Solve();
  double t_first = Now() - t0;
  std::string first = StatusLine(solver->status());
#ifdef EPS_GPU_ADAPTER
  if (dev) WriteTrace(out + "/trace_first.txt", dev->epochs());
#endif
  double t_second = 0;
  if (resolve) {
    MakeDir(out + "/x_first");
    for (const std::string& key : x.keys()) WriteVector(out + "/x_first/" + key, x(key));
    for (const std::string& id : ListDir(dir + "/reparam")) {
      Constant c;
      CHECK(google::protobuf::TextFormat::ParseFromString(ReadFile(dir + "/reparam/" + id), &c));
      solver->SetParameterValue(id, c);
    }
    t0 = Now();
    x = solver->Solve();
    t_second = Now() - t0;
  }
#ifdef EPS_GPU_ADAPTER
  if (dev) {
    WriteTrace(out + "/trace.txt", dev->epochs());
    WriteFile(out + "/describe.txt", dev->describe());
  }
#endif
  char buf[1024];
  snprintf(buf, sizeof(buf), "%s\nload_s %.6f\ntotal_s %.6f\nsecond_s %.6f\n",
           StatusLine(solver->status()).c_str(), t_load, t_first, t_second);
  WriteFile(out + "/status.txt", buf);
  WriteFile(out + "/status_first.txt", first + "\n");
  WriteSolution(out, x);
  return 0;
}

// Statement sequence of EvalProx, python/epopt/solvemodule.cc:205-236.
int RunProx(const std::string& dir) {
  Expression f_expr;
  CHECK(google::protobuf::TextFormat::ParseFromString(
      ReadFile(dir + "/f_expr.pbtxt"), &f_expr));
  const double lambda = std::stod(ReadFile(dir + "/lambda.txt"));
  DataMap data = LoadData(dir);
  CHECK_EQ(Expression::PROX_FUNCTION, f_expr.expression_type());

  AffineOperator H, A;
  for (int i = 0; i < f_expr.arg_size(); i++) {
    affine::BuildAffineOperator(
        f_expr.arg(i), data, affine::arg_key(i), &H.A, &H.b);
  }
  BlockVector v_in;
  int i = 0;
  for (const Expression* var_expr : GetVariables(f_expr)) {
    const std::string& var_id = var_expr->variable().variable_id();
    A.A(affine::constraint_key(i++), var_id) =
        (1 / sqrt(lambda)) * linear_map::Identity(GetDimension(*var_expr));
    std::string bytes = ReadFile(dir + "/v/" + var_id);
    v_in(var_id) = Eigen::Map<const Eigen::VectorXd>(
        reinterpret_cast<const double*>(bytes.data()),
        bytes.size() / sizeof(double));
  }
  BlockVector v = A.A * v_in;
  std::unique_ptr<ProxOperator> op = CreateProxOperator(
      f_expr.prox_function().prox_function_type(),
      f_expr.prox_function().epigraph());
  op->Init(ProxOperatorArg(f_expr.prox_function(), data, H, A));
  double t0 = Now();
  BlockVector x = op->Apply(v);
  double t_apply = Now() - t0;
  std::string out = dir + "/out_prox";
  MakeDir(out);
  char buf[64];
  snprintf(buf, sizeof(buf), "apply_s %.6f\n", t_apply);
  WriteFile(out + "/status.txt", buf);
  WriteSolution(out, x);
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  google::InitGoogleLogging(argv[0]);
  if (argc < 3) {
    fprintf(stderr, "usage: %s solve|trace|prox|resolve|gpu-solve|gpu-resolve <dir> [dump_iter]\n", argv[0]);
    return 2;
  }
  std::string mode = argv[1];
  std::string dir = argv[2];
  int dump_iter = argc > 3 ? atoi(argv[3]) : -1;
  if (mode == "solve") return RunSolve(dir, false, -1);
  if (mode == "trace") return RunSolve(dir, true, dump_iter);
  if (mode == "prox") return RunProx(dir);
  if (mode == "resolve" || mode == "gpu-solve" || mode == "gpu-resolve") return RunSolver(dir, mode);
  fprintf(stderr, "unknown mode %s\n", mode.c_str());
  return 2;
}
