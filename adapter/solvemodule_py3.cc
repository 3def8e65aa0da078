// _solve for CPython 3: the binding python/epopt/cvxpy_solver.py calls (`_solve.solve`, `_solve.eval_prox`),
// with the device solver behind it.
//
// The reference's binding (python/epopt/solvemodule.cc) is a CPython-2 extension (PyString_*, Py_InitModule) and
// cannot be built for the interpreter of this image.  This file is NOT a port of it line by line: it is the same
// contract -- bytes in, bytes out, `_solve.error` on a failed CHECK -- written against the CPython-3 API:
//
//   _solve.solve(problem: bytes, parameters: [(parameter_id: str, constant: bytes)], params: bytes,
//                data: {location: bytes}) -> (status: bytes, {variable_id: bytes})        solvemodule.cc:110-186
//   _solve.eval_prox(f_expr: bytes, lam: float, data: {location: bytes}, v: {variable_id: bytes})
//                -> {variable_id: bytes}                                                    solvemodule.cc:189-242
//
// CreateSolver (solvemodule.cc:74-87) picks ProxADMMGpuSolver (adapter/prox_admm_gpu.cc -> libepsgpu.so) for
// SolverParams.solver == PROX_ADMM unless EPSILON_GPU=0; PROX_ADMM_TWO_BLOCK and EPSILON_GPU=0 run the stock CPU
// classes, so one module serves both arms.  Warm start keeps the solver in a cache keyed by the problem bytes
// (solvemodule.cc:142-151).  Built by `make -C oracle py3` into oracle/_ref/_solve.so; test infrastructure here, the
// file a maintainer would drop next to solvemodule.cc upstream.
#define PY_SSIZE_T_CLEAN
#include <Python.h>

#include <setjmp.h>

#include <memory>
#include <string>
#include <unordered_map>

#include <glog/logging.h>

#include "epsilon/affine/affine.h"
#include "epsilon/algorithms/prox_admm.h"
#include "epsilon/algorithms/prox_admm_two_block.h"
#include "epsilon/expression.pb.h"
#include "epsilon/expression/expression_util.h"
#include "epsilon/prox/prox.h"
#include "epsilon/solver.pb.h"
#include "epsilon/solver_params.pb.h"
#include "epsilon/util/logging.h"
#include "epsilon/vector/vector_util.h"

#include "prox_admm_gpu.h"

namespace {

PyObject* g_error = nullptr;       // _solve.error
jmp_buf g_failure;                 // glog failure function -> longjmp (solvemodule.cc:17-19,245-248)
std::unordered_map<std::string, std::unique_ptr<Solver>> g_solver_cache;

void OnFailure() { longjmp(g_failure, 1); }

void LogToPython(const std::string& msg) { PySys_WriteStdout("%s\n", msg.c_str()); }

bool UseGpu(const SolverParams& params) {
  const char* e = getenv("EPSILON_GPU");
  return params.solver() == SolverParams::PROX_ADMM && !(e != nullptr && e[0] == '0');
}

// {str or bytes: bytes} -> DataMap; the strings are copied once, as WriteConstants does (solvemodule.cc:58-72)
bool ReadDataMap(PyObject* dict, DataMap* out) {
  if (!PyDict_Check(dict)) {
    PyErr_SetString(PyExc_TypeError, "data must be a dict {location: bytes}");
    return false;
  }
  Py_ssize_t pos = 0;
  PyObject *key, *value;
  while (PyDict_Next(dict, &pos, &key, &value)) {
    const char* k = PyUnicode_Check(key) ? PyUnicode_AsUTF8(key) : PyBytes_AsString(key);
    char* v = nullptr;
    Py_ssize_t len = 0;
    if (k == nullptr || PyBytes_AsStringAndSize(value, &v, &len) != 0) return false;
    out->emplace(k, std::string(v, len));
  }
  return true;
}

std::unique_ptr<Solver> MakeSolver(const Problem& problem, const DataMap& data, const SolverParams& params) {
  if (UseGpu(params)) return std::unique_ptr<Solver>(new ProxADMMGpuSolver(problem, data, params));
  if (params.solver() == SolverParams::PROX_ADMM)
    return std::unique_ptr<Solver>(new ProxADMMSolver(problem, data, params));
  if (params.solver() == SolverParams::PROX_ADMM_TWO_BLOCK)
    return std::unique_ptr<Solver>(new ProxADMMTwoBlockSolver(problem, data, params));
  LOG(FATAL) << "Unknown solver: " << params.solver();
  return nullptr;
}

PyObject* VectorsToDict(const BlockVector& x, const VariableSet& vars) {
  PyObject* out = PyDict_New();
  for (const Expression* expr : vars) {
    const std::string& id = expr->variable().variable_id();
    const Eigen::VectorXd v = x.Get(id, GetDimension(*expr));
    PyObject* val = PyBytes_FromStringAndSize(reinterpret_cast<const char*>(v.data()), v.rows() * sizeof(double));
    PyDict_SetItemString(out, id.c_str(), val);
    Py_DECREF(val);
  }
  return out;
}

}  // namespace

extern "C" {

// solvemodule.cc:110-186
static PyObject* Solve(PyObject*, PyObject* args) {
  const char *problem_str, *params_str;
  Py_ssize_t problem_len, params_len;
  PyObject *parameters, *data;
  if (!PyArg_ParseTuple(args, "y#Oy#O", &problem_str, &problem_len, &parameters, &params_str, &params_len, &data))
    return nullptr;
  Problem problem;
  SolverParams params;
  if (!problem.ParseFromArray(problem_str, (int)problem_len) || !params.ParseFromArray(params_str, (int)params_len)) {
    PyErr_SetString(g_error, "cannot parse Problem / SolverParams");
    return nullptr;
  }
  // the DataMap must outlive a cached solver (ProxADMMSolver keeps a reference to it): it lives with the cache entry
  static std::unordered_map<std::string, std::unique_ptr<DataMap>> data_cache;
  std::unique_ptr<DataMap> data_own(new DataMap());
  if (!ReadDataMap(data, data_own.get())) return nullptr;

  if (setjmp(g_failure)) {
    PyErr_SetString(g_error, "CHECK failed");   // destructors of the abandoned frame are skipped, as upstream
    return nullptr;
  }
  Solver* solver = nullptr;
  std::unique_ptr<Solver> solver_gc;
  DataMap* dm = data_own.get();
  if (params.warm_start()) {                                   // solvemodule.cc:142-151
    const std::string key(problem_str, problem_len);
    auto it = g_solver_cache.find(key);
    if (it == g_solver_cache.end()) {
      data_cache[key] = std::move(data_own);
      dm = data_cache[key].get();
      it = g_solver_cache.emplace(key, MakeSolver(problem, *dm, params)).first;
    } else {
      // new constants of this call (parameter values point at locations of the new dict)
      for (auto& kv : *data_own) (*data_cache[key])[kv.first] = kv.second;
    }
    solver = it->second.get();
  } else {
    solver_gc = MakeSolver(problem, *dm, params);
    solver = solver_gc.get();
  }
  // SetParameterValues (solvemodule.cc:89-106)
  PyObject* iter = PyObject_GetIter(parameters);
  if (iter == nullptr) return nullptr;
  while (PyObject* item = PyIter_Next(iter)) {
    PyObject* id = PyTuple_GetItem(item, 0);
    PyObject* val = PyTuple_GetItem(item, 1);
    char* buf = nullptr;
    Py_ssize_t len = 0;
    Constant constant;
    if (id == nullptr || val == nullptr || PyBytes_AsStringAndSize(val, &buf, &len) != 0 ||
        !constant.ParseFromArray(buf, (int)len)) {
      Py_DECREF(item);
      Py_DECREF(iter);
      if (!PyErr_Occurred()) PyErr_SetString(g_error, "cannot parse a parameter Constant");
      return nullptr;
    }
    solver->SetParameterValue(PyUnicode_Check(id) ? PyUnicode_AsUTF8(id) : PyBytes_AsString(id), constant);
    Py_DECREF(item);
  }
  Py_DECREF(iter);

  BlockVector x = solver->Solve();
  const std::string status = solver->status().SerializeAsString();
  PyObject* vars = VectorsToDict(x, GetVariables(problem));
  PyObject* ret = Py_BuildValue("y#O", status.data(), (Py_ssize_t)status.size(), vars);
  Py_DECREF(vars);
  return ret;
}

// solvemodule.cc:189-242
static PyObject* EvalProx(PyObject*, PyObject* args) {
  const char* expr_str;
  Py_ssize_t expr_len;
  double lambda;
  PyObject *data, *v_map;
  if (!PyArg_ParseTuple(args, "y#dOO", &expr_str, &expr_len, &lambda, &data, &v_map)) return nullptr;
  Expression f_expr;
  if (!f_expr.ParseFromArray(expr_str, (int)expr_len)) {
    PyErr_SetString(g_error, "cannot parse Expression");
    return nullptr;
  }
  DataMap data_map, v_bytes;
  if (!ReadDataMap(data, &data_map) || !ReadDataMap(v_map, &v_bytes)) return nullptr;
  if (setjmp(g_failure)) {
    PyErr_SetString(g_error, "CHECK failed");
    return nullptr;
  }
  CHECK_EQ(Expression::PROX_FUNCTION, f_expr.expression_type());
  BlockVector v;
  for (const Expression* var : GetVariables(f_expr)) {
    const std::string& id = var->variable().variable_id();
    auto it = v_bytes.find(id);
    CHECK(it != v_bytes.end()) << "no value for variable " << id;
    CHECK_EQ((size_t)GetDimension(*var) * sizeof(double), it->second.size());
    v(id) = Eigen::Map<const Eigen::VectorXd>(reinterpret_cast<const double*>(it->second.data()), GetDimension(*var));
  }
  // one prox function, no constraint: ProxADMMGpuSolver::EvalProx -> eps_prox_eval
  Problem problem;
  problem.mutable_objective()->set_expression_type(Expression::ADD);
  *problem.mutable_objective()->add_arg() = f_expr;
  SolverParams params;
  ProxADMMGpuSolver solver(problem, data_map, params);
  BlockVector x = solver.EvalProx(lambda, v);
  return VectorsToDict(x, GetVariables(f_expr));
}

static PyMethodDef kMethods[] = {
    {"solve", Solve, METH_VARARGS, "Solve a prox-affine problem (ProxADMM on the device)."},
    {"eval_prox", EvalProx, METH_VARARGS, "Evaluate one proximal operator on the device."},
    {nullptr, nullptr, 0, nullptr}};

static struct PyModuleDef kModule = {PyModuleDef_HEAD_INIT, "_solve", "Epsilon solver binding (CPython 3)", -1, kMethods};

PyMODINIT_FUNC PyInit__solve(void) {
  google::InitGoogleLogging("epopt");
  google::InstallFailureFunction(&OnFailure);
  SetVerboseLogger(&LogToPython);
  PyObject* m = PyModule_Create(&kModule);
  if (m == nullptr) return nullptr;
  g_error = PyErr_NewException("_solve.error", nullptr, nullptr);
  Py_INCREF(g_error);
  PyModule_AddObject(m, "error", g_error);
  return m;
}

}  // extern "C"
