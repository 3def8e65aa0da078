// ProxADMMGpuSolver -- the reference-side binding of libepsgpu.so (include/epsgpu.h).
//
// This is the file a maintainer of mfouda/epsilon adds as src/epsilon/algorithms/prox_admm_gpu.{h,cc}:
// a `Solver` subclass (src/epsilon/algorithms/solver.h:42-102) selected in CreateSolver
// (python/epopt/solvemodule.cc:74-87) in place of ProxADMMSolver.  It uses the reference's OWN front half
// of Init -- affine::BuildAffineOperator (src/epsilon/affine/affine.cc:121-129), GetVariables,
// GetDimension -- to obtain A_, b_ and every H_i, g_i as BlockMatrix / BlockVector, hands each block to
// the C ABI, and runs the whole iteration loop on the device.  epopt.solve(), the compiler, the protos
// and the returned SolverStatus + BlockVector are untouched.
//
// In THIS repository the file is compiled by oracle/Makefile against /root/reference (never copied
// there) into oracle/_ref/epsilon_ref_gpu, so that one harness runs the same problem.pbtxt through the
// stock ProxADMMSolver and through this class (tests/test_gpu_adapter.py).  It contains no solver math.
#ifndef EPSILON_ALGORITHMS_PROX_ADMM_GPU_H
#define EPSILON_ALGORITHMS_PROX_ADMM_GPU_H

#include <map>
#include <string>
#include <vector>

#include "epsilon/affine/affine.h"
#include "epsilon/algorithms/solver.h"
#include "epsilon/expression.pb.h"
#include "epsilon/solver_params.pb.h"
#include "epsilon/vector/block_matrix.h"
#include "epsilon/vector/block_vector.h"
#include "epsilon/vector/vector_util.h"

#include "epsgpu.h"

class ProxADMMGpuSolver final : public Solver {
 public:
  ProxADMMGpuSolver(const Problem& problem, const DataMap& data_map, const SolverParams& params);
  ~ProxADMMGpuSolver() override;
  BlockVector Solve() override;
  // EvalProx (python/epopt/solvemodule.cc:189-242): the problem holds exactly one prox function and no
  // constraint; returns argmin lambda*f(x) + 1/2 ||x - v||^2 over its variables.
  BlockVector EvalProx(double lambda, const BlockVector& v);

  // every ComputeResiduals result of the last Solve() (the reference keeps only the last one)
  const std::vector<SolverStatus>& epochs() const { return epochs_; }
  const eps_status& device_status() const { return device_status_; }
  const char* describe() const;

 private:
  // One affine expression of the problem after BuildAffineOperator: a constraint (prox < 0) or argument
  // `index` of prox function `prox`.
  struct Affine {
    int prox, index, rows;
    std::string row_key;
    BlockMatrix A;
    BlockVector b;
  };
  std::vector<Affine> Walk();
  void Build(const std::vector<Affine>& exprs);
  void BuildProblem(const std::vector<Affine>& exprs);
  bool SameMaps(const std::vector<Affine>& exprs) const;
  void RefreshConstants(const std::vector<Affine>& exprs);
  eps_map* Upload(const linear_map::LinearMap& L);
  void AddTerms(const Affine& e);
  void Release();
  static int OnEpoch(const eps_status* st, void* self);

  const DataMap& data_map_;
  SolverParams params_;
  eps_problem* prob_ = nullptr;
  eps_solver* solver_ = nullptr;
  std::map<std::string, int> var_index_;
  std::vector<Affine> built_;       // what the device solver was built from (for SetParameterValue)
  bool exprs_only_ = false;
  std::vector<SolverStatus> epochs_;
  eps_status device_status_;
};

#endif  // EPSILON_ALGORITHMS_PROX_ADMM_GPU_H
