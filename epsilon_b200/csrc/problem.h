// Prox-affine problem description held by the C ABI (eps_problem) -- the flattened form
// the reference reaches after affine::BuildAffineOperator (src/epsilon/affine/affine.cc:22-129).
#pragma once
#include "internal.h"

namespace eps {

struct Term {
  std::shared_ptr<MapData> map;  // null = identity
  bool trans = false;
  int var = -1;                  // >= 0: variable index, < 0: constant
  std::vector<double> constant;  // when var < 0
};

struct AffineExpr {
  int rows = 0;
  std::vector<Term> terms;
  int64_t global_rows = 0;  // > 0: these rows are this rank's shard of a larger argument
};

struct ProxFn {
  int type = 0;
  double alpha = 1.0;
  bool epigraph = false;
  double zone_m = 0.0;
  int sum_largest_k = 0;
  int soc_m = 0, soc_n = 0;
  int mat_m = 0, mat_n = 0;
  int axis = -1, ax_m = 0, ax_n = 0;   // has_axis / axis / arg_size(0) (vector_prox.cc:150-177)
  std::vector<double> q_alpha, q_beta;
  std::vector<AffineExpr> args;
};

struct VarInfo {
  std::string id;
  int size = 0;
};

}  // namespace eps

struct eps_problem {
  eps::Ctx* ctx = nullptr;
  std::vector<eps::VarInfo> vars;
  std::vector<eps::AffineExpr> constraints;
  std::vector<eps::ProxFn> prox;
};
