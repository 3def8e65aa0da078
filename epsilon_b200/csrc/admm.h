// Device ProxADMM solver (reference: src/epsilon/algorithms/prox_admm.{h,cc}).
#pragma once
#include "kkt.h"
#include "problem.h"
#include "prox.cuh"

namespace eps {

enum StageKind { ST_ELEM = 0, ST_NORM2 = 1, ST_TV = 2, ST_KKT = 3, ST_VEC = 4 };

// One f_i after Init.
struct ProxOp {
  int type = 0;
  std::vector<int> vars;        // variable indices, ordered by variable id (expression_util.cc:5-31)
  std::vector<int64_t> var_off; // offset of each inside x_i
  int64_t n = 0;                // total variable entries
  int64_t x_off = 0;            // into the x arena
  int64_t span_lo = 0, span_len = 0;  // rows of u this operator touches (contiguous cover)
  int64_t y_off = 0;            // into each y arena
  BlockOp A;                    // span_len x n, rows relative to span_lo
  // row-sharded constraints: A = A_sh (rows this rank holds a shard of) + A_rep (replicated rows)
  BlockOp A_sh, A_rep;
  bool a_sharded = false;       // replicated variables entering sharded constraints (KKT operators only)
  bool local = false;           // a sharded variable: the operator lives on this rank's rows alone
  bool simple_scalar = false;   // A = a*I on exactly the span
  double a = 0;
  std::unique_ptr<KktSolver> kkt;
  ElemSeg seg;                  // elementwise / block / tv operators
  std::vector<ElemSeg> axis_segs;   // NORM_2 with has_axis: one segment per column / row (replaces `seg` in the stage)
  // multi-argument / sort-based vector operators (ST_VEC): the general form of the VectorProx
  // scalar reduction, v = (beta/gamma) H A'w + g, x = (1/beta) H'(x' - g)  (vector_prox.cc:59-67,131-137)
  BlockOp Hop;
  int64_t rows_h = 0;
  double vec_bg = 0, vec_cinv = 0;
  int64_t g_off = -1;           // g in the constant arena
  // eps_solver_set_constant: rows of each argument inside the stacked g, the current g (host), and the
  // per-part scaling the operator applies to g when it stores it (SECOND_ORDER_CONE divides by at / ax)
  std::vector<int64_t> arg_off;
  std::vector<double> g_host;
  double g_div0 = 1.0, g_div1 = 1.0;
  int64_t g_split = 0;
  VecOpDesc vdesc;
  VecOpWork vwork;
  DevBuf<double> vbuf, vtmp;    // stacked argument (rows_h), A'w (n)
};

struct Stage {
  int kind = ST_ELEM;
  int type = 0;
  std::vector<int> members;     // indices into ops
  bool scalar_a = false;        // every member has a simple scalar A block
  // device tables
  DevBuf<ElemSeg> segs;
  DevBuf<SegPair> pairs;
  std::vector<SegPair> pairs_host;
  DevBuf<int64_t> prefix;
  int64_t total = 0;
  int max_len = 0;
  int n_segs = 0;               // entries of `segs` (more than members when an operator acts axis-wise)
  DevBuf<double> tv_work;
  unsigned tv_tag[2] = {0, 0};  // TV solver (tv1d_par.cu): last round tag of its slot maxima, rounds of the last call
};

}  // namespace eps

struct eps_solver {
  eps::Ctx* ctx = nullptr;
  eps_params params;
  std::vector<eps::VarInfo> vars;
  std::vector<int> var_owner;                 // prox index owning each variable (-1 none)
  std::vector<int64_t> var_x_off;             // absolute offset in the x arena
  std::vector<eps::ProxOp> ops;
  std::vector<std::unique_ptr<eps::Stage>> stages;
  int64_t m_total = 0;   // A_.m()
  int64_t n_total = 0;   // A_.n()
  int N = 0;
  bool has_b = false;
  double b_norm = 0;
  eps::DevBuf<double> u, b, x, consts, ybuf[2], tmp_m, diff_m, tmp_n, rhs_n, resid;
  double* h_resid = nullptr;  // pinned
  std::vector<double> b_host;               // constant terms of the constraints (eps_solver_set_constant)
  std::vector<double> consts_host;          // the constant arena as uploaded
  std::vector<int64_t> constraint_off;      // first row of every constraint in u-space
  // row-sharded constraints (eps_problem_constraint_sharded)
  bool any_sharded = false;
  std::vector<char> c_sharded;              // per constraint
  int64_t m_global = 0;                     // A_.m() over all ranks
  eps::DevBuf<double> resid_sh;             // partial sums of ComputeResidualsSharded
  std::vector<double> tb_hb_host;           // two-block: constants of the user's constraints
  int parity = 0;             // ybuf[parity] holds the current y
  int iter = 0;               // next iteration index (iter_ of the reference)
  bool started = false;
  eps_status status;
  std::string description;
  // graphs
  cudaGraphExec_t graph_exec[2] = {nullptr, nullptr};
  int64_t graph_nodes[2] = {0, 0};
  bool graph_ok = false;
  int eager_done[2] = {0, 0};

  // ---- ProxADMMTwoBlockSolver (prox_admm_two_block.cc): consensus form ----
  bool two_block = false;
  eps::BlockOp Hc;                       // the constraint matrix H over the x/z layout (m_c x n_all)
  std::unique_ptr<eps::KktSolver> proj;  // (H H') y = r of the ZeroProx projection (prox/zero.cc)
  int64_t m_c = 0, n_all = 0, tb_n = 0;  // rows of H; all variables; H.A.n() (constraint variables)
  bool has_mask = false;
  eps::DevBuf<double> zbuf[2], ub, zeros, hb, v_tmp, r_m, y_m, cmask;
  int zpar = 0;                          // zbuf[zpar] = z, zbuf[1 - zpar] = z_prev

  ~eps_solver();
  void Init(const eps_problem& p);
  void InitTwoBlock(const eps_problem& original);
  void RunStages(const double* yo, double* yn);  // x_i = prox_i(u) for every stage (+ fused dual updates)
  void TwoBlockIteration();
  void Iteration();          // one ADMM iteration, eager launches
  void RunIteration();       // graph replay when available
  void ComputeResiduals();
  void ComputeResidualsSharded();
  void ResetVariables();
  // Solver::SetParameterValue for constant vectors (solver.cc:109-116): a new constant term of
  // constraint `index` (prox < 0) or of argument `index` of prox function `prox`
  void SetConstant(int prox, int index, const double* values);
  void InvalidateGraphs();
};
