// Linear-equality (KKT) prox: SumSquareProx / AffineProx of the reference.
#pragma once
#include <memory>

#include "internal.h"

namespace eps {

// Solves  (2a H'H + A'A) x = A'v + c0,   c0 = -2a H'g - c
// which is what chol_.Solve(b_ + v).Select(var_keys_) returns for
//   SumSquareProx (src/epsilon/prox/sum_square.cc:10-40; alpha' = sqrt(2*alpha) scales H and g)
//   AffineProx    (src/epsilon/prox/affine.cc:19-49;  a = 0, c = alpha * linear term)
// The reference eliminates the KKT blocks with BlockCholesky
// (src/epsilon/vector/block_cholesky.cc:119-137); which Schur complement that produces
// (rows(H)^2 or n^2) is the WOODBURY / GRAM distinction here.
struct KktSolver {
  Ctx* ctx = nullptr;
  int mode = EPS_KKT_AUTO;
  int64_t n = 0, rows_h = 0, rows_a = 0;
  double alpha2 = 0;  // 2*alpha (0: no H term)
  bool h_sharded = false;
  BlockOp H, A;
  // Row-sharded constraints (multi-GPU): A = A_sh (blocks whose rows this rank holds a shard of: their
  // products A_sh'(A_sh p) and A_sh'v are summed over the ranks) + A_rep (rows every rank holds in full).
  bool a_sharded = false;
  BlockOp A_sh, A_rep;
  // The one sparse (SELL) block B of the fused CG step: M = fb_coef * s^2 B'B + diag(ddiag).  Either the
  // argument of a SUM_SQUARE (H, fb_coef = 2a: the lasso) or the data block of the constraints (A,
  // fb_coef = 1: logistic regression, SVM); fb_sharded: its rows are this rank's shard.
  Blk fb;
  double fb_coef = 0;
  bool fb_sharded = false;
  DevBuf<double> tb1;    // B p
  DevBuf<double> c0;
  // factors
  DevBuf<double> dinv;   // DIAGONAL: 1/diag(M); WOODBURY: 1/diag(A'A)
  DevBuf<double> ddiag;  // CG: diag(A'A) when A'A is diagonal
  bool a_diag = false;
  DevBuf<double> W;      // WOODBURY: rows_h^2, GRAM: n^2; lower-triangular inv(chol)
  // work vectors
  DevBuf<double> tn1, tn2, th1, th2, ta1;
  DevBuf<double> cg_r, cg_p, cg_w;
  DevBuf<double> scal;   // device scalars of the CG recurrences
  double* h_scal = nullptr;  // pinned mirror
  // CG without a host round trip per iteration: the tail kernel of iteration k publishes
  // (rr, converged, seq) into a host-mapped ring and raises `cg_done` on the device, which
  // turns the speculatively launched iteration k+1 into no-ops.
  struct CgResult { double rr; double converged; long long seq; long long pad; };
  CgResult* ring = nullptr;      // cudaHostAlloc(mapped), kCgRing entries
  long long cg_seq = 1;
  DevBuf<int> cg_done;
  DevBuf<double> cg_partials;
  int tail_grid = 0;
  bool fast_h = false;           // H = one SELL block, A'A diagonal: fused matvec
  bool fused_tail = false;       // plane reduction + exchange + CG update in one kernel (CgFusedTailKernel)
  // fused_tail, second generation (EPS_TAIL_V1=1 restores the first): the first-pass plane reduction rides in
  // the second SpMV pass (pre-phase + GridBarrier), the tail has one grid-wide synchronisation and an ordinary
  // (programmatic) launch, and the initial residual of a solve is carried over from the previous one
  bool tail_v2 = false;
  bool carry_enabled = true;     // false where one object serves several right-hand sides (Kronecker columns)
  bool carry_valid = false;      // cg_r / rhs_prev belong to the x the next Solve() is warm-started from
  int64_t carry_count = 0;
  DevBuf<double> rhs_prev;
  DevBuf<unsigned> cg_bar;       // GridBarrier state: [0,1] tail, [2,3] pre-phase of the second pass
  PeerBuf* peer = nullptr;       // row-sharded: the context's one-shot NVLink all-reduce buffers (not owned)
  unsigned long long pending_seq = 0;  // exchange the next consumer kernel has to complete (0: none)
  // H = (s I_k) (x) B with A'A repeating over the k blocks (matrix variables, A*X): the system
  // is block diagonal with k identical blocks; `inner` solves one block, Solve loops the columns
  std::unique_ptr<KktSolver> inner;
  int kron_k = 0;
  double cg_tol = 1e-12;
  int cg_max_iter = 500;
  int64_t cg_iters_total = 0;
  int64_t cg_unconverged = 0;    // inner solves that hit cg_max_iter above cg_tol (eps_status.cg_unconverged)
  int64_t gram_max_n = 60000;
  std::string description;

  ~KktSolver();
  // g: rows_h constants of the argument (may be empty = 0); c: n linear term (may be null)
  void Init(const std::vector<double>& g_host, const std::vector<double>& c_host);
  // c0 = -2a H'g - c for a new constant g of the argument (eps_solver_set_constant); the factors stay
  void SetConstant(const std::vector<double>& g_host);
  void ComputeC0(const std::vector<double>& g_host);
  std::vector<double> c_host_;   // linear term kept for SetConstant
  // x <- solution for rhs (both device, n).  x holds the warm start on entry (CG).
  void Solve(const double* rhs, double* x);
  void MatVec(const double* p, double* w, const int* skip = nullptr);  // w = (2a H'H + A'A) p
  // k systems at once (columns of rhs / x, leading dimension n): the factor and a dense H are
  // read once for all columns (DMMA GEMMs); returns false when this mode has no such path
  bool SolveMulti(const double* rhs, double* x, int k);
  DevBuf<double> mn1, mn2, mh1, mh2;  // work for SolveMulti: n*k, n*k, rows_h*k, rows_h*k
};

}  // namespace eps
