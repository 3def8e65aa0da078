/* epsgpu.h -- C ABI of libepsgpu.so: the B200 (sm_100a) implementation of Epsilon's
 * prox-affine ADMM hot path.
 *
 * Every entry point stands where one interface of the reference (mfouda/epsilon)
 * stands; the reference file:line each one replaces is given next to it.  The
 * signatures use plain pointers and sizes only.  All `const double*` / `double*`
 * buffers are HOST memory unless the name ends in `_dev`.
 *
 * Error model (reference: glog CHECK/LOG(FATAL) -> longjmp -> Python exception,
 * python/epopt/solvemodule.cc:158,184-186,245-263): no entry point throws, exits
 * or aborts.  Every function returns EPS_OK (0) or a negative code and leaves a
 * message retrievable with eps_last_error().  The C++ adapter turns a non-zero
 * return into LOG(FATAL) << eps_last_error() to keep the reference's behaviour.
 *
 * Threading (solvemodule.cc holds the GIL for the whole Solve): calls on one
 * context must be serialised by the caller.
 */
#ifndef EPSGPU_H_
#define EPSGPU_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EPS_OK 0
#define EPS_ERR_INVALID -1     /* bad argument / unsupported structure (a reference CHECK) */
#define EPS_ERR_CUDA -2        /* CUDA runtime failure */
#define EPS_ERR_NO_DEVICE -3   /* no sm_100 device: there is NO CPU fallback */
#define EPS_ERR_NCCL -4
#define EPS_ERR_NUMERIC -5     /* factorisation failed (not positive definite) */

typedef struct eps_ctx eps_ctx;
typedef struct eps_map eps_map;
typedef struct eps_problem eps_problem;
typedef struct eps_solver eps_solver;

/* ProxFunction.Type, proto/epsilon/expression.proto:123-170 (same numeric values). */
enum eps_prox_type {
  EPS_PROX_AFFINE = 1,
  EPS_PROX_CONSTANT = 2,
  EPS_PROX_ZERO = 10,
  EPS_PROX_SUM_SQUARE = 11,
  EPS_PROX_NON_NEGATIVE = 20,
  EPS_PROX_NORM_1 = 21,
  EPS_PROX_SUM_DEADZONE = 22,
  EPS_PROX_SUM_EXP = 23,
  EPS_PROX_SUM_HINGE = 24,
  EPS_PROX_SUM_INV_POS = 25,
  EPS_PROX_SUM_KL_DIV = 26,
  EPS_PROX_SUM_LOGISTIC = 27,
  EPS_PROX_SUM_NEG_ENTR = 28,
  EPS_PROX_SUM_NEG_LOG = 29,
  EPS_PROX_SUM_QUANTILE = 31,
  EPS_PROX_EXP = 32,
  EPS_PROX_LOG_SUM_EXP = 100,
  EPS_PROX_MAX = 101,
  EPS_PROX_NORM_2 = 102,
  EPS_PROX_SECOND_ORDER_CONE = 104,
  EPS_PROX_SUM_LARGEST = 105,
  EPS_PROX_TOTAL_VARIATION_1D = 106,
  EPS_PROX_LAMBDA_MAX = 200,
  EPS_PROX_NEG_LOG_DET = 202,
  EPS_PROX_NORM_NUCLEAR = 203,
  EPS_PROX_SEMIDEFINITE = 204
};

/* SolverStatus.State, proto/epsilon/solver.proto:4-16. */
enum eps_state {
  EPS_NOT_STARTED = 0,
  EPS_INITIALIZING = 1,
  EPS_RUNNING = 2,
  EPS_OPTIMAL = 3,
  EPS_MAX_ITERATIONS_REACHED = 4,
  EPS_ERROR = 5
};

/* SolverParams.Solver, proto/epsilon/solver_params.proto:63-67 (same numeric values).
 * PROX_ADMM           = ProxADMMSolver          (src/epsilon/algorithms/prox_admm.cc)
 * PROX_ADMM_TWO_BLOCK = ProxADMMTwoBlockSolver  (src/epsilon/algorithms/prox_admm_two_block.cc) */
enum eps_solver_kind { EPS_SOLVER_PROX_ADMM = 0, EPS_SOLVER_PROX_ADMM_TWO_BLOCK = 1 };

/* How the linear-equality (KKT) prox of SumSquare/Affine/Constant is solved.  The
 * reference always eliminates blocks with BlockCholesky in the order NextKey picks
 * (src/epsilon/vector/block_cholesky.cc:51-64,119-133); AUTO reproduces that choice
 * (Woodbury = it eliminated the arg block last, Gram = the variable block last) and
 * adds the matrix-free mode for shapes whose dense factor does not fit. */
enum eps_kkt_mode {
  EPS_KKT_AUTO = 0,
  EPS_KKT_DIAGONAL = 1,  /* every block scalar/diagonal: elementwise */
  EPS_KKT_WOODBURY = 2,  /* factor (I/(2a) + H D^-1 H')  (rows(H) x rows(H)) */
  EPS_KKT_GRAM = 3,      /* factor (2a H'H + A'A)         (n x n), explicit inverse */
  EPS_KKT_CG = 4         /* conjugate gradients on (2a H'H + A'A), warm-started */
};

/* ---- context ----------------------------------------------------------------- */

/* One context per process and GPU.  `stream` is a cudaStream_t to launch on, or
 * NULL to let the library create its own non-blocking stream. */
int eps_ctx_create(int device, void* stream, eps_ctx** out);
void eps_ctx_destroy(eps_ctx* ctx);
const char* eps_last_error(void);
/* Number of kernels of THIS library launched on the context so far (graph replays
 * count every kernel node they contain). */
int64_t eps_ctx_kernel_launches(const eps_ctx* ctx);
int64_t eps_ctx_bytes_allocated(const eps_ctx* ctx);
int eps_ctx_synchronize(eps_ctx* ctx);

/* Row-sharded multi-GPU (one process per GPU).  `unique_id` is the 128-byte
 * ncclUniqueId every rank received from rank 0 (the host exchanges it with
 * torch.distributed / MPI / a file; the library only consumes the bytes).
 * The reference has no counterpart (single process, single thread). */
int eps_nccl_unique_id(void* out_128_bytes);
int eps_ctx_init_comm(eps_ctx* ctx, int n_ranks, int rank, const void* unique_id);

/* ---- linear maps: LinearMapImpl, src/epsilon/linear/linear_map.h:33-57 --------- */

/* DenseMatrixImpl (dense_matrix_impl.cc:11-19): m x n, column-major f64, as decoded
 * by BuildMatrix (src/epsilon/vector/vector_util.cc:247-259). */
int eps_map_dense(eps_ctx* ctx, int m, int n, const double* colmajor, eps_map** out);
/* The same for a matrix that is ALREADY resident in the memory of the context's device (column-major f64):
 * the map borrows the buffer -- no copy; the caller keeps it alive until the map and every problem / solver
 * built on it are destroyed.  (An 80 GB data matrix produced on the device must not exist twice.) */
int eps_map_dense_dev(eps_ctx* ctx, int m, int n, double* colmajor_dev, eps_map** out);
/* SparseMatrixImpl from the wire layout of BuildSparseMatrix (vector_util.cc:261-281):
 * CSC, int32 colptr[n+1], int32 rowidx[nnz], f64 vals[nnz]. */
int eps_map_csc(eps_ctx* ctx, int m, int n, int64_t nnz, const int32_t* colptr,
                const int32_t* rowidx, const double* vals, eps_map** out);
/* DiagonalMatrixImpl (diagonal_matrix_impl.h:14-23). */
int eps_map_diagonal(eps_ctx* ctx, int n, const double* d, eps_map** out);
/* ScalarMatrixImpl (scalar_matrix_impl.h:14-31): alpha * I_n. */
int eps_map_scalar(eps_ctx* ctx, int n, double alpha, eps_map** out);
/* KroneckerProductImpl (kronecker_product_impl.cc:45-58): A (x) B; holds references. */
int eps_map_kronecker(eps_ctx* ctx, eps_map* a, eps_map* b, eps_map** out);
/* LinearMapImpl::Transpose(): a new map sharing storage (dense_matrix_impl.h:41-43). */
int eps_map_transpose(eps_map* map, eps_map** out);
/* LinearMapImpl::Inverse() (linear_map.h:49): scalar 1/alpha (scalar_matrix_impl.h:29-31), diagonal with
 * 0 kept 0 (diagonal_matrix_impl.cc:15-23), Kronecker inv(A) (x) inv(B) (kronecker_product_impl.h:37-39),
 * dense symmetric matrix as an explicit dense inverse (dense_matrix_impl.cc:21-30), sparse by way of
 * dense (sparse_matrix_impl.cc:60-78).  The dense factorisation is a Cholesky of M or -M on the device:
 * EPS_ERR_NUMERIC for an indefinite matrix. */
int eps_map_inverse(eps_map* map, eps_map** out);
int eps_map_dims(const eps_map* map, int* m, int* n);
/* LinearMapImpl::Apply(x) (linear_map.h:52): y = map * x; x has n, y has m entries. */
int eps_map_apply(eps_map* map, const double* x, double* y);
/* Same on device buffers already resident (no copies), on the context stream. */
int eps_map_apply_dev(eps_map* map, const double* x_dev, double* y_dev);
/* Time `reps` back-to-back device applies with CUDA events on the context's stream;
 * returns the mean milliseconds per launch and the algorithmic bytes one launch moves. */
int eps_map_time(eps_map* map, int reps, float* ms_per_apply, double* algorithmic_bytes);
void eps_map_destroy(eps_map* map);
/* y = A x for a dense column-major m x n matrix already on the device, WITHOUT any allocation: no
 * cudaMalloc / cudaFree / pool call, so it can be issued under CUDA graph capture or from a loop that
 * must not allocate.  No counterpart in the reference.
 * The kernels and the summation order are those of eps_map_apply_dev on an eps_map_dense (bit-identical
 * result); their column-split partial sums live in one static device array of
 * EPS_NOALLOC_SCRATCH_DOUBLES doubles (8 MiB, held by every GPU that loads the library until exit)
 * instead of the context's grown scratch.  EPS_ERR_INVALID when splits * m exceeds it, where
 * splits <= min(ceil(n / 64), 24 * SMs / ceil(m / 512)); m * ceil(n / 64) <= EPS_NOALLOC_SCRATCH_DOUBLES
 * always fits.  All calls share the array (it keeps the last call's partials; every entry read is
 * written first, nothing needs clearing): calls that overlap on one GPU -- two streams, two contexts --
 * race; calls on one stream are ordered.  Runs on the context's stream; a, x, y are device pointers. */
#define EPS_NOALLOC_SCRATCH_DOUBLES (1 << 20)
int eps_dense_apply_noalloc(eps_ctx* ctx, int m, int n, const double* a_colmajor_dev, const double* x_dev,
                            double* y_dev);

/* ---- prox-affine problem: what ProxADMMSolver::Init assembles ------------------
 *   minimize sum_i f_i(H_i x + g_i)   s.t.  sum_j A_j x_j + b = 0
 * AffineOperator {BlockMatrix A; BlockVector b} built by affine::BuildAffineOperator
 * (src/epsilon/affine/affine.cc:22-129): every term is `map * variable` (a block of
 * A) or `map * constant` (added into b).  Row keys "constraint:i" / "arg:i"
 * (affine.cc:131-140) are the integer indices used here. */
int eps_problem_create(eps_ctx* ctx, eps_problem** out);
void eps_problem_destroy(eps_problem* p);
int eps_problem_add_variable(eps_problem* p, const char* var_id, int size, int* var_index);
/* constraint i: INDICATOR(ZERO cone) over an affine expression of `rows` rows
 * (prox_admm.cc:25-43). */
int eps_problem_add_constraint(eps_problem* p, int rows, int* constraint_index);
/* objective term f_i (prox_admm.cc:45-94): type, ProxFunction.alpha, epigraph flag. */
int eps_problem_add_prox(eps_problem* p, int prox_type, double alpha, int epigraph,
                         int* prox_index);
/* scaled_zone_params (scaled_zone.cc:34-76): M for SUM_DEADZONE; per-element
 * alpha/beta (length n_q, or 1 = promoted) for SUM_QUANTILE. */
int eps_problem_prox_zone(eps_problem* p, int prox_index, double zone_m,
                          const double* q_alpha, const double* q_beta, int n_q);
/* sum_largest_params.k for SUM_LARGEST (sum_largest.cc:10-13). */
int eps_problem_prox_sum_largest(eps_problem* p, int prox_index, int k);
/* SECOND_ORDER_CONE: arguments (t in R^m, X in R^{m x n} column-major); m, n = arg_size(1)
 * (second_order_cone.cc:31-35). */
int eps_problem_prox_soc_dims(eps_problem* p, int prox_index, int m, int n);
/* Matrix functions (LAMBDA_MAX, NEG_LOG_DET, NORM_NUCLEAR, SEMIDEFINITE): the first argument is
 * mat(x, m, n), column-major; m, n = ProxFunction.arg_size(0).dim (src/epsilon/prox/ortho_invariant.cc:12-16). */
int eps_problem_prox_mat_dims(eps_problem* p, int prox_index, int m, int n);
/* ProxFunction.has_axis / axis (expression.proto:193-195; VectorProx::Apply, vector_prox.cc:150-177): the operator
 * acts on every column (axis 0) or every row (axis 1) of mat(arg, m, n), m, n = arg_size(0).  On the device:
 * NORM_2 (the row / column group norms of a matrix variable); other types return EPS_ERR_INVALID. */
int eps_problem_prox_axis(eps_problem* p, int prox_index, int axis, int m, int n);
int eps_problem_prox_add_arg(eps_problem* p, int prox_index, int rows, int* arg_index);
/* Row-sharded argument (multi-GPU): the rows given to this rank are its slice of an
 * argument with `global_rows` rows; products H'H p and H'g are all-reduced. */
int eps_problem_prox_arg_sharded(eps_problem* p, int prox_index, int arg_index,
                                 int64_t global_rows);
/* Row-sharded CONSTRAINT (multi-GPU, SURVEY section 8e: the "consensus blocks" of the north star): the rows
 * given to this rank are its slice of a constraint with `global_rows` rows -- for logistic regression or an
 * SVM the block -diag(y) A x - t = 0 whose row count is the number of samples.  Every rank holds its rows of
 * the data block, of the constant term, of u and y_i and of the variable that enters the block through a
 * scalar / diagonal map (t); variables that enter through a general map (x) are replicated.  The
 * linear-equality prox of the replicated variables sums A_r'(.) over the ranks (fused NVLink exchange per
 * CG step, or one all-reduce of the Gram matrix), ComputeResiduals all-reduces its partial sums in one
 * message (prox_admm.cc:141-147,178-217 on row blocks). */
int eps_problem_constraint_sharded(eps_problem* p, int constraint_index, int64_t global_rows);
/* Add `map * x[var_index]` (var_index >= 0) or `map * constant` (var_index < 0,
 * `constant` has map.n entries, or `rows` entries when map is NULL) to constraint
 * `constraint_index` (prox_index < 0) or to argument `arg_index` of prox `prox_index`.
 * map == NULL is the identity (a bare VARIABLE / CONSTANT node). */
int eps_problem_add_term(eps_problem* p, int prox_index, int arg_or_constraint_index,
                         eps_map* map, int var_index, const double* constant);

/* ---- solver: ProxADMMSolver, src/epsilon/algorithms/prox_admm.{h,cc} ------------ */

/* SolverParams, proto/epsilon/solver_params.proto:4-71 (live fields on this path)
 * plus device-side knobs.  Initialise with eps_params_default(). */
typedef struct eps_params {
  int32_t max_iterations;   /* 10000 */
  double rho;               /* 1 (anything else is rejected, prox_admm.cc:50) */
  double abs_tol;           /* 1e-4 */
  double rel_tol;           /* 1e-2 */
  int32_t epoch_iterations; /* 10 */
  int32_t warm_start;       /* 0 */
  int32_t kkt_mode;         /* eps_kkt_mode, AUTO */
  double cg_tol;            /* relative residual of the inner CG, 1e-12 */
  int32_t cg_max_iterations;/* 500 */
  int32_t use_graph;        /* capture the per-iteration kernel sequence in a CUDA graph */
  int32_t profile;          /* bracket kernel classes with CUDA events (disables graphs) */
  int64_t gram_max_n;       /* largest n factored densely in AUTO mode (60000) */
  int32_t solver;           /* eps_solver_kind: SolverParams.solver (solver_params.proto:63-67) */
} eps_params;
void eps_params_default(eps_params* p);

/* SolverStatus + Residuals, proto/epsilon/solver.proto:4-42, plus timings. */
typedef struct eps_status {
  int32_t state;
  int32_t num_iterations;   /* iter_ at the last ComputeResiduals (prox_admm.cc:216) */
  double r_norm, s_norm, epsilon_primal, epsilon_dual;
  int32_t iterations_run;   /* ADMM iterations executed by the last call */
  int32_t m, n, N;          /* A_.m(), A_.n(), number of prox operators */
  double init_ms;           /* host wall time of eps_solver_create */
  double loop_ms;           /* CUDA-event time of the iteration loop of the last call */
  int64_t cg_iterations;    /* inner CG iterations of the last call (all KKT solves) */
  int64_t cg_unconverged;   /* inner CG solves of the last call that stopped at cg_max_iterations
                               above cg_tol: the KKT prox was inexact there (0 in a healthy run) */
} eps_status;

/* Per-epoch callback: the reference offers status callbacks only at the end
 * (solver.cc:88-95); parity tests need every ComputeResiduals result.  Return
 * non-zero to stop (stop_callback_, solver.h:60-63). */
typedef int (*eps_trace_fn)(const eps_status* status, void* user);

/* ProxADMMSolver constructor + Init() (prox_admm.cc:14-129): uploads, analyses each
 * f_i (VectorProx scalar/diagonal reduction, vector_prox.cc:51-116; KKT structure,
 * sum_square.cc:12-29 / affine.cc:19-38) and factorises. */
int eps_solver_create(eps_problem* p, const eps_params* params, eps_solver** out);
void eps_solver_destroy(eps_solver* s);
/* ProxADMMSolver::Solve() loop (prox_admm.cc:131-169) incl. ComputeResiduals
 * (:178-217) every epoch_iterations; fills `status`. */
int eps_solver_solve(eps_solver* s, eps_status* status, eps_trace_fn trace, void* user);
/* Solver::SetParameterValue (src/epsilon/algorithms/solver.cc:38-61,109-116; called for every
 * parameter on every solve, python/epopt/solvemodule.cc:89-106,156) for constant VECTORS: replaces the
 * constant term of constraint `arg_or_constraint_index` (prox_index < 0; that many rows) or of argument
 * `arg_or_constraint_index` of prox function `prox_index` (the sum of the CONSTANT nodes of that affine
 * expression, as affine::BuildAffineOperator leaves it in b(row_key), affine.cc:51-69).  With
 * params.warm_start the next eps_solver_solve continues from the current iterate, which is the
 * warm-start path of solvemodule.cc:142-151 / prox_admm.cc:115-120 without the reference's
 * refactorisation.  A parameter inside a MATRIX needs a new solver (eps_solver_create). */
int eps_solver_set_constant(eps_solver* s, int prox_index, int arg_or_constraint_index,
                            const double* values);
/* Carries the iterate (u, x_i, y_i, y_prev; z, u of the two-block form) of `src` over to `dst`, a solver
 * created from a problem with the same variables and constraint rows (e.g. after a parameter inside a
 * matrix changed): with params.warm_start the next eps_solver_solve(dst) continues from it, as the
 * reference does when Init() re-runs on a live solver (prox_admm.cc:110-120 keeps x_, y_, u_). */
int eps_solver_copy_state(eps_solver* dst, eps_solver* src);
/* Run exactly `iterations` more ADMM iterations without stop test (benchmarking);
 * `elapsed_ms` = CUDA-event time on the context stream. */
int eps_solver_iterate(eps_solver* s, int iterations, float* elapsed_ms);
/* ComputeResiduals() on the current iterate. */
int eps_solver_residuals(eps_solver* s, eps_status* status);
/* GetSolution() (prox_admm.cc:171-176): value of one variable. */
int eps_solver_get_variable(eps_solver* s, const char* var_id, double* out);
/* Per-kernel-class CUDA-event totals gathered while params.profile != 0:
 * names[i] -> (launches, total ms, algorithmic bytes).  Returns the class count. */
int eps_solver_profile(eps_solver* s, int max_classes, const char** names,
                       int64_t* launches, double* total_ms, double* bytes);
/* Free-text description of what Init chose per prox operator (mode, sizes). */
const char* eps_solver_describe(eps_solver* s);

/* EvalProx (python/epopt/solvemodule.cc:189-242): the problem holds exactly one
 * prox function and no constraint; A = (1/sqrt(lambda)) I per variable.
 * v_concat / x_concat: the variables' values concatenated in variable-index order. */
int eps_prox_eval(eps_problem* p, double lambda, const double* v_concat, double* x_concat);

#ifdef __cplusplus
}
#endif
#endif /* EPSGPU_H_ */
