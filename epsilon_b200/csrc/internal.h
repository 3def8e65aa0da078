// Internal declarations of libepsgpu (host side, C++17).  Not installed.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/epsgpu.h"

namespace eps {

// ---- errors -------------------------------------------------------------------
struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
std::string Format(const char* fmt, ...);
void SetLastError(const std::string& msg);

#define EPS_CHECK(cond, ...)                                                     \
  do {                                                                           \
    if (!(cond))                                                                 \
      throw ::eps::Error(EPS_ERR_INVALID, std::string("check failed: " #cond " ") + \
                                              ::eps::Format(__VA_ARGS__));       \
  } while (0)

#define CUDA_CHECK(expr)                                                         \
  do {                                                                           \
    cudaError_t e_ = (expr);                                                     \
    if (e_ != cudaSuccess)                                                       \
      throw ::eps::Error(EPS_ERR_CUDA, ::eps::Format("%s:%d %s: %s", __FILE__,   \
                                                     __LINE__, #expr,            \
                                                     cudaGetErrorString(e_)));   \
  } while (0)

struct Ctx;
// Host (pageable) -> device copy ordered on the context stream; large buffers go through
// pinned staging buffers filled by several host threads (ctx.cu).
void UploadBytes(Ctx* ctx, void* dst, const void* src, size_t bytes);

// ---- device memory ----------------------------------------------------------------
// Device memory of a context: stream-ordered allocations from the device's default pool (kept
// across solves: release threshold = unlimited), cudaMalloc when EPS_NO_POOL is set.
bool CtxIsLive(const Ctx* ctx);  // false once eps_ctx_destroy has run (objects may outlive their context)
void* DevAlloc(Ctx* ctx, size_t bytes);
void DevFree(Ctx* ctx, void* p, size_t bytes);

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  Ctx* ctx = nullptr;
  bool borrowed = false;   // memory owned by the caller (eps_map_dense_dev): never freed here
  DevBuf() {}
  DevBuf(Ctx* c, size_t count) { alloc(c, count); }
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  DevBuf(DevBuf&& o) noexcept { *this = std::move(o); }
  DevBuf& operator=(DevBuf&& o) noexcept {
    if (this != &o) { release(); p = o.p; n = o.n; ctx = o.ctx; borrowed = o.borrowed; o.p = nullptr; o.n = 0; o.borrowed = false; }
    return *this;
  }
  ~DevBuf() { release(); }
  void alloc(Ctx* c, size_t count) {
    release();
    ctx = c; n = count;
    if (count == 0) return;
    p = static_cast<T*>(DevAlloc(c, count * sizeof(T)));
  }
  void release() {
    if (p && !borrowed) DevFree(ctx, p, n * sizeof(T));
    p = nullptr; n = 0; borrowed = false;
  }
  void borrow(Ctx* c, T* ptr, size_t count) {
    release();
    ctx = c; p = ptr; n = count; borrowed = true;
  }
  void zero() { if (n) CUDA_CHECK(cudaMemsetAsync(p, 0, n * sizeof(T), ctx->stream)); }
  void upload(const T* host, size_t count) {
    if (count) UploadBytes(ctx, p, host, count * sizeof(T));
  }
  void download(T* host, size_t count) const {
    if (count) CUDA_CHECK(cudaMemcpyAsync(host, p, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }
};

// ---- context --------------------------------------------------------------------
struct Comm;  // NCCL (dlopen'ed), comm.cu
struct PeerBuf;  // NVLink peer buffers of the one-shot all-reduce, comm.cu

// Kernel classes for the profile mode (CUDA events around each launch).
enum KClass {
  KC_SPMV_N = 0, KC_SPMV_T, KC_GEMV_N, KC_GEMV_T, KC_PROX, KC_VEC, KC_REDUCE,
  KC_KKT_DENSE, KC_TV, KC_ALLREDUCE, KC_OTHER, KC_COUNT
};
extern const char* const kKClassNames[KC_COUNT];

struct Ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  bool pooled = false;   // DevAlloc uses cudaMallocAsync on `stream`
  int sm_count = 148;
  int64_t launches = 0;        // kernels launched outside graphs + graph-node replays
  int64_t bytes_allocated = 0;
  bool capturing = false;      // inside stream capture: launches counted per capture
  int64_t capture_launches = 0;
  Comm* comm = nullptr;
  PeerBuf* peer = nullptr;     // owned; created collectively by PeerEnsure
  int n_ranks = 1, rank = 0;
  // profile mode
  bool profile = false;
  struct Prof { int64_t launches = 0; double ms = 0, bytes = 0; };
  Prof prof[KC_COUNT];
  cudaEvent_t prof_ev[2] = {nullptr, nullptr};
  // scratch owned by the context (grown on demand, never during graph capture)
  DevBuf<double> split_scratch, seg_scratch, red_partials, sell_scratch, sell_scratch2;
  DevBuf<unsigned int> red_counter;
  int64_t scratch_generation = 0;
  int* h_flag = nullptr;
  struct Staging;              // pinned upload staging (ctx.cu)
  Staging* staging = nullptr;
  ~Ctx();
  void count_launch(int n = 1) {
    if (capturing) capture_launches += n; else launches += n;
  }
};

// Brackets one kernel launch for the profile mode; no-op otherwise.
struct ProfScope {
  Ctx* ctx; int kc; double bytes;
  ProfScope(Ctx* c, int k, double b);
  ~ProfScope();
};

// ---- atomic linear maps (LinearMapImpl types, linear/linear_map.h:17-26) -----------
enum MapKind { MK_DENSE = 0, MK_SPARSE = 1, MK_DIAGONAL = 2, MK_SCALAR = 3, MK_KRONECKER = 4 };

// CSR on the device.  rowptr int64 when nnz does not fit int32.
struct Csr {
  int m = 0, n = 0;
  int64_t nnz = 0;
  DevBuf<int32_t> rowptr;  // m+1 (nnz < 2^31 per map / per shard is enforced)
  DevBuf<int32_t> col;
  DevBuf<double> val;
  int max_row_nnz = 0;
};

// Column-tiled SELL-32 storage (sell.cu).
// One warp's contiguous stream of slices inside a visit (stream layout, sell.cu).
struct SellStream {
  int64_t voff;      // first value of the stream (doubles)
  int64_t ioff;      // first 16-bit index of the stream
  int32_t ss_begin;  // first slice (stream order) and how many
  int32_t ss_cnt;
  int32_t q_total;   // 32-wide value rows in the stream
  int32_t pad;
};

struct Sell {
  bool built = false;
  int R = 0, Cn = 0, T = 0, TW = 0, n_slices = 0, n_cta = 0;
  int64_t nnz = 0, n_val = 0, n_idx = 0;
  // a CTA walks visits = (column tile, row range); inside a visit every one of the nw warps owns
  // one contiguous stream of slices
  int n_visits = 0, nw = 24;
  DevBuf<int32_t> sl_len, rowid;   // per slice, stream order: length; per lane: index into the planes
  DevBuf<int64_t> voff;            // only while building
  DevBuf<double> val;
  DevBuf<uint16_t> idx;
  DevBuf<int32_t> visit_tile, cta_visit;
  DevBuf<SellStream> streams;
};

struct MapData {
  Ctx* ctx = nullptr;
  int kind = MK_SCALAR;
  int m = 0, n = 0;       // of the stored (untransposed) matrix
  // dense: column-major m x n
  DevBuf<double> dense;
  // sparse: CSR of A and CSR of A' (= the CSC wire layout)
  Csr csr, csr_t;
  Sell sell, sell_t;  // built for large matrices; used by Apply instead of the CSR kernel
  // small host copy of the CSC arrays for block merging (only when nnz is small)
  std::vector<int32_t> h_colptr, h_rowidx;
  std::vector<double> h_vals;
  bool has_host = false;
  // diagonal
  DevBuf<double> diag;
  std::vector<double> h_diag;
  // scalar
  double alpha = 1.0;
  // kronecker
  std::shared_ptr<MapData> ka, kb;
  bool ka_trans = false, kb_trans = false;
};

}  // namespace eps

// Handle types of the C ABI.
struct eps_ctx { eps::Ctx c; };
struct eps_map {
  std::shared_ptr<eps::MapData> d;
  bool trans = false;
  int m() const { return trans ? d->n : d->m; }
  int n() const { return trans ? d->m : d->n; }
};

namespace eps {

// ---- kernels: maps.cu ---------------------------------------------------------------
// y = beta*y + alpha * op(A) x  for one atomic map.
void MapApply(Ctx* ctx, const MapData& A, bool trans, double alpha, const double* x,
              double beta, double* y);
double MapAlgorithmicBytes(const MapData& A, bool trans);
void KroneckerApply(Ctx* ctx, const MapData& K, bool trans, const double* x, double* y);
void CsrApply(Ctx* ctx, const Csr& A, double alpha, const double* x, double beta, double* y,
              int kclass);
void BuildCsrFromHost(Ctx* ctx, int m, int n, int64_t nnz, const int64_t* rowptr,
                      const int32_t* col, const double* val, Csr* out);
void BuildSell(Ctx* ctx, const Csr& A, Sell* S);
// One-shot all-reduce over NVLink peer memory (comm.cu): every rank PUSHES its n-vector into
// slot [parity][rank] of every peer's receive buffer and then raises flag [parity][rank] on
// that peer; the consumer waits for its G local flags and adds the G slots in rank order, so
// all ranks obtain bit-identical sums.  Passed to kernels by value.
constexpr int kMaxRanks = 8;
constexpr int kPeerMaxCtas = 256;  // CTAs of the fused exchange kernel (one per SM)
// "nothing has arrived yet": a signalling-NaN bit pattern no arithmetic produces (operations return the
// canonical quiet NaN), so a partial sum that is itself NaN still reads as "arrived"
constexpr unsigned long long kPeerSentinel = 0xFFF4DEADBEEFCAFEull;
struct PeerXchg {
  double* recv[kMaxRanks] = {};              // recv[r]: rank r's buffer [2][G][n], mapped here
  unsigned long long* flag[kMaxRanks] = {};  // flag[r]: rank r's flags [2][G]
  unsigned long long* cflag[kMaxRanks] = {}; // cflag[r]: rank r's per-CTA flags [2][G][kPeerMaxCtas]
  // flag-in-data packets (16 B per double: {lo32, seq32, hi32, seq32}; every 8-byte half carries its
  // own sequence number, so a packet needs no fence and no separate flag)
  uint4* ll1[kMaxRanks] = {};                // ll1[r]: rank r's reduce-scatter inbox [2][G][n]
  uint4* ll2[kMaxRanks] = {};                // ll2[r]: rank r's all-gather inbox [2][n]
  // sentinel inboxes: plain 8-byte doubles, a slot holds kPeerSentinel until its value lands and is put back
  // to the sentinel by the reader (half the NVLink bytes of the 16-byte packets)
  double* sn1[kMaxRanks] = {};               // sn1[r]: rank r's reduce-scatter inbox [2][G][n]
  double* sn2[kMaxRanks] = {};               // sn2[r]: rank r's all-gather inbox [2][n]
  int G = 0, rank = 0;
  int64_t n = 0;   // slot stride (capacity); vectors of up to n entries fit
};
struct PeerBuf {
  bool ready = false;
  PeerXchg x;
  void* block = nullptr;        // local allocation (flags + receive buffer)
  void* mapped[kMaxRanks] = {}; // peers' blocks opened through CUDA IPC
  unsigned* counter = nullptr;  // CTA arrival counter of the publishing kernel
  unsigned long long seq = 0;   // last sequence number handed out
};
// Collective over the communicator: the context's peer buffers with room for n-vectors, or
// nullptr when peer mapping is unavailable (the callers then fall back to ncclAllReduce).
PeerBuf* PeerEnsure(Ctx* ctx, int64_t n);
void PeerRelease(PeerBuf* pb);

// Optional fusions of a sparse apply (used by the CG loop of kkt.cu).
struct SpmvExtra {
  const PeerXchg* publish = nullptr;  // push the result to all ranks instead of writing y
  unsigned long long publish_seq = 0;
  unsigned* publish_counter = nullptr;
  const int* skip = nullptr;      // device flag: when non-zero the kernels return at once
  const double* ep_d = nullptr;   // y += ep_d .* ep_p after the alpha/beta update
  const double* ep_p = nullptr;
};
void SellApply(Ctx* ctx, const Sell& S, double alpha, const double* x, double beta, double* y, int kclass,
               const SpmvExtra* ex = nullptr);
// The stream kernel alone: returns the T x R partial planes (context scratch, valid until the
// next sparse apply); y = alpha * sum_t planes[t] is left to the caller's consumer kernel.
const double* SellApplyPlanes(Ctx* ctx, const Sell& S, const double* x, int kclass, const int* skip);
// Second pass of a matvec pair: x = alpha * sum_t prevP[t] is formed INSIDE the kernel (every CTA a slice,
// then a grid barrier on `bar`) before the x tiles are staged; returns this pass's planes (second buffer).
const double* SellApplyPlanesAfter(Ctx* ctx, const Sell& S, double* x, const double* prevP, int prevT, double alpha,
                                   unsigned* bar, int kclass, const int* skip);
void DenseGemv(Ctx* ctx, const double* A, int lda, int m, int n, bool trans, double alpha,
               const double* x, double beta, double* y, int tri = 0);

// ---- vector kernels: vecops.cu --------------------------------------------------------
void VecAxpby(Ctx* ctx, int64_t n, double a, const double* x, double b, const double* y, double* out);  // out = a x + b y
void VecScale(Ctx* ctx, int64_t n, double a, const double* x, double* out);                              // out = a x
void VecMulAdd(Ctx* ctx, int64_t n, const double* d, const double* x, double b, const double* y, double* out);  // out = d.*x + b y
void VecFill(Ctx* ctx, int64_t n, double a, double* out);
void VecCopy(Ctx* ctx, int64_t n, const double* x, double* out);
// out[0] (+)= sum x_i y_i  (deterministic two-stage reduction; out on device)
void VecDot(Ctx* ctx, int64_t n, const double* x, const double* y, double* out_dev, bool accumulate);
// scratch for two-stage reductions
struct ReduceScratch {
  double* partials;
  unsigned int* counter;
};
ReduceScratch GetReduceScratch(Ctx* ctx);
double* GrowScratch(Ctx* ctx, DevBuf<double>* buf, size_t count);
// EPS_STEP_TRACE=<k>: from the k-th sparse stream launch on, the next 9 kernels of the fused CG step (stream
// passes and tails) record per-CTA globaltimer stamps into one buffer WITHOUT any synchronisation in
// between; the table (us since the first stamp, min / median / max over the CTAs) goes to stderr afterwards.
// kind 0 = stream pass (4 stamps per CTA: start, x staged, first warp done, last warp done),
// kind 1 = fused tail (8 stamps per CTA).  Returns nullptr when not tracing this launch.
unsigned long long* StepTraceSlot(Ctx* ctx, int kind, int ctas, const char* label);
int* PinnedFlag(Ctx* ctx);  // one pinned host int for device->host round trips

// ---- composite block operator (BlockMatrix, vector/block_matrix.h) --------------------
struct Blk {
  int row_off = 0, col_off = 0;
  std::shared_ptr<MapData> map;
  bool trans = false;
  double scale = 1.0;
  int m() const { return trans ? map->n : map->m; }
  int n() const { return trans ? map->m : map->n; }
  bool diag_like() const { return map->kind == MK_SCALAR || map->kind == MK_DIAGONAL; }
};

struct BlockOp {
  Ctx* ctx = nullptr;
  int64_t rows = 0, cols = 0;
  std::vector<Blk> blks;
  // compiled form
  bool compiled = false;
  bool is_zero = false;
  // "simple scalar": exactly one scalar block covering [row_off, row_off+cols)
  bool simple_scalar = false;
  double simple_alpha = 0;
  int64_t simple_row_off = 0;
  // merged sparse-like blocks (CSR over rows x cols) and its transpose
  bool has_merged = false;
  Csr merged, merged_t;
  std::vector<Blk> heavy;  // blocks applied one by one (dense, large sparse, kronecker)
  void Compile();
  // y[0..rows) = op * x[0..cols)   (accumulate: y += )
  void Apply(const double* x, double* y, bool accumulate = false) const;
  // x[0..cols) = op' * y[0..rows)
  void ApplyT(const double* y, double* x, bool accumulate = false) const;
  // diag(op' op) when op' op is diagonal; false otherwise
  bool GramDiagonal(std::vector<double>* d) const;
  int64_t heavy_nnz() const;
};

// ---- dense factorisation kernels: dense.cu ---------------------------------------------
// C(n x n, ldc) (+)= alpha * A' A, A m x n column-major (lda).  fp64 tensor cores (DMMA).
void DenseGramTN(Ctx* ctx, const double* A, int64_t lda, int64_t m, int n, double alpha,
                 double* C, int64_t ldc, bool accumulate);
// Y(rows x k, ldy) = beta*Y + alpha * op(A) X(cols x k, ldx) for a few right-hand sides, on the fp64 tensor
// path (DMMA); op(A) = A (m x n) or A' (trans).  The matrix is read once for all k columns.
void DenseGemmRhs(Ctx* ctx, const double* A, int64_t lda, int m, int n, bool trans, double alpha,
                  const double* X, int64_t ldx, int k, double beta, double* Y, int64_t ldy);
// C(m x m) (+)= alpha * A diag(w) A', A m x n column-major; w may be null (ones).
void DenseGramNT(Ctx* ctx, const double* A, int64_t lda, int m, int64_t n, const double* w,
                 double alpha, double* C, int64_t ldc, bool accumulate);
// M (n x n SPD, lower part read, destroyed) -> W = inv(chol(M)), lower triangular with an
// explicit zero upper part, so that inv(M) = W' W.  Throws Error(EPS_ERR_NUMERIC) when a
// pivot is not positive.
void DenseCholInverseFactor(Ctx* ctx, double* M, int64_t ld, int n, double* W);
void DenseAddDiagonal(Ctx* ctx, double* M, int64_t ld, int n, const double* d, double alpha_scalar);
// Gram of a CSR block: M[col_off.., col_off..] += alpha * A' A
void CsrGramAtomic(Ctx* ctx, const Csr& A, double alpha, double* M, int64_t ld, int col_off);
void CsrDensify(Ctx* ctx, const Csr& A, double scale, double* out, int64_t ld, int row_off, int col_off, bool trans);

// ---- communication: comm.cu ---------------------------------------------------------------
void CommInit(Ctx* ctx, int n_ranks, int rank, const void* unique_id);
void CommUniqueId(void* out128);
void CommDestroy(Ctx* ctx);
void AllReduceSum(Ctx* ctx, double* buf, int64_t n);

}  // namespace eps
