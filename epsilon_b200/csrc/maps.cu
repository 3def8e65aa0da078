// Linear operators on the device: dense GEMV (N/T), CSR SpMV (A and A'), diagonal,
// scalar.  These back LinearMapImpl::Apply of the reference:
//   DenseMatrixImpl::Apply   src/epsilon/linear/dense_matrix_impl.cc:55-67 (dgemv_)
//   SparseMatrixImpl::Apply  src/epsilon/linear/sparse_matrix_impl.h:24    (Eigen CSC * x)
//   DiagonalMatrixImpl::Apply diagonal_matrix_impl.h:23, ScalarMatrixImpl::Apply scalar_matrix_impl.h:24
// All arithmetic is fp64 (LinearMapImpl::Scalar = double, linear_map.h:35); indices int32.
#include <chrono>
#include <mutex>

#include <cub/cub.cuh>

#include "device.cuh"
#include "internal.h"

namespace eps {

// =====================================================================================
// Dense GEMV, column-major.  HBM-bound: 8*m*n bytes of A per launch.
// =====================================================================================
namespace {

constexpr int kGemvBlock = 256;

// y_part[s][i] = sum_{j in split s} A[i,j] x[j].  Each thread owns ROWS consecutive-by-
// blockDim rows (coalesced across the warp), x chunk staged in shared memory.
constexpr int kGemvNChunk = 512;  // columns of x staged per pass
constexpr int kGemvNRows = 2;     // rows per thread
__global__ void __launch_bounds__(kGemvBlock)
GemvNKernel(const double* __restrict__ A, int64_t lda, int m, int n, int cols_per_split,
            const double* __restrict__ x, double* __restrict__ out, int64_t out_stride, int tri) {
  __shared__ double xs[kGemvNChunk];
  const int split = blockIdx.y;
  const int j0 = split * cols_per_split;
  int j1 = min(n, j0 + cols_per_split);
  // lower-triangular A: columns beyond this CTA's last row hold zeros
  if (tri) j1 = min(j1, (int)min((int64_t)m, (int64_t)(blockIdx.x + 1) * (kGemvBlock * kGemvNRows)));
  const int row0 = blockIdx.x * (kGemvBlock * kGemvNRows) + threadIdx.x;
  double acc[kGemvNRows];
#pragma unroll
  for (int r = 0; r < kGemvNRows; r++) acc[r] = 0.0;
  for (int jc = j0; jc < j1; jc += kGemvNChunk) {
    const int nc = min(kGemvNChunk, j1 - jc);
    __syncthreads();
    for (int t = threadIdx.x; t < nc; t += kGemvBlock) xs[t] = x[jc + t];
    __syncthreads();
    const double* Ac = A + (int64_t)jc * lda;
    int j = 0;
    for (; j + 8 <= nc; j += 8) {
      double a[kGemvNRows][8];
#pragma unroll
      for (int r = 0; r < kGemvNRows; r++) {
        const int row = row0 + r * kGemvBlock;
#pragma unroll
        for (int u = 0; u < 8; u++)
          a[r][u] = row < m ? LdStream(Ac + (int64_t)(j + u) * lda + row) : 0.0;
      }
#pragma unroll
      for (int r = 0; r < kGemvNRows; r++)
#pragma unroll
        for (int u = 0; u < 8; u++) acc[r] += a[r][u] * xs[j + u];
    }
    for (; j < nc; j++) {
#pragma unroll
      for (int r = 0; r < kGemvNRows; r++) {
        const int row = row0 + r * kGemvBlock;
        if (row < m) acc[r] += LdStream(Ac + (int64_t)j * lda + row) * xs[j];
      }
    }
  }
#pragma unroll
  for (int r = 0; r < kGemvNRows; r++) {
    const int row = row0 + r * kGemvBlock;
    if (row < m) out[(int64_t)split * out_stride + row] = acc[r];
  }
}

// y_part[s][j] = sum_{i in split s} A[i,j] x[i]  for groups of kGemvTCols columns.
// Work items (column group, row split) are walked with a grid-stride loop by a grid sized to
// the resident CTA slots, so there is no partial last wave (a 3.4-wave grid measured 75 % of
// the copy peak, profiles/r01_ncu_full_GemvTKernel_before.csv).  VEC2: 128-bit loads, two
// consecutive rows per thread (needs even lda / row offsets and 16-byte aligned A, x).
constexpr int kGemvTCols = 8;
template <bool VEC2>
__global__ void __launch_bounds__(kGemvBlock, 2)
GemvTKernel(const double* __restrict__ A, int64_t lda, int m, int n, int rows_per_split, int n_groups,
            int n_items, const double* __restrict__ x, double* __restrict__ out, int64_t out_stride,
            int tri) {
  __shared__ double smem[kGemvTCols * 32];
  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int jg = (item % n_groups) * kGemvTCols;
    const int split = item / n_groups;
    int i0 = split * rows_per_split;
    const int i1 = min(m, i0 + rows_per_split);
    if (tri) i0 = max(i0, jg & ~1);  // lower-triangular A: rows above the column group are zero
    double acc[kGemvTCols];
#pragma unroll
    for (int c = 0; c < kGemvTCols; c++) acc[c] = 0.0;
    const int ncol = min(kGemvTCols, n - jg);
    if (ncol == kGemvTCols && VEC2) {
      // rows i, i+1 (one double2) and i+512, i+513 per pass: 16 x 16 B in flight per thread
      int i = i0 + 2 * threadIdx.x;
      for (; i + 2 * kGemvBlock + 1 < i1; i += 4 * kGemvBlock) {
        const int i2 = i + 2 * kGemvBlock;
        double2 a0[kGemvTCols], a1[kGemvTCols];
#pragma unroll
        for (int c = 0; c < kGemvTCols; c++) {
          const double* col = A + (int64_t)(jg + c) * lda;
          a0[c] = LdStream2(col + i);
          a1[c] = LdStream2(col + i2);
        }
        const double2 x0 = *reinterpret_cast<const double2*>(x + i);
        const double2 x1 = *reinterpret_cast<const double2*>(x + i2);
#pragma unroll
        for (int c = 0; c < kGemvTCols; c++)
          acc[c] += (a0[c].x * x0.x + a0[c].y * x0.y) + (a1[c].x * x1.x + a1[c].y * x1.y);
      }
      for (; i < i1; i += 2 * kGemvBlock) {  // remainder passes, element-wise guards
        const bool two = i + 1 < i1;
        const double xa = x[i], xb = two ? x[i + 1] : 0.0;
#pragma unroll
        for (int c = 0; c < kGemvTCols; c++) {
          const double* col = A + (int64_t)(jg + c) * lda;
          acc[c] += col[i] * xa + (two ? col[i + 1] * xb : 0.0);
        }
      }
    } else if (ncol == kGemvTCols) {
      for (int i = i0 + threadIdx.x; i < i1; i += kGemvBlock * 2) {
        const int i2 = i + kGemvBlock;
        const bool v2 = i2 < i1;
        double a0[kGemvTCols], a1[kGemvTCols];
        const double xv0 = x[i];
        const double xv1 = v2 ? x[i2] : 0.0;
#pragma unroll
        for (int c = 0; c < kGemvTCols; c++) {
          const double* col = A + (int64_t)(jg + c) * lda;
          a0[c] = LdStream(col + i);
          a1[c] = v2 ? LdStream(col + i2) : 0.0;
        }
#pragma unroll
        for (int c = 0; c < kGemvTCols; c++) acc[c] += a0[c] * xv0 + a1[c] * xv1;
      }
    } else {
      for (int i = i0 + threadIdx.x; i < i1; i += kGemvBlock) {
        const double xv = x[i];
        for (int c = 0; c < ncol; c++) acc[c] += A[(int64_t)(jg + c) * lda + i] * xv;
      }
    }
    BlockSum<kGemvTCols>(acc, smem);
    if (threadIdx.x == 0) {
      for (int c = 0; c < ncol; c++) out[(int64_t)split * out_stride + jg + c] = acc[c];
    }
    __syncthreads();  // smem is reused by the next item
  }
}

// y = beta*y + alpha * sum_s part[s]
__global__ void __launch_bounds__(256)
SplitReduceKernel(int len, int splits, const double* __restrict__ part, int64_t stride,
                  double alpha, double beta, double* __restrict__ y) {
  int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= len) return;
  double s = 0.0;
  for (int k = 0; k < splits; k++) s += part[(int64_t)k * stride + i];
  y[i] = (beta == 0.0 ? 0.0 : beta * y[i]) + alpha * s;
}

double* SplitScratch(Ctx* ctx, size_t count) { return GrowScratch(ctx, &ctx->split_scratch, count); }

}  // namespace

void DenseGemv(Ctx* ctx, const double* A, int lda, int m, int n, bool trans, double alpha,
               const double* x, double beta, double* y, int tri) {
  if (m == 0 || n == 0) return;
  // enough CTAs for ~6+ waves (4 resident per SM): a grid of 2.1 waves runs as 3 and loses 30 %
  const int target = ctx->sm_count * 24;
  const double bytes = (tri ? 4.0 : 8.0) * m * n + 8.0 * m + 8.0 * n;
  if (!trans) {
    const int rb = (m + kGemvBlock * kGemvNRows - 1) / (kGemvBlock * kGemvNRows);
    int splits = std::max(1, std::min((n + 63) / 64, target / rb));
    int cps = (n + splits - 1) / splits;
    cps = (cps + 7) / 8 * 8;
    splits = (n + cps - 1) / cps;
    double* part = SplitScratch(ctx, (size_t)splits * m);
    {
      ProfScope ps(ctx, KC_GEMV_N, bytes);
      GemvNKernel<<<dim3(rb, splits), kGemvBlock, 0, ctx->stream>>>(A, lda, m, n, cps, x, part, m, tri);
      ctx->count_launch();
    }
    ProfScope ps(ctx, KC_VEC, 8.0 * splits * m);
    SplitReduceKernel<<<(m + 255) / 256, 256, 0, ctx->stream>>>(m, splits, part, m, alpha, beta, y);
    ctx->count_launch();
  } else {
    const int cg = (n + kGemvTCols - 1) / kGemvTCols;
    // items of ~4096 rows x 8 columns (256 KB): small against a CTA's share of the matrix, so
    // the grid-stride loop balances to a few per cent
    int splits = std::max(1, (m + 4095) / 4096);
    int rps = (m + splits - 1) / splits;
    rps = (rps + 1) & ~1;  // even row offsets keep the 128-bit loads aligned
    splits = (m + rps - 1) / rps;
    const int n_items = cg * splits;
    double* part = SplitScratch(ctx, (size_t)splits * n);
    const bool vec2 = (lda % 2 == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    static int slots_per_sm[2] = {0, 0};
    if (!slots_per_sm[vec2]) {
      int v = 0;
      if (vec2) CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, GemvTKernel<true>, kGemvBlock, 0));
      else CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, GemvTKernel<false>, kGemvBlock, 0));
      slots_per_sm[vec2] = std::max(1, v);
    }
    const int grid = std::min(n_items, ctx->sm_count * slots_per_sm[vec2]);
    {
      ProfScope ps(ctx, KC_GEMV_T, bytes);
      if (vec2)
        GemvTKernel<true><<<grid, kGemvBlock, 0, ctx->stream>>>(A, lda, m, n, rps, cg, n_items, x, part, n, tri);
      else
        GemvTKernel<false><<<grid, kGemvBlock, 0, ctx->stream>>>(A, lda, m, n, rps, cg, n_items, x, part, n, tri);
      ctx->count_launch();
    }
    ProfScope ps(ctx, KC_VEC, 8.0 * splits * n);
    SplitReduceKernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(n, splits, part, n, alpha, beta, y);
    ctx->count_launch();
  }
}

// =====================================================================================
// CSR SpMV: y = beta*y + alpha*A x.  T lanes cooperate on one row; 128-bit-friendly
// streaming loads of (col,val), x gathered through L1/L2.
// Algorithmic bytes: 12*nnz + 4*(m+1) + 8*n + 8*m   (SURVEY section 8d).
// =====================================================================================
namespace {

constexpr int kSpmvBlock = 256;

template <int T>
__global__ void __launch_bounds__(kSpmvBlock)
CsrSpmvKernel(int m, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
              const double* __restrict__ val, const double* __restrict__ x, double alpha,
              double beta, double* __restrict__ y) {
  constexpr int kRowsPerBlock = kSpmvBlock / T;
  const int lane = threadIdx.x % T;
  const int grp = threadIdx.x / T;
  for (int row = blockIdx.x * kRowsPerBlock + grp; row < m; row += gridDim.x * kRowsPerBlock) {
    const int start = __ldg(rowptr + row);
    const int end = __ldg(rowptr + row + 1);
    double acc = 0.0;
    int k = start + lane;
    // 4-way unrolled: indices and values first, then the dependent gathers.
    for (; k + 3 * T < end; k += 4 * T) {
      const int c0 = LdStreamI(col + k), c1 = LdStreamI(col + k + T);
      const int c2 = LdStreamI(col + k + 2 * T), c3 = LdStreamI(col + k + 3 * T);
      const double v0 = LdStream(val + k), v1 = LdStream(val + k + T);
      const double v2 = LdStream(val + k + 2 * T), v3 = LdStream(val + k + 3 * T);
      const double x0 = __ldg(x + c0), x1 = __ldg(x + c1), x2 = __ldg(x + c2), x3 = __ldg(x + c3);
      acc += v0 * x0;
      acc += v1 * x1;
      acc += v2 * x2;
      acc += v3 * x3;
    }
    for (; k < end; k += T) acc += LdStream(val + k) * __ldg(x + LdStreamI(col + k));
    acc = GroupSum<T>(acc);
    if (lane == 0) y[row] = (beta == 0.0 ? 0.0 : beta * y[row]) + alpha * acc;
  }
}

template <int T>
void LaunchCsr(Ctx* ctx, const Csr& A, double alpha, const double* x, double beta, double* y) {
  constexpr int rows_per_block = kSpmvBlock / T;
  int64_t blocks = ((int64_t)A.m + rows_per_block - 1) / rows_per_block;
  int64_t cap = (int64_t)ctx->sm_count * 8 * 4;
  int grid = (int)std::max<int64_t>(1, std::min(blocks, cap));
  CsrSpmvKernel<T><<<grid, kSpmvBlock, 0, ctx->stream>>>(A.m, A.rowptr.p, A.col.p, A.val.p, x,
                                                         alpha, beta, y);
  ctx->count_launch();
}

__global__ void ExpandColIdKernel(int n, int64_t nnz, const int32_t* __restrict__ colptr,
                                  int32_t* __restrict__ colid, int32_t* __restrict__ perm) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nnz) return;
  int lo = 0, hi = n;  // find j with colptr[j] <= k < colptr[j+1]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if ((int64_t)colptr[mid] <= k) lo = mid; else hi = mid;
  }
  colid[k] = lo;
  perm[k] = (int32_t)k;
}

__global__ void RowCountKernel(int64_t nnz, const int32_t* __restrict__ rowidx, int32_t* counts) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz) atomicAdd(&counts[rowidx[k]], 1);
}

__global__ void GatherKernel(int64_t nnz, const int32_t* __restrict__ perm,
                             const int32_t* __restrict__ colid, const double* __restrict__ vals,
                             int32_t* __restrict__ col_out, double* __restrict__ val_out) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nnz) return;
  int p = perm[k];
  col_out[k] = colid[p];
  val_out[k] = vals[p];
}

// Ingest validation of the wire CSC (the reference trusts Eigen's Map of the same arrays,
// vector_util.cc:261-281; here a malformed matrix must come back as EPS_ERR_INVALID, not as an
// out-of-bounds device access).  flag bit 0: colptr not monotone / row index outside [0, m);
// bit 1: row indices not ascending inside a column (legal, but the kernels need them sorted).
__global__ void CscValidateKernel(int m, int n, int64_t nnz, const int32_t* __restrict__ colptr,
                                  const int32_t* __restrict__ rowidx, int* __restrict__ flag) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n && colptr[k] > colptr[k + 1]) atomicOr(flag, 1);
  if (k >= nnz) return;
  const int r = rowidx[k];
  if (r < 0 || r >= m) { atomicOr(flag, 1); return; }
  if (k > 0 && rowidx[k - 1] > r) {
    // a descent is fine exactly at a column start: find the column of k
    int lo = 0, hi = n;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if ((int64_t)colptr[mid] <= k) lo = mid; else hi = mid;
    }
    if ((int64_t)colptr[lo] != k) atomicOr(flag, 2);
  }
}

__global__ void DiagScaleKernel(int n, const double* __restrict__ d, double alpha,
                                const double* __restrict__ x, double beta, double* __restrict__ y) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = (beta == 0.0 ? 0.0 : beta * y[i]) + alpha * (d ? d[i] : 1.0) * x[i];
}

}  // namespace

void CsrApply(Ctx* ctx, const Csr& A, double alpha, const double* x, double beta, double* y,
              int kclass) {
  if (A.m == 0) return;
  ProfScope ps(ctx, kclass, 12.0 * A.nnz + 4.0 * (A.m + 1) + 8.0 * A.n + 8.0 * A.m);
  double avg = A.m ? (double)A.nnz / A.m : 0;
  if (avg <= 3) LaunchCsr<2>(ctx, A, alpha, x, beta, y);
  else if (avg <= 8) LaunchCsr<4>(ctx, A, alpha, x, beta, y);
  else if (avg <= 24) LaunchCsr<8>(ctx, A, alpha, x, beta, y);
  else if (avg <= 64) LaunchCsr<16>(ctx, A, alpha, x, beta, y);
  else LaunchCsr<32>(ctx, A, alpha, x, beta, y);
}

void BuildCsrFromHost(Ctx* ctx, int m, int n, int64_t nnz, const int64_t* rowptr,
                      const int32_t* col, const double* val, Csr* out) {
  EPS_CHECK(nnz < (int64_t)2147483647, "nnz %lld does not fit int32 offsets", (long long)nnz);
  out->m = m; out->n = n; out->nnz = nnz;
  std::vector<int32_t> rp(m + 1);
  int maxr = 0;
  for (int i = 0; i <= m; i++) rp[i] = (int32_t)rowptr[i];
  for (int i = 0; i < m; i++) maxr = std::max(maxr, rp[i + 1] - rp[i]);
  out->max_row_nnz = maxr;
  out->rowptr.alloc(ctx, m + 1);
  out->col.alloc(ctx, nnz);
  out->val.alloc(ctx, nnz);
  out->rowptr.upload(rp.data(), m + 1);
  out->col.upload(col, nnz);
  out->val.upload(val, nnz);
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
}

// Returns after the wire CSC on the device has been checked; sorts the row indices of every column
// when they arrive unsorted.  Throws EPS_ERR_INVALID for a malformed matrix.
static void ValidateCscOnDevice(Ctx* ctx, Csr* At) {
  const int n = At->m, m = At->n;
  const int64_t nnz = At->nnz;
  DevBuf<int> flag(ctx, 1);
  flag.zero();
  const int64_t work = std::max<int64_t>(nnz, n);
  if (work == 0) return;
  CscValidateKernel<<<(unsigned)((work + 255) / 256), 256, 0, ctx->stream>>>(m, n, nnz, At->rowptr.p, At->col.p, flag.p);
  ctx->count_launch();
  int h = 0;
  flag.download(&h, 1);
  if (h & 1) throw Error(EPS_ERR_INVALID, Format("malformed CSC matrix (%d x %d, nnz %lld): colptr is not "
                                                 "non-decreasing or a row index lies outside [0, %d)", m, n,
                                                 (long long)nnz, m));
  if (h & 2) {
    DevBuf<int32_t> keys(ctx, nnz);
    DevBuf<double> vals(ctx, nnz);
    size_t bytes = 0;
    CUDA_CHECK(cub::DeviceSegmentedSort::SortPairs(nullptr, bytes, At->col.p, keys.p, At->val.p, vals.p, (int)nnz, n,
                                                   At->rowptr.p, At->rowptr.p + 1, ctx->stream));
    DevBuf<char> tmp(ctx, bytes);
    CUDA_CHECK(cub::DeviceSegmentedSort::SortPairs(tmp.p, bytes, At->col.p, keys.p, At->val.p, vals.p, (int)nnz, n,
                                                   At->rowptr.p, At->rowptr.p + 1, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    At->col = std::move(keys);
    At->val = std::move(vals);
  }
}

// CSC (= CSR of A') on the device -> CSR of A, by a stable radix sort on the row index
// (deterministic: within a row the columns stay ascending).
static void TransposeCsrOnDevice(Ctx* ctx, const Csr& At, Csr* A) {
  const int m = At.n, n = At.m;
  const int64_t nnz = At.nnz;
  A->m = m; A->n = n; A->nnz = nnz;
  A->rowptr.alloc(ctx, m + 1);
  A->col.alloc(ctx, nnz);
  A->val.alloc(ctx, nnz);
  A->rowptr.zero();
  if (nnz == 0) return;
  DevBuf<int32_t> colid(ctx, nnz), perm_in(ctx, nnz), perm_out(ctx, nnz), keys_out(ctx, nnz);
  const int tb = 256;
  const unsigned gb = (unsigned)((nnz + tb - 1) / tb);
  ExpandColIdKernel<<<gb, tb, 0, ctx->stream>>>(n, nnz, At.rowptr.p, colid.p, perm_in.p);
  int end_bit = 1;
  while ((1ll << end_bit) < (long long)m) end_bit++;
  size_t tmp_bytes = 0;
  CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, At.col.p, keys_out.p, perm_in.p,
                                             perm_out.p, (int)nnz, 0, end_bit, ctx->stream));
  DevBuf<char> tmp(ctx, tmp_bytes);
  CUDA_CHECK(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, At.col.p, keys_out.p, perm_in.p,
                                             perm_out.p, (int)nnz, 0, end_bit, ctx->stream));
  GatherKernel<<<gb, tb, 0, ctx->stream>>>(nnz, perm_out.p, colid.p, At.val.p, A->col.p, A->val.p);
  // row pointer: counts -> exclusive scan
  DevBuf<int32_t> counts(ctx, m + 1);
  counts.zero();
  RowCountKernel<<<gb, tb, 0, ctx->stream>>>(nnz, At.col.p, counts.p);
  size_t scan_bytes = 0;
  CUDA_CHECK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, counts.p, A->rowptr.p, m + 1, ctx->stream));
  DevBuf<char> tmp2(ctx, scan_bytes);
  CUDA_CHECK(cub::DeviceScan::ExclusiveSum(tmp2.p, scan_bytes, counts.p, A->rowptr.p, m + 1, ctx->stream));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  CUDA_CHECK(cudaGetLastError());
}

double MapAlgorithmicBytes(const MapData& A, bool trans) {
  switch (A.kind) {
    case MK_DENSE: return 8.0 * A.m * A.n + 8.0 * A.m + 8.0 * A.n;
    case MK_SPARSE: {
      const Csr& c = trans ? A.csr_t : A.csr;
      return 12.0 * c.nnz + 4.0 * (c.m + 1) + 8.0 * c.n + 8.0 * c.m;
    }
    case MK_DIAGONAL: return 24.0 * A.n;
    default: return 16.0 * A.n;
  }
}

void MapApply(Ctx* ctx, const MapData& A, bool trans, double alpha, const double* x, double beta,
              double* y) {
  switch (A.kind) {
    case MK_DENSE:
      DenseGemv(ctx, A.dense.p, A.m, A.m, A.n, trans, alpha, x, beta, y);
      break;
    case MK_SPARSE: {
      const Sell& S = trans ? A.sell_t : A.sell;
      if (S.built) SellApply(ctx, S, alpha, x, beta, y, trans ? KC_SPMV_T : KC_SPMV_N);
      else CsrApply(ctx, trans ? A.csr_t : A.csr, alpha, x, beta, y, trans ? KC_SPMV_T : KC_SPMV_N);
      break;
    }
    case MK_DIAGONAL:
    case MK_SCALAR: {
      if (A.n == 0) break;
      ProfScope ps(ctx, KC_VEC, 24.0 * A.n);
      const double a = A.kind == MK_SCALAR ? alpha * A.alpha : alpha;
      DiagScaleKernel<<<(A.n + 255) / 256, 256, 0, ctx->stream>>>(
          A.n, A.kind == MK_DIAGONAL ? A.diag.p : nullptr, a, x, beta, y);
      ctx->count_launch();
      break;
    }
    default:
      throw Error(EPS_ERR_INVALID, "map kind not applicable on device");
  }
}

}  // namespace eps

// =====================================================================================
// C ABI
// =====================================================================================
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

int eps_map_dense(eps_ctx* ctx, int m, int n, const double* colmajor, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(ctx && out && m >= 0 && n >= 0, "eps_map_dense arguments");
  EPS_CHECK(colmajor || (size_t)m * n == 0, "eps_map_dense: null data pointer");
  CUDA_CHECK(cudaSetDevice(ctx->c.device));
  auto d = std::make_shared<MapData>();
  d->ctx = &ctx->c; d->kind = MK_DENSE; d->m = m; d->n = n;
  d->dense.alloc(&ctx->c, (size_t)m * n);
  // chunked upload keeps the pageable staging bounded for 80 GB matrices
  const size_t total = (size_t)m * n, chunk = (size_t)1 << 27;
  for (size_t off = 0; off < total; off += chunk) {
    size_t cnt = std::min(chunk, total - off);
    CUDA_CHECK(cudaMemcpyAsync(d->dense.p + off, colmajor + off, cnt * sizeof(double),
                               cudaMemcpyHostToDevice, ctx->c.stream));
  }
  CUDA_CHECK(cudaStreamSynchronize(ctx->c.stream));
  auto* h = new eps_map();
  h->d = d;
  *out = h;
  EPS_API_END
}

int eps_map_dense_dev(eps_ctx* ctx, int m, int n, double* colmajor_dev, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(ctx && out && m >= 0 && n >= 0 && (colmajor_dev || (size_t)m * n == 0), "eps_map_dense_dev arguments");
  cudaPointerAttributes attr;
  CUDA_CHECK(cudaPointerGetAttributes(&attr, colmajor_dev));
  EPS_CHECK(attr.type == cudaMemoryTypeDevice && attr.device == ctx->c.device,
            "eps_map_dense_dev: the matrix must live in the memory of device %d", ctx->c.device);
  auto d = std::make_shared<MapData>();
  d->ctx = &ctx->c; d->kind = MK_DENSE; d->m = m; d->n = n;
  d->dense.borrow(&ctx->c, colmajor_dev, (size_t)m * n);
  auto* h = new eps_map();
  h->d = d;
  *out = h;
  EPS_API_END
}

int eps_map_csc(eps_ctx* ctx, int m, int n, int64_t nnz, const int32_t* colptr,
                const int32_t* rowidx, const double* vals, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(ctx && out && m >= 0 && n >= 0 && nnz >= 0, "eps_map_csc arguments");
  EPS_CHECK(colptr && (nnz == 0 || (rowidx && vals)), "eps_map_csc: null colptr / rowidx / vals");
  EPS_CHECK(nnz < (int64_t)2147483647, "nnz %lld does not fit int32 offsets (shard the rows)",
            (long long)nnz);
  EPS_CHECK(colptr[0] == 0 && colptr[n] == nnz, "colptr[0]=%d colptr[n]=%d nnz=%lld", colptr[0],
            colptr[n], (long long)nnz);
  CUDA_CHECK(cudaSetDevice(ctx->c.device));
  auto d = std::make_shared<MapData>();
  d->ctx = &ctx->c; d->kind = MK_SPARSE; d->m = m; d->n = n;
  // EPS_TIMING=1: wall time of the ingest phases on stderr
  static const bool timing = getenv("EPS_TIMING") != nullptr;
  auto now = [&]() { cudaStreamSynchronize(ctx->c.stream); return std::chrono::steady_clock::now(); };
  auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    return std::chrono::duration<double, std::milli>(b - a).count();
  };
  const auto t0 = timing ? now() : std::chrono::steady_clock::time_point();
  // CSR of A' is the wire layout itself.
  Csr& t = d->csr_t;
  t.m = n; t.n = m; t.nnz = nnz;
  t.rowptr.alloc(&ctx->c, n + 1);
  t.col.alloc(&ctx->c, nnz);
  t.val.alloc(&ctx->c, nnz);
  t.rowptr.upload(colptr, n + 1);
  t.col.upload(rowidx, nnz);
  t.val.upload(vals, nnz);
  ValidateCscOnDevice(&ctx->c, &t);   // synchronises; EPS_ERR_INVALID instead of a device fault
  const auto t1 = timing ? now() : t0;
  TransposeCsrOnDevice(&ctx->c, t, &d->csr);
  const auto t2 = timing ? now() : t0;
  const char* fmt = getenv("EPS_SPARSE_FORMAT");  // "csr" forces the plain CSR kernel
  const bool want_sell = fmt ? strcmp(fmt, "csr") != 0 : true;
  if (want_sell && (nnz >= ((int64_t)1 << 18) || (fmt && strcmp(fmt, "sell") == 0)) && m > 0 && n > 0) {
    BuildSell(&ctx->c, d->csr, &d->sell);
    BuildSell(&ctx->c, d->csr_t, &d->sell_t);
  }
  if (timing)
    fprintf(stderr, "[eps timing] map_csc %dx%d nnz=%lld: upload %.1f ms, transpose %.1f ms, sell x2 %.1f ms\n", m, n,
            (long long)nnz, ms(t0, t1), ms(t1, t2), ms(t2, now()));
  if (nnz <= (int64_t)4 << 20) {
    d->h_colptr.assign(colptr, colptr + n + 1);
    d->h_rowidx.assign(rowidx, rowidx + nnz);
    d->h_vals.assign(vals, vals + nnz);
    d->has_host = true;
  }
  auto* h = new eps_map();
  h->d = d;
  *out = h;
  EPS_API_END
}

int eps_map_diagonal(eps_ctx* ctx, int n, const double* dvals, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(ctx && out && n >= 0, "eps_map_diagonal arguments");
  EPS_CHECK(dvals || n == 0, "eps_map_diagonal: null data pointer");
  CUDA_CHECK(cudaSetDevice(ctx->c.device));
  auto d = std::make_shared<MapData>();
  d->ctx = &ctx->c; d->kind = MK_DIAGONAL; d->m = n; d->n = n;
  d->diag.alloc(&ctx->c, n);
  d->diag.upload(dvals, n);
  d->h_diag.assign(dvals, dvals + n);
  CUDA_CHECK(cudaStreamSynchronize(ctx->c.stream));
  auto* h = new eps_map();
  h->d = d;
  *out = h;
  EPS_API_END
}

int eps_map_scalar(eps_ctx* ctx, int n, double alpha, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(ctx && out && n >= 0, "eps_map_scalar arguments");
  auto d = std::make_shared<MapData>();
  d->ctx = &ctx->c; d->kind = MK_SCALAR; d->m = n; d->n = n; d->alpha = alpha;
  auto* h = new eps_map();
  h->d = d;
  *out = h;
  EPS_API_END
}

int eps_map_kronecker(eps_ctx* ctx, eps_map* a, eps_map* b, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(ctx && a && b && out, "eps_map_kronecker arguments");
  auto d = std::make_shared<MapData>();
  d->ctx = &ctx->c; d->kind = MK_KRONECKER;
  d->ka = a->d; d->kb = b->d; d->ka_trans = a->trans; d->kb_trans = b->trans;
  d->m = a->m() * b->m();
  d->n = a->n() * b->n();
  auto* h = new eps_map();
  h->d = d;
  *out = h;
  EPS_API_END
}

int eps_map_transpose(eps_map* map, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(map && out, "eps_map_transpose arguments");
  auto* h = new eps_map();
  h->d = map->d;
  h->trans = !map->trans;
  *out = h;
  EPS_API_END
}

// LinearMapImpl::Inverse() (linear_map.h:49).  What the reference does per type:
//   scalar    1/alpha                                    scalar_matrix_impl.h:29-31
//   diagonal  1/d_i, 0 where d_i == 0                    diagonal_matrix_impl.cc:15-23
//   Kronecker inv(A) (x) inv(B)                          kronecker_product_impl.h:37-39
//   dense     Eigen::LDLT of the (assumed symmetric) matrix, solve(Identity)   dense_matrix_impl.cc:21-30
//   sparse    scalar matrix -> scalar inverse, else densify and as dense       sparse_matrix_impl.cc:60-78
// Every dense matrix the solver inverts (BlockCholesky pivots, block_cholesky.cc:125) is symmetric
// definite (+ or -), so the device factorises M or -M by Cholesky, inv = W'W with W = inv(chol)
// (DMMA Gram kernel); an indefinite matrix is refused with EPS_ERR_NUMERIC.
static std::shared_ptr<MapData> InverseOf(Ctx* ctx, const MapData& A, bool trans) {
  EPS_CHECK(A.m == A.n || A.kind == MK_KRONECKER, "inverse of a %d x %d map", A.m, A.n);
  auto d = std::make_shared<MapData>();
  d->ctx = ctx; d->m = A.n; d->n = A.m;
  switch (A.kind) {
    case MK_SCALAR:
      d->kind = MK_SCALAR;
      d->alpha = 1 / A.alpha;
      return d;
    case MK_DIAGONAL: {
      d->kind = MK_DIAGONAL;
      d->h_diag.resize(A.n);
      for (int i = 0; i < A.n; i++) d->h_diag[i] = A.h_diag[i] != 0.0 ? 1 / A.h_diag[i] : 0.0;
      d->diag.alloc(ctx, A.n);
      d->diag.upload(d->h_diag.data(), A.n);
      CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      return d;
    }
    case MK_KRONECKER:
      d->kind = MK_KRONECKER;
      d->ka = InverseOf(ctx, *A.ka, A.ka_trans != trans);
      d->kb = InverseOf(ctx, *A.kb, A.kb_trans != trans);
      d->m = d->ka->m * d->kb->m;
      d->n = d->ka->n * d->kb->n;
      return d;
    case MK_SPARSE:
      if (A.has_host) {  // IsScalarMatrix (sparse_matrix_impl.cc:66-71)
        bool scalar = A.h_colptr[A.n] == A.n && A.n > 0;
        for (int j = 0; scalar && j < A.n; j++)
          scalar = A.h_colptr[j] == j && A.h_rowidx[j] == j && A.h_vals[j] == A.h_vals[0];
        if (scalar) {
          d->kind = MK_SCALAR;
          d->alpha = 1 / A.h_vals[0];
          return d;
        }
      }
      // fall through: densify
    case MK_DENSE: {
      const int n = A.n;
      EPS_CHECK(n <= 60000, "dense inverse of a %d x %d map does not fit", n, n);
      DevBuf<double> M(ctx, (size_t)n * n), W(ctx, (size_t)n * n);
      auto load = [&](double sign) {
        if (A.kind == MK_DENSE) {
          VecScale(ctx, (int64_t)n * n, sign, A.dense.p, M.p);
        } else {
          M.zero();
          CsrDensify(ctx, A.csr, sign, M.p, n, 0, 0, trans);
        }
      };
      double sign = 1.0;
      load(1.0);
      try {
        DenseCholInverseFactor(ctx, M.p, n, n, W.p);
      } catch (const Error& e) {
        if (e.code != EPS_ERR_NUMERIC) throw;
        sign = -1.0;
        load(-1.0);
        try {
          DenseCholInverseFactor(ctx, M.p, n, n, W.p);
        } catch (const Error& e2) {
          if (e2.code != EPS_ERR_NUMERIC) throw;
          throw Error(EPS_ERR_NUMERIC, Format("eps_map_inverse: the %d x %d matrix is neither positive nor negative "
                                              "definite (the reference's LDLT would accept it; the device path "
                                              "factorises by Cholesky)", n, n));
        }
      }
      d->kind = MK_DENSE;
      d->dense.alloc(ctx, (size_t)n * n);
      DenseGramTN(ctx, W.p, n, n, n, sign, d->dense.p, n, false);   // inv(M) = sign * W'W
      CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      return d;
    }
    default:
      throw Error(EPS_ERR_INVALID, "unknown map kind");
  }
}

int eps_map_inverse(eps_map* map, eps_map** out) {
  EPS_API_BEGIN
  EPS_CHECK(map && out, "eps_map_inverse arguments");
  Ctx* ctx = map->d->ctx;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  auto* h = new eps_map();
  try {
    h->d = InverseOf(ctx, *map->d, map->trans);
  } catch (...) {
    delete h;
    throw;
  }
  // the inverse of a symmetric matrix is symmetric; a Kronecker inverse was built for op(map) already
  h->trans = false;
  *out = h;
  EPS_API_END
}

int eps_map_dims(const eps_map* map, int* m, int* n) {
  EPS_API_BEGIN
  EPS_CHECK(map, "null map");
  if (m) *m = map->m();
  if (n) *n = map->n();
  EPS_API_END
}

int eps_map_apply_dev(eps_map* map, const double* x_dev, double* y_dev) {
  EPS_API_BEGIN
  EPS_CHECK(map && x_dev && y_dev, "eps_map_apply_dev arguments");
  Ctx* ctx = map->d->ctx;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  if (map->d->kind == MK_KRONECKER) KroneckerApply(ctx, *map->d, map->trans, x_dev, y_dev);
  else MapApply(ctx, *map->d, map->trans, 1.0, x_dev, 0.0, y_dev);
  CUDA_CHECK(cudaGetLastError());
  EPS_API_END
}

int eps_map_apply(eps_map* map, const double* x, double* y) {
  EPS_API_BEGIN
  EPS_CHECK(map && x && y, "eps_map_apply arguments");
  Ctx* ctx = map->d->ctx;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  DevBuf<double> dx(ctx, map->n()), dy(ctx, map->m());
  dx.upload(x, map->n());
  if (map->d->kind == MK_KRONECKER) KroneckerApply(ctx, *map->d, map->trans, dx.p, dy.p);
  else MapApply(ctx, *map->d, map->trans, 1.0, dx.p, 0.0, dy.p);
  CUDA_CHECK(cudaGetLastError());
  dy.download(y, map->m());
  EPS_API_END
}

int eps_map_time(eps_map* map, int reps, float* ms_per_apply, double* algorithmic_bytes) {
  EPS_API_BEGIN
  EPS_CHECK(map && reps > 0, "eps_map_time arguments");
  Ctx* ctx = map->d->ctx;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  DevBuf<double> dx(ctx, map->n()), dy(ctx, map->m());
  VecFill(ctx, map->n(), 1.0, dx.p);
  for (int i = 0; i < 3; i++) MapApply(ctx, *map->d, map->trans, 1.0, dx.p, 0.0, dy.p);
  cudaEvent_t e0, e1;
  CUDA_CHECK(cudaEventCreate(&e0));
  CUDA_CHECK(cudaEventCreate(&e1));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  CUDA_CHECK(cudaEventRecord(e0, ctx->stream));
  for (int i = 0; i < reps; i++) MapApply(ctx, *map->d, map->trans, 1.0, dx.p, 0.0, dy.p);
  CUDA_CHECK(cudaEventRecord(e1, ctx->stream));
  CUDA_CHECK(cudaEventSynchronize(e1));
  float ms = 0;
  CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (ms_per_apply) *ms_per_apply = ms / reps;
  if (algorithmic_bytes) *algorithmic_bytes = MapAlgorithmicBytes(*map->d, map->trans);
  EPS_API_END
}

void eps_map_destroy(eps_map* map) { delete map; }

}  // extern "C"

// =====================================================================================
// eps_dense_apply_noalloc: y = A x for a dense column-major A with the split partials in a static
// device array instead of the context's grown scratch.  No counterpart in the reference; added
// for callers that cannot allocate (CUDA graph capture, serving loops).
// Same two kernels, same split arithmetic and the same summation order as DenseGemv(trans = false),
// so the result is bit-identical to eps_map_apply_dev on an eps_map_dense of the same matrix.
// =====================================================================================
namespace {

// Lives in every GPU that loads the library until the process exits (8 MiB).  Shared by all calls:
// GemvNKernel overwrites every entry SplitReduceKernel reads, so nothing has to start at zero, but two
// calls that overlap on the GPU (different streams / contexts of one device) race on it.
static __device__ __align__(256) double g_noalloc_partials[EPS_NOALLOC_SCRATCH_DOUBLES];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
double* g_noalloc_ptr[kNoAllocMaxDevices] = {};   // device address of the array, per device ordinal

// In host code `g_noalloc_partials` is only a handle: the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
double* NoAllocScratch(int device) {
  EPS_CHECK(device >= 0 && device < kNoAllocMaxDevices, "eps_dense_apply_noalloc: device ordinal %d out of range", device);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[device] == nullptr) {
    void* p = nullptr;
    CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_partials));
    EPS_CHECK(p != nullptr, "eps_dense_apply_noalloc: no device address for the static scratch");
    g_noalloc_ptr[device] = static_cast<double*>(p);
  }
  return g_noalloc_ptr[device];
}

}  // namespace

extern "C" int eps_dense_apply_noalloc(eps_ctx* c, int m, int n, const double* a_colmajor_dev, const double* x_dev,
                                       double* y_dev) {
  EPS_API_BEGIN
  EPS_CHECK(c && a_colmajor_dev && x_dev && y_dev, "eps_dense_apply_noalloc arguments");
  EPS_CHECK(m > 0 && n > 0, "eps_dense_apply_noalloc: m = %d, n = %d must be positive", m, n);
  Ctx* ctx = &c->c;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  // the split of DenseGemv(trans = false), unchanged
  const int target = ctx->sm_count * 24;
  const int rb = (m + kGemvBlock * kGemvNRows - 1) / (kGemvBlock * kGemvNRows);
  int splits = std::max(1, std::min((n + 63) / 64, target / rb));
  int cps = (n + splits - 1) / splits;
  cps = (cps + 7) / 8 * 8;
  splits = (n + cps - 1) / cps;
  const size_t need = (size_t)splits * (size_t)m;
  if (need > (size_t)EPS_NOALLOC_SCRATCH_DOUBLES)
    throw Error(EPS_ERR_INVALID,
                Format("eps_dense_apply_noalloc: %d x %d needs %zu partial sums (%d column splits x %d rows), the static "
                       "scratch holds EPS_NOALLOC_SCRATCH_DOUBLES = %d; use eps_map_dense + eps_map_apply_dev",
                       m, n, need, splits, m, (int)EPS_NOALLOC_SCRATCH_DOUBLES));
  double* part = NoAllocScratch(ctx->device);
  GemvNKernel<<<dim3(rb, splits), kGemvBlock, 0, ctx->stream>>>(a_colmajor_dev, m, m, n, cps, x_dev, part, m, 0);
  SplitReduceKernel<<<(m + 255) / 256, 256, 0, ctx->stream>>>(m, splits, part, m, 1.0, 0.0, y_dev);
  ctx->count_launch(2);
  CUDA_CHECK(cudaGetLastError());
  EPS_API_END
}
