// Dense fp64 contractions for the one-time factorisations of the KKT prox
// (reference: Multiply_DenseMatrix_DenseMatrix -> dgemm_, linear/linear_map_multiply.cc:14-37;
//  DenseMatrixImpl::Inverse -> Eigen::LDLT + solve(I), linear/dense_matrix_impl.cc:21-30;
//  SparseMatrixImpl::Inverse densifies, linear/sparse_matrix_impl.cc:60-78).
//
// fp64 has no tcgen05 kind on Blackwell; the tensor path for doubles is DMMA
// (mma.sync.m8n8k4.f64), which is what GemmKernel issues.  This is the only place the
// path is a dense contraction.
#include "device.cuh"
#include "internal.h"

namespace eps {

namespace {

constexpr int BM = 64, BN = 64, BKK = 16, SPAD = 4;
constexpr int kGemmThreads = 128;

__device__ __forceinline__ void Dmma(double& d0, double& d1, double a, double b) {
  asm volatile(
      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(d0), "+d"(d1)
      : "d"(a), "d"(b));
}

// C[m + n*ldc] = beta*C + alpha * sum_k a(m,k) b(n,k)
//   a(m,k) = AK ? A[k + m*lda] : A[m + k*lda]     (times kscale[k] when given)
//   b(n,k) = BK ? B[k + n*ldb] : B[n + k*ldb]
// lower_only: tiles strictly above the diagonal are skipped (symmetric results).
// b_lower_tri: b(n,k) is zero for k < n (B lower triangular as a K x N matrix): the k loop
// starts at the tile's first column.
template <bool AK, bool BK>
__global__ void __launch_bounds__(kGemmThreads)
GemmKernel(int M, int N, int K, double alpha, const double* __restrict__ A, int64_t lda,
           const double* __restrict__ B, int64_t ldb, const double* __restrict__ kscale,
           double beta, double* __restrict__ C, int64_t ldc, int lower_only, int b_lower_tri,
           int kchunk = 0, int64_t cstride = 0) {
  // split-K (kchunk > 0): CTA z contracts k in [z*kchunk, (z+1)*kchunk) into its own plane of C
  const int Kfull = K;
  int kz0 = 0;
  if (kchunk > 0) {
    kz0 = blockIdx.z * kchunk;
    K = min(Kfull, kz0 + kchunk);
    C += (int64_t)blockIdx.z * cstride;
  }
  __shared__ double sA[BM][BKK + SPAD];
  __shared__ double sB[BN][BKK + SPAD];
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  if (lower_only && m0 + BM <= n0) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

  double ra[8], rb[8];
  auto load_tiles = [&](int k0) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      int m, k;
      if (AK) { k = tid & 15; m = (tid >> 4) + 8 * r; } else { m = tid & 63; k = (tid >> 6) + 2 * r; }
      const int gm = m0 + m, gk = k0 + k;
      double v = 0.0;
      if (gm < M && gk < K) {
        v = AK ? A[gk + (int64_t)gm * lda] : A[gm + (int64_t)gk * lda];
        if (kscale) v *= kscale[gk];
      }
      ra[r] = v;
    }
#pragma unroll
    for (int r = 0; r < 8; r++) {
      int n, k;
      if (BK) { k = tid & 15; n = (tid >> 4) + 8 * r; } else { n = tid & 63; k = (tid >> 6) + 2 * r; }
      const int gn = n0 + n, gk = k0 + k;
      rb[r] = (gn < N && gk < K) ? (BK ? B[gk + (int64_t)gn * ldb] : B[gn + (int64_t)gk * ldb]) : 0.0;
    }
  };
  auto store_tiles = [&]() {
#pragma unroll
    for (int r = 0; r < 8; r++) {
      int m, k;
      if (AK) { k = tid & 15; m = (tid >> 4) + 8 * r; } else { m = tid & 63; k = (tid >> 6) + 2 * r; }
      sA[m][k] = ra[r];
    }
#pragma unroll
    for (int r = 0; r < 8; r++) {
      int n, k;
      if (BK) { k = tid & 15; n = (tid >> 4) + 8 * r; } else { n = tid & 63; k = (tid >> 6) + 2 * r; }
      sB[n][k] = rb[r];
    }
  };

  int kbeg = kz0;
  if (b_lower_tri) kbeg = (n0 / BKK) * BKK;
  if (kbeg < K) {
    load_tiles(kbeg);
    for (int k0 = kbeg; k0 < K; k0 += BKK) {
      __syncthreads();
      store_tiles();
      __syncthreads();
      if (k0 + BKK < K) load_tiles(k0 + BKK);
#pragma unroll
      for (int kk = 0; kk < BKK; kk += 4) {
        double fa[4], fb[4];
#pragma unroll
        for (int i = 0; i < 4; i++) fa[i] = sA[wm + i * 8 + (lane >> 2)][kk + (lane & 3)];
#pragma unroll
        for (int j = 0; j < 4; j++) fb[j] = sB[wn + j * 8 + (lane >> 2)][kk + (lane & 3)];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 4; j++) Dmma(acc[i][j][0], acc[i][j][1], fa[i], fb[j]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int gm = m0 + wm + i * 8 + (lane >> 2);
        const int gn = n0 + wn + j * 8 + (lane & 3) * 2 + e;
        if (gm < M && gn < N) {
          double* c = C + gm + (int64_t)gn * ldc;
          *c = (beta == 0.0 ? 0.0 : beta * *c) + alpha * acc[i][j][e];
        }
      }
}

template <bool AK, bool BK>
void Gemm(Ctx* ctx, int M, int N, int K, double alpha, const double* A, int64_t lda,
          const double* B, int64_t ldb, const double* kscale, double beta, double* C, int64_t ldc,
          bool lower_only, bool b_lower_tri) {
  if (M <= 0 || N <= 0) return;
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN);
  ProfScope ps(ctx, KC_KKT_DENSE, 0);
  GemmKernel<AK, BK><<<grid, kGemmThreads, 0, ctx->stream>>>(M, N, K, alpha, A, lda, B, ldb, kscale,
                                                              beta, C, ldc, lower_only ? 1 : 0,
                                                              b_lower_tri ? 1 : 0);
  ctx->count_launch();
}

// C upper <- C lower (after a lower_only product)
__global__ void SymmetrizeKernel(int n, double* C, int64_t ld) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y * blockDim.y + threadIdx.y;
  if (i < n && j < n && i > j) C[j + (int64_t)i * ld] = C[i + (int64_t)j * ld];
}

constexpr int NB = 64;

// Cholesky of one diagonal block (kb <= 64) in shared memory; writes L (upper part zeroed)
// back and inv(L) (kb x kb, leading dimension NB) to `linv`.  flag != 0 on a bad pivot.
__global__ void __launch_bounds__(256) PotrfDiagKernel(double* A, int64_t ld, int kb, double* linv, int* flag) {
  __shared__ double s[NB][NB + 1];
  const int tid = threadIdx.x;
  for (int e = tid; e < NB * NB; e += 256) {
    int i = e % NB, j = e / NB;
    s[i][j] = (i < kb && j < kb) ? A[i + (int64_t)j * ld] : (i == j ? 1.0 : 0.0);
  }
  __syncthreads();
  for (int j = 0; j < kb; j++) {
    double d = s[j][j];
    if (!(d > 0.0)) {
      if (tid == 0) atomicExch(flag, 1);
      d = 1.0;
    }
    d = sqrt(d);
    __syncthreads();
    if (tid == 0) s[j][j] = d;
    for (int i = j + 1 + tid; i < kb; i += 256) s[i][j] /= d;
    __syncthreads();
    // trailing update of the lower triangle
    const int rem = kb - j - 1;
    for (int e = tid; e < rem * rem; e += 256) {
      int i = j + 1 + e % rem, k = j + 1 + e / rem;
      if (i >= k) s[i][k] -= s[i][j] * s[k][j];
    }
    __syncthreads();
  }
  // inverse of L by forward substitution, one column per thread
  if (tid < kb) {
    const int c = tid;
    double w[NB];
    for (int i = 0; i < c; i++) w[i] = 0.0;
    w[c] = 1.0 / s[c][c];
    for (int i = c + 1; i < kb; i++) {
      double t = 0.0;
      for (int k = c; k < i; k++) t += s[i][k] * w[k];
      w[i] = -t / s[i][i];
    }
    for (int i = 0; i < kb; i++) linv[i + c * NB] = w[i];
  }
  __syncthreads();
  for (int e = tid; e < kb * kb; e += 256) {
    int i = e % kb, j = e / kb;
    A[i + (int64_t)j * ld] = i >= j ? s[i][j] : 0.0;
  }
}

__global__ void Copy2DKernel(int m, int n, const double* __restrict__ src, int64_t lds,
                             double* __restrict__ dst, int64_t ldd, double scale) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y;
  if (i < m && j < n) dst[i + (int64_t)j * ldd] = scale * src[i + (int64_t)j * lds];
}

void Copy2D(Ctx* ctx, int m, int n, const double* src, int64_t lds, double* dst, int64_t ldd,
            double scale = 1.0) {
  if (m <= 0 || n <= 0) return;
  Copy2DKernel<<<dim3((m + 255) / 256, n), 256, 0, ctx->stream>>>(m, n, src, lds, dst, ldd, scale);
  ctx->count_launch();
}

__global__ void AddDiagKernel(int n, double* M, int64_t ld, const double* d, double a) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) M[i + (int64_t)i * ld] += (d ? d[i] : 0.0) + a;
}

// One thread per row of A: M[c1,c2] += alpha * v1 v2 over the row's pairs.
__global__ void CsrGramKernel(int m, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                              const double* __restrict__ val, double alpha, double* M, int64_t ld,
                              int col_off) {
  int row = blockIdx.x;
  if (row >= m) return;
  const int s = rowptr[row], e = rowptr[row + 1], len = e - s;
  for (int p = threadIdx.x; p < len * len; p += blockDim.x) {
    int a = s + p % len, b = s + p / len;
    atomicAdd(&M[(col_off + col[a]) + (int64_t)(col_off + col[b]) * ld], alpha * val[a] * val[b]);
  }
}

__global__ void CsrDensifyKernel(int m, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                 const double* __restrict__ val, double scale, double* out, int64_t ld,
                                 int row_off, int col_off, int trans) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= m) return;
  for (int k = rowptr[row]; k < rowptr[row + 1]; k++) {
    int r = row, c = col[k];
    if (trans) { int t = r; r = c; c = t; }
    atomicAdd(&out[(row_off + r) + (int64_t)(col_off + c) * ld], scale * val[k]);
  }
}

}  // namespace

void DenseGramTN(Ctx* ctx, const double* A, int64_t lda, int64_t m, int n, double alpha, double* C,
                 int64_t ldc, bool accumulate) {
  EPS_CHECK(m < ((int64_t)1 << 31), "Gram K dimension %lld", (long long)m);
  Gemm<true, true>(ctx, n, n, (int)m, alpha, A, lda, A, lda, nullptr, accumulate ? 1.0 : 0.0, C, ldc,
                   true, false);
  dim3 tb(16, 16);
  SymmetrizeKernel<<<dim3((n + 15) / 16, (n + 15) / 16), tb, 0, ctx->stream>>>(n, C, ldc);
  ctx->count_launch();
}

void DenseGramNT(Ctx* ctx, const double* A, int64_t lda, int m, int64_t n, const double* w, double alpha,
                 double* C, int64_t ldc, bool accumulate) {
  EPS_CHECK(n < ((int64_t)1 << 31), "Gram K dimension %lld", (long long)n);
  Gemm<false, false>(ctx, m, m, (int)n, alpha, A, lda, A, lda, w, accumulate ? 1.0 : 0.0, C, ldc, true,
                     false);
  dim3 tb(16, 16);
  SymmetrizeKernel<<<dim3((m + 15) / 16, (m + 15) / 16), tb, 0, ctx->stream>>>(m, C, ldc);
  ctx->count_launch();
}

__global__ void __launch_bounds__(256)
GemmSplitReduceKernel(int64_t len, int rows, int splits, const double* __restrict__ part, int64_t stride,
                      double alpha, double beta, double* __restrict__ Y, int64_t ldy) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= len) return;
  double s = 0.0;
  for (int z = 0; z < splits; z++) s += part[(int64_t)z * stride + i];
  double* y = Y + (i % rows) + (i / rows) * ldy;
  *y = (beta == 0.0 ? 0.0 : beta * *y) + alpha * s;
}

void DenseGemmRhs(Ctx* ctx, const double* A, int64_t lda, int m, int n, bool trans, double alpha,
                  const double* X, int64_t ldx, int k, double beta, double* Y, int64_t ldy) {
  // C[i + j ldy] = sum_kk a(i, kk) x(j, kk): a(i,kk) = A[i + kk lda] (AK = false) or A[kk + i lda] (A', AK = true);
  // x(j,kk) = X[kk + j ldx] (BK = true).  With a few columns the output has only rows/64 tiles: the
  // contraction is split over gridDim.z so that ~2 CTAs per SM stream the matrix, partial planes are
  // added in a fixed order.
  const int M = trans ? n : m, K = trans ? m : n;
  if (M <= 0 || k <= 0) return;
  const int tiles = ((M + BM - 1) / BM) * ((k + BN - 1) / BN);
  int splits = std::max(1, std::min((2 * ctx->sm_count + tiles - 1) / tiles, (K + 4 * BKK - 1) / (4 * BKK)));
  int kchunk = (K + splits - 1) / splits;
  kchunk = (kchunk + BKK - 1) / BKK * BKK;
  splits = (K + kchunk - 1) / kchunk;
  if (splits <= 1) {
    if (!trans) Gemm<false, true>(ctx, m, k, n, alpha, A, lda, X, ldx, nullptr, beta, Y, ldy, false, false);
    else Gemm<true, true>(ctx, n, k, m, alpha, A, lda, X, ldx, nullptr, beta, Y, ldy, false, false);
    return;
  }
  const int64_t plane = (int64_t)M * k;
  double* part = GrowScratch(ctx, &ctx->split_scratch, (size_t)plane * splits);
  dim3 grid((M + BM - 1) / BM, (k + BN - 1) / BN, splits);
  {
    ProfScope ps(ctx, KC_KKT_DENSE, 8.0 * m * n);
    if (!trans)
      GemmKernel<false, true><<<grid, kGemmThreads, 0, ctx->stream>>>(M, k, K, 1.0, A, lda, X, ldx, nullptr, 0.0, part, M,
                                                                       0, 0, kchunk, plane);
    else
      GemmKernel<true, true><<<grid, kGemmThreads, 0, ctx->stream>>>(M, k, K, 1.0, A, lda, X, ldx, nullptr, 0.0, part, M,
                                                                      0, 0, kchunk, plane);
    ctx->count_launch();
  }
  ProfScope ps(ctx, KC_VEC, 8.0 * plane * splits);
  GemmSplitReduceKernel<<<(unsigned)((plane + 255) / 256), 256, 0, ctx->stream>>>(plane, M, splits, part, plane, alpha,
                                                                                 beta, Y, ldy);
  ctx->count_launch();
}

void DenseAddDiagonal(Ctx* ctx, double* M, int64_t ld, int n, const double* d, double a) {
  if (n <= 0) return;
  AddDiagKernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(n, M, ld, d, a);
  ctx->count_launch();
}

void CsrGramAtomic(Ctx* ctx, const Csr& A, double alpha, double* M, int64_t ld, int col_off) {
  if (A.m == 0) return;
  CsrGramKernel<<<A.m, 128, 0, ctx->stream>>>(A.m, A.rowptr.p, A.col.p, A.val.p, alpha, M, ld, col_off);
  ctx->count_launch();
}

void CsrDensify(Ctx* ctx, const Csr& A, double scale, double* out, int64_t ld, int row_off, int col_off,
                bool trans) {
  if (A.m == 0) return;
  CsrDensifyKernel<<<(A.m + 127) / 128, 128, 0, ctx->stream>>>(A.m, A.rowptr.p, A.col.p, A.val.p, scale,
                                                                out, ld, row_off, col_off, trans ? 1 : 0);
  ctx->count_launch();
}

// M (n x n SPD, lower part read, destroyed) -> W = inv(chol(M)) lower triangular, so that
// inv(M) = W' W.  W must hold n*n doubles (leading dimension n).
void DenseCholInverseFactor(Ctx* ctx, double* M, int64_t ld, int n, double* W) {
  if (n == 0) return;
  const int nblk = (n + NB - 1) / NB;
  DevBuf<double> linv(ctx, (size_t)nblk * NB * NB);
  DevBuf<double> panel(ctx, (size_t)n * NB);
  DevBuf<int> flag(ctx, 1);
  flag.zero();
  for (int b = 0; b < nblk; b++) {
    const int k0 = b * NB, kb = std::min(NB, n - k0), rem = n - k0 - kb;
    double* Akk = M + k0 + (int64_t)k0 * ld;
    double* Lkk_inv = linv.p + (size_t)b * NB * NB;
    PotrfDiagKernel<<<1, 256, 0, ctx->stream>>>(Akk, ld, kb, Lkk_inv, flag.p);
    ctx->count_launch();
    if (rem > 0) {
      double* P = M + (k0 + kb) + (int64_t)k0 * ld;
      // L_panel = P * inv(L_kk)'
      Gemm<false, false>(ctx, rem, kb, kb, 1.0, P, ld, Lkk_inv, NB, nullptr, 0.0, panel.p, rem, false, false);
      Copy2D(ctx, rem, kb, panel.p, rem, P, ld);
      double* A22 = M + (k0 + kb) + (int64_t)(k0 + kb) * ld;
      Gemm<false, false>(ctx, rem, rem, kb, -1.0, P, ld, P, ld, nullptr, 1.0, A22, ld, true, false);
    }
  }
  int h_flag = 0;
  CUDA_CHECK(cudaMemcpyAsync(&h_flag, flag.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  if (h_flag) throw Error(EPS_ERR_NUMERIC, "Cholesky: matrix is not positive definite");
  // W = inv(L) by block forward substitution
  CUDA_CHECK(cudaMemsetAsync(W, 0, (size_t)n * n * sizeof(double), ctx->stream));
  for (int b = 0; b < nblk; b++) {
    const int i0 = b * NB, ib = std::min(NB, n - i0);
    double* Lii_inv = linv.p + (size_t)b * NB * NB;
    Copy2D(ctx, ib, ib, Lii_inv, NB, W + i0 + (int64_t)i0 * n, n);
    if (i0 > 0) {
      // T = L[i, 0:i0] * W[0:i0, 0:i0]
      Gemm<false, true>(ctx, ib, i0, i0, 1.0, M + i0, ld, W, n, nullptr, 0.0, panel.p, ib, false, true);
      // W[i, 0:i0] = -inv(L_ii) * T
      Gemm<false, true>(ctx, ib, i0, ib, -1.0, Lii_inv, NB, panel.p, ib, nullptr, 0.0, W + i0, n, false, false);
    }
  }
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  CUDA_CHECK(cudaGetLastError());
}

}  // namespace eps
