// Composite block operator on flat device slabs: the device analogue of the reference's
// BlockMatrix (src/epsilon/vector/block_matrix.{h,cc}): y = A*x (block_matrix.cc:155-168),
// Transpose (:58-67).  String keys become fixed row/column offsets decided at Init; a
// missing block is an implicit zero exactly as in the reference.
#include <algorithm>

#include "device.cuh"
#include "internal.h"

namespace eps {

namespace {

struct Triplet { int r, c; double v; };

void BuildCsrFromTriplets(Ctx* ctx, int m, int n, std::vector<Triplet>& t, Csr* out) {
  std::vector<int64_t> rowptr(m + 1, 0);
  for (auto& e : t) rowptr[e.r + 1]++;
  for (int i = 0; i < m; i++) rowptr[i + 1] += rowptr[i];
  std::vector<int64_t> cursor(rowptr.begin(), rowptr.end() - 1);
  std::vector<int32_t> col(t.size());
  std::vector<double> val(t.size());
  // stable fill keeps block order inside a row; then order by column for determinism
  for (auto& e : t) {
    int64_t p = cursor[e.r]++;
    col[p] = e.c;
    val[p] = e.v;
  }
  for (int i = 0; i < m; i++) {
    int64_t a = rowptr[i], b = rowptr[i + 1];
    if (b - a > 1) {
      std::vector<std::pair<int32_t, double>> tmp(b - a);
      for (int64_t k = a; k < b; k++) tmp[k - a] = {col[k], val[k]};
      std::stable_sort(tmp.begin(), tmp.end(),
                       [](const auto& x, const auto& y) { return x.first < y.first; });
      for (int64_t k = a; k < b; k++) { col[k] = tmp[k - a].first; val[k] = tmp[k - a].second; }
    }
  }
  BuildCsrFromHost(ctx, m, n, (int64_t)t.size(), rowptr.data(), col.data(), val.data(), out);
}

void BlockTriplets(const Blk& b, std::vector<Triplet>* out) {
  const MapData& M = *b.map;
  if (M.kind == MK_SCALAR) {
    for (int i = 0; i < M.n; i++) out->push_back({b.row_off + i, b.col_off + i, b.scale * M.alpha});
  } else if (M.kind == MK_DIAGONAL) {
    for (int i = 0; i < M.n; i++) out->push_back({b.row_off + i, b.col_off + i, b.scale * M.h_diag[i]});
  } else {
    for (int j = 0; j < M.n; j++)
      for (int k = M.h_colptr[j]; k < M.h_colptr[j + 1]; k++) {
        int r = M.h_rowidx[k], c = j;
        if (b.trans) std::swap(r, c);
        out->push_back({b.row_off + r, b.col_off + c, b.scale * M.h_vals[k]});
      }
  }
}

__global__ void TransposeKernel(int m, int n, const double* __restrict__ in, double* __restrict__ out) {
  // in: m x n column-major -> out: n x m column-major
  __shared__ double tile[32][33];
  int bx = blockIdx.x * 32, by = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int r = bx + threadIdx.x, c = by + j;
    if (r < m && c < n) tile[j][threadIdx.x] = in[(int64_t)c * m + r];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int c = by + threadIdx.x, r = bx + j;
    if (r < m && c < n) out[(int64_t)r * n + c] = tile[threadIdx.x][j];
  }
}

}  // namespace

void BlockOp::Compile() {
  compiled = true;
  heavy.clear();
  has_merged = false;
  simple_scalar = false;
  is_zero = blks.empty();
  if (is_zero) return;
  for (auto& b : blks) {
    EPS_CHECK(b.row_off >= 0 && b.row_off + b.m() <= rows && b.col_off >= 0 &&
                  b.col_off + b.n() <= cols,
              "block (%d,%d) %dx%d outside %lldx%lld", b.row_off, b.col_off, b.m(), b.n(),
              (long long)rows, (long long)cols);
  }
  if (blks.size() == 1 && blks[0].map->kind == MK_SCALAR && blks[0].col_off == 0 &&
      blks[0].n() == cols) {
    simple_scalar = true;
    simple_alpha = blks[0].scale * blks[0].map->alpha;
    simple_row_off = blks[0].row_off;
  }
  std::vector<const Blk*> light;
  for (auto& b : blks) {
    bool l = b.diag_like() || (b.map->kind == MK_SPARSE && b.map->has_host);
    if (l) light.push_back(&b); else heavy.push_back(b);
  }
  if (light.size() >= 2) {
    std::vector<Triplet> t, tt;
    for (auto* b : light) BlockTriplets(*b, &t);
    tt.reserve(t.size());
    for (auto& e : t) tt.push_back({e.c, e.r, e.v});
    BuildCsrFromTriplets(ctx, (int)rows, (int)cols, t, &merged);
    BuildCsrFromTriplets(ctx, (int)cols, (int)rows, tt, &merged_t);
    has_merged = true;
  } else {
    for (auto* b : light) heavy.push_back(*b);
  }
}

int64_t BlockOp::heavy_nnz() const {
  int64_t t = 0;
  for (auto& b : blks) {
    if (b.map->kind == MK_DENSE) t += (int64_t)b.map->m * b.map->n;
    else if (b.map->kind == MK_SPARSE) t += b.map->csr.nnz;
  }
  return t;
}

void KroneckerApply(Ctx* ctx, const MapData& K, bool trans, const double* x, double* y);

static void ApplyBlock(Ctx* ctx, const Blk& b, bool transpose, double beta, const double* in,
                       double* out) {
  const bool t = b.trans != transpose;
  if (b.map->kind == MK_KRONECKER) {
    EPS_CHECK(beta == 0.0, "accumulating Kronecker block");
    KroneckerApply(ctx, *b.map, t, in, out);
    return;
  }
  MapApply(ctx, *b.map, t, b.scale, in, beta, out);
}

void BlockOp::Apply(const double* x, double* y, bool accumulate) const {
  EPS_CHECK(compiled, "BlockOp not compiled");
  if (rows == 0) return;
  bool direct = !accumulate && !has_merged && heavy.size() == 1 && heavy[0].row_off == 0 &&
                heavy[0].m() == rows;
  if (!accumulate && !direct && !(has_merged && heavy.empty()))
    CUDA_CHECK(cudaMemsetAsync(y, 0, rows * sizeof(double), ctx->stream));
  if (has_merged) {
    double beta = (accumulate || !heavy.empty()) ? 1.0 : 0.0;
    CsrApply(ctx, merged, 1.0, x, beta, y, KC_SPMV_N);
  }
  for (auto& b : heavy) ApplyBlock(ctx, b, false, direct ? 0.0 : 1.0, x + b.col_off, y + b.row_off);
}

void BlockOp::ApplyT(const double* y, double* x, bool accumulate) const {
  EPS_CHECK(compiled, "BlockOp not compiled");
  if (cols == 0) return;
  bool direct = !accumulate && !has_merged && heavy.size() == 1 && heavy[0].col_off == 0 &&
                heavy[0].n() == cols;
  if (!accumulate && !direct && !(has_merged && heavy.empty()))
    CUDA_CHECK(cudaMemsetAsync(x, 0, cols * sizeof(double), ctx->stream));
  if (has_merged) {
    double beta = (accumulate || !heavy.empty()) ? 1.0 : 0.0;
    CsrApply(ctx, merged_t, 1.0, y, beta, x, KC_SPMV_T);
  }
  for (auto& b : heavy) ApplyBlock(ctx, b, true, direct ? 0.0 : 1.0, y + b.row_off, x + b.col_off);
}

bool BlockOp::GramDiagonal(std::vector<double>* d) const {
  d->assign(cols, 0.0);
  // sum duplicates at the same (row_off, col_off) first (BlockMatrix::InsertOrAdd)
  std::map<std::pair<int, int>, std::vector<double>> acc;
  for (auto& b : blks) {
    if (!b.diag_like()) return false;
    auto& v = acc[{b.row_off, b.col_off}];
    if (v.empty()) v.assign(b.n(), 0.0);
    if ((int)v.size() != b.n()) return false;
    for (int i = 0; i < b.n(); i++)
      v[i] += b.scale * (b.map->kind == MK_SCALAR ? b.map->alpha : b.map->h_diag[i]);
  }
  // off-diagonal terms appear iff two distinct blocks share rows
  std::vector<std::pair<int, int>> ranges;
  for (auto& kv : acc) ranges.push_back({kv.first.first, kv.first.first + (int)kv.second.size()});
  std::sort(ranges.begin(), ranges.end());
  for (size_t i = 1; i < ranges.size(); i++)
    if (ranges[i].first < ranges[i - 1].second) return false;
  for (auto& kv : acc)
    for (size_t i = 0; i < kv.second.size(); i++) (*d)[kv.first.second + i] += kv.second[i] * kv.second[i];
  return true;
}

// (A (x) B) x = vec(B X A'), X = reshape(x, [B.n, A.n]) column-major
// (reference: KroneckerProductImpl::Apply, kronecker_product_impl.cc:45-58).
void KroneckerApply(Ctx* ctx, const MapData& K, bool trans, const double* x, double* y) {
  const MapData& A = *K.ka;
  const MapData& B = *K.kb;
  const bool at = K.ka_trans != trans, bt = K.kb_trans != trans;
  EPS_CHECK(A.kind != MK_KRONECKER && B.kind != MK_KRONECKER, "nested Kronecker products");
  const int am = at ? A.n : A.m, an = at ? A.m : A.n;
  const int bm = bt ? B.n : B.m, bn = bt ? B.m : B.n;
  if (A.kind == MK_SCALAR) {
    // (s I_k) (x) B: column c of the result is s * B x_c -- the matrix-variable product A*X of
    // the compiler (transforms/linear.py); no transposes, no temporaries
    for (int c = 0; c < an; c++)
      MapApply(ctx, B, bt, A.alpha, x + (size_t)c * bn, 0.0, y + (size_t)c * bm);
    return;
  }
  DevBuf<double> T(ctx, (size_t)bm * an), Tt(ctx, (size_t)an * bm), Yt(ctx, (size_t)am * bm);
  for (int c = 0; c < an; c++) MapApply(ctx, B, bt, 1.0, x + (size_t)c * bn, 0.0, T.p + (size_t)c * bm);
  dim3 tb(32, 8);
  TransposeKernel<<<dim3((bm + 31) / 32, (an + 31) / 32), tb, 0, ctx->stream>>>(bm, an, T.p, Tt.p);
  ctx->count_launch();
  for (int i = 0; i < bm; i++) MapApply(ctx, A, at, 1.0, Tt.p + (size_t)i * an, 0.0, Yt.p + (size_t)i * am);
  TransposeKernel<<<dim3((am + 31) / 32, (bm + 31) / 32), tb, 0, ctx->stream>>>(am, bm, Yt.p, y);
  ctx->count_launch();
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
}

}  // namespace eps
