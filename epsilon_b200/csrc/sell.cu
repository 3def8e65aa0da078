// Column-tiled sliced-ELL SpMV with the x tile staged in shared memory.
//
// Why: with plain CSR the gather of x goes through L1/L2 in 32-byte sectors for 8 useful
// bytes; ncu on the 1M x 100k lasso shows DRAM at 35 % while L1/L2 throughput sits at
// 82-88 % (profiles/r01_spmv_csr_baseline.txt).  Here the columns are cut into tiles of at
// most kMaxTile entries so that the matching slice of x (<= 200 KB) lives in shared memory
// (staged by one TMA bulk copy), the rows of each (tile, row-range) work item are sorted by
// length and packed 32 at a time column-major (SELL-32-sigma) so that every load of the
// (index, value) stream is a full coalesced line, and column indices shrink to 16 bits
// (10 bytes per non-zero instead of 12).  Each work item writes one partial sum per row
// and tile; a second kernel adds the tiles in a fixed order (deterministic, no atomics).
//
// Backs SparseMatrixImpl::Apply (src/epsilon/linear/sparse_matrix_impl.h:24) and its
// transpose for large matrices; results differ from the reference only by summation order.
#include <cub/cub.cuh>

#include "device.cuh"
#include "internal.h"
#include "peer.cuh"

namespace eps {

namespace {

constexpr int kMaxTile = 25600;      // doubles of x per tile (200 KB of shared memory)
constexpr int kSellThreads = 1024;

__global__ void SegBoundsKernel(int R, int T, int TW, const int32_t* __restrict__ rowptr,
                                const int32_t* __restrict__ col, int32_t* __restrict__ segstart,
                                int32_t* __restrict__ seglen) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= R) return;
  const int s = rowptr[row], e = rowptr[row + 1];
  int prev = s;
  for (int t = 0; t < T; t++) {
    // first position with col >= (t+1)*TW
    const int bound = (t + 1) * TW;
    int lo = prev, hi = e;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (col[mid] < bound) lo = mid + 1; else hi = mid;
    }
    segstart[(size_t)t * R + row] = prev;
    seglen[(size_t)t * R + row] = lo - prev;
    prev = lo;
  }
}

__global__ void SellKeysKernel(int R, int T, int RR, int rows_per, const int32_t* __restrict__ seglen,
                               uint32_t* __restrict__ keys, int32_t* __restrict__ vals) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)T * R) return;
  const int t = (int)(i / R), row = (int)(i % R);
  const uint32_t item = (uint32_t)(t * RR + row / rows_per);
  keys[i] = (item << 16) | (uint32_t)(0xFFFF - seglen[i]);
  vals[i] = (int32_t)i;
}

__global__ void SliceLenKernel(int R, int T, int RR, int rows_per, const uint32_t* __restrict__ skeys,
                               int32_t* __restrict__ sl_len, int64_t* __restrict__ vsize,
                               int64_t* __restrict__ isize) {
  const int spi = rows_per / 32;
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= T * RR * spi) return;
  const int item = s / spi, j = s % spi;
  const int t = item / RR, rr = item % RR;
  const int first = rr * rows_per + 32 * j;
  int len = 0;
  if (first < R && first < (rr + 1) * rows_per) len = 0xFFFF - (int)(skeys[(size_t)t * R + first] & 0xFFFF);
  sl_len[s] = len;
  vsize[s] = 32 * (int64_t)len;
  isize[s] = 32 * (int64_t)((len + 3) / 4 * 4);
}

// one warp per slice
__global__ void SellFillKernel(int R, int T, int RR, int rows_per, int TW, int n_slices,
                               const int32_t* __restrict__ svals, const int32_t* __restrict__ segstart,
                               const int32_t* __restrict__ seglen, const int32_t* __restrict__ ccol,
                               const double* __restrict__ cval, const int32_t* __restrict__ sl_len,
                               const int64_t* __restrict__ voff, const int64_t* __restrict__ ioff,
                               double* __restrict__ val, uint16_t* __restrict__ idx,
                               int32_t* __restrict__ rowid) {
  const int s = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (s >= n_slices) return;
  const int spi = rows_per / 32;
  const int item = s / spi, j = s % spi;
  const int t = item / RR, rr = item % RR;
  const int pos = rr * rows_per + 32 * j + lane;
  int row = -1, start = 0, mylen = 0;
  if (pos < R && pos < (rr + 1) * rows_per) {
    const int i = svals[(size_t)t * R + pos];
    row = i - t * R;
    start = segstart[i];
    mylen = seglen[i];
  }
  rowid[(size_t)s * 32 + lane] = row;
  const int len = sl_len[s];
  const int64_t vo = voff[s], io = ioff[s];
  const int base = t * TW;
  for (int k = 0; k < len; k++) {
    const bool in = k < mylen;
    val[vo + (int64_t)k * 32 + lane] = in ? cval[start + k] : 0.0;
    idx[io + (int64_t)(k >> 2) * 128 + lane * 4 + (k & 3)] = in ? (uint16_t)(ccol[start + k] - base) : (uint16_t)0;
  }
  for (int k = len; k < (len + 3) / 4 * 4; k++) idx[io + (int64_t)(k >> 2) * 128 + lane * 4 + (k & 3)] = 0;
}

__device__ __forceinline__ uint32_t SmemAddr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// P[tile][row] = sum over the row's non-zeros inside the tile.
// Persistent: CTA g owns the contiguous work items [cta_begin[g], cta_begin[g+1]) -- a static
// partition balanced by stored non-zeros at build time (deterministic, no atomics).  Items
// are ordered tile-major, so a CTA restages its x tile only when the tile changes.
__global__ void __launch_bounds__(kSellThreads, 1)
SellSpmvKernel(int R, int Cn, int TW, int RR, int spi, const int32_t* __restrict__ cta_begin,
               const int32_t* __restrict__ sl_len,
               const int64_t* __restrict__ voff, const int64_t* __restrict__ ioff,
               const double* __restrict__ val, const uint16_t* __restrict__ idx,
               const int32_t* __restrict__ rowid, const double* __restrict__ x,
               double* __restrict__ P, int use_tma, const int* __restrict__ skip) {
  extern __shared__ __align__(128) double xs[];
  __shared__ __align__(8) unsigned long long bar;
  if (skip && *skip) return;  // speculative launch after the CG loop converged (kkt.cu)
  const int item_lo = cta_begin[blockIdx.x], item_hi = cta_begin[blockIdx.x + 1];
  if (item_lo >= item_hi) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t bar_a = SmemAddr(&bar);
  if (use_tma) {
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
  }
  int cur_t = -1;
  uint32_t phase = 0;
  for (int item = item_lo; item < item_hi; item++) {
    const int t = item / RR;
    if (t != cur_t) {
      const int c0 = t * TW;
      const int cw = min(TW, Cn - c0);
      __syncthreads();  // every warp is done reading the previous tile (and sees the barrier init)
      if (use_tma) {
        // one TMA bulk copy of the x tile (cp.async.bulk, completion on an mbarrier)
        const uint32_t bytes = (uint32_t)cw * 8u;
        if (threadIdx.x == 0) {
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(bytes) : "memory");
          asm volatile(
              "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
              ::"r"(SmemAddr(xs)), "l"(x + c0), "r"(bytes), "r"(bar_a) : "memory");
        }
        uint32_t done = 0;
        while (!done) {
          asm volatile(
              "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
              : "=r"(done) : "r"(bar_a), "r"(phase) : "memory");
        }
        phase ^= 1u;
      } else {
        for (int i = threadIdx.x; i < cw; i += kSellThreads) xs[i] = x[c0 + i];
        __syncthreads();
      }
      cur_t = t;
    }
    const int s0 = item * spi;
    double* Pt = P + (size_t)t * R;
    for (int s = s0 + warp; s < s0 + spi; s += kSellThreads / 32) {
      const int len = sl_len[s];
      const int row = rowid[(size_t)s * 32 + lane];
      const double* v = val + voff[s] + lane;
      const uint16_t* ix = idx + ioff[s] + lane * 4;
      double acc0 = 0.0, acc1 = 0.0;
      int k = 0;
      for (; k + 8 <= len; k += 8) {
        const uint2 qa = *reinterpret_cast<const uint2*>(ix + (size_t)(k >> 2) * 128);
        const uint2 qb = *reinterpret_cast<const uint2*>(ix + (size_t)((k >> 2) + 1) * 128);
        const double v0 = LdStream(v + (size_t)(k + 0) * 32), v1 = LdStream(v + (size_t)(k + 1) * 32);
        const double v2 = LdStream(v + (size_t)(k + 2) * 32), v3 = LdStream(v + (size_t)(k + 3) * 32);
        const double v4 = LdStream(v + (size_t)(k + 4) * 32), v5 = LdStream(v + (size_t)(k + 5) * 32);
        const double v6 = LdStream(v + (size_t)(k + 6) * 32), v7 = LdStream(v + (size_t)(k + 7) * 32);
        acc0 += v0 * xs[qa.x & 0xFFFF];
        acc1 += v1 * xs[qa.x >> 16];
        acc0 += v2 * xs[qa.y & 0xFFFF];
        acc1 += v3 * xs[qa.y >> 16];
        acc0 += v4 * xs[qb.x & 0xFFFF];
        acc1 += v5 * xs[qb.x >> 16];
        acc0 += v6 * xs[qb.y & 0xFFFF];
        acc1 += v7 * xs[qb.y >> 16];
      }
      for (; k < len; k += 4) {
        const uint2 q = *reinterpret_cast<const uint2*>(ix + (size_t)(k >> 2) * 128);
        const int rem = len - k;
        acc0 += LdStream(v + (size_t)k * 32) * xs[q.x & 0xFFFF];
        if (rem > 1) acc1 += LdStream(v + (size_t)(k + 1) * 32) * xs[q.x >> 16];
        if (rem > 2) acc0 += LdStream(v + (size_t)(k + 2) * 32) * xs[q.y & 0xFFFF];
        if (rem > 3) acc1 += LdStream(v + (size_t)(k + 3) * 32) * xs[q.y >> 16];
      }
      if (row >= 0) Pt[row] = acc0 + acc1;
    }
  }
}

__global__ void __launch_bounds__(256)
SellReduceKernel(int R, int T, const double* __restrict__ P, double alpha, double beta,
                 double* __restrict__ y, const int* __restrict__ skip,
                 const double* __restrict__ ep_d, const double* __restrict__ ep_p) {
  if (skip && *skip) return;
  const int row = blockIdx.x * 256 + threadIdx.x;
  if (row >= R) return;
  double s = 0.0;
  for (int t = 0; t < T; t++) s += P[(size_t)t * R + row];
  double out = (beta == 0.0 ? 0.0 : beta * y[row]) + alpha * s;
  if (ep_d) out += ep_d[row] * ep_p[row];  // fused diagonal term of the KKT operator
  y[row] = out;
}

// Same reduction, but the result is this rank's contribution to a cross-GPU sum: it is
// pushed over NVLink into slot [parity][rank] of every rank's receive buffer (P2P stores),
// fused with the reduction so that the transfer overlaps it CTA by CTA.
__global__ void __launch_bounds__(256)
SellReducePublishKernel(int R, int T, const double* __restrict__ P, double alpha, PeerXchg px,
                        unsigned long long seq, unsigned* counter) {
  const int row = blockIdx.x * 256 + threadIdx.x;
  if (row < R) {
    double s = 0.0;
    for (int t = 0; t < T; t++) s += P[(size_t)t * R + row];
    const double out = alpha * s;
    const size_t slot = ((size_t)(seq & 1ull) * px.G + px.rank) * px.n + row;
    for (int k = 0; k < px.G; k++) {
      const int r = (px.rank + k) % px.G;  // start with the local copy, spread the peers
      px.recv[r][slot] = out;
    }
  }
  PeerPublishFinish(px, seq, counter);
}

}  // namespace

void BuildSell(Ctx* ctx, const Csr& A, Sell* S) {
  const int R = A.m, Cn = A.n;
  S->R = R; S->Cn = Cn; S->nnz = A.nnz;
  const int T = (Cn + kMaxTile - 1) / kMaxTile;
  int TW = (Cn + T - 1) / T;
  TW = (TW + 1) / 2 * 2;  // keep tile starts 16-byte aligned for the bulk copy
  S->T = T; S->TW = TW;
  // about 8 work items per CTA so that the static partition below balances to ~1/8 of an
  // item, but at least 1024 rows (one 32-row slice per warp) per item
  const int64_t want_items = (int64_t)ctx->sm_count * 8;
  int rows_per = (int)(((int64_t)R * T + want_items - 1) / want_items);
  rows_per = std::max(1024, rows_per);
  rows_per = std::min(rows_per, std::max(32, R));
  rows_per = (rows_per + 31) / 32 * 32;
  int RR = (R + rows_per - 1) / rows_per;
  S->RR = RR; S->rows_per = rows_per;
  const int spi = rows_per / 32;
  const int n_slices = T * RR * spi;
  S->n_slices = n_slices;
  const size_t TR = (size_t)T * R;
  EPS_CHECK(T * RR < 65536, "too many SELL work items (%d)", T * RR);
  DevBuf<int32_t> segstart(ctx, TR), seglen(ctx, TR), vals_in(ctx, TR), vals_out(ctx, TR);
  DevBuf<uint32_t> keys_in(ctx, TR), keys_out(ctx, TR);
  SegBoundsKernel<<<(R + 255) / 256, 256, 0, ctx->stream>>>(R, T, TW, A.rowptr.p, A.col.p, segstart.p, seglen.p);
  SellKeysKernel<<<(unsigned)((TR + 255) / 256), 256, 0, ctx->stream>>>(R, T, RR, rows_per, seglen.p, keys_in.p, vals_in.p);
  size_t tmp_bytes = 0;
  CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_in.p, keys_out.p, vals_in.p, vals_out.p,
                                             (int)TR, 0, 32, ctx->stream));
  DevBuf<char> tmp(ctx, tmp_bytes);
  CUDA_CHECK(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys_in.p, keys_out.p, vals_in.p, vals_out.p,
                                             (int)TR, 0, 32, ctx->stream));
  S->sl_len.alloc(ctx, n_slices);
  S->voff.alloc(ctx, n_slices + 1);
  S->ioff.alloc(ctx, n_slices + 1);
  DevBuf<int64_t> vsize(ctx, n_slices + 1), isize(ctx, n_slices + 1);
  vsize.zero();
  isize.zero();
  SliceLenKernel<<<(n_slices + 255) / 256, 256, 0, ctx->stream>>>(R, T, RR, rows_per, keys_out.p, S->sl_len.p,
                                                                   vsize.p, isize.p);
  size_t scan_bytes = 0;
  CUDA_CHECK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, vsize.p, S->voff.p, n_slices + 1, ctx->stream));
  DevBuf<char> tmp2(ctx, scan_bytes);
  CUDA_CHECK(cub::DeviceScan::ExclusiveSum(tmp2.p, scan_bytes, vsize.p, S->voff.p, n_slices + 1, ctx->stream));
  CUDA_CHECK(cub::DeviceScan::ExclusiveSum(tmp2.p, scan_bytes, isize.p, S->ioff.p, n_slices + 1, ctx->stream));
  int64_t totals[2];
  CUDA_CHECK(cudaMemcpyAsync(&totals[0], S->voff.p + n_slices, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_CHECK(cudaMemcpyAsync(&totals[1], S->ioff.p + n_slices, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  S->n_val = totals[0];
  S->n_idx = totals[1];
  S->val.alloc(ctx, std::max<int64_t>(S->n_val, 1));
  S->idx.alloc(ctx, std::max<int64_t>(S->n_idx, 4));
  S->rowid.alloc(ctx, (size_t)n_slices * 32);
  {
    // static partition of the tile-major item list over the CTAs, balanced by stored bytes
    // (10 per non-zero) plus the x-tile staging a CTA pays when it enters a new tile
    const int items = T * RR;
    std::vector<int64_t> hv(n_slices + 1);
    CUDA_CHECK(cudaMemcpyAsync(hv.data(), S->voff.p, (n_slices + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    std::vector<double> cost(items + 1, 0.0);
    for (int i = 0; i < items; i++) {
      double c = 10.0 * (double)(hv[(size_t)(i + 1) * spi] - hv[(size_t)i * spi]) + 12.0 * rows_per;
      if (i % RR == 0) c += 8.0 * TW;
      cost[i + 1] = cost[i] + c;
    }
    const int G = std::max(1, std::min(items, ctx->sm_count));
    std::vector<int32_t> begin(G + 1, items);
    begin[0] = 0;
    int it = 0;
    for (int g = 1; g < G; g++) {
      const double target = cost[items] * g / G;
      while (it < items && cost[it + 1] <= target) it++;
      // closer boundary of the two around the target
      int cut = it;
      if (it < items && (target - cost[it]) > (cost[it + 1] - target)) cut = it + 1;
      begin[g] = std::max(cut, begin[g - 1]);
    }
    begin[G] = items;
    S->n_cta = G;
    S->cta_begin.alloc(ctx, G + 1);
    S->cta_begin.upload(begin.data(), G + 1);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }
  SellFillKernel<<<(n_slices + 7) / 8, 256, 0, ctx->stream>>>(
      R, T, RR, rows_per, TW, n_slices, vals_out.p, segstart.p, seglen.p, A.col.p, A.val.p, S->sl_len.p,
      S->voff.p, S->ioff.p, S->val.p, S->idx.p, S->rowid.p);
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  CUDA_CHECK(cudaGetLastError());
  static bool attr_set = false;
  if (!attr_set) {
    CUDA_CHECK(cudaFuncSetAttribute(SellSpmvKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    kMaxTile * 8 + 64));
    attr_set = true;
  }
  S->built = true;
}

void SellApply(Ctx* ctx, const Sell& S, double alpha, const double* x, double beta, double* y, int kclass,
               const SpmvExtra* ex) {
  const int* skip = ex ? ex->skip : nullptr;
  if (S.R == 0) return;
  double* P = GrowScratch(ctx, &ctx->sell_scratch, (size_t)S.T * S.R);
  {
    ProfScope ps(ctx, kclass, 12.0 * S.nnz + 4.0 * (S.R + 1) + 8.0 * S.Cn + 8.0 * S.R);
    // bulk copies need 16-byte aligned addresses and sizes: TW is even, so only the base
    // pointer and the width of the last tile can break that
    const int last = S.Cn - (S.T - 1) * S.TW;
    const int use_tma = ((reinterpret_cast<uintptr_t>(x) & 15) == 0 && last % 2 == 0) ? 1 : 0;
    SellSpmvKernel<<<S.n_cta, kSellThreads, (size_t)S.TW * 8 + 64, ctx->stream>>>(
        S.R, S.Cn, S.TW, S.RR, S.rows_per / 32, S.cta_begin.p, S.sl_len.p, S.voff.p, S.ioff.p, S.val.p, S.idx.p,
        S.rowid.p, x, P, use_tma, skip);
    ctx->count_launch();
  }
  if (ex && ex->publish) {
    EPS_CHECK(beta == 0.0 && ex->publish->n >= S.R, "publishing apply: beta must be 0 and the vector must fit the slots");
    ProfScope ps(ctx, KC_ALLREDUCE, 8.0 * S.T * S.R + 8.0 * S.R * ex->publish->G);
    SellReducePublishKernel<<<(S.R + 255) / 256, 256, 0, ctx->stream>>>(S.R, S.T, P, alpha, *ex->publish,
                                                                         ex->publish_seq, ex->publish_counter);
    ctx->count_launch();
    return;
  }
  ProfScope ps(ctx, KC_VEC, 8.0 * S.T * S.R + 8.0 * S.R);
  SellReduceKernel<<<(S.R + 255) / 256, 256, 0, ctx->stream>>>(S.R, S.T, P, alpha, beta, y, skip,
                                                                ex ? ex->ep_d : nullptr, ex ? ex->ep_p : nullptr);
  ctx->count_launch();
}

}  // namespace eps
