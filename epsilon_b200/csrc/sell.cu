// Column-tiled sliced-ELL SpMV in stream order, with the x tile staged in shared memory.
//
// Why: with plain CSR the gather of x goes through L1/L2 in 32-byte sectors for 8 useful
// bytes; ncu on the 1M x 100k lasso shows DRAM at 35 % while L1/L2 throughput sits at
// 82-88 %.  Here the columns are cut into tiles of at most kMaxTile entries so that the
// matching slice of x (<= 200 KB) lives in shared memory (staged by one TMA bulk copy), rows
// are packed 32 at a time column-major (SELL-32-sigma) so that every load of the (index,
// value) stream is a full coalesced line, and column indices shrink to 16 bits (10 bytes per
// non-zero instead of 12).  A visit writes one partial sum per row and tile into the planes
// P[tile][row]; a second kernel (or the fused CG update, kkt.cu) adds the tiles in a fixed
// order (deterministic, no atomics).
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

__device__ __forceinline__ uint32_t SmemAddr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__global__ void __launch_bounds__(256)
SellReduceKernel(int R, int T, const double* __restrict__ P, double alpha, double beta,
                 double* __restrict__ y, const int* __restrict__ skip,
                 const double* __restrict__ ep_d, const double* __restrict__ ep_p) {
  GridDepLaunch();  // the next SpMV pass may start prefetching its matrix streams
  GridDepWait();    // launched with programmatic serialization behind the SpMV pass that writes P
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


// ===== stream layout ========================================================================
// Work is cut at (tile, row) granularity: the tile-major list of (tile, row) pairs is split
// into n_cta contiguous ranges of equal cost; the part of a range inside one tile is a
// "visit".  Inside a visit the rows are sorted by length, packed into 32-row slices, and the
// slices are dealt to the CTA's warps in serpentine order; the slices of one warp are stored
// back to back, so a warp reads ONE contiguous stream of values and of indices per visit and
// can keep two batches of loads in flight across slice boundaries (an earlier kernel that
// walked (tile, row-range) items slice by slice drained its loads at the end of every
// 25-entry slice: long-scoreboard stalls, 77 % of the HBM peak,
// profiles/r01_ncu_full_SellSpmvKernel_v1_before.csv).


struct VisitTab {
  const int32_t* lin0;        // [n_visits+1] first (tile,row) pair of the visit, tile-major
  const int32_t* slice_base;  // [n_visits+1] first slice of the visit (visit-major numbering)
  int n_visits;
};

__host__ __device__ __forceinline__ int WarpPrefix(int n_sl, int w, int NW) {
  // slices that warps < w receive when n_sl slices are dealt to NW warps in serpentine order
  const int full = n_sl / NW, rem = n_sl % NW;
  const int extra = (full & 1) == 0 ? (w < rem ? w : rem) : (w - (NW - rem) > 0 ? w - (NW - rem) : 0);
  return w * full + extra;
}

// slice j of a visit with n_sl slices -> position in stream order, relative to the visit
__host__ __device__ __forceinline__ int StreamPos(int n_sl, int j, int NW) {
  const int round = j / NW, pos = j % NW;
  const int w = (round & 1) == 0 ? pos : NW - 1 - pos;
  return WarpPrefix(n_sl, w, NW) + round;
}

__device__ __forceinline__ int UpperBound(const int32_t* a, int n, int key) {
  // first index with a[i] > key
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (a[mid] <= key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// Streamed bytes of one (tile,row) pair; the x-tile staging a CTA pays when it enters a tile
// is charged to the first row of the tile (a CTA that starts inside a tile pays it as well, so
// every CTA carries one staging more than its share shows: equal for all, it cancels).
struct CostOp {
  const int32_t* seglen;
  int64_t R, stage;
  __host__ __device__ int64_t operator()(int64_t i) const {
    return 10 * (int64_t)seglen[i] + 12 + (i % R == 0 ? stage : 0);
  }
};

__global__ void CutKernel(int64_t TR, int G, const int64_t* __restrict__ csum, int32_t* __restrict__ cut) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g > G) return;
  if (g == 0) { cut[0] = 0; return; }
  if (g == G) { cut[G] = (int32_t)TR; return; }
  const int64_t total = csum[TR - 1];
  const int64_t target = (int64_t)((double)total * g / G);
  int64_t lo = 0, hi = TR;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (csum[mid] < target) lo = mid + 1; else hi = mid;
  }
  cut[g] = (int32_t)lo;
}

// Cut positions for explicit cost targets (no-span assignment of CTAs to tiles): cut[g] = first index whose
// running cost reaches target[g], or forced[g] when that is >= 0 (tile boundaries).
__global__ void CutTargetsKernel(int64_t TR, int G, const int64_t* __restrict__ csum, const int64_t* __restrict__ target,
                                 const int64_t* __restrict__ forced, int32_t* __restrict__ cut) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g > G) return;
  if (forced[g] >= 0) { cut[g] = (int32_t)forced[g]; return; }
  const int64_t t = target[g];
  int64_t lo = 0, hi = TR;
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (csum[mid] < t) lo = mid + 1; else hi = mid;
  }
  cut[g] = (int32_t)lo;
}

// Rows are sorted by length inside windows of kSortWindow consecutive rows of a visit, not
// over the whole visit: a warp's slices then walk the visit's rows window by window, so the
// 8-byte partial sums a CTA scatters into P fall into a few 32 KB windows at a time and complete
// their 32-byte sectors in L2 before eviction.  (Sorted over a 108k-row visit -- logreg 2M x 200k,
// 128 MB of partial planes -- the half-written sectors were evicted by the matrix stream:
// 73 % of the HBM peak instead of 87 %.)
constexpr int kSortWindow = 4096;

__global__ void VisitKeysKernel(int64_t TR, VisitTab vt, const int32_t* __restrict__ seglen,
                                uint64_t* __restrict__ keys, int32_t* __restrict__ vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= TR) return;
  const int v = UpperBound(vt.lin0, vt.n_visits, (int)i) - 1;
  const uint64_t window = (uint64_t)(((int)i - vt.lin0[v]) / kSortWindow);
  keys[i] = ((uint64_t)v << 40) | (window << 16) | (uint64_t)(0xFFFF - seglen[i]);
  vals[i] = (int32_t)i;
}

__global__ void StreamSliceLenKernel(int n_slices, int NW, VisitTab vt, const uint64_t* __restrict__ skeys,
                                     int32_t* __restrict__ sl_len, int64_t* __restrict__ vsize) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slices) return;
  const int v = UpperBound(vt.slice_base, vt.n_visits, s) - 1;
  const int j = s - vt.slice_base[v];
  const int n_sl = vt.slice_base[v + 1] - vt.slice_base[v];
  const int len = 0xFFFF - (int)(skeys[vt.lin0[v] + 32 * j] & 0xFFFFull);
  const int ss = vt.slice_base[v] + StreamPos(n_sl, j, NW);
  sl_len[ss] = len;
  vsize[ss] = 32 * (int64_t)len;
}

// one warp per slice
__global__ void StreamFillKernel(int R, int TW, int n_slices, int NW, VisitTab vt, const int32_t* __restrict__ visit_tile,
                                 const SellStream* __restrict__ streams, const int32_t* __restrict__ svals,
                                 const int32_t* __restrict__ segstart, const int32_t* __restrict__ seglen,
                                 const int32_t* __restrict__ ccol, const double* __restrict__ cval,
                                 const int32_t* __restrict__ sl_len, const int64_t* __restrict__ voff,
                                 double* __restrict__ val, uint16_t* __restrict__ idx,
                                 int32_t* __restrict__ rowid) {
  const int s = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (s >= n_slices) return;
  const int v = UpperBound(vt.slice_base, vt.n_visits, s) - 1;
  const int j = s - vt.slice_base[v];
  const int n_sl = vt.slice_base[v + 1] - vt.slice_base[v];
  const int round = j / NW, pos5 = j % NW;
  const int w = (round & 1) == 0 ? pos5 : NW - 1 - pos5;
  const int ss = vt.slice_base[v] + WarpPrefix(n_sl, w, NW) + round;
  const int t = visit_tile[v];
  const int pos = vt.lin0[v] + 32 * j + lane;
  int row = -1, start = 0, mylen = 0;
  if (pos < vt.lin0[v + 1]) {
    const int i = svals[pos];
    row = i - t * R;
    start = segstart[i];
    mylen = seglen[i];
  }
  rowid[(size_t)ss * 32 + lane] = row < 0 ? -1 : t * R + row;
  const int len = sl_len[ss];
  const SellStream sd = streams[v * NW + w];
  const int64_t vo = voff[ss];
  const int q0 = (int)((vo - sd.voff) >> 5);
  const int base = t * TW;
  for (int k = 0; k < len; k++) {
    const bool in = k < mylen;
    const int q = q0 + k;
    val[vo + (int64_t)k * 32 + lane] = in ? cval[start + k] : 0.0;
    idx[sd.ioff + (int64_t)(q >> 2) * 128 + lane * 4 + (q & 3)] = in ? (uint16_t)(ccol[start + k] - base) : (uint16_t)0;
  }
}

// The x tile as several bulk copies in flight at once (all completing on the same mbarrier): one 200 KB request
// is walked by the copy engine front to back, several smaller ones overlap their L2 round trips.
// (experiment knob EPS_SELL_TMA_CHUNK: bytes per request, 0 = one request)
__device__ __forceinline__ void StageTile(uint32_t dst, const double* src, uint32_t bytes, uint32_t bar_a, uint32_t chunk) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(bytes) : "memory");
  if (chunk == 0 || chunk >= bytes) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar_a) : "memory");
    return;
  }
  for (uint32_t off = 0; off < bytes; off += chunk) {
    const uint32_t len = min(chunk, bytes - off);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst + off), "l"(reinterpret_cast<const char*>(src) + off), "r"(len), "r"(bar_a) : "memory");
  }
}

__device__ __forceinline__ unsigned long long GlobalTimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ uint2 LdStreamU2(const uint16_t* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}

// KB 32-wide value rows per batch of loads
template <int KB>
struct SellBatch {
  double v[KB];
  uint2 ix[KB / 4];
};

template <bool GUARD, int KB>
__device__ __forceinline__ void LoadBatch(SellBatch<KB>& b, const double* v, const uint16_t* ix, int q0, int Q) {
  constexpr int kBatch = KB;
#pragma unroll
  for (int j = 0; j < kBatch / 4; j++) {
    if (!GUARD || q0 + 4 * j < Q) b.ix[j] = LdStreamU2(ix + (size_t)((q0 >> 2) + j) * 128);
    else b.ix[j] = make_uint2(0u, 0u);
  }
#pragma unroll
  for (int j = 0; j < kBatch; j++) {
    if (!GUARD || q0 + j < Q) b.v[j] = LdStream(v + (size_t)(q0 + j) * 32);
    else b.v[j] = 0.0;
  }
}

// Per-warp cursor over the slices of a stream: which row the running sum belongs to.
// rowid holds the index into P (tile * R + row), -1 for the padding lanes of a visit's last slice.
struct SliceCursor {
  const int32_t* sl_len;
  const int32_t* rowid;
  double* P;
  int cur, end;          // current slice and one past the last one of the stream
  int rem;               // value rows left in the current slice
  int row_cur, len_next, row_next;
  double acc;
  __device__ __forceinline__ void Prefetch() {
    if (cur + 1 < end) {
      len_next = sl_len[cur + 1];
      row_next = rowid[(size_t)(cur + 1) * 32 + (threadIdx.x & 31)];
    } else {
      len_next = 0x7FFFFFFF;
      row_next = -1;
    }
  }
  __device__ __forceinline__ void Start(int begin, int cnt) {
    cur = begin; end = begin + cnt; acc = 0.0;
    rem = sl_len[begin];
    row_cur = rowid[(size_t)begin * 32 + (threadIdx.x & 31)];
    Prefetch();
  }
  __device__ __forceinline__ void Next() {
    if (row_cur >= 0) P[row_cur] = acc;
    acc = 0.0;
    cur++;
    rem = len_next; row_cur = row_next;
    Prefetch();
  }
  __device__ __forceinline__ void SkipEmpty() {
    while (cur < end && rem == 0) Next();
  }
  __device__ __forceinline__ void Add(double v, double x) {
    acc += v * x;
    if (--rem == 0) { Next(); SkipEmpty(); }
  }
};

template <bool GUARD, int KB>
__device__ __forceinline__ void ConsumeBatch(const SellBatch<KB>& b, const double* xs, SliceCursor& c, int q0, int Q) {
  constexpr int kBatch = KB;
#pragma unroll
  for (int j = 0; j < kBatch; j++) {
    const uint32_t word = (j & 2) ? b.ix[j >> 2].y : b.ix[j >> 2].x;
    const uint32_t col = (j & 1) ? (word >> 16) : (word & 0xFFFF);
    if (!GUARD || q0 + j < Q) c.Add(b.v[j], xs[col]);
  }
}

template <int NW, int KB, bool TRACE>
__global__ void __launch_bounds__(NW * 32, 1)
SellStreamKernel(int R, int Cn, int TW, const int32_t* __restrict__ cta_visit,
                 const int32_t* __restrict__ visit_tile, const SellStream* __restrict__ streams,
                 const int32_t* __restrict__ sl_len, const double* __restrict__ val,
                 const uint16_t* __restrict__ idx, const int32_t* __restrict__ rowid,
                 const double* __restrict__ x, double* __restrict__ P, int use_tma,
                 const int* __restrict__ skip, unsigned long long* __restrict__ trace,
                 const double* __restrict__ preP, int preT, double preAlpha, double* __restrict__ preY,
                 unsigned* __restrict__ gbar) {
  // Optional pre-phase (preP != nullptr): x itself is not in memory yet -- it is alpha * the sum of the preT
  // partial planes preP[t][0..Cn) the PREVIOUS pass of the CG matvec left behind.  Every CTA adds a fixed
  // slice of the planes into preY (== x), the grid meets at a barrier, and only then are the x tiles staged.
  // This replaces the separate plane-reduction launch between the two passes (SellReduceKernel).
  constexpr int kBatch = KB;
  extern __shared__ __align__(128) double xs[];
  __shared__ __align__(8) unsigned long long bar;
  // Launched with programmatic stream serialization (SellLaunch): everything up to GridDepWait() only
  // touches the matrix streams, which the preceding kernel does not write, and may run while that
  // kernel drains; x, P and the skip flag are read / written after it.
  if (TRACE && threadIdx.x == 0) trace[blockIdx.x * 4 + 0] = GlobalTimer();
  GridDepLaunch();  // the plane reduction behind this pass is set up while the streams drain
  const int v_lo = cta_visit[blockIdx.x], v_hi = cta_visit[blockIdx.x + 1];
  auto pre_phase = [&]() {
    int per = (Cn + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo = min(Cn, (int)blockIdx.x * per), hi = min(Cn, lo + per);
    // four rows per thread and round: the 4 x preT plane loads are independent, so a thread keeps
    // 16 of them in flight (one row at a time this phase was latency-bound: 24 us for 1M rows)
    constexpr int kStep = NW * 32;
    for (int i = lo + (int)threadIdx.x; i < hi; i += 4 * kStep) {
      const int i1 = i + kStep, i2 = i + 2 * kStep, i3 = i + 3 * kStep;
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll 4
      for (int t = 0; t < preT; t++) {
        const double* row = preP + (size_t)t * Cn;
        const double a0 = __ldcg(row + i);
        const double a1 = i1 < hi ? __ldcg(row + i1) : 0.0;
        const double a2 = i2 < hi ? __ldcg(row + i2) : 0.0;
        const double a3 = i3 < hi ? __ldcg(row + i3) : 0.0;
        s0 += a0; s1 += a1; s2 += a2; s3 += a3;
      }
      preY[i] = preAlpha * s0;
      if (i1 < hi) preY[i1] = preAlpha * s1;
      if (i2 < hi) preY[i2] = preAlpha * s2;
      if (i3 < hi) preY[i3] = preAlpha * s3;
    }
    // the tiles of preY are read by bulk copies (async proxy) on other SMs: order the generic-proxy stores
    // before them on both sides of the barrier
    asm volatile("fence.proxy.async.global;" ::: "memory");
    GridBarrier(gbar);
    asm volatile("fence.proxy.async.global;" ::: "memory");
  };
  if (v_lo >= v_hi) {
    // a CTA without visits still owes its slice of the pre-phase and its arrival at the barrier
    if (preP) {
      GridDepWait();
      if (!(skip && *skip)) pre_phase();
    }
    return;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t bar_a = SmemAddr(&bar);
  if (use_tma && threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  int cur_t = -1;
  uint32_t phase = 0;
  for (int vi = v_lo; vi < v_hi; vi++) {
    const int t = visit_tile[vi];
    const bool restage = t != cur_t;
    const int c0 = t * TW;
    const int cw = min(TW, Cn - c0);
    if (restage) {
      // every warp is done reading the previous tile; nothing to wait for on the first visit
      if (vi != v_lo) __syncthreads();
      if (vi != v_lo && use_tma && threadIdx.x == 0) {
        // one TMA bulk copy of the x tile (cp.async.bulk, completion on an mbarrier), issued
        // before anything else so that it overlaps the first loads of the streams
        StageTile(SmemAddr(xs), x + c0, (uint32_t)cw * 8u, bar_a, (uint32_t)(use_tma >> 1) << 4);
      }
    }
    // this warp's stream: two batches of loads in flight before the x tile is needed
    const SellStream sd = streams[vi * NW + warp];
    const int Q = sd.q_total;
    const double* v = val + sd.voff + lane;
    const uint16_t* ix = idx + sd.ioff + lane * 4;
    SellBatch<KB> ba, bb;
    if (Q >= 2 * kBatch) {
      LoadBatch<false>(ba, v, ix, 0, Q);
      LoadBatch<false>(bb, v, ix, kBatch, Q);
    } else {
      LoadBatch<true>(ba, v, ix, 0, Q);
      LoadBatch<true>(bb, v, ix, kBatch, Q);
    }
    SliceCursor c;
    c.sl_len = sl_len; c.rowid = rowid; c.P = P;
    c.cur = c.end = 0;
    if (sd.ss_cnt > 0) c.Start(sd.ss_begin, sd.ss_cnt);
    if (vi == v_lo) {
      // first visit: the loads above are in flight; now order behind the producer of x
      GridDepWait();
      if (skip && *skip) return;  // speculative launch after the CG loop converged (kkt.cu)
      if (preP) pre_phase();
      if (use_tma && threadIdx.x == 0) {
        StageTile(SmemAddr(xs), x + c0, (uint32_t)cw * 8u, bar_a, (uint32_t)(use_tma >> 1) << 4);
      }
    }
    if (restage) {
      if (use_tma) {
        if (vi == v_lo) __syncthreads();  // the barrier init is visible to every thread
        uint32_t done = 0;
        while (!done) {
          asm volatile(
              "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
              : "=r"(done) : "r"(bar_a), "r"(phase) : "memory");
        }
        phase ^= 1u;
      } else {
        for (int i = threadIdx.x; i < cw; i += NW * 32) xs[i] = __ldcg(x + c0 + i);  // may be written in-kernel (pre-phase)
        __syncthreads();
      }
      cur_t = t;
      if (TRACE && threadIdx.x == 0) trace[blockIdx.x * 4 + 1] = GlobalTimer();
    }
    if (sd.ss_cnt == 0) continue;
    c.SkipEmpty();
    int q = 0;
    // full batches need no guards
    while (q + 4 * kBatch <= Q) {
      ConsumeBatch<false>(ba, xs, c, q, Q);
      LoadBatch<false>(ba, v, ix, q + 2 * kBatch, Q);
      ConsumeBatch<false>(bb, xs, c, q + kBatch, Q);
      LoadBatch<false>(bb, v, ix, q + 3 * kBatch, Q);
      q += 2 * kBatch;
    }
    while (q < Q) {
      ConsumeBatch<true>(ba, xs, c, q, Q);
      LoadBatch<true>(ba, v, ix, q + 2 * kBatch, Q);
      ConsumeBatch<true>(bb, xs, c, q + kBatch, Q);
      LoadBatch<true>(bb, v, ix, q + 3 * kBatch, Q);
      q += 2 * kBatch;
    }
    // rows after the last stored entry (empty in this tile)
    while (c.cur < c.end) c.Next();
    if (TRACE && lane == 0) {
      const unsigned long long now = GlobalTimer();
      atomicMin(&trace[blockIdx.x * 4 + 2], now);  // first warp done
      atomicMax(&trace[blockIdx.x * 4 + 3], now);  // last warp done
    }
  }
}

}  // namespace


static void BuildSellStream(Ctx* ctx, const Csr& A, Sell* S) {
  const int R = A.m, Cn = A.n;
  // 24 warps: 80 registers per thread hold two batches of loads without spills; measured on
  // B200 (1M x 100k, density 1e-3): 96.4 % of the HBM copy peak vs 96.1 % (32 warps, spills),
  // 93.8 % (20), 86.2 % (16)
  const int NW = getenv("EPS_SELL_WARPS") ? atoi(getenv("EPS_SELL_WARPS")) : 24;
  EPS_CHECK(NW == 16 || NW == 20 || NW == 24 || NW == 32, "EPS_SELL_WARPS must be 16, 20, 24 or 32");
  S->nw = NW;
  S->R = R; S->Cn = Cn; S->nnz = A.nnz;
  // Tile width.  200 KB tiles (25 600 doubles) leave the SM 27 KB of L1 for the slice descriptors and the
  // partial-sum stores; 160 KB tiles measured 5.7 % faster on the 1M x 100k matrix (184.4 vs 174.4 it/s) although
  // they add a plane.  On small matrices the 8*TW bytes every CTA stages before it can start weigh more: the
  // 125k x 100k shard of an 8-GPU run is fastest at 100 KB (473.7 vs 454.4 it/s at 160 KB).  So: staging at most
  // ~10 % of the bytes a CTA streams, between 12 800 and 20 000 doubles.  EPS_SELL_TILE overrides.
  int max_tile = (int)std::min<double>(20000.0, std::max<double>(12800.0, 0.1 * 10.0 * (double)A.nnz / ctx->sm_count / 8.0));
  if (getenv("EPS_SELL_TILE")) max_tile = std::min(kMaxTile, std::max(256, atoi(getenv("EPS_SELL_TILE"))));
  const int T = (Cn + max_tile - 1) / max_tile;
  int TW = (Cn + T - 1) / T;
  TW = (TW + 1) / 2 * 2;  // keep tile starts 16-byte aligned for the bulk copy
  S->T = T; S->TW = TW;
  const int64_t TR = (int64_t)T * R;
  EPS_CHECK(TR < ((int64_t)1 << 31) - 64, "too many (tile,row) pairs for the SELL build (%lld)", (long long)TR);
  DevBuf<int32_t> segstart(ctx, TR), seglen(ctx, TR);
  SegBoundsKernel<<<(R + 255) / 256, 256, 0, ctx->stream>>>(R, T, TW, A.rowptr.p, A.col.p, segstart.p, seglen.p);

  // equal-cost cut of the tile-major (tile,row) list over the CTAs
  const int G = (int)std::max<int64_t>(1, std::min<int64_t>(ctx->sm_count, (TR + 31) / 32));
  std::vector<int32_t> cut(G + 1);
  {
    DevBuf<int64_t> csum(ctx, TR);
    cub::CountingInputIterator<int64_t> cnt(0);
    cub::TransformInputIterator<int64_t, CostOp, cub::CountingInputIterator<int64_t>> cost(
        cnt, CostOp{seglen.p, (int64_t)R, getenv("EPS_SELL_NOSTAGE") ? (int64_t)0 : (int64_t)8 * TW});
    size_t bytes = 0;
    CUDA_CHECK(cub::DeviceScan::InclusiveSum(nullptr, bytes, cost, csum.p, (int)TR, ctx->stream));
    DevBuf<char> tmp(ctx, bytes);
    CUDA_CHECK(cub::DeviceScan::InclusiveSum(tmp.p, bytes, cost, csum.p, (int)TR, ctx->stream));
    DevBuf<int32_t> dcut(ctx, G + 1);
    // With few tiles every CTA gets rows of ONE tile (whole CTAs per tile, largest-remainder rounding): a range
    // that spills a handful of rows into the next tile pays a full restaging of x for them and finishes last
    // (8-GPU shard: the slowest CTA ended 4 us after the median one, and every rank waits for the slowest CTA of
    // the slowest rank).  With many tiles (T > G/8) whole CTAs per tile would cost more balance than restaging.
    const bool no_span = T > 1 && T <= G / 8 && !getenv("EPS_SELL_SPAN");
    if (no_span) {
      std::vector<int64_t> tile_end(T);
      for (int t = 0; t < T; t++)
        CUDA_CHECK(cudaMemcpyAsync(&tile_end[t], csum.p + ((int64_t)(t + 1) * R - 1), sizeof(int64_t),
                                   cudaMemcpyDeviceToHost, ctx->stream));
      CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      std::vector<double> cost_t(T);
      double total = 0;
      for (int t = 0; t < T; t++) { cost_t[t] = (double)(tile_end[t] - (t ? tile_end[t - 1] : 0)); total += cost_t[t]; }
      std::vector<int> share(T, 1);
      int left = G - T;
      std::vector<std::pair<double, int>> frac;
      for (int t = 0; t < T; t++) {
        const double want = total > 0 ? cost_t[t] / total * G : (double)G / T;
        const int extra = std::max(0, (int)std::floor(want) - 1);
        share[t] += std::min(extra, left);
        left -= std::min(extra, left);
        frac.push_back({want - std::floor(want), t});
      }
      std::sort(frac.begin(), frac.end(), [](const auto& a, const auto& b) { return a.first > b.first || (a.first == b.first && a.second < b.second); });
      for (size_t k = 0; left > 0; k = (k + 1) % frac.size()) { share[frac[k].second]++; left--; }
      std::vector<int64_t> target(G + 1, 0), forced(G + 1, -1);
      int g = 0;
      for (int t = 0; t < T; t++) {
        const int64_t base = t ? tile_end[t - 1] : 0;
        forced[g] = (int64_t)t * R;
        for (int j = 1; j < share[t]; j++) target[g + j] = base + (int64_t)(cost_t[t] * j / share[t]);
        g += share[t];
      }
      forced[G] = TR;
      DevBuf<int64_t> dt(ctx, G + 1), df(ctx, G + 1);
      dt.upload(target.data(), G + 1);
      df.upload(forced.data(), G + 1);
      CutTargetsKernel<<<(G + 1 + 255) / 256, 256, 0, ctx->stream>>>(TR, G, csum.p, dt.p, df.p, dcut.p);
    } else {
      CutKernel<<<(G + 1 + 255) / 256, 256, 0, ctx->stream>>>(TR, G, csum.p, dcut.p);
    }
    dcut.download(cut.data(), G + 1);
    for (int g = 1; g <= G; g++) cut[g] = std::max(cut[g], cut[g - 1]);   // monotone whatever the rounding did
  }
  // visits: the part of each CTA's range inside one tile
  std::vector<int32_t> lin0, slice_base, vtile, cta_visit(G + 1);
  int n_slices = 0;
  for (int g = 0; g < G; g++) {
    cta_visit[g] = (int)vtile.size();
    int64_t c0 = cut[g];
    const int64_t c1 = std::max<int64_t>(cut[g + 1], c0);
    while (c0 < c1) {
      const int t = (int)(c0 / R);
      const int64_t e = std::min<int64_t>(c1, (int64_t)(t + 1) * R);
      lin0.push_back((int32_t)c0);
      slice_base.push_back(n_slices);
      vtile.push_back(t);
      n_slices += (int)((e - c0 + 31) / 32);
      c0 = e;
    }
  }
  const int NV = (int)vtile.size();
  cta_visit[G] = NV;
  lin0.push_back((int32_t)TR);
  slice_base.push_back(n_slices);
  EPS_CHECK(NV < (1 << 23), "too many SELL visits (%d)", NV);
  S->n_cta = G; S->n_visits = NV; S->n_slices = n_slices;
  DevBuf<int32_t> d_lin0(ctx, NV + 1), d_sbase(ctx, NV + 1);
  d_lin0.upload(lin0.data(), NV + 1);
  d_sbase.upload(slice_base.data(), NV + 1);
  S->visit_tile.alloc(ctx, std::max(NV, 1));
  S->visit_tile.upload(vtile.data(), NV);
  S->cta_visit.alloc(ctx, G + 1);
  S->cta_visit.upload(cta_visit.data(), G + 1);
  const VisitTab vt{d_lin0.p, d_sbase.p, NV};

  // sort the rows of every visit by length (descending, stable)
  DevBuf<int32_t> vals_in(ctx, TR), vals_out(ctx, TR);
  DevBuf<uint64_t> keys_in(ctx, TR), keys_out(ctx, TR);
  VisitKeysKernel<<<(unsigned)((TR + 255) / 256), 256, 0, ctx->stream>>>(TR, vt, seglen.p, keys_in.p, vals_in.p);
  {
    int end_bit = 41;
    while (end_bit < 64 && ((int64_t)1 << (end_bit - 40)) <= NV) end_bit++;
    size_t bytes = 0;
    CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys_in.p, keys_out.p, vals_in.p, vals_out.p,
                                               (int)TR, 0, end_bit, ctx->stream));
    DevBuf<char> tmp(ctx, bytes);
    CUDA_CHECK(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, keys_in.p, keys_out.p, vals_in.p, vals_out.p,
                                               (int)TR, 0, end_bit, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }
  // slice lengths in stream order and their value offsets
  S->sl_len.alloc(ctx, n_slices + 1);
  S->voff.alloc(ctx, n_slices + 1);
  std::vector<int64_t> hv(n_slices + 1);
  {
    DevBuf<int64_t> vsize(ctx, n_slices + 1);
    vsize.zero();
    StreamSliceLenKernel<<<(n_slices + 255) / 256, 256, 0, ctx->stream>>>(n_slices, NW, vt, keys_out.p, S->sl_len.p, vsize.p);
    size_t bytes = 0;
    CUDA_CHECK(cub::DeviceScan::ExclusiveSum(nullptr, bytes, vsize.p, S->voff.p, n_slices + 1, ctx->stream));
    DevBuf<char> tmp(ctx, bytes);
    CUDA_CHECK(cub::DeviceScan::ExclusiveSum(tmp.p, bytes, vsize.p, S->voff.p, n_slices + 1, ctx->stream));
    S->voff.download(hv.data(), n_slices + 1);
  }
  // one stream per (visit, warp)
  std::vector<SellStream> st((size_t)std::max(NV, 1) * NW);
  int64_t ioff = 0;
  for (int v = 0; v < NV; v++) {
    const int n_sl = slice_base[v + 1] - slice_base[v];
    for (int w = 0; w < NW; w++) {
      SellStream& d = st[(size_t)v * NW + w];
      d.ss_begin = slice_base[v] + WarpPrefix(n_sl, w, NW);
      d.ss_cnt = WarpPrefix(n_sl, w + 1, NW) - WarpPrefix(n_sl, w, NW);
      d.voff = hv[d.ss_begin];
      d.q_total = (int32_t)((hv[d.ss_begin + d.ss_cnt] - hv[d.ss_begin]) / 32);
      d.ioff = ioff;
      d.pad = 0;
      ioff += (int64_t)((d.q_total + 3) / 4) * 128;
    }
  }
  S->streams.alloc(ctx, st.size());
  S->streams.upload(st.data(), st.size());
  S->n_val = hv[n_slices];
  S->n_idx = ioff;
  S->val.alloc(ctx, std::max<int64_t>(S->n_val, 1));
  S->idx.alloc(ctx, std::max<int64_t>(S->n_idx, 4));
  S->idx.zero();
  S->rowid.alloc(ctx, (size_t)std::max(n_slices, 1) * 32);
  StreamFillKernel<<<(n_slices + 7) / 8, 256, 0, ctx->stream>>>(
      R, TW, n_slices, NW, vt, S->visit_tile.p, S->streams.p, vals_out.p, segstart.p, seglen.p, A.col.p, A.val.p,
      S->sl_len.p, S->voff.p, S->val.p, S->idx.p, S->rowid.p);
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  CUDA_CHECK(cudaGetLastError());
  S->voff.release();  // only needed while building
  static bool attr_set = false;
  if (!attr_set) {
    const int smem = kMaxTile * 8 + 64;
#define EPS_SELL_ATTR(NW) \
    CUDA_CHECK(cudaFuncSetAttribute(SellStreamKernel<NW, 8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
    CUDA_CHECK(cudaFuncSetAttribute(SellStreamKernel<NW, 8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    EPS_SELL_ATTR(16) EPS_SELL_ATTR(20) EPS_SELL_ATTR(24) EPS_SELL_ATTR(32)
#undef EPS_SELL_ATTR
    attr_set = true;
  }
  S->built = true;
}

void BuildSell(Ctx* ctx, const Csr& A, Sell* S) { BuildSellStream(ctx, A, S); }

struct SellPre {
  const double* P = nullptr;  // planes of the previous pass, T x Cn of THIS pass
  int T = 0;
  double alpha = 1.0;
  double* y = nullptr;        // receives alpha * sum_t P[t]; must be the x of this pass
  unsigned* bar = nullptr;    // two zero-initialised words (GridBarrier)
};

static double* SellLaunch(Ctx* ctx, const Sell& S, const double* x, int kclass, const int* skip,
                          const SellPre* pre = nullptr, bool second_buffer = false) {
  double* P = GrowScratch(ctx, second_buffer ? &ctx->sell_scratch2 : &ctx->sell_scratch, (size_t)S.T * S.R);
  {
    ProfScope ps(ctx, kclass, 12.0 * S.nnz + 4.0 * (S.R + 1) + 8.0 * S.Cn + 8.0 * S.R);
    // bulk copies need 16-byte aligned addresses and sizes: TW is even, so only the base
    // pointer and the width of the last tile can break that
    const int last = S.Cn - (S.T - 1) * S.TW;
    static const int tma_chunk = getenv("EPS_SELL_TMA_CHUNK") ? atoi(getenv("EPS_SELL_TMA_CHUNK")) : 0;
    // bit 0: stage by bulk copies; bits 1..: bytes per request / 16 (0 = the whole tile in one request)
    const int use_tma = ((reinterpret_cast<uintptr_t>(x) & 15) == 0 && last % 2 == 0) ? (1 | ((tma_chunk / 16) << 1)) : 0;
    {
    // EPS_SELL_TRACE=<launch number>: per-CTA timeline of that launch on stderr (diagnostics)
    static const int trace_at = getenv("EPS_SELL_TRACE") ? atoi(getenv("EPS_SELL_TRACE")) : -1;
    static int launch_no = 0;
    static DevBuf<unsigned long long> tbuf;
    unsigned long long* trace = nullptr;
    bool step_trace = false;
    if (unsigned long long* st = StepTraceSlot(ctx, 0, S.n_cta, pre ? "pass 2 (+pre-phase)" : "pass 1")) {
      trace = st;
      step_trace = true;
    } else if (trace_at >= 0 && launch_no++ == trace_at) {
      tbuf.alloc(ctx, (size_t)S.n_cta * 4);
      std::vector<unsigned long long> init((size_t)S.n_cta * 4, 0ull);
      for (int g = 0; g < S.n_cta; g++) init[g * 4 + 2] = ~0ull;
      tbuf.upload(init.data(), init.size());
      CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      trace = tbuf.p;
    }
    using KernT = void (*)(int, int, int, const int32_t*, const int32_t*, const SellStream*, const int32_t*,
                           const double*, const uint16_t*, const int32_t*, const double*, double*, int, const int*,
                           unsigned long long*, const double*, int, double, double*, unsigned*);
    KernT kern = nullptr;
    switch (S.nw) {
      case 16: kern = trace ? SellStreamKernel<16, 8, true> : SellStreamKernel<16, 8, false>; break;
      case 20: kern = trace ? SellStreamKernel<20, 8, true> : SellStreamKernel<20, 8, false>; break;
      case 24: kern = trace ? SellStreamKernel<24, 8, true> : SellStreamKernel<24, 8, false>; break;
      default: kern = trace ? SellStreamKernel<32, 8, true> : SellStreamKernel<32, 8, false>; break;
    }
    // programmatic stream serialization: the prologue (matrix prefetch) overlaps the tail of the
    // kernel that produces x (EPS_NO_PDL=1: plain launch)
    static const bool pdl = getenv("EPS_NO_PDL") == nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(S.n_cta);
    cfg.blockDim = dim3(S.nw * 32);
    cfg.dynamicSmemBytes = (size_t)S.TW * 8 + 64;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    const int32_t *a_cv = S.cta_visit.p, *a_vt = S.visit_tile.p, *a_sl = S.sl_len.p, *a_ri = S.rowid.p;
    const SellStream* a_st = S.streams.p;
    const double* a_val = S.val.p;
    const uint16_t* a_idx = S.idx.p;
    const double* pre_P = pre ? pre->P : nullptr;
    const int pre_T = pre ? pre->T : 0;
    const double pre_alpha = pre ? pre->alpha : 1.0;
    double* pre_y = pre ? pre->y : nullptr;
    unsigned* pre_bar = pre ? pre->bar : nullptr;
    if (pre) EPS_CHECK(pre->y == x && pre->bar && S.n_cta <= ctx->sm_count, "pre-phase of the stream kernel: bad arguments");
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, S.R, S.Cn, S.TW, a_cv, a_vt, a_st, a_sl, a_val, a_idx, a_ri, x, P,
                                  use_tma, skip, trace, pre_P, pre_T, pre_alpha, pre_y, pre_bar));
    if (trace && !step_trace) {
      std::vector<unsigned long long> h((size_t)S.n_cta * 4);
      tbuf.download(h.data(), h.size());
      unsigned long long t0 = ~0ull;
      for (int g = 0; g < S.n_cta; g++) if (h[g * 4]) t0 = std::min(t0, h[g * 4]);
      std::vector<double> a, b, c, d;
      for (int g = 0; g < S.n_cta; g++) {
        if (!h[g * 4]) continue;
        a.push_back((h[g * 4] - t0) * 1e-3);
        b.push_back((h[g * 4 + 1] - t0) * 1e-3);
        c.push_back((h[g * 4 + 2] - t0) * 1e-3);
        d.push_back((h[g * 4 + 3] - t0) * 1e-3);
      }
      auto q = [](std::vector<double>& v, double f) { std::sort(v.begin(), v.end()); return v.empty() ? 0.0 : v[(size_t)(f * (v.size() - 1))]; };
      fprintf(stderr, "[sell trace] R=%d Cn=%d T=%d ctas=%zu us since first CTA start: start min/med/max %.1f %.1f %.1f | "
              "x staged %.1f %.1f %.1f | first warp done %.1f %.1f %.1f | last warp done %.1f %.1f %.1f\n",
              S.R, S.Cn, S.T, a.size(), q(a, 0), q(a, .5), q(a, 1), q(b, 0), q(b, .5), q(b, 1), q(c, 0), q(c, .5), q(c, 1),
              q(d, 0), q(d, .5), q(d, 1));
    }
    }
    ctx->count_launch();
  }
  return P;
}

const double* SellApplyPlanes(Ctx* ctx, const Sell& S, const double* x, int kclass, const int* skip) {
  EPS_CHECK(S.R > 0, "empty sparse map");
  return SellLaunch(ctx, S, x, kclass, skip);
}

const double* SellApplyPlanesAfter(Ctx* ctx, const Sell& S, double* x, const double* prevP, int prevT, double alpha,
                                   unsigned* bar, int kclass, const int* skip) {
  EPS_CHECK(S.R > 0, "empty sparse map");
  SellPre pre;
  pre.P = prevP; pre.T = prevT; pre.alpha = alpha; pre.y = x; pre.bar = bar;
  return SellLaunch(ctx, S, x, kclass, skip, &pre, true);
}

void SellApply(Ctx* ctx, const Sell& S, double alpha, const double* x, double beta, double* y, int kclass,
               const SpmvExtra* ex) {
  const int* skip = ex ? ex->skip : nullptr;
  if (S.R == 0) return;
  double* P = SellLaunch(ctx, S, x, kclass, skip);
  if (ex && ex->publish) {
    EPS_CHECK(beta == 0.0 && ex->publish->n >= S.R, "publishing apply: beta must be 0 and the vector must fit the slots");
    ProfScope ps(ctx, KC_ALLREDUCE, 8.0 * S.T * S.R + 8.0 * S.R * ex->publish->G);
    SellReducePublishKernel<<<(S.R + 255) / 256, 256, 0, ctx->stream>>>(S.R, S.T, P, alpha, *ex->publish,
                                                                         ex->publish_seq, ex->publish_counter);
    ctx->count_launch();
    return;
  }
  ProfScope ps(ctx, KC_VEC, 8.0 * S.T * S.R + 8.0 * S.R);
  static const bool pdl = getenv("EPS_NO_PDL") == nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((S.R + 255) / 256);
  cfg.blockDim = dim3(256);
  cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  const double* Pc = P;
  const double *ep_d = ex ? ex->ep_d : nullptr, *ep_p = ex ? ex->ep_p : nullptr;
  CUDA_CHECK(cudaLaunchKernelEx(&cfg, SellReduceKernel, S.R, S.T, Pc, alpha, beta, y, skip, ep_d, ep_p));
  ctx->count_launch();
}

}  // namespace eps
