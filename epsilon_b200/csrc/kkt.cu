// KKT prox solve object: DIAGONAL / WOODBURY / GRAM / CG.  See kkt.h.
#include <algorithm>

#include <cooperative_groups.h>

#include "device.cuh"
#include "kkt.h"
#include "peer.cuh"
#include "prox.cuh"

namespace eps {

void AddBlock(Ctx* ctx, int nb, const double* G, int64_t ldg, double* M, int64_t ldm);

namespace {

constexpr int kBlock = 256;
constexpr int kUnroll = 4;

__device__ __forceinline__ unsigned long long GlobalTimerNs() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

constexpr int kCgRing = 64;
constexpr int kTailBlock = 512;

// Sum of `count` per-CTA partials in a fixed order; identical in every CTA.
__device__ __forceinline__ double SumPartials(const double* partials, int count, double* smem) {
  double v[1] = {0.0};
  for (int i = threadIdx.x; i < count; i += blockDim.x) v[0] += __ldcg(&partials[i]);
  BlockSum<1>(v, smem);
  __shared__ double total_;
  __syncthreads();
  if (threadIdx.x == 0) total_ = v[0];
  __syncthreads();
  return total_;
}

// r = rhs - w; p = r; rr = r.r; thresh = tol*sqrt(rhs.rhs); publishes result `seq`.
// scal: [0] rr  [1] thresh
__global__ void __launch_bounds__(kBlock)
CgStartKernel(int64_t n, const double* __restrict__ rhs, const double* __restrict__ w,
              double* __restrict__ r, double* __restrict__ p, double tol, double* scal,
              double* partials, unsigned* counter, int* done, KktSolver::CgResult* ring,
              long long seq, PeerXchg px, unsigned long long xseq, const double* __restrict__ dg,
              const double* __restrict__ x0) {
  __shared__ double smem[64];
  double acc[2] = {0.0, 0.0};
  const int64_t stride = (int64_t)gridDim.x * kBlock;
  if (xseq) PeerWait(px, xseq);  // w = sum over ranks of the published parts + d.*x0
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += stride) {
    const double b = rhs[i];
    const double wi = xseq ? PeerSum(px, xseq, i) + dg[i] * x0[i] : w[i];
    const double t = b - wi;
    r[i] = t;
    p[i] = t;
    acc[0] += t * t;
    acc[1] += b * b;
  }
  if (GridSumLast<2>(acc, smem, partials, counter)) {
    // rhs == 0 (exact solution 0): measure against the initial residual instead of a zero threshold
    const double thresh = tol * sqrt(acc[1] > 0.0 ? acc[1] : acc[0]);
    const int conv = sqrt(acc[0]) <= thresh ? 1 : 0;
    scal[0] = acc[0];
    scal[1] = thresh;
    *done = conv;
    KktSolver::CgResult* e = &ring[seq % kCgRing];
    e->rr = acc[0];
    e->converged = conv;
    __threadfence_system();
    *reinterpret_cast<volatile long long*>(&e->seq) = seq;
  }
}

// One CG update after w = M p, as one cooperative kernel:
//   a = rr / p.w;  x += a p;  r -= a w;  rr' = r.r;  p = r + (rr'/rr) p
// and the publication of (rr', converged) for the host.  No-op once `done` is raised.
__global__ void __launch_bounds__(kTailBlock)
CgTailKernel(int64_t n, double* __restrict__ w, double* __restrict__ p, double* __restrict__ x,
             double* __restrict__ r, double* scal, double* partials, int* done,
             KktSolver::CgResult* ring, long long seq, PeerXchg px, unsigned long long xseq,
             const double* __restrict__ dg) {
  namespace cg = cooperative_groups;
  __shared__ double smem[64];
  if (*done) {  // uniform: only ever written by the last phase of an earlier launch
    // the exchange of the speculative iteration is still consumed, so that a peer can never
    // run two exchanges ahead of this rank (double-buffered slots)
    if (xseq) PeerWait(px, xseq);
    return;
  }
  cg::grid_group grid = cg::this_grid();
  const int G = gridDim.x;
  const int64_t stride = (int64_t)G * kTailBlock;
  const int64_t first = (int64_t)blockIdx.x * kTailBlock + threadIdx.x;
  const double rr = scal[0];
  const double thresh = scal[1];
  double v[1] = {0.0};
  if (xseq) {
    // fused receive side of the all-reduce: w = sum_ranks(published) + d.*p
    PeerWait(px, xseq);
    for (int64_t i = first; i < n; i += stride) {
      const double pi = p[i];
      const double wi = PeerSum(px, xseq, i) + dg[i] * pi;
      w[i] = wi;
      v[0] += pi * wi;
    }
  } else {
    for (int64_t i = first; i < n; i += stride) v[0] += p[i] * w[i];
  }
  BlockSum<1>(v, smem);
  if (threadIdx.x == 0) partials[blockIdx.x] = v[0];
  grid.sync();
  const double pw = SumPartials(partials, G, smem);
  const double a = pw != 0.0 ? rr / pw : 0.0;
  v[0] = 0.0;
  for (int64_t i = first; i < n; i += stride) {
    x[i] += a * p[i];
    const double t = r[i] - a * w[i];
    r[i] = t;
    v[0] += t * t;
  }
  BlockSum<1>(v, smem);
  if (threadIdx.x == 0) partials[G + blockIdx.x] = v[0];
  grid.sync();
  const double rr_new = SumPartials(partials + G, G, smem);
  const double b = rr != 0.0 ? rr_new / rr : 0.0;
  for (int64_t i = first; i < n; i += stride) p[i] = r[i] + b * p[i];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const int conv = sqrt(rr_new) <= thresh ? 1 : 0;
    scal[0] = rr_new;
    if (conv) *done = 1;
    KktSolver::CgResult* e = &ring[seq % kCgRing];
    e->rr = rr_new;
    e->converged = conv;
    __threadfence_system();
    *reinterpret_cast<volatile long long*>(&e->seq) = seq;
  }
}

// The reduction of the second SpMV pass, the cross-GPU exchange and the CG update as ONE
// cooperative kernel (one CTA per SM, every CTA owns the same contiguous slice of the n-vector
// on every rank):
//   part_i  = alpha * sum_t planes[t][i]                     (this rank's rows of H'(H p))
//   G > 1 : part is pushed into slot [parity][rank] of every rank over NVLink (P2P stores);
//           the CTA then raises ITS flag [parity][rank][cta] on every rank and waits for the G
//           flags of the same slice -- no grid-wide or device-wide wait for the exchange, no
//           arrival counter: a slice is summed (in rank order, identical bits everywhere) as soon
//           as its G parts have landed
//   w = sum + d.*p ; then the CG update of CgTailKernel.
// Once `done` is raised only the flag handshake is executed, so that no rank can run two
// exchanges ahead of a reader of the double-buffered slots.
__global__ void __launch_bounds__(kTailBlock)
CgFusedTailKernel(int64_t n, int T, const double* __restrict__ planes, double alpha,
                  double* __restrict__ w, double* __restrict__ p, double* __restrict__ x,
                  double* __restrict__ r, double* scal, double* partials, int* done,
                  KktSolver::CgResult* ring, long long seq, PeerXchg px, unsigned long long xseq,
                  const double* __restrict__ dg, unsigned long long* __restrict__ trace) {
  namespace cg = cooperative_groups;
  __shared__ double smem[64];
  const int G = gridDim.x;
  const int c = blockIdx.x;
#define EPS_TT(k) if (trace && threadIdx.x == 0) trace[c * 8 + (k)] = GlobalTimerNs()
  EPS_TT(0);
  GridDepLaunch();  // the next SpMV pass may start prefetching its matrix streams (sell.cu)
  GridDepWait();    // no-op unless launched with programmatic serialization behind the SpMV pass
  const int parity = (int)(xseq & 1ull);
  const bool finished = *done != 0;  // uniform: only ever written by the last phase of an earlier launch
  // slice of this CTA (even length so that the rows stay 16-byte aligned)
  int64_t per = (n + G - 1) / G;
  per = (per + 1) & ~(int64_t)1;
  const int64_t lo = min(n, (int64_t)c * per), hi = min(n, lo + per);
  const bool ll = px.G > 1 && px.ll1[0] != nullptr;
  const unsigned seq32 = (unsigned)xseq;
  if (ll) {
    // two-shot exchange with flag-in-data packets: CTA c's slice is owned by rank c % G.  A thread
    // keeps the same elements in every step, so there is no barrier, no fence and no flag:
    //   (1) partial -> owner's inbox; (2) owner: wait for the G partials, add them in rank order,
    //   result -> every rank's inbox; (3) everybody reads the result while forming w (below).
    if (!finished) {
      const int owner = c % px.G;
      for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
        double s = 0.0;
        for (int t = 0; t < T; t++) s += planes[(size_t)t * n + i];
        LlStore(px.ll1[owner] + ((size_t)parity * px.G + px.rank) * px.n + i, alpha * s, seq32);
      }
      EPS_TT(1);
      if (px.rank == owner) {
        for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
          double sum = 0.0;
          for (int q = 0; q < px.G; q++)
            sum += LlLoad(px.ll1[px.rank] + ((size_t)parity * px.G + q) * px.n + i, seq32);
          for (int k = 0; k < px.G; k++)
            LlStore(px.ll2[(px.rank + k) % px.G] + (size_t)parity * px.n + i, sum, seq32);
        }
      }
      EPS_TT(2);
    }
  } else if (px.G > 1) {
    if (!finished) {
      for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
        double s = 0.0;
        for (int t = 0; t < T; t++) s += planes[(size_t)t * n + i];
        const double out = alpha * s;
        const size_t slot = ((size_t)parity * px.G + px.rank) * px.n + i;
        for (int k = 0; k < px.G; k++) {
          const int q = (px.rank + k) % px.G;  // start with the local copy, spread the peers
          px.recv[q][slot] = out;
        }
      }
    }
    __syncthreads();
    EPS_TT(1);
    if (threadIdx.x == 0) __threadfence_system();  // cumulative: orders the CTA's remote stores
    __syncthreads();
    EPS_TT(2);
    if (threadIdx.x < px.G)
      StReleaseSys(px.cflag[threadIdx.x] + ((size_t)parity * px.G + px.rank) * kPeerMaxCtas + c, xseq);
    if (threadIdx.x < px.G) {
      const unsigned long long* f = px.cflag[px.rank] + ((size_t)parity * px.G + threadIdx.x) * kPeerMaxCtas + c;
      while (LdAcquireSys(f) < xseq) {}
    }
    __syncthreads();
    EPS_TT(3);
  }
  if (finished) return;
  cg::grid_group grid = cg::this_grid();
  const double rr = scal[0];
  const double thresh = scal[1];
  double v[1] = {0.0};
  if (ll) {
    for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
      const double pi = p[i];
      const double wi = LlLoad(px.ll2[px.rank] + (size_t)parity * px.n + i, seq32) + dg[i] * pi;
      w[i] = wi;
      v[0] += pi * wi;
    }
    EPS_TT(3);
  } else if (px.G > 1) {
    for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
      const double pi = p[i];
      const double wi = PeerSum(px, xseq, i) + dg[i] * pi;
      w[i] = wi;
      v[0] += pi * wi;
    }
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
      double s = 0.0;
      for (int t = 0; t < T; t++) s += planes[(size_t)t * n + i];
      const double pi = p[i];
      const double wi = alpha * s + dg[i] * pi;
      w[i] = wi;
      v[0] += pi * wi;
    }
  }
  BlockSum<1>(v, smem);
  if (threadIdx.x == 0) partials[blockIdx.x] = v[0];
  EPS_TT(4);
  grid.sync();
  EPS_TT(5);
  const double pw = SumPartials(partials, G, smem);
  const double a = pw != 0.0 ? rr / pw : 0.0;
  v[0] = 0.0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
    x[i] += a * p[i];
    const double t = r[i] - a * w[i];
    r[i] = t;
    v[0] += t * t;
  }
  BlockSum<1>(v, smem);
  if (threadIdx.x == 0) partials[G + blockIdx.x] = v[0];
  grid.sync();
  const double rr_new = SumPartials(partials + G, G, smem);
  const double b = rr != 0.0 ? rr_new / rr : 0.0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) p[i] = r[i] + b * p[i];
  EPS_TT(6);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const int conv = sqrt(rr_new) <= thresh ? 1 : 0;
    scal[0] = rr_new;
    if (conv) *done = 1;
    KktSolver::CgResult* e = &ring[seq % kCgRing];
    e->rr = rr_new;
    e->converged = conv;
    __threadfence_system();
    *reinterpret_cast<volatile long long*>(&e->seq) = seq;
  }
#undef EPS_TT
}

// r0 without a matvec: the previous solve left r_prev = rhs_prev - M x_prev for the x that is the warm
// start of this one, so r0 = rhs - M x_prev = r_prev + (rhs - rhs_prev).  (One product with M = two passes over
// the matrix saved per ADMM iteration; the true residual is recomputed every few solves to bound the drift
// of the recurrence.)  Otherwise as CgStartKernel; also leaves rhs in rhs_prev for the next solve.
__global__ void __launch_bounds__(kBlock)
CgCarryStartKernel(int64_t n, const double* __restrict__ rhs, double* __restrict__ rhs_prev,
                   double* __restrict__ r, double* __restrict__ p, double tol, double* scal, double* partials,
                   unsigned* counter, int* done, KktSolver::CgResult* ring, long long seq) {
  __shared__ double smem[64];
  double acc[2] = {0.0, 0.0};
  const int64_t stride = (int64_t)gridDim.x * kBlock;
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += stride) {
    const double b = rhs[i];
    const double t = r[i] + (b - rhs_prev[i]);
    rhs_prev[i] = b;
    r[i] = t;
    p[i] = t;
    acc[0] += t * t;
    acc[1] += b * b;
  }
  if (GridSumLast<2>(acc, smem, partials, counter)) {
    const double thresh = tol * sqrt(acc[1] > 0.0 ? acc[1] : acc[0]);
    const int conv = sqrt(acc[0]) <= thresh ? 1 : 0;
    scal[0] = acc[0];
    scal[1] = thresh;
    *done = conv;
    KktSolver::CgResult* e = &ring[seq % kCgRing];
    e->rr = acc[0];
    e->converged = conv;
    __threadfence_system();
    *reinterpret_cast<volatile long long*>(&e->seq) = seq;
  }
}

// CgFusedTailKernel with ONE grid-wide synchronisation and no cooperative launch.
//   The four sums p.w, r.w, w.w, r.r are formed in the same pass that builds w; after the barrier every CTA knows
//       a = rr / p.w,   rr' = ||r - a w||^2 = rr - 2a r.w + a^2 w.w,   b = rr' / rr
//   and updates x += a p, r -= a w, p = r + b p in one sweep (the classical recurrence needs r.r of the UPDATED
//   r, i.e. a second synchronisation).  rr is recomputed from r every step, so the expansion never feeds on
//   itself: its error is a few ulp of rr, far below the 1e-2..1e-1 by which a CG step shrinks rr here.
//   The barrier is GridBarrier (device.cuh): the launch is an ordinary one with programmatic stream
//   serialization, so the kernel's launch latency hides behind the drain of the second SpMV pass.
__global__ void __launch_bounds__(kTailBlock)
CgFusedTail2Kernel(int64_t n, int T, const double* __restrict__ planes, double alpha,
                   double* __restrict__ w, double* __restrict__ p, double* __restrict__ x,
                   double* __restrict__ r, double* scal, double* partials, int* done,
                   KktSolver::CgResult* ring, long long seq, PeerXchg px, unsigned long long xseq,
                   const double* __restrict__ dg, unsigned* bar, unsigned long long* __restrict__ trace) {
  __shared__ double smem[128];
  const int G = gridDim.x;
  const int c = blockIdx.x;
#define EPS_TT(k) if (trace && threadIdx.x == 0) trace[c * 8 + (k)] = GlobalTimerNs()
  EPS_TT(0);
  GridDepLaunch();
  GridDepWait();
  const int parity = (int)(xseq & 1ull);
  const bool finished = *done != 0;  // uniform: written only by the last phase of an earlier launch
  int64_t per = (n + G - 1) / G;
  per = (per + 1) & ~(int64_t)1;
  const int64_t lo = min(n, (int64_t)c * per), hi = min(n, lo + per);
  const bool ll = px.G > 1 && px.ll1[0] != nullptr;
  const bool sn = ll && px.sn1[0] != nullptr;   // same two-shot exchange, 8-byte sentinel slots instead of 16-byte packets
  const unsigned seq32 = (unsigned)xseq;
  if (sn) {
    if (!finished) {
      for (int64_t i = lo + threadIdx.x; i < hi; i += 2 * kTailBlock) {
        const int64_t j = i + kTailBlock;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll 8
        for (int t = 0; t < T; t++) {
          s0 += __ldcg(planes + (size_t)t * n + i);
          if (j < hi) s1 += __ldcg(planes + (size_t)t * n + j);
        }
        SnStore(px.sn1[(int)((i >> 6) % px.G)] + ((size_t)parity * px.G + px.rank) * px.n + i, alpha * s0);
        if (j < hi) SnStore(px.sn1[(int)((j >> 6) % px.G)] + ((size_t)parity * px.G + px.rank) * px.n + j, alpha * s1);
      }
      EPS_TT(1);
      const int64_t ch_lo = lo >> 6, ch_hi = (hi + 63) >> 6;
      int64_t first = ch_lo + ((px.rank - ch_lo % px.G) + px.G) % px.G;
      for (int64_t k = threadIdx.x;; k += kTailBlock) {
        const int64_t ch = first + (k >> 6) * px.G;
        if (ch >= ch_hi) break;
        const int64_t i = (ch << 6) + (k & 63);
        if (i < lo || i >= hi) continue;
        const double sum = SnTakeRanks(px.sn1[px.rank] + (size_t)parity * px.G * px.n + i, (size_t)px.n, px.G);
        for (int q = 0; q < px.G; q++)
          SnStore(px.sn2[(px.rank + q) % px.G] + (size_t)parity * px.n + i, sum);
      }
      EPS_TT(2);
    }
  } else if (ll) {
    // two-shot exchange with flag-in-data packets.  Ownership is by 64-element chunk, chunk k belongs to rank
    // k % G, so EVERY CTA of every rank owns ~1/G of its own slice: all 148 CTAs take part in the owner step
    // (with whole slices owned by rank c % G, 7 of 8 CTAs idled while the owners walked their slice twice).
    if (!finished) {
      for (int64_t i = lo + threadIdx.x; i < hi; i += 2 * kTailBlock) {
        const int64_t j = i + kTailBlock;
        double s0 = 0.0, s1 = 0.0;
#pragma unroll 8
        for (int t = 0; t < T; t++) {
          s0 += __ldcg(planes + (size_t)t * n + i);
          if (j < hi) s1 += __ldcg(planes + (size_t)t * n + j);
        }
        LlStore(px.ll1[(int)((i >> 6) % px.G)] + ((size_t)parity * px.G + px.rank) * px.n + i, alpha * s0, seq32);
        if (j < hi) LlStore(px.ll1[(int)((j >> 6) % px.G)] + ((size_t)parity * px.G + px.rank) * px.n + j, alpha * s1, seq32);
      }
      EPS_TT(1);
      // owner step: the chunks of this slice that belong to this rank
      const int64_t ch_lo = lo >> 6, ch_hi = (hi + 63) >> 6;
      int64_t first = ch_lo + ((px.rank - ch_lo % px.G) + px.G) % px.G;
      for (int64_t k = threadIdx.x;; k += kTailBlock) {
        const int64_t ch = first + (k >> 6) * px.G;
        if (ch >= ch_hi) break;
        const int64_t i = (ch << 6) + (k & 63);
        if (i < lo || i >= hi) continue;
        const double sum = LlSumRanks(px.ll1[px.rank] + (size_t)parity * px.G * px.n + i, (size_t)px.n, px.G, seq32);
        for (int q = 0; q < px.G; q++)
          LlStore(px.ll2[(px.rank + q) % px.G] + (size_t)parity * px.n + i, sum, seq32);
      }
      EPS_TT(2);
    }
  } else if (px.G > 1) {
    if (!finished) {
      for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
        double s = 0.0;
#pragma unroll 8
        for (int t = 0; t < T; t++) s += __ldcg(planes + (size_t)t * n + i);
        const double out = alpha * s;
        const size_t slot = ((size_t)parity * px.G + px.rank) * px.n + i;
        for (int k = 0; k < px.G; k++) px.recv[(px.rank + k) % px.G][slot] = out;
      }
    }
    __syncthreads();
    EPS_TT(1);
    if (threadIdx.x == 0) __threadfence_system();
    __syncthreads();
    EPS_TT(2);
    if (threadIdx.x < px.G)
      StReleaseSys(px.cflag[threadIdx.x] + ((size_t)parity * px.G + px.rank) * kPeerMaxCtas + c, xseq);
    if (threadIdx.x < px.G) {
      const unsigned long long* f = px.cflag[px.rank] + ((size_t)parity * px.G + threadIdx.x) * kPeerMaxCtas + c;
      while (LdAcquireSys(f) < xseq) {}
    }
    __syncthreads();
  }
  if (finished) return;
  const double thresh = scal[1];
  // A thread owns elements i0 = lo + tid and i1 = i0 + kTailBlock (slices of up to 2 * kTailBlock elements, i.e.
  // n <= 1024 * gridDim: everything of both elements is loaded up front -- one L2 round trip instead of one per
  // operand -- and p, r, w stay in registers across the barrier); longer slices take the generic loop below.
  const bool two = hi - lo <= 2 * kTailBlock;
  double v[4] = {0.0, 0.0, 0.0, 0.0};
  double pk[2] = {0.0, 0.0}, rk[2] = {0.0, 0.0}, wk[2] = {0.0, 0.0};
  if (two) {
    const int64_t i0 = lo + threadIdx.x, i1 = i0 + kTailBlock;
    const bool h0 = i0 < hi, h1 = i1 < hi;
    double dk[2] = {0.0, 0.0}, sk[2] = {0.0, 0.0};
    if (h0) { pk[0] = p[i0]; rk[0] = r[i0]; dk[0] = dg[i0]; }
    if (h1) { pk[1] = p[i1]; rk[1] = r[i1]; dk[1] = dg[i1]; }
    if (sn) {
      if (h0) sk[0] = SnTake(px.sn2[px.rank] + (size_t)parity * px.n + i0);
      if (h1) sk[1] = SnTake(px.sn2[px.rank] + (size_t)parity * px.n + i1);
    } else if (ll) {
      if (h0) sk[0] = LlLoad(px.ll2[px.rank] + (size_t)parity * px.n + i0, seq32);
      if (h1) sk[1] = LlLoad(px.ll2[px.rank] + (size_t)parity * px.n + i1, seq32);
    } else if (px.G > 1) {
      if (h0) sk[0] = PeerSum(px, xseq, i0);
      if (h1) sk[1] = PeerSum(px, xseq, i1);
    } else {
#pragma unroll 8
      for (int t = 0; t < T; t++) {
        if (h0) sk[0] += __ldcg(planes + (size_t)t * n + i0);
        if (h1) sk[1] += __ldcg(planes + (size_t)t * n + i1);
      }
      sk[0] *= alpha; sk[1] *= alpha;
    }
#pragma unroll
    for (int k = 0; k < 2; k++) {
      wk[k] = sk[k] + dk[k] * pk[k];
      v[0] += pk[k] * wk[k];
      v[1] += rk[k] * wk[k];
      v[2] += wk[k] * wk[k];
      v[3] += rk[k] * rk[k];
    }
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
      const double pi = p[i], ri = r[i];
      double wi;
      if (sn) {
        wi = SnTake(px.sn2[px.rank] + (size_t)parity * px.n + i) + dg[i] * pi;
      } else if (ll) {
        wi = LlLoad(px.ll2[px.rank] + (size_t)parity * px.n + i, seq32) + dg[i] * pi;
      } else if (px.G > 1) {
        wi = PeerSum(px, xseq, i) + dg[i] * pi;
      } else {
        double s = 0.0;
#pragma unroll 8
        for (int t = 0; t < T; t++) s += __ldcg(planes + (size_t)t * n + i);
        wi = alpha * s + dg[i] * pi;
      }
      w[i] = wi;
      v[0] += pi * wi;
      v[1] += ri * wi;
      v[2] += wi * wi;
      v[3] += ri * ri;
    }
  }
  EPS_TT(3);
  // x of both elements: requested now, needed after the barrier
  double xk[2] = {0.0, 0.0};
  if (two) {
    const int64_t i0 = lo + threadIdx.x, i1 = i0 + kTailBlock;
    if (i0 < hi) xk[0] = x[i0];
    if (i1 < hi) xk[1] = x[i1];
  }
  BlockSum<4>(v, smem);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 4; k++) partials[k * G + c] = v[k];
  }
  EPS_TT(4);
  GridBarrier(bar);
  EPS_TT(5);
  // every CTA adds the G partials of the four sums in the same order: identical a, b everywhere
  double t4[4] = {0.0, 0.0, 0.0, 0.0};
  for (int i = threadIdx.x; i < G; i += kTailBlock) {
#pragma unroll
    for (int k = 0; k < 4; k++) t4[k] += __ldcg(&partials[k * G + i]);
  }
  BlockSum<4>(t4, smem);
  __shared__ double tot_[4];
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 4; k++) tot_[k] = t4[k];
  }
  __syncthreads();
  const double pw = tot_[0], rw = tot_[1], ww = tot_[2], rr = tot_[3];
  const double a = pw != 0.0 ? rr / pw : 0.0;
  const double rr_new = fmax(rr - 2.0 * a * rw + a * a * ww, 0.0);
  const double b = rr != 0.0 ? rr_new / rr : 0.0;
  if (two) {
    const int64_t i0 = lo + threadIdx.x, i1 = i0 + kTailBlock;
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const int64_t i = k ? i1 : i0;
      if (i < hi) {
        x[i] = xk[k] + a * pk[k];
        const double t = rk[k] - a * wk[k];
        r[i] = t;
        p[i] = t + b * pk[k];
      }
    }
  } else {
    for (int64_t i = lo + threadIdx.x; i < hi; i += kTailBlock) {
      const double pi = p[i];
      x[i] += a * pi;
      const double t = r[i] - a * w[i];
      r[i] = t;
      p[i] = t + b * pi;
    }
  }
  EPS_TT(6);
  if (c == 0 && threadIdx.x == 0) {
    const int conv = sqrt(rr_new) <= thresh ? 1 : 0;
    scal[0] = rr_new;
    if (conv) *done = 1;
    KktSolver::CgResult* e = &ring[seq % kCgRing];
    e->rr = rr_new;
    e->converged = conv;
    __threadfence_system();
    *reinterpret_cast<volatile long long*>(&e->seq) = seq;
  }
#undef EPS_TT
}

__global__ void InvertKernel(int64_t n, const double* __restrict__ d, double* __restrict__ out, int* bad) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double v = d[i];
  if (!(v > 0.0)) atomicExch(bad, 1);
  out[i] = 1.0 / v;
}

bool RowDisjoint(const BlockOp& op) {
  std::vector<std::pair<int, int>> r;
  for (auto& b : op.blks) r.push_back({b.row_off, b.row_off + b.m()});
  std::sort(r.begin(), r.end());
  for (size_t i = 1; i < r.size(); i++)
    if (r[i].first < r[i - 1].second) return false;
  return true;
}

// out (rows x cols, column-major) = dense copy of op
void Densify(Ctx* ctx, const BlockOp& op, double* out) {
  CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)op.rows * op.cols * sizeof(double), ctx->stream));
  for (auto& b : op.blks) {
    const MapData& M = *b.map;
    if (M.kind == MK_SPARSE) {
      CsrDensify(ctx, M.csr, b.scale, out, op.rows, b.row_off, b.col_off, b.trans);
    } else {
      // apply the block to unit vectors column by column (dense / diagonal / scalar; small only)
      DevBuf<double> e(ctx, b.n());
      for (int j = 0; j < b.n(); j++) {
        e.zero();
        VecFill(ctx, 1, 1.0, e.p + j);
        // accumulate: several blocks may share rows
        DevBuf<double> col(ctx, b.m());
        MapApply(ctx, M, b.trans, b.scale, e.p, 0.0, col.p);
        VecAxpby(ctx, b.m(), 1.0, col.p, 1.0, out + b.row_off + (size_t)(b.col_off + j) * op.rows,
                 out + b.row_off + (size_t)(b.col_off + j) * op.rows);
        CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      }
    }
  }
}

// Mmat (n x n) += weight * op' op
void AccumulateGram(Ctx* ctx, const BlockOp& op, double weight, double* Mmat, int64_t n) {
  if (op.blks.empty()) return;
  if (RowDisjoint(op)) {
    for (auto& b : op.blks) {
      const MapData& M = *b.map;
      const double wgt = weight * b.scale * b.scale;
      double* Mb = Mmat + b.col_off + (size_t)b.col_off * n;
      if (M.kind == MK_DENSE) {
        DevBuf<double> G(ctx, (size_t)b.n() * b.n());
        if (!b.trans) DenseGramTN(ctx, M.dense.p, M.m, M.m, M.n, wgt, G.p, b.n(), false);
        else DenseGramNT(ctx, M.dense.p, M.m, M.m, M.n, nullptr, wgt, G.p, b.n(), false);
        // add G into the (col_off, col_off) block of Mmat
        AddBlock(ctx, b.n(), G.p, b.n(), Mb, n);
        CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      } else if (M.kind == MK_SPARSE) {
        CsrGramAtomic(ctx, b.trans ? M.csr_t : M.csr, wgt, Mmat, n, b.col_off);
      } else if (M.kind == MK_SCALAR) {
        DenseAddDiagonal(ctx, Mb, n, b.n(), nullptr, wgt * M.alpha * M.alpha);
      } else if (M.kind == MK_DIAGONAL) {
        std::vector<double> d2(M.n);
        for (int i = 0; i < M.n; i++) d2[i] = wgt * M.h_diag[i] * M.h_diag[i];
        DevBuf<double> dd(ctx, M.n);
        dd.upload(d2.data(), M.n);
        DenseAddDiagonal(ctx, Mb, n, b.n(), dd.p, 0.0);
        CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      } else {
        throw Error(EPS_ERR_INVALID, "Gram of a Kronecker block is not supported");
      }
    }
  } else {
    EPS_CHECK((double)op.rows * op.cols < 2e9, "row-overlapping blocks too large to densify (%lld x %lld)",
              (long long)op.rows, (long long)op.cols);
    DevBuf<double> D(ctx, (size_t)op.rows * op.cols);
    Densify(ctx, op, D.p);
    DenseGramTN(ctx, D.p, op.rows, op.rows, (int)op.cols, weight, Mmat, n, true);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }
}

__global__ void AddBlockKernel(int nb, const double* __restrict__ G, int64_t ldg, double* M, int64_t ldm) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y;
  if (i < nb && j < nb) M[i + (int64_t)j * ldm] += G[i + (int64_t)j * ldg];
}

}  // namespace

void AddBlock(Ctx* ctx, int nb, const double* G, int64_t ldg, double* M, int64_t ldm) {
  if (nb <= 0) return;
  AddBlockKernel<<<dim3((nb + 255) / 256, nb), 256, 0, ctx->stream>>>(nb, G, ldg, M, ldm);
  ctx->count_launch();
}

KktSolver::~KktSolver() {
  if (h_scal) cudaFreeHost(h_scal);
  if (ring) cudaFreeHost(ring);
}

// c0 = -2a H'g - c
void KktSolver::ComputeC0(const std::vector<double>& g_host) {
  if (!c0.p) c0.alloc(ctx, n);
  c0.zero();
  if (alpha2 != 0.0 && !g_host.empty() && rows_h > 0) {
    EPS_CHECK((int64_t)g_host.size() == rows_h, "constant of the argument has %zu entries, expected %lld", g_host.size(),
              (long long)rows_h);
    DevBuf<double> g(ctx, rows_h);
    g.upload(g_host.data(), rows_h);
    H.ApplyT(g.p, c0.p);
    if (h_sharded) AllReduceSum(ctx, c0.p, n);
    VecScale(ctx, n, -alpha2, c0.p, c0.p);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }
  if (!c_host_.empty()) {
    DevBuf<double> c(ctx, n);
    c.upload(c_host_.data(), n);
    VecAxpby(ctx, n, 1.0, c0.p, -1.0, c.p, c0.p);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }
}

void KktSolver::SetConstant(const std::vector<double>& g_host) { ComputeC0(g_host); }

void KktSolver::Init(const std::vector<double>& g_host, const std::vector<double>& c_host) {
  EPS_CHECK(n > 0, "KKT prox without variables");
  H.Compile();
  A.Compile();
  tn1.alloc(ctx, n);
  tn2.alloc(ctx, n);
  th1.alloc(ctx, std::max<int64_t>(rows_h, 1));
  th2.alloc(ctx, std::max<int64_t>(rows_h, 1));
  ta1.alloc(ctx, std::max<int64_t>(rows_a, 1));
  scal.alloc(ctx, 8);
  scal.zero();
  CUDA_CHECK(cudaMallocHost(&h_scal, 8 * sizeof(double)));

  c_host_ = c_host;
  ComputeC0(g_host);
  // ---- Kronecker structure: k identical diagonal blocks ----
  if (alpha2 != 0.0 && H.blks.size() == 1 && H.blks[0].map->kind == MK_KRONECKER && !H.blks[0].trans &&
      H.blks[0].map->ka->kind == MK_SCALAR && !h_sharded && H.blks[0].row_off == 0 && H.blks[0].col_off == 0 &&
      H.blks[0].m() == rows_h && H.blks[0].n() == n && !getenv("EPS_NO_KRON_KKT")) {
    const Blk& hb = H.blks[0];
    const MapData& K = *hb.map;
    const int k = K.ka->n;
    const int64_t nb = n / k, mb = rows_h / k;
    std::vector<double> da_all;
    bool rep = A.GramDiagonal(&da_all);
    for (int64_t j = 0; rep && j < n; j++) rep = da_all[j] == da_all[j % nb];
    if (rep && k >= 1) {
      inner = std::make_unique<KktSolver>();
      inner->ctx = ctx;
      inner->mode = mode;
      inner->n = nb;
      inner->rows_h = mb;
      inner->rows_a = nb;
      inner->alpha2 = alpha2;
      inner->cg_tol = cg_tol;
      inner->cg_max_iter = cg_max_iter;
      inner->gram_max_n = gram_max_n;
      inner->H.ctx = ctx; inner->H.rows = mb; inner->H.cols = nb;
      Blk ib;
      ib.map = K.kb;
      ib.trans = K.kb_trans;
      ib.scale = hb.scale * K.ka->alpha;
      ib.row_off = 0; ib.col_off = 0;
      inner->H.blks.push_back(ib);
      // A'A of one block as a diagonal map sqrt(d)
      auto dm = std::make_shared<MapData>();
      dm->ctx = ctx; dm->kind = MK_DIAGONAL; dm->m = dm->n = (int)nb;
      dm->h_diag.resize(nb);
      for (int64_t j = 0; j < nb; j++) dm->h_diag[j] = std::sqrt(da_all[j]);
      dm->diag.alloc(ctx, nb);
      dm->diag.upload(dm->h_diag.data(), nb);
      CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
      Blk ab;
      ab.map = dm; ab.trans = false; ab.scale = 1.0; ab.row_off = 0; ab.col_off = 0;
      inner->A.ctx = ctx; inner->A.rows = nb; inner->A.cols = nb;
      inner->A.blks.push_back(ab);
      inner->carry_enabled = k == 1;
      inner->Init(std::vector<double>(), std::vector<double>());
      kron_k = k;
      mode = inner->mode;
      description = Format("kron x%d of [%s]", k, inner->description.c_str());
      return;
    }
  }

  std::vector<double> dh, da;
  const bool h_diag = (alpha2 == 0.0 || H.blks.empty()) ? (dh.assign(n, 0.0), true) : H.GramDiagonal(&dh);
  if (a_sharded) {
    A_sh.Compile();
    A_rep.Compile();
    EPS_CHECK(mode == EPS_KKT_AUTO || mode == EPS_KKT_GRAM || mode == EPS_KKT_CG,
              "row-sharded constraints need kkt mode AUTO, GRAM or CG");
  }
  a_diag = !a_sharded && A.GramDiagonal(&da);
  bool a_pos = a_diag;
  if (a_diag) for (double v : da) if (!(v > 0.0)) { a_pos = false; break; }

  if (mode == EPS_KKT_AUTO) {
    if (h_diag && a_diag && !h_sharded) mode = EPS_KKT_DIAGONAL;
    else if (a_pos && rows_h > 0 && rows_h < n && rows_h <= gram_max_n && !h_sharded) mode = EPS_KKT_WOODBURY;
    else if (n <= gram_max_n) mode = EPS_KKT_GRAM;
    else mode = EPS_KKT_CG;
    // the dense Gram of row-overlapping blocks goes through a densified copy: matrix-free
    // when that copy would not fit
    if (mode == EPS_KKT_GRAM &&
        ((alpha2 != 0.0 && !RowDisjoint(H) && (double)H.rows * H.cols >= 2e9) ||
         (!RowDisjoint(A) && (double)A.rows * A.cols >= 2e9)))
      mode = EPS_KKT_CG;
  }

  if (mode == EPS_KKT_DIAGONAL) {
    EPS_CHECK(h_diag && a_diag && !h_sharded, "DIAGONAL mode needs scalar/diagonal blocks");
    std::vector<double> di(n);
    for (int64_t i = 0; i < n; i++) {
      const double d = alpha2 * dh[i] + da[i];
      EPS_CHECK(d > 0.0, "singular diagonal KKT system at %lld", (long long)i);
      di[i] = 1.0 / d;
    }
    dinv.alloc(ctx, n);
    dinv.upload(di.data(), n);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    description = Format("kkt=DIAGONAL n=%lld", (long long)n);
  } else if (mode == EPS_KKT_WOODBURY) {
    EPS_CHECK(a_pos && rows_h > 0 && !h_sharded, "WOODBURY mode needs diagonal positive A'A and an H block");
    std::vector<double> di(n);
    for (int64_t i = 0; i < n; i++) di[i] = 1.0 / da[i];
    dinv.alloc(ctx, n);
    dinv.upload(di.data(), n);
    // S = I/(2a) + H D^-1 H'
    DevBuf<double> S(ctx, (size_t)rows_h * rows_h);
    const bool single_dense = H.blks.size() == 1 && H.blks[0].map->kind == MK_DENSE &&
                              !H.blks[0].trans && H.blks[0].row_off == 0 && H.blks[0].col_off == 0 &&
                              H.blks[0].m() == rows_h && H.blks[0].n() == n;
    if (single_dense) {
      const Blk& b = H.blks[0];
      DenseGramNT(ctx, b.map->dense.p, b.map->m, (int)rows_h, n, dinv.p, b.scale * b.scale, S.p, rows_h, false);
    } else {
      EPS_CHECK((double)rows_h * n < 2e9, "H too large to densify for WOODBURY (%lld x %lld)",
                (long long)rows_h, (long long)n);
      DevBuf<double> D(ctx, (size_t)rows_h * n);
      Densify(ctx, H, D.p);
      DenseGramNT(ctx, D.p, rows_h, (int)rows_h, n, dinv.p, 1.0, S.p, rows_h, false);
      CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    }
    DenseAddDiagonal(ctx, S.p, rows_h, (int)rows_h, nullptr, 1.0 / alpha2);
    W.alloc(ctx, (size_t)rows_h * rows_h);
    DenseCholInverseFactor(ctx, S.p, rows_h, (int)rows_h, W.p);
    description = Format("kkt=WOODBURY n=%lld factor=%lldx%lld", (long long)n, (long long)rows_h, (long long)rows_h);
  } else if (mode == EPS_KKT_GRAM) {
    DevBuf<double> Mmat(ctx, (size_t)n * n);
    Mmat.zero();
    // the parts whose rows are sharded first, one all-reduce for them, then the replicated parts
    if (alpha2 != 0.0 && h_sharded) AccumulateGram(ctx, H, alpha2, Mmat.p, n);
    if (a_sharded) AccumulateGram(ctx, A_sh, 1.0, Mmat.p, n);
    if (h_sharded || a_sharded) AllReduceSum(ctx, Mmat.p, n * n);
    if (alpha2 != 0.0 && !h_sharded) AccumulateGram(ctx, H, alpha2, Mmat.p, n);
    AccumulateGram(ctx, a_sharded ? A_rep : A, 1.0, Mmat.p, n);
    W.alloc(ctx, (size_t)n * n);
    DenseCholInverseFactor(ctx, Mmat.p, n, (int)n, W.p);
    description = Format("kkt=GRAM n=%lld factor=%lldx%lld%s", (long long)n, (long long)n, (long long)n,
                         (h_sharded || a_sharded) ? " sharded (Gram all-reduced)" : "");
  } else if (mode == EPS_KKT_CG) {
    if (a_diag) {
      ddiag.alloc(ctx, n);
      ddiag.upload(da.data(), n);
    }
    cg_r.alloc(ctx, n);
    cg_p.alloc(ctx, n);
    cg_w.alloc(ctx, n);
    cg_done.alloc(ctx, 1);
    cg_done.zero();
    CUDA_CHECK(cudaHostAlloc(&ring, kCgRing * sizeof(CgResult), cudaHostAllocMapped));
    memset(ring, 0, kCgRing * sizeof(CgResult));
    {
      int per_sm = 0;
      CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, CgTailKernel, kTailBlock, 0));
      EPS_CHECK(per_sm >= 1, "CgTailKernel does not fit an SM");
      const int64_t want = (n + kTailBlock - 1) / kTailBlock;
      tail_grid = (int)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)ctx->sm_count * per_sm));
      tail_grid = std::min(tail_grid, ctx->sm_count);
      cg_partials.alloc(ctx, 4 * (size_t)tail_grid);
    }
    // fused matvec: M = c B'B + diag(d) with ONE SELL-packed sparse block B over all of x.
    //   (i)  B is the argument of the SUM_SQUARE (H) and A'A is diagonal           -- the lasso
    //   (ii) B is the data block of the constraints (A), every other block of A and all of H are
    //        scalar / diagonal                                                      -- logistic regression, SVM
    auto sell_block = [&](const Blk& b) {
      return b.map->kind == MK_SPARSE && b.col_off == 0 && b.n() == n && b.map->sell.built && b.map->sell_t.built;
    };
    fast_h = a_diag && alpha2 != 0.0 && H.blks.size() == 1 && sell_block(H.blks[0]) && H.blks[0].row_off == 0 &&
             H.blks[0].m() == rows_h;
    if (fast_h) {
      fb = H.blks[0];
      fb_coef = alpha2;
      fb_sharded = h_sharded;
    } else if (h_diag && !h_sharded) {
      const BlockOp& src = a_sharded ? A_sh : A;
      int n_sparse = 0;
      for (auto& b : src.blks) if (!b.diag_like()) n_sparse++;
      BlockOp rest;
      rest.ctx = ctx; rest.rows = A.rows; rest.cols = n;
      for (auto& b : A.blks) if (b.diag_like()) rest.blks.push_back(b);
      std::vector<double> dr;
      bool ok = n_sparse == 1 && (!a_sharded || A_sh.blks.size() == 1) && (int)A.blks.size() == (int)rest.blks.size() + 1;
      if (ok) for (auto& b : src.blks) if (!b.diag_like()) ok = sell_block(b), fb = b;
      if (ok) for (auto& b : A_rep.blks) ok = ok && b.diag_like();
      if (ok && rest.GramDiagonal(&dr)) {
        fast_h = true;
        fb_coef = 1.0;
        fb_sharded = a_sharded;
        da.assign(n, 0.0);
        for (int64_t j = 0; j < n; j++) da[j] = alpha2 * dh[j] + dr[j];
        ddiag.alloc(ctx, n);
        ddiag.upload(da.data(), n);
      }
    }
    if (fast_h) tb1.alloc(ctx, std::max<int64_t>(fb.m(), 1));
    if (fast_h && fb_sharded) peer = PeerEnsure(ctx, n);
    fused_tail = fast_h && (!fb_sharded || peer) && tail_grid <= kPeerMaxCtas && !getenv("EPS_NO_FUSED_TAIL");
    if (fused_tail) {
      int per_sm = 0;
      CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, CgFusedTailKernel, kTailBlock, 0));
      EPS_CHECK(per_sm >= 1, "CgFusedTailKernel does not fit an SM");
      tail_v2 = !getenv("EPS_TAIL_V1") && fb.map->sell.n_cta <= ctx->sm_count &&
                fb.map->sell_t.n_cta <= ctx->sm_count;
      if (tail_v2) {
        rhs_prev.alloc(ctx, n);
        cg_bar.alloc(ctx, 4);
        cg_bar.zero();
      }
    }
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    description = Format("kkt=CG n=%lld rows_h=%lld tol=%.1e%s%s%s", (long long)n, (long long)rows_h, cg_tol,
                         (h_sharded || a_sharded) ? " sharded" : "",
                         fast_h ? (fb_coef == alpha2 && alpha2 != 0.0 ? " fused-matvec" : " fused-matvec(constraint block)") : "",
                         peer ? " peer-allreduce" : "");
    if (fused_tail) description += tail_v2 ? " fused-tail(one-barrier, pre-phase, residual-carry)" : " fused-tail";
  } else {
    throw Error(EPS_ERR_INVALID, Format("unknown kkt mode %d", mode));
  }
}

void KktSolver::MatVec(const double* p, double* w, const int* skip) {
  if (fast_h) {
    // w = 2a s^2 H'(H p) + d.*p with the diagonal term fused into the reduction of the second
    // pass.  Row-sharded: every rank contributes its rows' part, rank 0 alone adds d.*p, and
    // the all-reduce completes the sum (identical bits on every rank).
    const Blk& b = fb;
    const MapData& M = *b.map;
    SpmvExtra e1, e2;
    e1.skip = skip;
    e2.skip = skip;
    pending_seq = 0;
    if (peer) {
      // push this rank's part to every rank; the kernel that consumes w (CgStart / CgTail)
      // completes the sum and adds d.*p.  Never skipped: see CgTailKernel.
      e2.publish = &peer->x;
      e2.publish_seq = pending_seq = ++peer->seq;
      e2.publish_counter = peer->counter;
    } else if (!fb_sharded || ctx->rank == 0) {
      e2.ep_d = ddiag.p;
      e2.ep_p = p;
    }
    SellApply(ctx, b.trans ? M.sell_t : M.sell, b.scale, p, 0.0, tb1.p, b.trans ? KC_SPMV_T : KC_SPMV_N, &e1);
    SellApply(ctx, b.trans ? M.sell : M.sell_t, fb_coef * b.scale, tb1.p, 0.0, w,
              b.trans ? KC_SPMV_N : KC_SPMV_T, &e2);
    if (fb_sharded && !peer) AllReduceSum(ctx, w, n);
    return;
  }
  bool have = false;
  if (alpha2 != 0.0 && !H.blks.empty()) {
    H.Apply(p, th1.p);
    H.ApplyT(th1.p, w);
    if (h_sharded) AllReduceSum(ctx, w, n);
    have = true;
  }
  if (a_diag) {
    if (have) VecMulAdd(ctx, n, ddiag.p, p, alpha2, w, w);
    else VecMulAdd(ctx, n, ddiag.p, p, 0.0, nullptr, w);
  } else if (a_sharded) {
    // sharded rows: this rank's part of A_sh'(A_sh p), summed over the ranks; then the replicated rows
    A_sh.Apply(p, ta1.p);
    A_sh.ApplyT(ta1.p, tn2.p);
    AllReduceSum(ctx, tn2.p, n);
    if (have) VecAxpby(ctx, n, alpha2, w, 1.0, tn2.p, w);
    else VecCopy(ctx, n, tn2.p, w);
    if (!A_rep.blks.empty()) {
      A_rep.Apply(p, ta1.p);
      A_rep.ApplyT(ta1.p, w, true);
    }
  } else {
    if (have) VecScale(ctx, n, alpha2, w, w);
    A.Apply(p, ta1.p);
    A.ApplyT(ta1.p, w, have);
  }
}

bool KktSolver::SolveMulti(const double* rhs, double* x, int k) {
  if (mode == EPS_KKT_GRAM) {
    if (mn1.n < (size_t)n * k) mn1.alloc(ctx, (size_t)n * k);
    DenseGemmRhs(ctx, W.p, n, (int)n, (int)n, false, 1.0, rhs, n, k, 0.0, mn1.p, n);
    DenseGemmRhs(ctx, W.p, n, (int)n, (int)n, true, 1.0, mn1.p, n, k, 0.0, x, n);
    return true;
  }
  if (mode == EPS_KKT_WOODBURY) {
    const bool single_dense = H.blks.size() == 1 && H.blks[0].map->kind == MK_DENSE && H.blks[0].row_off == 0 &&
                              H.blks[0].col_off == 0 && H.blks[0].m() == rows_h && H.blks[0].n() == n;
    if (!single_dense) return false;
    const Blk& b = H.blks[0];
    const MapData& M = *b.map;
    if (mn1.n < (size_t)n * k) { mn1.alloc(ctx, (size_t)n * k); mn2.alloc(ctx, (size_t)n * k); }
    if (mh1.n < (size_t)rows_h * k) { mh1.alloc(ctx, (size_t)rows_h * k); mh2.alloc(ctx, (size_t)rows_h * k); }
    // X = D^-1 R - D^-1 H' S^-1 H D^-1 R,  S^-1 = W'W   (the single-column path of Solve, k columns at once)
    DiagMulti(ctx, n, k, dinv.p, rhs, 0.0, nullptr, mn1.p);
    DenseGemmRhs(ctx, M.dense.p, M.m, M.m, M.n, b.trans, b.scale, mn1.p, n, k, 0.0, mh1.p, rows_h);
    DenseGemmRhs(ctx, W.p, rows_h, (int)rows_h, (int)rows_h, false, 1.0, mh1.p, rows_h, k, 0.0, mh2.p, rows_h);
    DenseGemmRhs(ctx, W.p, rows_h, (int)rows_h, (int)rows_h, true, 1.0, mh2.p, rows_h, k, 0.0, mh1.p, rows_h);
    DenseGemmRhs(ctx, M.dense.p, M.m, M.m, M.n, !b.trans, b.scale, mh1.p, rows_h, k, 0.0, mn2.p, n);
    DiagMulti(ctx, n, k, dinv.p, mn2.p, -1.0, mn1.p, mn2.p);   // d.*(H'z) - tn1 = -x
    VecScale(ctx, n * k, -1.0, mn2.p, x);
    return true;
  }
  return false;
}

void KktSolver::Solve(const double* rhs, double* x) {
  if (inner) {
    const int64_t nb = n / kron_k;
    const int64_t before = inner->cg_iters_total;
    static const bool no_multi = getenv("EPS_NO_KRON_GEMM") != nullptr;
    if (!no_multi && kron_k > 1 && inner->SolveMulti(rhs, x, kron_k)) return;
    for (int c = 0; c < kron_k; c++) inner->Solve(rhs + (size_t)c * nb, x + (size_t)c * nb);
    cg_iters_total += inner->cg_iters_total - before;
    cg_unconverged = inner->cg_unconverged;
    return;
  }
  switch (mode) {
    case EPS_KKT_DIAGONAL:
      KktDiagSolve(ctx, n, dinv.p, rhs, x);
      break;
    case EPS_KKT_WOODBURY: {
      // x = D^-1 r - D^-1 H' S^-1 H D^-1 r,  S^-1 = W'W
      KktDiagSolve(ctx, n, dinv.p, rhs, tn1.p);
      H.Apply(tn1.p, th1.p);
      DenseGemv(ctx, W.p, (int)rows_h, (int)rows_h, (int)rows_h, false, 1.0, th1.p, 0.0, th2.p, 1);
      DenseGemv(ctx, W.p, (int)rows_h, (int)rows_h, (int)rows_h, true, 1.0, th2.p, 0.0, th1.p, 1);
      H.ApplyT(th1.p, tn2.p);
      VecMulAdd(ctx, n, dinv.p, tn2.p, -1.0, tn1.p, tn2.p);  // tn2 = dinv.*(H'z) - tn1 = -x
      VecScale(ctx, n, -1.0, tn2.p, x);
      break;
    }
    case EPS_KKT_GRAM:
      DenseGemv(ctx, W.p, (int)n, (int)n, (int)n, false, 1.0, rhs, 0.0, tn1.p, 1);
      DenseGemv(ctx, W.p, (int)n, (int)n, (int)n, true, 1.0, tn1.p, 0.0, x, 1);
      break;
    case EPS_KKT_CG: {
      // Conjugate gradients, warm-started from x.  The host never waits for the iteration it
      // has just launched: it launches iteration k, then reads the published result of
      // iteration k-1 from the host-mapped ring, so the device always has work queued.  When
      // that result says "converged", iteration k was already a no-op (cg_done is raised on
      // the device by the kernel that detected convergence).
      ReduceScratch rs = GetReduceScratch(ctx);
      const int grid = std::min(1024, GridFor(n, kBlock, kUnroll, ctx->sm_count, 4));
      const long long base = cg_seq;
      auto wait_result = [&](long long seq) -> const CgResult& {
        volatile CgResult* e = &ring[seq % kCgRing];
        long long spins = 0;
        while (e->seq != seq) {
          if ((++spins & 0xFFFFF) == 0) {  // surface device faults instead of spinning forever
            cudaError_t q = cudaStreamQuery(ctx->stream);
            if (q != cudaSuccess && q != cudaErrorNotReady) CUDA_CHECK(q);
            if (q == cudaSuccess && e->seq != seq)
              throw Error(EPS_ERR_CUDA, "CG result was never published");
          }
        }
        return *const_cast<const CgResult*>(e);
      };
      static const bool no_carry = getenv("EPS_NO_CARRY") != nullptr;
      // every 8th solve recomputes the true residual (bounds the drift of the carried recurrence)
      const bool carry = tail_v2 && carry_enabled && carry_valid && !no_carry && (carry_count % 8) != 0;
      if (carry) {
        ProfScope ps(ctx, KC_VEC, 48.0 * n);
        CgCarryStartKernel<<<grid, kBlock, 0, ctx->stream>>>(n, rhs, rhs_prev.p, cg_r.p, cg_p.p, cg_tol, scal.p,
                                                             rs.partials, rs.counter, cg_done.p, ring, base);
        ctx->count_launch();
      } else {
        MatVec(x, cg_w.p, nullptr);
        CgStartKernel<<<grid, kBlock, 0, ctx->stream>>>(n, rhs, cg_w.p, cg_r.p, cg_p.p, cg_tol, scal.p,
                                                         rs.partials, rs.counter, cg_done.p, ring, base,
                                                         peer ? peer->x : PeerXchg(), pending_seq, ddiag.p, x);
        ctx->count_launch();
        if (tail_v2) VecCopy(ctx, n, rhs, rhs_prev.p);
      }
      if (tail_v2) carry_count++;
      int it = 0;
      bool converged = false;
      for (; it < cg_max_iter; it++) {
        // profile mode times every launch with its own event pair: no speculation there, so
        // that no-op launches do not enter the per-class statistics
        if (ctx->profile && wait_result(base + it).converged != 0.0) { converged = true; break; }
        if (fused_tail) {
          // first pass with its reduction, second pass as partial planes, then ONE kernel for
          // the plane reduction, the cross-GPU exchange and the CG update
          const Blk& b = fb;
          const MapData& M = *b.map;
          const Sell& S1 = b.trans ? M.sell_t : M.sell;
          const Sell& S2 = b.trans ? M.sell : M.sell_t;
          const double* planes = nullptr;
          if (tail_v2) {
            // first pass as planes; their reduction rides in the second pass (pre-phase + grid barrier)
            const double* p1 = SellApplyPlanes(ctx, S1, cg_p.p, b.trans ? KC_SPMV_T : KC_SPMV_N, cg_done.p);
            planes = SellApplyPlanesAfter(ctx, S2, tb1.p, p1, S1.T, b.scale, cg_bar.p + 2,
                                          b.trans ? KC_SPMV_N : KC_SPMV_T, cg_done.p);
          } else {
            SpmvExtra e1;
            e1.skip = cg_done.p;
            SellApply(ctx, S1, b.scale, cg_p.p, 0.0, tb1.p, b.trans ? KC_SPMV_T : KC_SPMV_N, &e1);
            planes = SellApplyPlanes(ctx, S2, tb1.p, b.trans ? KC_SPMV_N : KC_SPMV_T, cg_done.p);
          }
          {
          ProfScope ps(ctx, peer ? KC_ALLREDUCE : KC_VEC, 8.0 * S2.T * n + 72.0 * n);
          int64_t nn = n;
          int T = S2.T;
          double alpha = fb_coef * b.scale;
          double* wp = cg_w.p;
          double *pp = cg_p.p, *xp = x, *rp = cg_r.p, *sp = scal.p, *part = cg_partials.p;
          int* dp = cg_done.p;
          CgResult* rg = ring;
          long long seq = base + it + 1;
          PeerXchg pxv = peer ? peer->x : PeerXchg();
          unsigned long long xs = peer ? ++peer->seq : 0ull;
          const double* dgp = ddiag.p;
          // EPS_TAIL_TRACE=<launch number>: per-CTA phase timeline of that launch on stderr (rank 0)
          static const int trace_at = getenv("EPS_TAIL_TRACE") ? atoi(getenv("EPS_TAIL_TRACE")) : -1;
          static int launch_no = 0;
          static DevBuf<unsigned long long> tbuf;
          unsigned long long* trace = nullptr;
          if (trace_at >= 0) {
            if (launch_no == 0) {  // allocated up front: no synchronisation next to the traced launch
              tbuf.alloc(ctx, (size_t)tail_grid * 8);
              tbuf.zero();
            }
            if (launch_no++ == trace_at) trace = tbuf.p;
          }
          bool step_trace = false;
          if (tail_v2) {
            if (unsigned long long* st = StepTraceSlot(ctx, 1, tail_grid, "tail")) {
              trace = st;
              step_trace = true;
            }
            // ordinary launch with programmatic stream serialization; one GridBarrier inside
            if (cg_partials.n < 4 * (size_t)tail_grid) cg_partials.alloc(ctx, 4 * (size_t)tail_grid);
            part = cg_partials.p;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(tail_grid);
            cfg.blockDim = dim3(kTailBlock);
            cfg.stream = ctx->stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            static const bool pdl = getenv("EPS_NO_PDL") == nullptr;
            cfg.numAttrs = pdl ? 1 : 0;
            unsigned* barp = cg_bar.p;
            CUDA_CHECK(cudaLaunchKernelEx(&cfg, CgFusedTail2Kernel, nn, T, planes, alpha, wp, pp, xp, rp, sp, part, dp,
                                          rg, seq, pxv, xs, dgp, barp, trace));
            ctx->count_launch();
          } else {
          void* args[] = {&nn, &T, &planes, &alpha, &wp, &pp, &xp, &rp, &sp, &part, &dp, &rg, &seq, &pxv, &xs, &dgp, &trace};
          // EPS_TAIL_PDL=1: cooperative + programmatic serialization (set-up overlaps the SpMV drain)
          static int tail_pdl = getenv("EPS_TAIL_PDL") ? atoi(getenv("EPS_TAIL_PDL")) : 0;
          bool launched = false;
          if (tail_pdl) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(tail_grid);
            cfg.blockDim = dim3(kTailBlock);
            cfg.stream = ctx->stream;
            cudaLaunchAttribute attr[2];
            attr[0].id = cudaLaunchAttributeCooperative;
            attr[0].val.cooperative = 1;
            attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[1].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 2;
            cudaError_t e = cudaLaunchKernelExC(&cfg, (const void*)CgFusedTailKernel, args);
            if (e == cudaSuccess) {
              launched = true;
            } else {
              (void)cudaGetLastError();
              fprintf(stderr, "[eps] cooperative + programmatic launch refused (%s): plain cooperative launch\n",
                      cudaGetErrorString(e));
              tail_pdl = 0;
            }
          }
          if (!launched)
            CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)CgFusedTailKernel, dim3(tail_grid),
                                                   dim3(kTailBlock), args, 0, ctx->stream));
          ctx->count_launch();
          }
          if (trace && !step_trace) {
            std::vector<unsigned long long> h((size_t)tail_grid * 8);
            tbuf.download(h.data(), h.size());
            unsigned long long t0 = ~0ull;
            for (int g = 0; g < tail_grid; g++) t0 = std::min(t0, h[g * 8]);
            std::string line = Format("[tail trace] rank %d G=%d T=%d n=%lld us since first CTA (median/max):", ctx->rank,
                                      peer ? peer->x.G : 1, T, (long long)n);
            const char* names[7] = {"start", "pushed", "fenced", "peers-in", "dot", "sync1", "end"};
            for (int k = 0; k < 7; k++) {
              std::vector<double> v;
              for (int g = 0; g < tail_grid; g++) if (h[g * 8 + k]) v.push_back((h[g * 8 + k] - t0) * 1e-3);
              if (v.empty()) continue;
              std::sort(v.begin(), v.end());
              line += Format(" %s %.1f/%.1f", names[k], v[v.size() / 2], v.back());
            }
            fprintf(stderr, "%s\n", line.c_str());
          }
          }
          if (wait_result(base + it).converged != 0.0) { converged = true; break; }
          continue;
        }
        MatVec(cg_p.p, cg_w.p, cg_done.p);
        {
          ProfScope ps(ctx, KC_VEC, 72.0 * n);
          int64_t nn = n;
          double* wp = cg_w.p;
          double *pp = cg_p.p, *xp = x, *rp = cg_r.p, *sp = scal.p, *part = cg_partials.p;
          int* dp = cg_done.p;
          CgResult* rg = ring;
          long long seq = base + it + 1;
          PeerXchg pxv = peer ? peer->x : PeerXchg();
          unsigned long long xs = pending_seq;
          const double* dgp = ddiag.p;
          void* args[] = {&nn, &wp, &pp, &xp, &rp, &sp, &part, &dp, &rg, &seq, &pxv, &xs, &dgp};
          CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)CgTailKernel, dim3(tail_grid),
                                                 dim3(kTailBlock), args, 0, ctx->stream));
          ctx->count_launch();
        }
        if (wait_result(base + it).converged != 0.0) { converged = true; break; }
      }
      if (!converged) {
        // it == cg_max_iter: drain the last iteration; an inexact KKT solve is counted, never silent
        if (wait_result(base + it).converged == 0.0) cg_unconverged++;
      }
      cg_seq = base + cg_max_iter + 2;
      cg_iters_total += it;
      carry_valid = tail_v2;   // cg_r is the residual of the x just returned
      break;
    }
    default:
      throw Error(EPS_ERR_INVALID, "KKT solver not initialised");
  }
}

}  // namespace eps
