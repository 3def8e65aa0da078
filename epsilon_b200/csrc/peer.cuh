// Device side of the one-shot peer all-reduce (see PeerXchg in internal.h, host side in comm.cu).
#pragma once
#include "device.cuh"
#include "internal.h"

namespace eps {

// Publisher, called by every thread of every CTA after its last store into the peers' slots:
// the last CTA to arrive raises flag [parity][rank] = seq on every rank.
__device__ __forceinline__ void PeerPublishFinish(const PeerXchg& px, unsigned long long seq,
                                                  unsigned* counter) {
  // one system-scope fence per CTA: the barrier orders every thread's remote stores before
  // thread 0's fence (fences are cumulative), which orders them before the counter update
  __syncthreads();
  __shared__ bool last_;
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned prev = atomicAdd(counter, 1u);
    last_ = prev == gridDim.x - 1;
  }
  __syncthreads();
  if (!last_) return;
  __threadfence_system();
  if (threadIdx.x < px.G) {
    const int parity = (int)(seq & 1ull);
    StReleaseSys(px.flag[threadIdx.x] + parity * px.G + px.rank, seq);
  }
  if (threadIdx.x == 0) *counter = 0;
}

// Consumer: block until all G ranks have published `seq`.  Called by every thread of the CTA.
__device__ __forceinline__ void PeerWait(const PeerXchg& px, unsigned long long seq) {
  if (threadIdx.x < px.G) {
    const unsigned long long* f = px.flag[px.rank] + (int)(seq & 1ull) * px.G + threadIdx.x;
    while (LdAcquireSys(f) < seq) {}
  }
  __syncthreads();
}

// Sum of the G published slots at element i, in rank order (identical on every rank).
// __ldcg: the slots are rewritten by peers every other exchange, L1 must not serve them.
__device__ __forceinline__ double PeerSum(const PeerXchg& px, unsigned long long seq, int64_t i) {
  const double* base = px.recv[px.rank] + (size_t)(seq & 1ull) * px.G * px.n + i;
  double s = 0.0;
  for (int r = 0; r < px.G; r++) s += __ldcg(base + (size_t)r * px.n);
  return s;
}

// ---- flag-in-data packets -----------------------------------------------------------------
__device__ __forceinline__ void LlStore(uint4* p, double v, unsigned seq) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"((unsigned)b), "r"(seq),
               "r"((unsigned)(b >> 32)), "r"(seq) : "memory");
}

// Spins until both halves carry `seq`.
__device__ __forceinline__ double LlLoad(const uint4* p, unsigned seq) {
  unsigned a, fa, b, fb;
  do {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(fa), "=r"(b), "=r"(fb) : "l"(p) : "memory");
  } while (fa != seq || fb != seq);
  return __longlong_as_double((long long)(((unsigned long long)b << 32) | a));
}

// ---- sentinel slots ------------------------------------------------------------------------------
__device__ __forceinline__ void SnStore(double* p, double v) {
  asm volatile("st.volatile.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ unsigned long long SnPeek(const double* p) {
  unsigned long long v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void SnReset(double* p) {
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(kPeerSentinel) : "memory");
}
// Waits for the slot, returns its value and hands the slot back (sentinel) for the exchange after next.
__device__ __forceinline__ double SnTake(double* p) {
  unsigned long long v;
  do { v = SnPeek(p); } while (v == kPeerSentinel);
  SnReset(p);
  return __longlong_as_double((long long)v);
}
// Sum in rank order of the G slots base[q * stride]; all loads of a poll are in flight together.
__device__ __forceinline__ double SnTakeRanks(double* base, size_t stride, int G) {
  unsigned long long v[kMaxRanks];
  bool ok;
  do {
    ok = true;
#pragma unroll
    for (int q = 0; q < kMaxRanks; q++)
      if (q < G) v[q] = SnPeek(base + (size_t)q * stride);
#pragma unroll
    for (int q = 0; q < kMaxRanks; q++)
      if (q < G) ok = ok && v[q] != kPeerSentinel;
  } while (!ok);
  double s = 0.0;
#pragma unroll
  for (int q = 0; q < kMaxRanks; q++)
    if (q < G) {
      s += __longlong_as_double((long long)v[q]);
      SnReset(base + (size_t)q * stride);
    }
  return s;
}

__device__ __forceinline__ uint4 LlPeek(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double LlValue(const uint4& v) {
  return __longlong_as_double((long long)(((unsigned long long)v.z << 32) | v.x));
}

// Sum, in rank order, of the G packets base[q * stride], q < G: all G loads are issued before any of them is
// tested, so a poll costs one L2 round trip instead of G (one packet at a time the owner step of the exchange
// took 8 dependent round trips per element: ~6 us at G = 8).
__device__ __forceinline__ double LlSumRanks(const uint4* base, size_t stride, int G, unsigned seq) {
  uint4 v[kMaxRanks];
  bool ok;
  do {
    ok = true;
#pragma unroll
    for (int q = 0; q < kMaxRanks; q++)
      if (q < G) v[q] = LlPeek(base + (size_t)q * stride);
#pragma unroll
    for (int q = 0; q < kMaxRanks; q++)
      if (q < G) ok = ok && v[q].y == seq && v[q].w == seq;
  } while (!ok);
  double s = 0.0;
#pragma unroll
  for (int q = 0; q < kMaxRanks; q++)
    if (q < G) s += LlValue(v[q]);
  return s;
}

}  // namespace eps
