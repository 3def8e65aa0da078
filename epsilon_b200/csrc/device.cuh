// Device-side helpers shared by the kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace eps {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ double WarpSum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// Sum over `W` consecutive lanes (W a power of two <= 32); every lane of the group gets it.
template <int W>
__device__ __forceinline__ double GroupSum(double v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// Programmatic dependent launch (PTX griddepcontrol): a kernel launched with
// cudaLaunchAttributeProgrammaticStreamSerialization may start while its predecessor in the stream is
// still running; GridDepWait() returns once that predecessor has completed and its writes are visible
// (a no-op for a normal launch).  GridDepLaunch() in the predecessor lets the dependent CTAs be
// scheduled as soon as resources free up instead of at its very end.
__device__ __forceinline__ void GridDepWait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void GridDepLaunch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Block-wide sum of NV values per thread; result valid in thread 0.  smem: NV*32 doubles.
template <int NV>
__device__ __forceinline__ void BlockSum(double (&v)[NV], double* smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; k++) v[k] = WarpSum(v[k]);
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) smem[k * 32 + warp] = v[k];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) {
      double t = lane < nwarp ? smem[k * 32 + lane] : 0.0;
      v[k] = WarpSum(t);
    }
  }
}

// Grid-wide deterministic finish: every CTA deposits NV partial sums; the last CTA to
// arrive adds them in CTA order and writes/accumulates out[k].  `partials` holds
// NV*gridDim.x doubles, `counter` one zero-initialised unsigned that is reset on exit.
// op: 0 = overwrite, 1 = accumulate into out.
template <int NV>
__device__ __forceinline__ void GridSumFinish(double (&v)[NV], double* smem, double* partials,
                                              unsigned* counter, double* out, int accumulate) {
  BlockSum<NV>(v, smem);
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) partials[k * gridDim.x + blockIdx.x] = v[k];
    __threadfence();
    unsigned prev = atomicAdd(counter, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    double acc[NV];
#pragma unroll
    for (int k = 0; k < NV; k++) {
      double t = 0;
      for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x)
        t += __ldcg(&partials[k * gridDim.x + i]);
      acc[k] = t;
    }
    BlockSum<NV>(acc, smem);
    if (threadIdx.x == 0) {
#pragma unroll
      for (int k = 0; k < NV; k++) out[k] = accumulate ? out[k] + acc[k] : acc[k];
      *counter = 0;
    }
  }
}

// Like GridSumFinish but leaves the totals in v[] of thread 0 of the LAST CTA to arrive and
// returns true there (false everywhere else), so that the caller can append scalar work.
template <int NV>
__device__ __forceinline__ bool GridSumLast(double (&v)[NV], double* smem, double* partials,
                                            unsigned* counter) {
  BlockSum<NV>(v, smem);
  __shared__ bool is_last_;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) partials[k * gridDim.x + blockIdx.x] = v[k];
    __threadfence();
    unsigned prev = atomicAdd(counter, 1u);
    is_last_ = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last_) return false;
  __threadfence();
  double acc[NV];
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double t = 0;
    for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) t += __ldcg(&partials[k * gridDim.x + i]);
    acc[k] = t;
  }
  BlockSum<NV>(acc, smem);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) v[k] = acc[k];
    *counter = 0;
    return true;
  }
  return false;
}

// Streaming 128-bit loads that do not pollute L1 (data read exactly once).
__device__ __forceinline__ double2 LdStream2(const double* p) {
  double2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];"
               : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ double LdStream(const double* p) {
  double r;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ int4 LdStreamI4(const int32_t* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ int2 LdStreamI2(const int32_t* p) {
  int2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0, %1}, [%2];"
               : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ int LdStreamI(const int32_t* p) {
  int r;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

__device__ __forceinline__ unsigned long long LdAcquireSys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void StReleaseSys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Grid-wide barrier for kernels whose CTAs are all co-resident (grid <= number of SMs, one CTA per SM
// fits next to whatever the programmatic launch overlaps with): bar[0] counts arrivals, bar[1] is the
// generation.  Unlike cooperative_groups::grid_group::sync() it does not need a cooperative launch, so
// the kernel can be launched with programmatic stream serialization (its prologue overlaps the drain
// of its predecessor).  Both words start at zero; the barrier leaves bar[0] at zero again.
__device__ __forceinline__ void GridBarrier(unsigned* bar) {
  __syncthreads();
  if (threadIdx.x == 0) {
    volatile unsigned* gen = bar + 1;
    const unsigned g = *gen;
    __threadfence();
    if (atomicAdd(bar, 1u) == gridDim.x - 1) {
      bar[0] = 0;
      __threadfence();
      atomicAdd(bar + 1, 1u);
    } else {
      while (*gen == g) {}
    }
    __threadfence();
  }
  __syncthreads();
}

inline int GridFor(int64_t n, int block, int per_thread, int sm_count, int max_waves = 8) {
  int64_t g = (n + (int64_t)block * per_thread - 1) / ((int64_t)block * per_thread);
  int64_t cap = (int64_t)sm_count * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace eps
