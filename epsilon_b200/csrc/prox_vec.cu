// Multi-argument and sort-based proximal operators on the stacked, reduced argument vector
// V = [arg0 | arg1 | ...] (VectorProx::ApplyVector of the reference after PreProcessInput,
// src/epsilon/prox/vector_prox.cc:131-137):
//   MaxProx / MaxEpigraph            src/epsilon/prox/max.cc:8-41, 46-85
//   SumSquareEpigraph                src/epsilon/prox/sum_square.cc:42-57 (+ LargestRealCubicRoot, newton.cc:293-323)
//   ScaledZoneEpigraph               src/epsilon/prox/scaled_zone.cc:152-279 (NORM_1, SUM_HINGE, SUM_DEADZONE,
//                                    SUM_QUANTILE with constant alpha/beta)
//   SumKLDivProx                     src/epsilon/prox/sum_kl_div.cc:6-67
// The reference sorts (or quick-selects) on one thread and accumulates sequentially; here the
// thresholds come from a device radix sort + prefix sums + a parallel search for the first
// index at which the reference's loop would stop.  Results differ from the reference by the
// summation order of the prefix sums only (~1e-16 relative).  No host synchronisation: every
// decision stays in device scalars, so the stage can sit inside a captured iteration.
#include <cub/cub.cuh>
#include <thrust/complex.h>

#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

constexpr int kBlock = 256;

// scal layout (device doubles)
enum { SC_NORM2 = 0, SC_T = 1, SC_LAM = 2, SC_FVAL = 3, SC_SKIP = 4, SC_COUNT = 8 };

inline int Grid(int64_t n) { return (int)std::max<int64_t>(1, std::min<int64_t>((n + kBlock - 1) / kBlock, 148 * 8)); }

// ---- SUM_KL_DIV: per-element Newton on yhat (sum_kl_div.cc:6-52) -------------------------
__global__ void KlDivKernel(int64_t n, double lam, double* __restrict__ V) {
  const double eps = 1e-13;
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    const double u = V[i], v = V[n + i];
    if (fabs(u) < eps * eps && fabs(v) < eps * eps) continue;  // x = u, y = v
    double yhat = fmax((0.5 + lam - v) / lam, eps);
    for (int it = 0; it < 1000; it++) {
      const double f = lam * yhat * yhat + (v - lam) * yhat - u + lam * log(yhat);
      const double F = 2 * lam * yhat + (v - lam) + lam / yhat;
      const double res = f / F;
      if (fabs(res) < eps || (yhat <= eps * 2 && res > 0) || (lam * yhat + v - lam <= eps * 2 && res > 0)) break;
      yhat = yhat - res;
      if (yhat < eps) yhat = eps;
      if (lam * yhat + v - lam < eps) yhat = (eps + lam - v) / lam;
    }
    const double y = lam * yhat + v - lam;
    V[n + i] = y;
    V[i] = y * yhat;
  }
}

// ---- SUM_SQUARE epigraph ------------------------------------------------------------------
// Durand-Kerner from the reference's starting points (newton.cc:293-323).
__device__ double LargestRealCubicRoot(double b, double c, double d) {
  typedef thrust::complex<double> cd;
  const double eps = 1e-12;
  cd p(0.4, 0.9), q = p * p, r = p * p * p;
  for (int it = 0; it < 100; it++) {
    const cd fp = p * p * p + b * p * p + c * p + d;
    const cd np = p - fp / ((p - q) * (p - r));
    const cd fq = q * q * q + b * q * q + c * q + d;
    const cd nq = q - fq / ((q - p) * (q - r));
    const cd fr = r * r * r + b * r * r + c * r + d;
    const cd nr = r - fr / ((r - p) * (r - q));
    if (thrust::abs(fp) < eps && thrust::abs(fq) < eps && thrust::abs(fr) < eps) break;
    p = np; q = nq; r = nr;
  }
  double m = -1e41;
  if (fabs(p.imag()) < eps && p.real() > m) m = p.real();
  if (fabs(q.imag()) < eps && q.real() > m) m = q.real();
  if (fabs(r.imag()) < eps && r.real() > m) m = r.real();
  return m;
}

// scal[SC_NORM2] = |u|^2 on entry; lam = max(root, 0); t = s + lam
__global__ void SumSquareEpiRootKernel(int64_t n, double* __restrict__ V, double* __restrict__ scal) {
  const double s = V[n];
  double lam = LargestRealCubicRoot(1 + s, 0.25 + s, (s - scal[SC_NORM2]) / 4);
  if (lam < 0) lam = 0;
  scal[SC_LAM] = lam;
  V[n] = s + lam;
}

__global__ void SumSquareEpiScaleKernel(int64_t n, double* __restrict__ V, const double* __restrict__ scal) {
  const double lam = scal[SC_LAM];
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock)
    V[i] = V[i] / (1 + 2 * lam);
}

// ---- common -------------------------------------------------------------------------------
// scal[SC_T] = s (the epigraph variable), idx[0] = "not found"
__global__ void PrepKernel(int64_t n0, int epigraph, const double* __restrict__ V, double* __restrict__ scal,
                           int* __restrict__ idx) {
  if (epigraph) scal[SC_T] = V[n0];
  idx[0] = 0x7FFFFFFF;
}

// ---- MAX ----------------------------------------------------------------------------------
// Y sorted descending, S its inclusive prefix sums.  The reference loop (max.cc:24-30) stops at
// the first i with y_i * i < -lam + sum_{j<i} y_j; epigraph (max.cc:66-72): (1+i)(y_i - s) <
// sum_{j<i} (y_j - s).
__global__ void MaxBreakKernel(int64_t n, int epigraph, double lam, const double* __restrict__ Y,
                               const double* __restrict__ S, const double* __restrict__ scal, int* __restrict__ idx) {
  const double s = epigraph ? scal[SC_T] : 0.0;
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    const double before = i > 0 ? S[i - 1] : 0.0;
    bool stop;
    if (epigraph) stop = (1.0 + (double)i) * (Y[i] - s) < before - (double)i * s;
    else stop = Y[i] * (double)i < -lam + before;
    if (stop) atomicMin(idx, (int)i);
  }
}

__global__ void MaxApplyKernel(int64_t n, int epigraph, double lam, double* __restrict__ V,
                               const double* __restrict__ Y, const double* __restrict__ S,
                               const double* __restrict__ scal, const int* __restrict__ idx) {
  const int64_t k = min((int64_t)idx[0], n);
  const double sk = k > 0 ? S[k - 1] : 0.0;
  if (epigraph) {
    const double s = scal[SC_T];
    if (s >= Y[0]) return;  // already inside: x = v, t = s
    const double delta = k > 0 ? (sk - (double)k * s) / (1.0 + (double)k) : 0.0;
    const double t = s + delta;
    for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
      const double v = V[i];
      V[i] = v - fmax(0.0, v - t);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) V[n] = t;
  } else {
    const double t = k > 0 ? (-lam + sk) / (double)k : 0.0;
    for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock)
      V[i] = fmin(V[i], t);
  }
}

// ---- scaled-zone epigraph -------------------------------------------------------------------
__device__ __forceinline__ double ScaledZone1(double x, double lam, double M, double a, double b) {
  if (fabs(x) <= M) return x;
  if (x > M + lam * a) return x - lam * a;
  if (x < -M - lam * b) return x + lam * b;
  return x > 0 ? M : -M;
}

// key = |(x -+ M) / slope|, w2 = slope^2, contribution to f; filtered entries get key = w2 = 0
// (scaled_zone.cc:172-187)
__global__ void ZoneKeyKernel(int64_t n, double M, double a, double b, const double* __restrict__ V,
                              double* __restrict__ key, double* __restrict__ w2, double* __restrict__ fc) {
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    const double y = V[i];
    double k = 0.0, w = 0.0, f = 0.0;
    if (!(fabs(y) <= M || (y > 0 && a == 0) || (y < 0 && b == 0))) {
      if (y > M) f = a * (y - M);
      else if (y < -M) f = b * (-y - M);
      if (y > 0) { k = (y - M) / a; w = a * a; }
      else { k = -(y + M) / b; w = b * b; }
    }
    key[i] = k; w2[i] = w; fc[i] = f;
  }
}

__global__ void MulKernel(int64_t n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock)
    out[i] = a[i] * b[i];
}

// lam_k = (-s + sum_{j<=k} w2_j key_j) / (1 + sum_{j<=k} w2_j) is the root iff lam_k >= key_{k+1}
__global__ void ZoneBreakKernel(int64_t n, const double* __restrict__ key, const double* __restrict__ SA,
                                const double* __restrict__ SD, const double* __restrict__ scal, int* __restrict__ idx) {
  const double s = scal[SC_T];
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    const double lam = (-s + SA[i]) / (SD[i] + 1.0);
    const double nxt = i + 1 < n ? key[i + 1] : 0.0;
    if (lam >= nxt) atomicMin(idx, (int)i);
  }
}

__global__ void ZoneApplyKernel(int64_t n, double M, double a, double b, double* __restrict__ V,
                                const double* __restrict__ SA, const double* __restrict__ SD,
                                const double* __restrict__ scal, const int* __restrict__ idx) {
  const double s = scal[SC_T];
  if (scal[SC_FVAL] <= s) return;  // f(v) <= s: the point is its own projection (scaled_zone.cc:189-193)
  const int64_t k = min((int64_t)idx[0], n - 1);
  const double lam = (-s + SA[k]) / (SD[k] + 1.0);
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock)
    V[i] = ScaledZone1(V[i], lam, M, a, b);
  if (blockIdx.x == 0 && threadIdx.x == 0) V[n] = s + lam;
}

}  // namespace

bool IsVecOpType(int type, bool epigraph) {
  if (epigraph)
    return type == EPS_PROX_SUM_SQUARE || type == EPS_PROX_MAX || type == EPS_PROX_NORM_1 ||
           type == EPS_PROX_SUM_HINGE || type == EPS_PROX_SUM_DEADZONE || type == EPS_PROX_SUM_QUANTILE;
  return type == EPS_PROX_MAX || type == EPS_PROX_SUM_KL_DIV;
}

static bool NeedsSort(const VecOpDesc& d) {
  return d.type == EPS_PROX_MAX || (d.epigraph && d.type != EPS_PROX_SUM_SQUARE);
}

void VecOpInit(Ctx* ctx, const VecOpDesc& d, VecOpWork* w) {
  EPS_CHECK(d.n0 > 0 && d.n0 < ((int64_t)1 << 30), "vector operator over %lld entries", (long long)d.n0);
  w->scal.alloc(ctx, SC_COUNT);
  w->scal.zero();
  w->idx.alloc(ctx, 4);
  if (!NeedsSort(d)) return;
  const int n = (int)d.n0;
  w->keys.alloc(ctx, n); w->keys2.alloc(ctx, n);
  w->scan1.alloc(ctx, n);
  size_t bytes = 0, b2 = 0;
  if (d.type == EPS_PROX_MAX) {
    CUDA_CHECK(cub::DeviceRadixSort::SortKeysDescending(nullptr, bytes, w->keys.p, w->keys2.p, n, 0, 64, ctx->stream));
  } else {
    w->vals.alloc(ctx, n); w->vals2.alloc(ctx, n); w->scan2.alloc(ctx, n);
    CUDA_CHECK(cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, w->keys.p, w->keys2.p, w->vals.p,
                                                         w->vals2.p, n, 0, 64, ctx->stream));
    CUDA_CHECK(cub::DeviceReduce::Sum(nullptr, b2, w->scan2.p, w->scal.p, n, ctx->stream));
    bytes = std::max(bytes, b2);
  }
  CUDA_CHECK(cub::DeviceScan::InclusiveSum(nullptr, b2, w->keys2.p, w->scan1.p, n, ctx->stream));
  bytes = std::max(bytes, b2);
  w->tmp.alloc(ctx, bytes + 16);
}

void VecOpApply(Ctx* ctx, const VecOpDesc& d, double* V, VecOpWork* w) {
  const int64_t n = d.n0;
  cudaStream_t st = ctx->stream;
  if (d.type == EPS_PROX_SUM_KL_DIV && !d.epigraph) {
    KlDivKernel<<<Grid(n), kBlock, 0, st>>>(n, d.lam, V);
    ctx->count_launch();
    return;
  }
  if (d.type == EPS_PROX_SUM_SQUARE && d.epigraph) {
    VecDot(ctx, n, V, V, w->scal.p + SC_NORM2, false);
    SumSquareEpiRootKernel<<<1, 1, 0, st>>>(n, V, w->scal.p);
    SumSquareEpiScaleKernel<<<Grid(n), kBlock, 0, st>>>(n, V, w->scal.p);
    ctx->count_launch(2);
    return;
  }
  size_t bytes = w->tmp.n;
  const int ni = (int)n;
  PrepKernel<<<1, 1, 0, st>>>(n, d.epigraph ? 1 : 0, V, w->scal.p, w->idx.p);
  if (d.type == EPS_PROX_MAX) {
    CUDA_CHECK(cub::DeviceRadixSort::SortKeysDescending(w->tmp.p, bytes, V, w->keys2.p, ni, 0, 64, st));
    bytes = w->tmp.n;
    CUDA_CHECK(cub::DeviceScan::InclusiveSum(w->tmp.p, bytes, w->keys2.p, w->scan1.p, ni, st));
    MaxBreakKernel<<<Grid(n), kBlock, 0, st>>>(n, d.epigraph ? 1 : 0, d.lam, w->keys2.p, w->scan1.p, w->scal.p, w->idx.p);
    MaxApplyKernel<<<Grid(n), kBlock, 0, st>>>(n, d.epigraph ? 1 : 0, d.lam, V, w->keys2.p, w->scan1.p, w->scal.p, w->idx.p);
    ctx->count_launch(5);
    return;
  }
  // scaled-zone epigraph
  ZoneKeyKernel<<<Grid(n), kBlock, 0, st>>>(n, d.zone_m, d.zone_alpha, d.zone_beta, V, w->keys.p, w->vals.p, w->scan2.p);
  CUDA_CHECK(cub::DeviceReduce::Sum(w->tmp.p, bytes, w->scan2.p, w->scal.p + SC_FVAL, ni, st));
  bytes = w->tmp.n;
  CUDA_CHECK(cub::DeviceRadixSort::SortPairsDescending(w->tmp.p, bytes, w->keys.p, w->keys2.p, w->vals.p, w->vals2.p,
                                                       ni, 0, 64, st));
  MulKernel<<<Grid(n), kBlock, 0, st>>>(n, w->keys2.p, w->vals2.p, w->keys.p);
  bytes = w->tmp.n;
  CUDA_CHECK(cub::DeviceScan::InclusiveSum(w->tmp.p, bytes, w->keys.p, w->scan1.p, ni, st));
  bytes = w->tmp.n;
  CUDA_CHECK(cub::DeviceScan::InclusiveSum(w->tmp.p, bytes, w->vals2.p, w->scan2.p, ni, st));
  ZoneBreakKernel<<<Grid(n), kBlock, 0, st>>>(n, w->keys2.p, w->scan1.p, w->scan2.p, w->scal.p, w->idx.p);
  ZoneApplyKernel<<<Grid(n), kBlock, 0, st>>>(n, d.zone_m, d.zone_alpha, d.zone_beta, V, w->scan1.p, w->scan2.p,
                                              w->scal.p, w->idx.p);
  ctx->count_launch(9);
}

}  // namespace eps
