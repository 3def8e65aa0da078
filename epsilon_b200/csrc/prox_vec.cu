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
#include <cooperative_groups.h>
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
__device__ __forceinline__ void KlProxElem(double lam, double u, double v, double* x, double* y) {
  const double eps = 1e-13;
  if (fabs(u) < eps * eps && fabs(v) < eps * eps) { *x = u; *y = v; return; }
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
  *y = lam * yhat + v - lam;
  *x = *y * yhat;
}

__global__ void KlDivKernel(int64_t n, double lam, double* __restrict__ V) {
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    double x, y;
    KlProxElem(lam, V[i], V[n + i], &x, &y);
    V[i] = x;
    V[n + i] = y;
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

// ---- SUM_LARGEST (sum_largest.cc:17-58) -------------------------------------------------------
// The reference sweeps two pointers over the sorted values: i adds y_i ("enters the window
// [q, q + lam)"), j adds -(y_j - lam) ("leaves it, saturated"); the larger of y_i and y_j - lam
// goes first.  That is a merge of two descending lists: here the 2n events are sorted once
// (stable, enters first on ties), prefix sums give the state before every event, and the first
// event at which the reference's loop would stop is found in parallel.
__global__ void LargestEventsKernel(int64_t n, double lam, const double* __restrict__ Y, double* __restrict__ E,
                                    double* __restrict__ tag) {
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    E[i] = Y[i];
    tag[i] = 1.0;          // enter: +value, +1 inside
    E[n + i] = Y[i] - lam;
    tag[n + i] = -1.0;     // leave: -value, -1 inside
  }
}

// acc contribution tag*E -> C; enters (tag > 0) -> ENT
__global__ void LargestContribKernel(int64_t n2, const double* __restrict__ E, const double* __restrict__ tag,
                                     double* __restrict__ C, double* __restrict__ ENT) {
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n2; i += (int64_t)gridDim.x * kBlock) {
    C[i] = tag[i] * E[i];
    ENT[i] = tag[i] > 0 ? 1.0 : 0.0;
  }
}

__global__ void LargestBreakKernel(int64_t n, double lam, double klam, const double* __restrict__ Y,
                                   const double* __restrict__ ACC, const double* __restrict__ INS,
                                   const double* __restrict__ ENT, int* __restrict__ idx) {
  const int64_t n2 = 2 * n;
  for (int64_t e = (int64_t)blockIdx.x * kBlock + threadIdx.x; e < n2; e += (int64_t)gridDim.x * kBlock) {
    const double acc = -klam + (e > 0 ? ACC[e - 1] : 0.0);
    const double ins = e > 0 ? INS[e - 1] : 0.0;
    const int64_t i = e > 0 ? (int64_t)ENT[e - 1] : 0;
    const int64_t j = e - i;
    bool stop = !(i < n && j < n);
    if (!stop) stop = Y[i] * ins <= acc && (Y[j] - lam) * ins <= acc;
    if (stop) atomicMin(idx, (int)e);
  }
}

__global__ void LargestApplyKernel(int64_t n, double lam, double klam, double* __restrict__ V,
                                   const double* __restrict__ ACC, const double* __restrict__ INS,
                                   const int* __restrict__ idx) {
  const int64_t e = min((int64_t)idx[0], 2 * n);
  const double q = e > 0 ? (-klam + ACC[e - 1]) / INS[e - 1] : 0.0;
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    const double v = V[i];
    V[i] = v - fmax(0.0, fmin(lam, v - q));
  }
}

// ---- SECOND_ORDER_CONE (second_order_cone.cc:54-70): rows of [t X] onto ||X_i|| <= beta t_i ----
// V = [t (m) | X (m x n, column-major)]; one thread per row, coalesced across rows.
__global__ void SocProjectKernel(int64_t m, int64_t n, double beta, double* __restrict__ V) {
  const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
  if (i >= m) return;
  double* X = V + m;
  double ss = 0.0;
  for (int64_t j = 0; j < n; j++) { const double x = X[i + j * m]; ss += x * x; }
  const double vn = sqrt(ss);
  const double beta2 = beta * beta;
  double t = V[i];
  double alpha = (1 / (beta2 + 1)) * (beta2 + beta * t / vn);
  if (isnan(alpha) || alpha > 1) {
    alpha = 1;
  } else if (alpha < 0) {
    alpha = 0;
    t = 0;
  } else {
    t = (1 / beta) * alpha * vn;
  }
  V[i] = t;
  for (int64_t j = 0; j < n; j++) X[i + j * m] = alpha * X[i + j * m];
}

// ---- EXP epigraph (exp.cc:12-78): elementwise projection onto {exp(x_i) <= t_i} ---------------
// V = [v (n) | s (n)] -> [x | t].  The reference iterates all elements together until the largest
// residual is below 1e-12; here every element stops on its own residuals (same Newton steps, a
// converged element would only move by rounding afterwards).
__global__ void ExpEpiKernel(int64_t n, double* __restrict__ V) {
  const double eps = 1e-12;
  for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (int64_t)gridDim.x * kBlock) {
    const double v = V[i], s = V[n + i];
    if (exp(v) <= s) continue;  // already feasible (exp.cc:68-74)
    double x = v, t = s, l = 1.0;
    for (int it = 0; it < 100; it++) {
      const double ex = exp(x);
      const double r0 = (x - v) + l * ex, r1 = t - s - l, r2 = ex - t;
      if (fabs(r0) < eps && fabs(r1) < eps && fabs(r2) < eps) break;
      const double h = 1 + l * ex, d = ex;
      const double dl = (-d * r0 + h * (r1 + r2)) / (d * d + h);
      const double dx = -(d * dl + r0) / h;
      const double dt = dl - r1;
      x += dx; t += dt; l += dl;
    }
    V[i] = x;
    V[n + i] = t;
  }
}

// ---- NewtonEpigraph (newton.cc:108-185): SUM_EXP, SUM_LOGISTIC, SUM_INV_POS epigraphs --------
// Newton on (x, t, lam) for the projection onto {f(x) <= t}; the KKT matrix is an arrowhead, so a
// step needs sum_i g_i^2 / (1 + lam h_i) and every line-search trial the norm of the residual.
// One cooperative kernel runs the whole iteration: a thread owns the same elements in every pass
// (x lives in V, trial points in `xn`), the scalars (t, lam, theta, ...) are recomputed
// identically by every CTA from per-CTA partial sums added in a fixed order after a grid sync.
template <int TYPE>
struct EpiFn {
  static __device__ __forceinline__ double f(double x) {
    if (TYPE == EPS_PROX_SUM_EXP) return exp(x);
    if (TYPE == EPS_PROX_SUM_LOGISTIC) return log(1 + exp(x));
    return 1 / x;
  }
  static __device__ __forceinline__ double grad(double x) {
    if (TYPE == EPS_PROX_SUM_EXP) return exp(x);
    if (TYPE == EPS_PROX_SUM_LOGISTIC) { const double e = exp(x); return e / (1 + e); }
    return -1 / (x * x);
  }
  static __device__ __forceinline__ double hess(double x) {
    if (TYPE == EPS_PROX_SUM_EXP) return exp(x);
    if (TYPE == EPS_PROX_SUM_LOGISTIC) { const double e = exp(x); return e / ((1 + e) * (1 + e)); }
    return 2 / (x * x * x);
  }
  static __device__ __forceinline__ double proj(double x) {
    return TYPE == EPS_PROX_SUM_INV_POS ? fmax(x, 1e-6) : x;
  }
};

constexpr int kEpiBlock = 256;
constexpr int kGridSumsMaxK = 4;

// Grid-wide sums of K values per thread, identical in every thread of every CTA on return.
template <int K>
__device__ __forceinline__ void GridSums(double (&v)[K], double* partials, int& toggle, double* smem) {
  namespace cg = cooperative_groups;
  static_assert(K <= kGridSumsMaxK, "partials hold kGridSumsMaxK planes per buffer");
  const int G = gridDim.x;
  double* P = partials + (size_t)toggle * kGridSumsMaxK * G;   // same two buffers whatever K (no overlap between calls)
  toggle ^= 1;
  BlockSum<K>(v, smem);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < K; k++) P[k * G + blockIdx.x] = v[k];
  }
  cg::this_grid().sync();
  double acc[K];
#pragma unroll
  for (int k = 0; k < K; k++) {
    double t = 0.0;
    for (int i = threadIdx.x; i < G; i += blockDim.x) t += __ldcg(&P[k * G + i]);
    acc[k] = t;
  }
  BlockSum<K>(acc, smem);
  __shared__ double total_[K];
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < K; k++) total_[k] = acc[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; k++) v[k] = total_[k];
  __syncthreads();
}

template <int TYPE>
__global__ void __launch_bounds__(kEpiBlock)
NewtonEpiKernel(int64_t n, double* __restrict__ V, const double* __restrict__ v0, double* __restrict__ xn,
                double* __restrict__ partials) {
  __shared__ double smem[3 * 32];
  typedef EpiFn<TYPE> F;
  const int64_t first = (int64_t)blockIdx.x * kEpiBlock + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kEpiBlock;
  const double s = V[n];
  const double eps = fmax(1e-12, 1e-10 / (double)n);
  int toggle = 0;
  // x = proj_feasible(v); the easy case: v feasible and f(v) <= s
  {
    double a[2] = {0.0, 0.0};
    for (int64_t i = first; i < n; i += stride) {
      const double vi = v0[i], xi = F::proj(vi);
      V[i] = xi;
      a[0] += (vi - xi) * (vi - xi);
      a[1] += F::f(xi);
    }
    GridSums<2>(a, partials, toggle, smem);
    if (sqrt(a[0]) < eps && a[1] <= s) {
      for (int64_t i = first; i < n; i += stride) V[i] = v0[i];
      return;
    }
  }
  double t = s, lam = 1.0;
  for (int iter = 0; iter < 100; iter++) {
    // residual at (x, t, lam) and the Newton step of the arrowhead system
    double a[3] = {0.0, 0.0, 0.0};
    for (int64_t i = first; i < n; i += stride) {
      const double xi = V[i];
      const double g = xi - v0[i] + lam * F::grad(xi);
      a[0] += g * g;
      a[1] += g * (g / (1.0 + lam * F::hess(xi)));
      a[2] += F::f(xi);
    }
    GridSums<3>(a, partials, toggle, smem);
    const double gn = t - s - lam, gn1 = a[2] - t, nt_res = a[1];
    double x_res = sqrt(a[0] + gn * gn + gn1 * gn1);
    const double scale = (-nt_res + gn + gn1) / (nt_res + 1);
    const double step_t = gn - scale, step_l = -scale;
    double theta = 1.0;
    while (theta > eps) {
      const double nt = t - theta * step_t;
      double nlam = lam - theta * step_l;
      if (nlam < eps) nlam = eps;
      double b[2] = {0.0, 0.0};
      for (int64_t i = first; i < n; i += stride) {
        const double xi = V[i];
        const double g = xi - v0[i] + lam * F::grad(xi);
        const double nx = F::proj(xi - theta * ((1 + scale) * (g / (1.0 + lam * F::hess(xi)))));
        xn[i] = nx;
        const double r = nx - v0[i] + nlam * F::grad(nx);
        b[0] += r * r;
        b[1] += F::f(nx);
      }
      GridSums<2>(b, partials, toggle, smem);
      const double r1 = nt - s - nlam, r2 = b[1] - nt;
      const double nx_res = sqrt(b[0] + r1 * r1 + r2 * r2);
      if (nx_res <= (1 - 0.001 * theta) * x_res) {
        for (int64_t i = first; i < n; i += stride) V[i] = xn[i];
        t = nt; lam = nlam; x_res = nx_res;
        break;
      }
      theta *= 0.5;
    }
    if (x_res < eps) break;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) V[n] = t;
}

// ---- implicit epigraph operators: scalar Newton on the multiplier lambda around the plain prox ---
// Each is one cooperative kernel; lambda and the stopping decisions are recomputed identically by
// every CTA from the grid-wide sums.

// SumNegLogEpigraph (sum_neg_log.cc:41-88): g(lam) = -sum log x_i(lam) - lam - s with the closed-form prox.
__global__ void __launch_bounds__(kEpiBlock)
NegLogEpiKernel(int64_t n, double* __restrict__ V, double* __restrict__ partials) {
  __shared__ double smem[3 * 32];
  const int64_t first = (int64_t)blockIdx.x * kEpiBlock + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kEpiBlock;
  const double s = V[n];
  int toggle = 0;
  double lam = 1.0;
  for (int iter = 0; iter < 1000; iter++) {
    double a[2] = {0.0, 0.0};
    for (int64_t i = first; i < n; i += stride) {
      const double v = V[i], z = sqrt(v * v + 4 * lam);
      if (v >= 0) {
        a[0] -= log((v + z) / 2);
        a[1] += 1.0 / (v * (v + z) / 2 + 2 * lam);
      } else {
        a[0] += log((-v + z) / (2 * lam));
        a[1] += 1.0 / (v * 2 * lam / (-v + z) + 2 * lam);
      }
    }
    GridSums<2>(a, partials, toggle, smem);
    const double g = -lam - s + a[0];
    double h = -1.0 - a[1];
    if (fabs(g) <= 1e-10) break;
    if (h >= -1e-10) h = -1e-10;
    lam -= g / h;
    if (lam <= 1e-10) lam = 1e-10;
  }
  for (int64_t i = first; i < n; i += stride) {
    const double v = V[i], z = sqrt(v * v + 4 * lam);
    V[i] = v >= 0 ? (v + z) / 2 : 2 * lam / (-v + z);   // ApplyNegLogProx, sum_neg_log.cc:10-26
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) V[n] = s + lam;
}

// SumKLDivEpigraph (sum_kl_div.cc:68-118): arguments [u (n) | v (n) | s]
__global__ void __launch_bounds__(kEpiBlock)
KlDivEpiKernel(int64_t n, double* __restrict__ V, double* __restrict__ partials) {
  __shared__ double smem[3 * 32];
  const int64_t first = (int64_t)blockIdx.x * kEpiBlock + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kEpiBlock;
  const double s = V[2 * n], eps = 1e-10;
  int toggle = 0;
  double lam = 1.0;
  for (int iter = 0; iter < 100; iter++) {
    double a[2] = {0.0, 0.0};
    for (int64_t i = first; i < n; i += stride) {
      double x, y;
      KlProxElem(lam, V[i], V[n + i], &x, &y);
      a[0] += x * log(x / y) - x + y;
      const double g0 = log(x / y), g1 = -x / y + 1;
      const double h00 = 1 + lam / x, h01 = -lam / y, h11 = 1 + lam * x / (y * y);
      a[1] += (g0 * (h11 * g0 - h01 * g1) + g1 * (h00 * g1 - h01 * g0)) / (h00 * h11 - h01 * h01);
    }
    GridSums<2>(a, partials, toggle, smem);
    const double glam = -s - lam + a[0], hlam = -1.0 - a[1];
    if (fabs(glam) < eps || (lam <= eps * 2 && glam / hlam > 0)) break;
    lam = lam - glam / hlam;
    if (lam < eps) lam = eps;
  }
  for (int64_t i = first; i < n; i += stride) {
    double x, y;
    KlProxElem(lam, V[i], V[n + i], &x, &y);
    V[i] = x;
    V[n + i] = y;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) V[2 * n] = s + lam;
}

// ImplicitNewtonEpigraph (newton.cc:192-238) over an elementwise function, with ApplyNewtonProx
// (newton.cc:47-91, global backtracking on the residual norm) as the inner solve.
struct NegEntrFn {
  static __device__ __forceinline__ double f(double x) { return x <= 0 ? 0.0 : x * log(x); }  // sum_neg_entr.cc:12-21
  static __device__ __forceinline__ double grad(double x) { return 1 + log(x); }
  static __device__ __forceinline__ double hess(double x) { return 1 / x; }
  static __device__ __forceinline__ double proj(double x) { return fmax(x, 1e-6); }
};

template <class F>
__device__ void NewtonProxGrid(int64_t n, double lam, double eps, double* __restrict__ X,
                               const double* __restrict__ v0, double* __restrict__ xn, double* partials,
                               int& toggle, double* smem) {
  const int64_t first = (int64_t)blockIdx.x * kEpiBlock + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kEpiBlock;
  for (int64_t i = first; i < n; i += stride) X[i] = F::proj(v0[i]);
  for (int iter = 0; iter < 100; iter++) {
    double a[1] = {0.0};
    for (int64_t i = first; i < n; i += stride) {
      const double xi = X[i], g = xi - v0[i] + lam * F::grad(xi);
      a[0] += g * g;
    }
    GridSums<1>(a, partials, toggle, smem);
    double x_res = sqrt(a[0]), theta = 1.0;
    while (theta > eps) {
      double b[1] = {0.0};
      for (int64_t i = first; i < n; i += stride) {
        const double xi = X[i], g = xi - v0[i] + lam * F::grad(xi);
        const double nx = F::proj(xi - theta * (g / (1.0 + lam * F::hess(xi))));
        xn[i] = nx;
        const double r = nx - v0[i] + lam * F::grad(nx);
        b[0] += r * r;
      }
      GridSums<1>(b, partials, toggle, smem);
      const double nx_res = sqrt(b[0]);
      if (nx_res <= (1 - 0.001 * theta) * x_res) {
        for (int64_t i = first; i < n; i += stride) X[i] = xn[i];
        x_res = nx_res;
        break;
      }
      theta *= 0.5;
    }
    if (x_res < eps) break;
  }
}

template <class F>
__global__ void __launch_bounds__(kEpiBlock)
ImplicitNewtonEpiKernel(int64_t n, double* __restrict__ V, const double* __restrict__ v0, double* __restrict__ xn,
                        double* __restrict__ partials) {
  __shared__ double smem[3 * 32];
  const int64_t first = (int64_t)blockIdx.x * kEpiBlock + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kEpiBlock;
  const double s = V[n];
  const double eps = fmax(1e-12, 1e-10 / (double)n);
  int toggle = 0;
  {
    double a[1] = {0.0};
    for (int64_t i = first; i < n; i += stride) {
      const double xi = F::proj(v0[i]);
      V[i] = xi;
      a[0] += F::f(xi);
    }
    GridSums<1>(a, partials, toggle, smem);
    if (a[0] <= s) return;   // x = proj_feasible(v), t = s
  }
  double lam = 1.0;
  for (int iter = 0; iter < 100; iter++) {
    NewtonProxGrid<F>(n, lam, eps, V, v0, xn, partials, toggle, smem);
    double a[2] = {0.0, 0.0};
    for (int64_t i = first; i < n; i += stride) {
      const double xi = V[i], gx = F::grad(xi);
      a[0] += F::f(xi);
      a[1] += gx / (1.0 + lam * F::hess(xi)) * gx;
    }
    GridSums<2>(a, partials, toggle, smem);
    const double glam = a[0] - lam - s, hlam = -a[1] - 1;
    if (fabs(glam) < 1e-10) break;
    lam = lam - glam / hlam;
    if (lam < 0) lam = 1e-6;
  }
  NewtonProxGrid<F>(n, lam, eps, V, v0, xn, partials, toggle, smem);
  if (blockIdx.x == 0 && threadIdx.x == 0) V[n] = s + lam;
}

// ---- LOG_SUM_EXP (log_sum_exp.cc): NewtonProx / NewtonEpigraph with a full Hessian ---------------
// f = log sum exp(x_i), grad = w = softmax(x), (I + lam (diag(w) - w w'))^{-1} g by Sherman-Morrison
// (log_sum_exp.cc:41-64).  Every evaluation point needs max(x) and sum exp(x - max) over the grid.
__device__ __forceinline__ double GridMax(double v, double* partials, int& toggle, double* smem) {
  namespace cg = cooperative_groups;
  const int G = gridDim.x;
  double* P = partials + (size_t)toggle * kGridSumsMaxK * G;
  toggle ^= 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
  __syncthreads();
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double m = smem[0];
    for (int w = 1; w < kEpiBlock / 32; w++) m = fmax(m, smem[w]);
    P[blockIdx.x] = m;
  }
  cg::this_grid().sync();
  double m = -INFINITY;
  for (int i = lane; i < G; i += 32) m = fmax(m, __ldcg(&P[i]));
  for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
  return m;
}

// max and sum exp(x - max) of X[0..n)
__device__ __forceinline__ void LseStats(int64_t n, const double* __restrict__ X, double* mx, double* z,
                                         double* partials, int& toggle, double* smem) {
  const int64_t first = (int64_t)blockIdx.x * kEpiBlock + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kEpiBlock;
  double m = -INFINITY;
  for (int64_t i = first; i < n; i += stride) m = fmax(m, X[i]);
  m = GridMax(m, partials, toggle, smem);
  double a[1] = {0.0};
  for (int64_t i = first; i < n; i += stride) a[0] += exp(X[i] - m);
  GridSums<1>(a, partials, toggle, smem);
  *mx = m;
  *z = a[0];
}

template <bool EPI>
__global__ void __launch_bounds__(kEpiBlock)
LogSumExpKernel(int64_t n, double lam_prox, double* __restrict__ V, const double* __restrict__ v0,
                double* __restrict__ xn, double* __restrict__ dx, double* __restrict__ partials) {
  __shared__ double smem[4 * 32];
  const int64_t first = (int64_t)blockIdx.x * kEpiBlock + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kEpiBlock;
  const double s = EPI ? V[n] : 0.0;
  const double eps = fmax(1e-12, 1e-10 / (double)n);
  int toggle = 0;
  double mx, z;
  LseStats(n, V, &mx, &z, partials, toggle, smem);   // x = v
  if (EPI && mx + log(z) <= s) return;               // easy case: already inside
  double t = s, lam = EPI ? 1.0 : lam_prox;
  for (int iter = 0; iter < 100; iter++) {
    // residual g = x - v + lam w and the sums of hess_inv; (mx, z) describe the current x
    double a[4] = {0.0, 0.0, 0.0, 0.0};
    for (int64_t i = first; i < n; i += stride) {
      const double xi = V[i], w = exp(xi - mx) / z, g = xi - v0[i] + lam * w, d = 1 + lam * w;
      a[0] += g * g;
      a[1] += w * w / d;
      a[2] += g * w / d;
      a[3] += g * g / d;
    }
    GridSums<4>(a, partials, toggle, smem);
    const double sc = lam * a[2] / (1 - lam * a[1]);
    double mul = 1.0, step_t = 0.0, step_l = 0.0, x_res = sqrt(a[0]);
    if (EPI) {
      const double gn = t - s - lam, gn1 = mx + log(z) - t;
      const double nt_res = a[3] + sc * a[2];
      const double scale = (-nt_res + gn + gn1) / (nt_res + 1);
      mul = 1 + scale; step_t = gn - scale; step_l = -scale;
      x_res = sqrt(a[0] + gn * gn + gn1 * gn1);
    }
    for (int64_t i = first; i < n; i += stride) {
      const double xi = V[i], w = exp(xi - mx) / z, g = xi - v0[i] + lam * w, d = 1 + lam * w;
      dx[i] = mul * (g / d + w / d * sc);
    }
    double theta = 1.0;
    while (theta > eps) {
      const double nt = t - theta * step_t;
      double nlam = lam - theta * step_l;
      if (EPI && nlam < eps) nlam = eps;
      for (int64_t i = first; i < n; i += stride) xn[i] = V[i] - theta * dx[i];
      double nmx, nz;
      LseStats(n, xn, &nmx, &nz, partials, toggle, smem);
      double b[1] = {0.0};
      for (int64_t i = first; i < n; i += stride) {
        const double r = xn[i] - v0[i] + nlam * (exp(xn[i] - nmx) / nz);
        b[0] += r * r;
      }
      GridSums<1>(b, partials, toggle, smem);
      double nx_res = sqrt(b[0]);
      if (EPI) {
        const double r1 = nt - s - nlam, r2 = nmx + log(nz) - nt;
        nx_res = sqrt(b[0] + r1 * r1 + r2 * r2);
      }
      if (nx_res <= (1 - 0.001 * theta) * x_res) {
        for (int64_t i = first; i < n; i += stride) V[i] = xn[i];
        t = nt; lam = nlam; x_res = nx_res; mx = nmx; z = nz;
        break;
      }
      theta *= 0.5;
    }
    if (x_res < eps) break;
  }
  if (EPI && blockIdx.x == 0 && threadIdx.x == 0) V[n] = t;
}

// ---- BisectionEpigraph around SumLargestProx (newton.cc:240-286, sum_largest.cc:60-85) ---------
// sum of the k largest: Y sorted descending
__global__ void TopKSumKernel(int k, const double* __restrict__ Y, double* __restrict__ out) {
  double t = 0.0;
  for (int i = 0; i < k; i++) t += Y[i];
  *out = t;
}

__global__ void SetScalarKernel(double* __restrict__ p, double v) { *p = v; }

}  // namespace

bool IsVecOpType(int type, bool epigraph) {
  if (IsMatOpType(type)) return !epigraph || type != EPS_PROX_SEMIDEFINITE;
  if (epigraph)
    return type == EPS_PROX_SUM_SQUARE || type == EPS_PROX_MAX || type == EPS_PROX_NORM_1 ||
           type == EPS_PROX_SUM_HINGE || type == EPS_PROX_SUM_DEADZONE || type == EPS_PROX_SUM_QUANTILE ||
           type == EPS_PROX_SUM_EXP || type == EPS_PROX_SUM_LOGISTIC || type == EPS_PROX_SUM_INV_POS ||
           type == EPS_PROX_EXP || type == EPS_PROX_SUM_NEG_LOG || type == EPS_PROX_SUM_NEG_ENTR ||
           type == EPS_PROX_SUM_KL_DIV || type == EPS_PROX_LOG_SUM_EXP || type == EPS_PROX_SUM_LARGEST;
  return type == EPS_PROX_MAX || type == EPS_PROX_SUM_KL_DIV || type == EPS_PROX_SUM_LARGEST ||
         type == EPS_PROX_SECOND_ORDER_CONE || type == EPS_PROX_LOG_SUM_EXP;
}

static bool IsNewtonEpi(const VecOpDesc& d) {
  return d.epigraph &&
         (d.type == EPS_PROX_SUM_EXP || d.type == EPS_PROX_SUM_LOGISTIC || d.type == EPS_PROX_SUM_INV_POS);
}

// one cooperative kernel iterating on grid-wide sums (work: copy of v, trial point, step)
static bool IsGridNewton(const VecOpDesc& d) {
  return IsNewtonEpi(d) || d.type == EPS_PROX_LOG_SUM_EXP ||
         (d.epigraph && (d.type == EPS_PROX_SUM_NEG_LOG || d.type == EPS_PROX_SUM_NEG_ENTR || d.type == EPS_PROX_SUM_KL_DIV));
}

static bool NeedsSort(const VecOpDesc& d) {
  return d.type == EPS_PROX_MAX || d.type == EPS_PROX_SUM_LARGEST ||
         (d.epigraph && d.type != EPS_PROX_SUM_SQUARE && d.type != EPS_PROX_EXP && !IsGridNewton(d));
}

static int NewtonEpiGrid(Ctx* ctx, int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ctx->sm_count, (n + kEpiBlock - 1) / kEpiBlock));
}

static void SumLargestProxRun(Ctx* ctx, const VecOpDesc& d, double lam, double* V, VecOpWork* w) {
    const int64_t n = d.n0;
    const int ni = (int)n;
    cudaStream_t st = ctx->stream;
    size_t bytes = w->tmp.n;
    const double klam = (double)d.k * lam;
    const int n2 = 2 * ni;
    CUDA_CHECK(cub::DeviceRadixSort::SortKeysDescending(w->tmp.p, bytes, V, w->keys2.p, ni, 0, 64, st));
    LargestEventsKernel<<<Grid(n), kBlock, 0, st>>>(n, lam, w->keys2.p, w->keys.p, w->vals.p);
    bytes = w->tmp.n;
    CUDA_CHECK(cub::DeviceRadixSort::SortPairsDescending(w->tmp.p, bytes, w->keys.p, w->ev.p, w->vals.p, w->vals2.p,
                                                         n2, 0, 64, st));
    // keys <- signed contributions, vals <- enter indicator (the unsorted inputs are dead)
    LargestContribKernel<<<Grid(2 * n), kBlock, 0, st>>>(2 * n, w->ev.p, w->vals2.p, w->keys.p, w->vals.p);
    bytes = w->tmp.n;
    CUDA_CHECK(cub::DeviceScan::InclusiveSum(w->tmp.p, bytes, w->keys.p, w->scan1.p, n2, st));   // ACC
    bytes = w->tmp.n;
    CUDA_CHECK(cub::DeviceScan::InclusiveSum(w->tmp.p, bytes, w->vals2.p, w->scan2.p, n2, st));  // INS
    bytes = w->tmp.n;
    CUDA_CHECK(cub::DeviceScan::InclusiveSum(w->tmp.p, bytes, w->vals.p, w->ent.p, n2, st));     // ENT
    LargestBreakKernel<<<Grid(2 * n), kBlock, 0, st>>>(n, lam, klam, w->keys2.p, w->scan1.p, w->scan2.p, w->ent.p,
                                                       w->idx.p);
    LargestApplyKernel<<<Grid(n), kBlock, 0, st>>>(n, lam, klam, V, w->scan1.p, w->scan2.p, w->idx.p);
    ctx->count_launch(10);
}

// BisectionEpigraph::ApplyVector (newton.cc:240-286): the doubling / halving decisions depend on the
// sign of g = f(prox_lam(v)) - lam - s, one scalar read back per round (the only host-synchronous
// operator; lam runs over dyadic rationals, so the rounds are those of the reference).
static void SumLargestEpigraphRun(Ctx* ctx, const VecOpDesc& d, double* V, VecOpWork* w) {
  const int64_t n = d.n0;
  const int ni = (int)n;
  cudaStream_t st = ctx->stream;
  EPS_CHECK(d.k >= 0 && d.k <= n, "SUM_LARGEST: k = %d of %lld", d.k, (long long)n);
  auto eval = [&](const double* X) {   // SumLargestProx::Eval, sum_largest.cc:60-73
    size_t bytes = w->tmp.n;
    CUDA_CHECK(cub::DeviceRadixSort::SortKeysDescending(w->tmp.p, bytes, X, w->keys2.p, ni, 0, 64, st));
    TopKSumKernel<<<1, 1, 0, st>>>(d.k, w->keys2.p, w->scal.p + SC_FVAL);
    ctx->count_launch(2);
    double f = 0.0;
    CUDA_CHECK(cudaMemcpyAsync(&f, w->scal.p + SC_FVAL, sizeof(double), cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    return f;
  };
  double s = 0.0;
  CUDA_CHECK(cudaMemcpyAsync(&s, V + n, sizeof(double), cudaMemcpyDeviceToHost, st));
  if (eval(V) <= s) return;   // x = v, t = s
  VecCopy(ctx, n, V, w->v0.p);
  double lam = 1.0, upper = 1.0, lower = 0.0, t = NAN;
  bool upper_fixed = false;
  for (int iter = 0; iter < 100; iter++) {
    if (iter > 0) VecCopy(ctx, n, w->v0.p, V);
    PrepKernel<<<1, 1, 0, st>>>(n, 0, V, w->scal.p, w->idx.p);
    SumLargestProxRun(ctx, d, lam, V, w);
    const double g = eval(V) - (lam + s);
    if (fabs(g) <= 1e-5) { t = lam + s; break; }
    if (g > 0 && !upper_fixed) { lam *= 2; upper = lam; }
    else if (g > 0) { lower = lam; lam = (lam + upper) / 2; }
    else { upper = lam; lam = (lam + lower) / 2; upper_fixed = true; }
  }
  SetScalarKernel<<<1, 1, 0, st>>>(V + n, t);   // left unset by the reference after 100 rounds: NaN here
  ctx->count_launch();
}

void VecOpInit(Ctx* ctx, const VecOpDesc& d, VecOpWork* w) {
  EPS_CHECK(d.n0 > 0 && d.n0 < ((int64_t)1 << 30), "vector operator over %lld entries", (long long)d.n0);
  w->scal.alloc(ctx, SC_COUNT);
  w->scal.zero();
  w->idx.alloc(ctx, 4);
  if (IsMatOpType(d.type)) {
    MatOpInit(ctx, d, w);
    return;
  }
  if (IsGridNewton(d)) {
    w->keys.alloc(ctx, d.n0);    // copy of v
    w->keys2.alloc(ctx, d.n0);   // trial point of the line search
    if (d.type == EPS_PROX_LOG_SUM_EXP) w->vals.alloc(ctx, d.n0);   // Newton step
    w->scan1.alloc(ctx, (size_t)2 * kGridSumsMaxK * NewtonEpiGrid(ctx, d.n0));
    return;
  }
  if (!NeedsSort(d)) return;
  const int n = (int)d.n0;
  size_t bytes = 0, b2 = 0;
  if (d.type == EPS_PROX_SUM_LARGEST) {
    // keys2 = sorted values (n) | keys/vals = events and tags (2n) | keys sorted -> vals2.. see VecOpApply
    w->keys2.alloc(ctx, n);
    w->keys.alloc(ctx, 2 * (size_t)n); w->vals.alloc(ctx, 2 * (size_t)n);
    w->vals2.alloc(ctx, 2 * (size_t)n); w->scan1.alloc(ctx, 2 * (size_t)n); w->scan2.alloc(ctx, 2 * (size_t)n);
    w->ev.alloc(ctx, 2 * (size_t)n); w->ent.alloc(ctx, 2 * (size_t)n);
    if (d.epigraph) w->v0.alloc(ctx, n);
    CUDA_CHECK(cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, w->keys.p, w->ev.p, w->vals.p, w->vals2.p,
                                                         2 * n, 0, 64, ctx->stream));
    CUDA_CHECK(cub::DeviceScan::InclusiveSum(nullptr, b2, w->keys.p, w->scan1.p, 2 * n, ctx->stream));
    bytes = std::max(bytes, b2);
    w->tmp.alloc(ctx, bytes + 16);
    return;
  }
  w->keys.alloc(ctx, n); w->keys2.alloc(ctx, n);
  w->scan1.alloc(ctx, n);
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
  if (IsMatOpType(d.type)) {
    MatOpApply(ctx, d, V, w);
    return;
  }
  if (d.type == EPS_PROX_EXP) {
    ExpEpiKernel<<<Grid(n), kBlock, 0, st>>>(n, V);
    ctx->count_launch();
    return;
  }
  if (d.type == EPS_PROX_SECOND_ORDER_CONE) {
    SocProjectKernel<<<Grid(n), kBlock, 0, st>>>(n, d.soc_n, d.soc_a, V);
    ctx->count_launch();
    return;
  }
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
  if (IsGridNewton(d) && !IsNewtonEpi(d)) {
    int64_t nn = n;
    double* Vp = V;
    const double* v0 = w->keys.p;
    double *xn = w->keys2.p, *dx = w->vals.p, *part = w->scan1.p;
    double lam = d.lam;
    const dim3 grid(NewtonEpiGrid(ctx, n)), block(kEpiBlock);
    if (d.type == EPS_PROX_SUM_NEG_LOG) {
      void* args[] = {&nn, &Vp, &part};
      CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)NegLogEpiKernel, grid, block, args, 0, st));
    } else if (d.type == EPS_PROX_SUM_KL_DIV) {
      void* args[] = {&nn, &Vp, &part};
      CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)KlDivEpiKernel, grid, block, args, 0, st));
    } else if (d.type == EPS_PROX_SUM_NEG_ENTR) {
      VecCopy(ctx, n, V, w->keys.p);
      void* args[] = {&nn, &Vp, &v0, &xn, &part};
      CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)ImplicitNewtonEpiKernel<NegEntrFn>, grid, block, args, 0, st));
    } else {
      VecCopy(ctx, n, V, w->keys.p);
      void* args[] = {&nn, &lam, &Vp, &v0, &xn, &dx, &part};
      const void* fn = d.epigraph ? (const void*)LogSumExpKernel<true> : (const void*)LogSumExpKernel<false>;
      CUDA_CHECK(cudaLaunchCooperativeKernel(fn, grid, block, args, 0, st));
    }
    ctx->count_launch();
    return;
  }
  if (IsNewtonEpi(d)) {
    VecCopy(ctx, n, V, w->keys.p);
    int64_t nn = n;
    double* Vp = V;
    const double* v0 = w->keys.p;
    double *xn = w->keys2.p, *part = w->scan1.p;
    void* args[] = {&nn, &Vp, &v0, &xn, &part};
    const void* fn = d.type == EPS_PROX_SUM_EXP ? (const void*)NewtonEpiKernel<EPS_PROX_SUM_EXP>
                     : d.type == EPS_PROX_SUM_LOGISTIC ? (const void*)NewtonEpiKernel<EPS_PROX_SUM_LOGISTIC>
                                                       : (const void*)NewtonEpiKernel<EPS_PROX_SUM_INV_POS>;
    CUDA_CHECK(cudaLaunchCooperativeKernel(fn, dim3(NewtonEpiGrid(ctx, n)), dim3(kEpiBlock), args, 0, st));
    ctx->count_launch();
    return;
  }
  size_t bytes = w->tmp.n;
  const int ni = (int)n;
  PrepKernel<<<1, 1, 0, st>>>(n, d.epigraph ? 1 : 0, V, w->scal.p, w->idx.p);
  if (d.type == EPS_PROX_SUM_LARGEST && !d.epigraph) {
    SumLargestProxRun(ctx, d, d.lam, V, w);
    return;
  }
  if (d.type == EPS_PROX_SUM_LARGEST) {
    SumLargestEpigraphRun(ctx, d, V, w);
    return;
  }
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
