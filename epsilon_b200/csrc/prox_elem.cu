// Fused prox kernels: prox + y = A x + dual update in one pass over HBM.
// Compiled with -fmad=false so that the elementwise arithmetic rounds exactly like the
// reference's x86-64 build (no FMA contraction).
//
// Reference operators restated here (each on the reduced form of VectorProx,
// src/epsilon/prox/vector_prox.cc:51-70,131-137):
//   ScaledZoneProx   src/epsilon/prox/scaled_zone.cc:78-104  (NORM_1, SUM_HINGE, SUM_DEADZONE, SUM_QUANTILE)
//   NonNegativeProx  src/epsilon/prox/non_negative.cc:5-9
//   Norm2Prox        src/epsilon/prox/norm_2.cc:6-17
//   NewtonProx+Logistic  src/epsilon/prox/newton.cc:47-91, sum_logistic.cc:8-32
//   TotalVariation1DProx src/epsilon/prox/total_variation_1d.cc:12-24 -> Johnson's DP (glmgen tf_dp.c:43-170)
#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

constexpr int kBlock = 256;
constexpr int kUnroll = 4;

__device__ __forceinline__ int FindSeg(const int64_t* __restrict__ prefix, int nseg, int64_t e) {
  int lo = 0, hi = nseg;  // prefix[lo] <= e < prefix[lo+1]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (prefix[mid] <= e) lo = mid; else hi = mid;
  }
  return lo;
}

// scaled_zone.cc:78-104 with C = 0.
__device__ __forceinline__ double ScaledZone(double x, double lam, double M, double a, double b) {
  if (fabs(x) <= M) return x;
  if (x > M + lam * a) return x - lam * a;
  if (x < -M - lam * b) return x + lam * b;
  return x > 0 ? M : -M;
}

// Solve f(x) = x - v + lam*sigma(x) = 0 (newton.cc:10-16 ProxResidual with Logistic::gradf) to working precision.
//
// The reference runs a damped Newton from x = v (newton.cc:47-91); in fp64 on the device that is ~7 evaluations of
// exp and ~14 divisions per element, and the kernel sat at 17 % of the HBM roofline on the fp64 pipe.  Three stages:
//   1. fp32, cheap (MUFU ex2 / rcp): an UNDAMPED Newton that cannot overshoot.  f is increasing, convex on x <= 0
//      and concave on x >= 0, and the root has the sign of v - lam/2; started on the side where the tangent
//      under-shoots (f >= 0 in the convex half, f <= 0 in the concave half) Newton converges monotonically.
//      Leaves x0 with |x0 - x*| ~ 1e-7 (1 + |x*|).
//   2. fp64: ONE exp and two divisions at x0, then a Newton step and a chord step in which sigma at the stepped
//      point comes from its Taylor expansion around x0 (the step d is ~1e-6, so the d^3 remainder is below 1e-17).
//   3. accepted when the cubic remainder of the two steps is below rounding and the second step is as small as
//      quadratic convergence predicts; otherwise -- a stage-1 start that
//      did not converge in its 12 steps (huge lam: Newton walks an exp tail one unit per step) -- the lanes fall
//      into the damped Newton of the reference (backtracking constants of newton.cc:63-77) from where they stand.
// All lanes of the warp iterate together and leave when every lane has converged.
// single MUFU instructions (the __frcp_rn / __fdividef / __expf intrinsics carry range and rounding fix-ups that
// tripled the instruction count of the fp32 stage; none of it matters for a starting point)
__device__ __forceinline__ float RcpApprox(float y) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y));
  return r;
}
__device__ __forceinline__ float Ex2Approx(float y) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y));
  return r;
}

// 1 / y for 1 <= y < 1e38 from the fp32 reciprocal and STEPS Newton refinements (relative error 1e-7 -> 1e-14 ->
// below 1 ulp): 2 conversions + 2 fused multiply-adds per step instead of the ~30 instructions of an IEEE division.
template <int STEPS>
__device__ __forceinline__ double RcpSeeded(double y) {
  double r = (double)RcpApprox((float)y);
#pragma unroll
  for (int i = 0; i < STEPS; i++) r = fma(r, fma(-y, r, 1.0), r);
  return r;
}

__device__ __forceinline__ double LogisticProx(double v, double lam, bool active) {
  auto sigma = [](double t) {
    if (t >= 0) { double e = exp(-t); return 1.0 / (1.0 + e); }
    double e = exp(t);
    return e / (1.0 + e);
  };
  double x = v;
  bool done = !active;
  if (active && lam > 0.0 && fabs(v) < 1e30 && lam < 1e30) {
    // ---- stage 1
    const float vf = (float)v, lf = (float)lam;
    // valid starts: x = v (root < 0) / x = v - lam (root > 0), and the root of the tangent model at 0,
    // sigma ~ 1/2 + x/4, which lies on the right side by convexity; the closer of the two
    const float xlin = (vf - 0.5f * lf) * RcpApprox(1.0f + 0.25f * lf);
    float xf = v < 0.5 * lam ? fminf(vf, xlin) : fmaxf(vf - lf, xlin);
#pragma unroll 1
    for (int it = 0; it < 12; it++) {
      const float e = Ex2Approx(-1.4426950408889634f * fabsf(xf));
      const float r = RcpApprox(1.0f + e);
      const float sg = xf >= 0.0f ? r : e * r;
      const float dx = (xf - vf + lf * sg) * RcpApprox(1.0f + lf * (e * r * r));
      xf -= dx;
      // warp-uniform exit: a converged lane sits on its fixed point while the others finish
      if (__all_sync(kFull, fabsf(dx) <= 2e-7f * (1.0f + fabsf(xf)))) break;
    }
    // ---- stage 2
    const double x0 = (double)xf;
    const double e = exp(-fabs(x0));
    const double r = RcpSeeded<2>(1.0 + e);
    const double s0 = x0 >= 0.0 ? r : e * r;
    const double h0 = e * r * r;                               // sigma' = sigma (1 - sigma), without the cancellation
    const double inv = RcpSeeded<1>(fma(lam, h0, 1.0));        // 1 / f'(x0) to 1e-14, kept for the second (chord)
    const double d0 = -(fma(lam, s0, x0 - v)) * inv;           // step: f' moves by ~lam sigma'' d0 / f' ~ 1e-6 relative
    const double x1 = x0 + d0;
    const double s1 = fma(h0 * d0, fma(0.5 * d0, fma(-2.0, s0, 1.0), 1.0), s0);
    const double d1 = -(fma(lam, s1, x1 - v)) * inv;
    // ---- stage 3: what the two steps leave behind is the Taylor remainder of sigma (its third derivative is
    // bounded by sigma' ~ h0 over the step) and the cube of the Newton error, both <= ~lam h0 |d0|^3 / (1 + lam h0)
    // in x.  Held below 1e-16 max(1, |x0|); where sigma is saturated (h0 -> 0) any |d0| <= 1 passes.
    const bool small0 = fabs(d0) <= 1.0 &&
                        lam * h0 * fabs(d0 * d0 * d0) <= 1e-16 * fmax(1.0, fabs(x0)) * fma(lam, h0, 1.0);
    if (small0 && fabs(d1) <= 1e-3 * fabs(d0) + 1e-300) {
      x = x1 + d1;
      done = true;
    } else if (small0 || fabs(x1 - v) <= lam) {
      x = x1;                                                  // still a better start than v
    }
  }
  if (__all_sync(kFull, done)) return x;
  double s = sigma(x);
  double g = x - v + lam * s;
  for (int it = 0; it < 100; it++) {
    if (__all_sync(kFull, done)) break;
    if (!done) {
      const double h = s * (1.0 - s);
      const double dx = g / (1.0 + lam * h);
      double theta = 1.0;
      bool moved = false;
      while (theta > 1e-12) {
        const double nx = x - theta * dx;
        const double ns = sigma(nx);
        const double ng = nx - v + lam * ns;
        if (fabs(ng) <= (1.0 - 0.001 * theta) * fabs(g)) {
          x = nx; s = ns; g = ng; moved = true;
          break;
        }
        theta *= 0.5;
      }
      // converged to working precision, or no further decrease is representable
      if (!moved || fabs(g) <= 2.3e-16 * fmax(1.0, fabs(v))) done = true;
    }
  }
  return x;
}

// The other ElemwiseSmoothFunction operators of the reference (NewtonProx, newton.cc:47-107):
//   SUM_EXP      f = exp(x)         sum_exp.cc:10-36
//   SUM_NEG_ENTR f = x log x        sum_neg_entr.cc:10-42   (proj_feasible: x >= 1e-6)
//   SUM_INV_POS  f = 1/x            sum_inv_pos.cc:10-41    (proj_feasible: x >= 1e-6)
// Same per-element damped Newton as LogisticProx: solve x - v + lam f'(x) = 0.
template <int TYPE>
struct Smooth {
  static __device__ __forceinline__ double grad(double x) {
    if (TYPE == EPS_PROX_SUM_EXP) return exp(x);
    if (TYPE == EPS_PROX_SUM_NEG_ENTR) return 1.0 + log(x);
    return -1.0 / (x * x);
  }
  static __device__ __forceinline__ double hess(double x) {
    if (TYPE == EPS_PROX_SUM_EXP) return exp(x);
    if (TYPE == EPS_PROX_SUM_NEG_ENTR) return 1.0 / x;
    return 2.0 / (x * x * x);
  }
  static __device__ __forceinline__ double proj(double x) {
    return TYPE == EPS_PROX_SUM_EXP ? x : fmax(x, 1e-6);
  }
};

template <int TYPE>
__device__ __forceinline__ double NewtonProxElem(double v, double lam, bool active) {
  double x = Smooth<TYPE>::proj(v);
  // exp overflows at the reference's starting point x = v for large v (it then returns NaN);
  // start right of the root but where exp is finite: x + lam e^x = v  =>  x <= log(v/lam)
  if (TYPE == EPS_PROX_SUM_EXP && lam > 0.0 && v > 1.0) x = fmin(v, log(v / lam) + 1.0);
  double g = x - v + lam * Smooth<TYPE>::grad(x);
  bool done = !active || lam == 0.0;
  for (int it = 0; it < 100; it++) {
    if (__all_sync(kFull, done)) break;
    if (!done) {
      const double dx = g / (1.0 + lam * Smooth<TYPE>::hess(x));
      double theta = 1.0;
      bool moved = false;
      while (theta > 1e-12) {
        const double nx = Smooth<TYPE>::proj(x - theta * dx);
        const double ng = nx - v + lam * Smooth<TYPE>::grad(nx);
        if (fabs(ng) <= (1.0 - 0.001 * theta) * fabs(g)) {
          moved = nx != x;
          x = nx; g = ng;
          break;
        }
        theta *= 0.5;
      }
      if (!moved || fabs(g) <= 2.3e-16 * fmax(1.0, fmax(fabs(v), fabs(x)))) done = true;
    }
  }
  return x;
}

// SumNegLogProx, closed form (sum_neg_log.cc:10-26): root of x - v - lam/x = 0, x > 0
__device__ __forceinline__ double NegLogProx(double v, double lam) {
  if (v >= 0) return (v + sqrt(v * v + 4 * lam)) / 2;
  return 2 * lam / (-v + sqrt(v * v + 4 * lam));
}

template <int TYPE>
__global__ void __launch_bounds__(kBlock)
ElemProxKernel(const ElemSeg* __restrict__ segs, const int64_t* __restrict__ prefix, int nseg,
               int64_t total, const double* __restrict__ consts, double* __restrict__ u,
               const double* __restrict__ y_old, double* __restrict__ y_new, double* __restrict__ x) {
  const int64_t stride = (int64_t)gridDim.x * kBlock * kUnroll;
  for (int64_t base = (int64_t)blockIdx.x * kBlock * kUnroll + threadIdx.x;
       base - threadIdx.x < total; base += stride) {
#pragma unroll
    for (int k = 0; k < kUnroll; k++) {
      const int64_t e = base + (int64_t)k * kBlock;
      const bool active = e < total;
      double w = 0, v = 0, g = 0, lam = 0, xo = 0, dcoef = 0;
      int64_t j = 0;
      ElemSeg sg;
      if (active) {
        const int si = nseg == 1 ? 0 : FindSeg(prefix, nseg, e);
        sg = segs[si];
        j = e - prefix[si];
        w = u[sg.u_off + j] + y_old[sg.y_off + j];
        g = sg.g_off >= 0 ? consts[sg.g_off + j] : 0.0;
        // InitDiagonal (vector_prox.cc:72-116): B, C, D, lambda per element
        if (sg.d_off >= 0) {
          sg.Bs = consts[sg.d_off + j];
          sg.Cs = consts[sg.d_off + sg.len + j];
          dcoef = consts[sg.d_off + 2 * (int64_t)sg.len + j];
          sg.lam = consts[sg.d_off + 3 * (int64_t)sg.len + j];
        }
        v = sg.Bs * w + g;
        lam = sg.lam;
      }
      if (TYPE == EPS_PROX_SUM_LOGISTIC) {
        xo = LogisticProx(v, lam, active);  // warp-cooperative: every lane takes part
      } else if (TYPE == EPS_PROX_SUM_EXP || TYPE == EPS_PROX_SUM_NEG_ENTR || TYPE == EPS_PROX_SUM_INV_POS) {
        xo = NewtonProxElem<TYPE>(v, lam, active);
      } else if (TYPE == EPS_PROX_SUM_NEG_LOG) {
        if (active) xo = NegLogProx(v, lam);
      } else if (active) {
        if (TYPE == EPS_PROX_NON_NEGATIVE) {
          xo = fmax(v, 0.0);
        } else if (TYPE == EPS_PROX_SUM_QUANTILE) {
          const double a = sg.q_off >= 0 ? consts[sg.q_off + j] : sg.zone_alpha;
          const double b = sg.q_off >= 0 ? consts[sg.q_off + sg.len + j] : sg.zone_beta;
          xo = ScaledZone(v, lam, 0.0, a, b);
        } else {
          xo = ScaledZone(v, lam, sg.zone_m, sg.zone_alpha, sg.zone_beta);
        }
      }
      if (active) {
        double xv = sg.Cs * (xo - g);
        if (sg.d_off >= 0) xv = xv + dcoef * w;     // PostProcessOutput: C (x' - g) + D v
        const double yn = sg.As * xv;
        x[sg.x_off + j] = xv;
        y_new[sg.y_off + j] = yn;
        u[sg.u_off + j] = w - yn;
      }
    }
  }
}

// One prox value (everything after v = Bs*w + g has been formed); LOGISTIC / NEWTON are warp-cooperative: every
// lane of the warp must make the call.
template <int TYPE>
__device__ __forceinline__ double ProxValue(const ElemSeg& sg, const double* __restrict__ consts, int64_t j,
                                            double v, bool active) {
  if (TYPE == EPS_PROX_SUM_LOGISTIC) return LogisticProx(v, sg.lam, active);
  if (TYPE == EPS_PROX_SUM_EXP || TYPE == EPS_PROX_SUM_NEG_ENTR || TYPE == EPS_PROX_SUM_INV_POS)
    return NewtonProxElem<TYPE>(v, sg.lam, active);
  if (!active) return 0.0;
  if (TYPE == EPS_PROX_SUM_NEG_LOG) return NegLogProx(v, sg.lam);
  if (TYPE == EPS_PROX_NON_NEGATIVE) return fmax(v, 0.0);
  if (TYPE == EPS_PROX_SUM_QUANTILE) {
    const double a = sg.q_off >= 0 ? consts[sg.q_off + j] : sg.zone_alpha;
    const double b = sg.q_off >= 0 ? consts[sg.q_off + sg.len + j] : sg.zone_beta;
    return ScaledZone(v, sg.lam, 0.0, a, b);
  }
  return ScaledZone(v, sg.lam, sg.zone_m, sg.zone_alpha, sg.zone_beta);
}

// Single-segment stage (the large vectors: t of a logistic regression / SVM, x of a lasso): no segment search, the
// segment descriptor travels as a kernel argument, and every access is a 128-bit load or store of two consecutive
// elements (u, y_old [, g] in; x, y_new, u out).  Needs even offsets (16-byte alignment); an odd last element is
// finished by the first warp of CTA 0.
constexpr int kVecPairs = 2;   // double2 pairs per thread and pass
template <int TYPE>
__global__ void __launch_bounds__(kBlock)
ElemProxVec2Kernel(ElemSeg sg, const double* __restrict__ consts, double* __restrict__ u,
                   const double* __restrict__ y_old, double* __restrict__ y_new, double* __restrict__ x) {
  const int64_t n2 = sg.len / 2;
  double2* u2 = reinterpret_cast<double2*>(u + sg.u_off);
  const double2* yo2 = reinterpret_cast<const double2*>(y_old + sg.y_off);
  double2* yn2 = reinterpret_cast<double2*>(y_new + sg.y_off);
  double2* x2 = reinterpret_cast<double2*>(x + sg.x_off);
  const double2* g2 = sg.g_off >= 0 ? reinterpret_cast<const double2*>(consts + sg.g_off) : nullptr;
  const int64_t stride = (int64_t)gridDim.x * kBlock * kVecPairs;
  for (int64_t base = (int64_t)blockIdx.x * kBlock * kVecPairs + threadIdx.x; base - threadIdx.x < n2; base += stride) {
    double2 w[kVecPairs], g[kVecPairs];
    bool act[kVecPairs];
#pragma unroll
    for (int k = 0; k < kVecPairs; k++) {
      const int64_t p = base + (int64_t)k * kBlock;
      act[k] = p < n2;
      w[k] = make_double2(0.0, 0.0);
      g[k] = make_double2(0.0, 0.0);
      if (act[k]) {
        const double2 a = u2[p], b = LdStream2(reinterpret_cast<const double*>(yo2 + p));
        w[k] = make_double2(a.x + b.x, a.y + b.y);
        if (g2) g[k] = g2[p];
      }
    }
#pragma unroll
    for (int k = 0; k < kVecPairs; k++) {
      const int64_t p = base + (int64_t)k * kBlock;
      const double xa = ProxValue<TYPE>(sg, consts, 2 * p, sg.Bs * w[k].x + g[k].x, act[k]);
      const double xb = ProxValue<TYPE>(sg, consts, 2 * p + 1, sg.Bs * w[k].y + g[k].y, act[k]);
      if (act[k]) {
        const double2 xv = make_double2(sg.Cs * (xa - g[k].x), sg.Cs * (xb - g[k].y));
        const double2 yn = make_double2(sg.As * xv.x, sg.As * xv.y);
        x2[p] = xv;
        yn2[p] = yn;
        u2[p] = make_double2(w[k].x - yn.x, w[k].y - yn.y);
      }
    }
  }
  if ((sg.len & 1) && blockIdx.x == 0 && threadIdx.x < 32) {
    const int64_t j = sg.len - 1;
    const bool active = threadIdx.x == 0;
    double w = 0.0, g = 0.0;
    if (active) {
      w = u[sg.u_off + j] + y_old[sg.y_off + j];
      g = sg.g_off >= 0 ? consts[sg.g_off + j] : 0.0;
    }
    const double xo = ProxValue<TYPE>(sg, consts, j, sg.Bs * w + g, active);
    if (active) {
      const double xv = sg.Cs * (xo - g);
      const double yn = sg.As * xv;
      x[sg.x_off + j] = xv;
      y_new[sg.y_off + j] = yn;
      u[sg.u_off + j] = w - yn;
    }
  }
}

}  // namespace

void ElemProxStepSingle(Ctx* ctx, const ElemSeg& sg, const double* consts, double* u, const double* y_old,
                        double* y_new, double* x) {
  ProfScope ps(ctx, KC_PROX, 40.0 * sg.len);
  const int grid = GridFor(sg.len / 2 + 1, kBlock, kVecPairs, ctx->sm_count);
#define LAUNCHV(T) ElemProxVec2Kernel<T><<<grid, kBlock, 0, ctx->stream>>>(sg, consts, u, y_old, y_new, x)
  switch (sg.type) {
    case EPS_PROX_NON_NEGATIVE: LAUNCHV(EPS_PROX_NON_NEGATIVE); break;
    case EPS_PROX_SUM_QUANTILE: LAUNCHV(EPS_PROX_SUM_QUANTILE); break;
    case EPS_PROX_SUM_LOGISTIC: LAUNCHV(EPS_PROX_SUM_LOGISTIC); break;
    case EPS_PROX_SUM_EXP: LAUNCHV(EPS_PROX_SUM_EXP); break;
    case EPS_PROX_SUM_NEG_ENTR: LAUNCHV(EPS_PROX_SUM_NEG_ENTR); break;
    case EPS_PROX_SUM_INV_POS: LAUNCHV(EPS_PROX_SUM_INV_POS); break;
    case EPS_PROX_SUM_NEG_LOG: LAUNCHV(EPS_PROX_SUM_NEG_LOG); break;
    case EPS_PROX_NORM_1:
    case EPS_PROX_SUM_HINGE:
    case EPS_PROX_SUM_DEADZONE: LAUNCHV(EPS_PROX_NORM_1); break;
    default: throw Error(EPS_ERR_INVALID, Format("prox type %d is not elementwise", sg.type));
  }
#undef LAUNCHV
  ctx->count_launch();
}

void ElemProxStep(Ctx* ctx, const ElemSeg* segs, const int64_t* prefix, int nseg, int64_t total,
                  int type, const double* consts, double* u, const double* y_old, double* y_new,
                  double* x) {
  if (total <= 0) return;
  ProfScope ps(ctx, KC_PROX, 40.0 * total);
  const int grid = GridFor(total, kBlock, kUnroll, ctx->sm_count);
#define LAUNCH(T)                                                                              \
  ElemProxKernel<T><<<grid, kBlock, 0, ctx->stream>>>(segs, prefix, nseg, total, consts, u, y_old, \
                                                      y_new, x)
  switch (type) {
    case EPS_PROX_NON_NEGATIVE: LAUNCH(EPS_PROX_NON_NEGATIVE); break;
    case EPS_PROX_SUM_QUANTILE: LAUNCH(EPS_PROX_SUM_QUANTILE); break;
    case EPS_PROX_SUM_LOGISTIC: LAUNCH(EPS_PROX_SUM_LOGISTIC); break;
    case EPS_PROX_SUM_EXP: LAUNCH(EPS_PROX_SUM_EXP); break;
    case EPS_PROX_SUM_NEG_ENTR: LAUNCH(EPS_PROX_SUM_NEG_ENTR); break;
    case EPS_PROX_SUM_INV_POS: LAUNCH(EPS_PROX_SUM_INV_POS); break;
    case EPS_PROX_SUM_NEG_LOG: LAUNCH(EPS_PROX_SUM_NEG_LOG); break;
    case EPS_PROX_NORM_1:
    case EPS_PROX_SUM_HINGE:
    case EPS_PROX_SUM_DEADZONE: LAUNCH(EPS_PROX_NORM_1); break;
    default: throw Error(EPS_ERR_INVALID, Format("prox type %d is not elementwise", type));
  }
#undef LAUNCH
  ctx->count_launch();
}

}  // namespace eps
