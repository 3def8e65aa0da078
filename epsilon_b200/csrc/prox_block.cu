// Block prox (NORM_2), segment utilities for the ADMM bookkeeping, the elementwise
// pre/post kernels of the KKT prox, and the 1-D total-variation prox.
// Compiled with -fmad=false (see prox_elem.cu).
#include <algorithm>

#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

// ---- NORM_2: x = (1 - lam/||v||) v if ||v|| >= lam else 0   (norm_2.cc:6-17) -----------
// GROUP threads per segment (32: one warp, 256: one CTA).
template <int GROUP>
__global__ void __launch_bounds__(256)
Norm2Kernel(const ElemSeg* __restrict__ segs, int nseg, const double* __restrict__ consts,
            double* __restrict__ u, const double* __restrict__ y_old, double* __restrict__ y_new,
            double* __restrict__ x) {
  __shared__ double smem[32];
  const int per_block = 256 / GROUP;
  const int seg_id = blockIdx.x * per_block + threadIdx.x / GROUP;
  const int t = threadIdx.x % GROUP;
  const bool valid = seg_id < nseg;
  ElemSeg sg;
  if (valid) sg = segs[seg_id];
  const int len = valid ? sg.len : 0;
  double ss[1] = {0.0};
  const int64_t st = valid ? sg.stride : 1;   // 1: contiguous block / column; rows of a matrix argument: its row count
  for (int j = t; j < len; j += GROUP) {
    const double w = u[sg.u_off + j * st] + y_old[sg.y_off + j * st];
    const double g = sg.g_off >= 0 ? consts[sg.g_off + j * st] : 0.0;
    const double v = sg.Bs * w + g;
    ss[0] += v * v;
  }
  double total;
  if (GROUP == 32) {
    total = WarpSum(ss[0]);
  } else {
    BlockSum<1>(ss, smem);
    __shared__ double bcast;
    if (threadIdx.x == 0) bcast = ss[0];
    __syncthreads();
    total = bcast;
  }
  if (!valid) return;
  const double nrm = sqrt(total);
  const bool keep = nrm >= sg.lam;
  const double scale = keep ? 1.0 - sg.lam / nrm : 0.0;
  for (int j = t; j < len; j += GROUP) {
    const double w = u[sg.u_off + j * st] + y_old[sg.y_off + j * st];
    const double g = sg.g_off >= 0 ? consts[sg.g_off + j * st] : 0.0;
    const double v = sg.Bs * w + g;
    const double xo = keep ? scale * v : 0.0;
    const double xv = sg.Cs * (xo - g);
    const double yn = sg.As * xv;
    x[sg.x_off + j * st] = xv;
    y_new[sg.y_off + j * st] = yn;
    u[sg.u_off + j * st] = w - yn;
  }
}

__device__ __forceinline__ int FindPair(const int64_t* __restrict__ prefix, int nseg, int64_t e) {
  int lo = 0, hi = nseg;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (prefix[mid] <= e) lo = mid; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(256)
SegAxpyKernel(const SegPair* __restrict__ pairs, const int64_t* __restrict__ prefix, int nseg,
              int64_t total, double* __restrict__ dst, const double* __restrict__ s1, double c1,
              const double* __restrict__ s2, double c2) {
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += stride) {
    const int si = nseg == 1 ? 0 : FindPair(prefix, nseg, e);
    const SegPair p = pairs[si];
    const int64_t j = e - prefix[si];
    double add = c1 * s1[p.b_off + j];
    if (s2) add += c2 * s2[p.b_off + j];
    dst[p.a_off + j] += add;
  }
}

// One warp per segment: segsum[s] = sum_j v[off + j]^2
__global__ void __launch_bounds__(256)
SegSumSqKernel(const SegPair* __restrict__ pairs, int nseg, const double* __restrict__ v, int use_b,
               double* __restrict__ segsum) {
  const int seg = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (seg >= nseg) return;
  const SegPair p = pairs[seg];
  const int64_t off = use_b ? p.b_off : p.a_off;
  double s = 0.0;
  for (int j = threadIdx.x & 31; j < p.len; j += 32) {
    const double t = v[off + j];
    s += t * t;
  }
  s = WarpSum(s);
  if ((threadIdx.x & 31) == 0) segsum[seg] = s;
}

// out_sum += sum_s coef_s*segsum[s]; out_max = max(out_max, max_s segsum[s])   (one CTA)
__global__ void __launch_bounds__(256)
SegFinishKernel(const SegPair* __restrict__ pairs, int nseg, const double* __restrict__ segsum,
                double* out_sum, double* out_max) {
  __shared__ double smem[64];
  double acc[1] = {0.0};
  double mx = 0.0;
  for (int s = threadIdx.x; s < nseg; s += 256) {
    const double t = segsum[s];
    acc[0] += pairs[s].coef * t;
    mx = fmax(mx, t);
  }
  BlockSum<1>(acc, smem);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(kFull, mx, o));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) smem[32 + (threadIdx.x >> 5)] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 0; w < 8; w++) mx = fmax(mx, smem[32 + w]);
    if (out_sum) *out_sum += acc[0];
    if (out_max) *out_max = fmax(*out_max, mx);
  }
}

template <typename F>
__global__ void __launch_bounds__(256) Elem1Kernel(int64_t n, F f) {
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) f(i);
}

template <typename F>
void LaunchElem1(Ctx* ctx, int64_t n, int kc, double bytes, F f) {
  if (n <= 0) return;
  ProfScope ps(ctx, kc, bytes);
  Elem1Kernel<<<GridFor(n, 256, 2, ctx->sm_count), 256, 0, ctx->stream>>>(n, f);
  ctx->count_launch();
}

}  // namespace

void Norm2ProxStep(Ctx* ctx, const ElemSeg* segs, int nseg, int max_len, const double* consts,
                   double* u, const double* y_old, double* y_new, double* x) {
  if (nseg <= 0) return;
  ProfScope ps(ctx, KC_PROX, 0);
  if (max_len <= 1024) {
    Norm2Kernel<32><<<(nseg + 7) / 8, 256, 0, ctx->stream>>>(segs, nseg, consts, u, y_old, y_new, x);
  } else {
    Norm2Kernel<256><<<nseg, 256, 0, ctx->stream>>>(segs, nseg, consts, u, y_old, y_new, x);
  }
  ctx->count_launch();
}

void SegAxpy(Ctx* ctx, const SegPair* pairs, const int64_t* prefix, int nseg, int64_t total,
             double* dst, const double* s1, double c1, const double* s2, double c2) {
  if (total <= 0) return;
  ProfScope ps(ctx, KC_VEC, (s2 ? 32.0 : 24.0) * total);
  SegAxpyKernel<<<GridFor(total, 256, 2, ctx->sm_count), 256, 0, ctx->stream>>>(
      pairs, prefix, nseg, total, dst, s1, c1, s2, c2);
  ctx->count_launch();
}

void SegSumSq(Ctx* ctx, const SegPair* pairs, const SegPair* pairs_host, int nseg, const double* v,
              double* out_sum, const double* w, double* out_max) {
  if (nseg <= 0) return;
  DevBuf<double>* buf = &ctx->seg_scratch;
  GrowScratch(ctx, buf, std::max(nseg, 1024));
  for (int pass = 0; pass < 2; pass++) {
    const double* src = pass == 0 ? v : w;
    double* os = pass == 0 ? out_sum : nullptr;
    double* om = pass == 0 ? nullptr : out_max;
    if (!src || (!os && !om)) continue;
    ProfScope ps(ctx, KC_REDUCE, 0);
    if (nseg == 1) {
      // long single segment: full-grid deterministic dot product
      const SegPair p = pairs_host[0];
      const double* base = src + (pass == 0 ? p.a_off : p.b_off);
      VecDot(ctx, p.len, base, base, buf->p, false);
    } else {
      SegSumSqKernel<<<(nseg + 7) / 8, 256, 0, ctx->stream>>>(pairs, nseg, src, pass, buf->p);
      ctx->count_launch();
    }
    SegFinishKernel<<<1, 256, 0, ctx->stream>>>(pairs, nseg, buf->p, os, om);
    ctx->count_launch();
  }
}

void KktScalarPre(Ctx* ctx, int64_t n, double a, const double* c0, double* u, const double* y_old,
                  double* rhs) {
  LaunchElem1(ctx, n, KC_VEC, 40.0 * n, [=] __device__(int64_t i) {
    const double w = u[i] + y_old[i];
    u[i] = w;
    rhs[i] = a * w + c0[i];
  });
}

void KktScalarPost(Ctx* ctx, int64_t n, double a, const double* x, double* u, double* y_new) {
  LaunchElem1(ctx, n, KC_VEC, 32.0 * n, [=] __device__(int64_t i) {
    const double yn = a * x[i];
    y_new[i] = yn;
    u[i] = u[i] - yn;
  });
}

// KktScalarPre + the DIAGONAL solve + KktScalarPost in one pass (56n bytes instead of 96n); same operations in
// the same order, so the result is bit-identical to the three kernels.
void KktScalarDiagStep(Ctx* ctx, int64_t n, double a, const double* c0, const double* dinv, double* u,
                       const double* y_old, double* y_new, double* x) {
  LaunchElem1(ctx, n, KC_VEC, 56.0 * n, [=] __device__(int64_t i) {
    const double w = u[i] + y_old[i];
    const double rhs = a * w + c0[i];
    const double xv = dinv[i] * rhs;
    const double yn = a * xv;
    x[i] = xv;
    y_new[i] = yn;
    u[i] = w - yn;
  });
}

namespace {
__global__ void __launch_bounds__(256)
TwoBlockDualKernel(int64_t n, const double* __restrict__ x, const double* __restrict__ v,
                   const double* __restrict__ t, const double* __restrict__ mask,
                   const double* __restrict__ z_prev, double* __restrict__ z, double* __restrict__ u,
                   double* partials, unsigned* counter, double* sums) {
  __shared__ double smem[5 * 32];
  double acc[5] = {0, 0, 0, 0, 0};
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    double zi = v[i] - (t ? t[i] : 0.0);
    if (mask) zi *= mask[i];
    const double xi = x[i];
    const double d = xi - zi;
    const double ui = u[i] + d;
    const double dz = zi - z_prev[i];
    z[i] = zi;
    u[i] = ui;
    acc[0] += d * d;
    acc[1] += dz * dz;
    acc[2] += xi * xi;
    acc[3] += zi * zi;
    acc[4] += ui * ui;
  }
  GridSumFinish<5>(acc, smem, partials, counter, sums, 0);
}
}  // namespace

void TwoBlockDualStep(Ctx* ctx, int64_t n, const double* x, const double* v, const double* t,
                      const double* mask, const double* z_prev, double* z, double* u, double* sums) {
  if (n <= 0) return;
  ReduceScratch rs = GetReduceScratch(ctx);
  const int grid = std::min(512, GridFor(n, 256, 4, ctx->sm_count, 4));
  ProfScope ps(ctx, KC_VEC, 56.0 * n);
  TwoBlockDualKernel<<<grid, 256, 0, ctx->stream>>>(n, x, v, t, mask, z_prev, z, u, rs.partials, rs.counter, sums);
  ctx->count_launch();
}

void DiagMulti(Ctx* ctx, int64_t n, int k, const double* d, const double* in, double b, const double* y, double* out) {
  LaunchElem1(ctx, n * k, KC_VEC, 32.0 * n * k, [=] __device__(int64_t i) {
    out[i] = d[i % n] * in[i] + (y ? b * y[i] : 0.0);
  });
}

void KktDiagSolve(Ctx* ctx, int64_t n, const double* dinv, const double* rhs, double* x) {
  LaunchElem1(ctx, n, KC_VEC, 24.0 * n, [=] __device__(int64_t i) { x[i] = dinv[i] * rhs[i]; });
}

}  // namespace eps
