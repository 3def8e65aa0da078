// 1-D total variation prox:  argmin_b 0.5*||b - v||^2 + lam * sum_k |b[k+1] - b[k]|.
// Reference: TotalVariation1DProx::ApplyVector (src/epsilon/prox/total_variation_1d.cc:12-24)
// calls glmgen's tf_dp (third_party/glmgen/c_lib/glmgen/src/tf/tf_dp.c:43-170), N. Johnson's
// dynamic program ("A dynamic programming algorithm for the fused lasso and L0-segmentation",
// JCGS 2013).  The algorithm is restated here from the paper's recurrences:
//
//   forward  : d_k(b) = (b - v_k) + clip(d_{k-1}(b), -lam, +lam) is the derivative of the
//              k-th message; it is piecewise linear and kept as a deque of knots, each knot
//              carrying the slope/intercept increment that becomes active to its right.
//              lo_k / hi_k are the arguments where d_k = -lam / +lam.
//   backward : b_{n-1} = root of d_{n-1};  b_k = clamp(b_{k+1}, lo_k, hi_k).
//
// The forward recursion is inherently sequential (every step consumes the deque the
// previous one left), so this first version runs it on ONE thread; it is correct and
// matches the reference to rounding but is far from the HBM roofline.  The
// backward clamp chain and the pre/post affine maps are ordinary parallel kernels.
// Compiled with -fmad=false: the slope/intercept sums then round exactly as on the CPU.
#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

__global__ void TvForwardKernel(int n, const double* __restrict__ v, double lam, double* knot,
                                double* slope, double* icpt, double* lo, double* hi, double* beta) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (n == 1 || lam == 0.0) {
    for (int i = 0; i < n; i++) beta[i] = v[i];
    if (n > 1) for (int i = 0; i < n - 1; i++) { lo[i] = v[i]; hi[i] = v[i]; }
    return;
  }
  // deque occupies [left, right] inside arrays of 2n entries, starting in the middle
  int left = n - 1, right = n;
  lo[0] = v[0] - lam;
  hi[0] = v[0] + lam;
  knot[left] = lo[0];
  knot[right] = hi[0];
  slope[left] = 1.0;   icpt[left] = lam - v[0];
  slope[right] = -1.0; icpt[right] = lam + v[0];
  // line active left of the first knot / (negated) right of the last knot for element k
  double s_first = 1.0, c_first = -v[1] - lam;
  double s_last = -1.0, c_last = v[1] - lam;
  for (int k = 1; k < n - 1; k++) {
    // walk up from the left end until the derivative exceeds -lam
    double s = s_first, c = c_first;
    int p = left;
    for (; p <= right; p++) {
      if (s * knot[p] + c > -lam) break;
      s += slope[p];
      c += icpt[p];
    }
    // walk down from the right end until the derivative drops below +lam
    double t = s_last, d = c_last;
    int q = right;
    for (; q >= p; q--) {
      if (-t * knot[q] - d < lam) break;
      t += slope[q];
      d += icpt[q];
    }
    const double new_lo = (-lam - c) / s;
    const double new_hi = (lam + d) / (-t);
    lo[k] = new_lo;
    hi[k] = new_hi;
    left = p - 1;
    right = q + 1;
    knot[left] = new_lo;
    knot[right] = new_hi;
    slope[left] = s;  icpt[left] = c + lam;
    slope[right] = t; icpt[right] = d + lam;
    const double vn = v[k + 1];
    s_first = 1.0;  c_first = -vn - lam;
    s_last = -1.0;  c_last = vn - lam;
  }
  // last element: zero of the derivative
  double s = s_first, c = c_first;
  for (int p = left; p <= right; p++) {
    if (s * knot[p] + c > 0.0) break;
    s += slope[p];
    c += icpt[p];
  }
  beta[n - 1] = -c / s;
}

// Backward chain b_k = clamp(b_{k+1}, lo_k, hi_k), k = n-2 .. 0.  A clamp is the map
// z -> min(max(z, L), H); the composition of clamps is again such a map, so the chain is a
// suffix scan over (L, H) pairs: compose(outer=(L1,H1), inner=(L2,H2)) = (clamp(L2; L1,H1), clamp(H2; L1,H1))
// evaluated as z -> min(max(z, L), H) with L <= H maintained.  One CTA scans a tile of
// kTile elements; tiles are chained through a small sequential pass.
constexpr int kTvBlock = 256;
constexpr int kTvPer = 16;
constexpr int kTvTile = kTvBlock * kTvPer;

struct Clamp { double L, H; };
__device__ __forceinline__ double ApplyClamp(const Clamp& f, double z) {
  // tf_dp.c:157-162: if (z > hi) hi; else if (z < lo) lo; else z
  return z > f.H ? f.H : (z < f.L ? f.L : z);
}
// (outer o inner)(z) = outer(inner(z))
__device__ __forceinline__ Clamp Compose(const Clamp& outer, const Clamp& inner) {
  Clamp r;
  r.L = ApplyClamp(outer, inner.L);
  r.H = ApplyClamp(outer, inner.H);
  return r;
}

// Pass 1: per tile, the composed map  F_tile = f_{first} o ... o f_{last}  (first = lowest k).
__global__ void __launch_bounds__(kTvBlock)
TvTileComposeKernel(int n1, const double* __restrict__ lo, const double* __restrict__ hi, Clamp* tiles) {
  __shared__ Clamp sm[kTvBlock];
  const int base = blockIdx.x * kTvTile + threadIdx.x * kTvPer;
  Clamp f = {-INFINITY, INFINITY};
  // thread-local: f = f_base o f_{base+1} o ... (apply higher k first)
  for (int i = kTvPer - 1; i >= 0; i--) {
    const int k = base + i;
    if (k < n1) { Clamp g = {lo[k], hi[k]}; f = Compose(g, f); }
  }
  sm[threadIdx.x] = f;
  __syncthreads();
  // tree: result for thread t = f_t o f_{t+1} o ... o f_{last}
  for (int off = 1; off < kTvBlock; off <<= 1) {
    Clamp mine = sm[threadIdx.x];
    Clamp other = {-INFINITY, INFINITY};
    const bool has = threadIdx.x + off < kTvBlock;
    if (has) other = sm[threadIdx.x + off];
    __syncthreads();
    if (has) sm[threadIdx.x] = Compose(mine, other);
    __syncthreads();
  }
  if (threadIdx.x == 0) tiles[blockIdx.x] = sm[0];
}

// Pass 2 (one thread): value entering each tile from the right.
__global__ void TvTileChainKernel(int ntiles, const Clamp* tiles, const double* beta_last, double* tile_in) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double z = *beta_last;
  for (int t = ntiles - 1; t >= 0; t--) {
    tile_in[t] = z;
    z = ApplyClamp(tiles[t], z);
  }
}

// Pass 3: inside a tile, b_k = (f_k o ... o f_{tile_last})(tile_in).
__global__ void __launch_bounds__(kTvBlock)
TvTileApplyKernel(int n1, const double* __restrict__ lo, const double* __restrict__ hi,
                  const double* __restrict__ tile_in, double* __restrict__ beta) {
  __shared__ Clamp sm[kTvBlock];
  const int base = blockIdx.x * kTvTile + threadIdx.x * kTvPer;
  Clamp f = {-INFINITY, INFINITY};
  for (int i = kTvPer - 1; i >= 0; i--) {
    const int k = base + i;
    if (k < n1) { Clamp g = {lo[k], hi[k]}; f = Compose(g, f); }
  }
  sm[threadIdx.x] = f;
  __syncthreads();
  for (int off = 1; off < kTvBlock; off <<= 1) {
    Clamp mine = sm[threadIdx.x];
    Clamp other = {-INFINITY, INFINITY};
    const bool has = threadIdx.x + off < kTvBlock;
    if (has) other = sm[threadIdx.x + off];
    __syncthreads();
    if (has) sm[threadIdx.x] = Compose(mine, other);
    __syncthreads();
  }
  // value entering this thread's chunk from the right
  double z = tile_in[blockIdx.x];
  if (threadIdx.x + 1 < kTvBlock) z = ApplyClamp(sm[threadIdx.x + 1], z);
  for (int i = kTvPer - 1; i >= 0; i--) {
    const int k = base + i;
    if (k < n1) {
      Clamp g = {lo[k], hi[k]};
      z = ApplyClamp(g, z);
      beta[k] = z;
    }
  }
}

__global__ void __launch_bounds__(256)
TvPreKernel(ElemSeg sg, const double* __restrict__ consts, double* __restrict__ u,
            const double* __restrict__ y_old, double* __restrict__ v) {
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < sg.len; j += stride) {
    const double w = u[sg.u_off + j] + y_old[sg.y_off + j];
    u[sg.u_off + j] = w;
    const double g = sg.g_off >= 0 ? consts[sg.g_off + j] : 0.0;
    v[j] = sg.Bs * w + g;
  }
}

__global__ void __launch_bounds__(256)
TvPostKernel(ElemSeg sg, const double* __restrict__ consts, const double* __restrict__ beta,
             double* __restrict__ u, double* __restrict__ y_new, double* __restrict__ x) {
  const int64_t stride = (int64_t)gridDim.x * 256;
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < sg.len; j += stride) {
    const double g = sg.g_off >= 0 ? consts[sg.g_off + j] : 0.0;
    const double xv = sg.Cs * (beta[j] - g);
    const double yn = sg.As * xv;
    x[sg.x_off + j] = xv;
    y_new[sg.y_off + j] = yn;
    u[sg.u_off + j] = u[sg.u_off + j] - yn;
  }
}

}  // namespace

// tv1d_par.cu
size_t Tv1dParallelBytes(int64_t n);
int Tv1dParallelSolve(Ctx* ctx, int n, const double* v, double lam, double* beta, char* scratch,
                      int* h_changed);

namespace {
// Below this length the sequential dynamic program on one thread beats ~15 rounds of
// launches + host round trips of the parallel solver.
constexpr int kTvParallelMin = 16384;
size_t DpDoubles(int64_t len) {
  const int64_t ntiles = (len + kTvTile - 1) / kTvTile + 1;
  return (size_t)(8 * len + 4 * ntiles + 16);
}
}  // namespace

// work layout (doubles): v[n] | beta[n] | scratch of the dynamic program | scratch of the
// parallel solver (whose first part, the active set, persists between calls: warm start)
size_t Tv1dWorkDoubles(int64_t len) {
  return (size_t)(2 * len) + DpDoubles(len) + (len >= kTvParallelMin ? Tv1dParallelBytes(len) / 8 + 32 : 0);
}

void Tv1dProxStep(Ctx* ctx, const ElemSeg& seg, const double* consts, double* u, const double* y_old,
                  double* y_new, double* x, double* work) {
  const int n = seg.len;
  if (n <= 0) return;
  double* v = work;
  double* beta = v + n;
  double* knot = beta + n;
  double* slope = knot + 2 * (size_t)n;
  double* icpt = slope + 2 * (size_t)n;
  double* lo = icpt + 2 * (size_t)n;
  double* hi = lo + n;
  const int n1 = n - 1;
  const int ntiles = (n1 + kTvTile - 1) / kTvTile;
  double* tile_in = hi + n;
  Clamp* tiles = reinterpret_cast<Clamp*>(tile_in + ntiles + 1);
  char* par = reinterpret_cast<char*>(work + 2 * (size_t)n + DpDoubles(n));
  par += (256 - (reinterpret_cast<uintptr_t>(par) & 255)) & 255;
  const int grid = GridFor(n, 256, 2, ctx->sm_count);
  {
    ProfScope ps(ctx, KC_VEC, 32.0 * n);
    TvPreKernel<<<grid, 256, 0, ctx->stream>>>(seg, consts, u, y_old, v);
    ctx->count_launch();
  }
  bool solved = false;
  if (n >= kTvParallelMin && seg.lam > 0.0 && !getenv("EPS_TV_SEQUENTIAL")) {
    EPS_CHECK(!ctx->capturing, "the parallel TV solver cannot be captured in a CUDA graph");
    solved = Tv1dParallelSolve(ctx, n, v, seg.lam, beta, par, PinnedFlag(ctx)) >= 0;
  }
  if (!solved) {
    // sequential dynamic program (exact; also the fallback should the rounds above not settle)
    ProfScope ps(ctx, KC_TV, 16.0 * n);
    TvForwardKernel<<<1, 32, 0, ctx->stream>>>(n, v, seg.lam, knot, slope, icpt, lo, hi, beta);
    ctx->count_launch();
    if (n1 > 0 && seg.lam != 0.0) {
      TvTileComposeKernel<<<ntiles, kTvBlock, 0, ctx->stream>>>(n1, lo, hi, tiles);
      TvTileChainKernel<<<1, 32, 0, ctx->stream>>>(ntiles, tiles, beta + n - 1, tile_in);
      TvTileApplyKernel<<<ntiles, kTvBlock, 0, ctx->stream>>>(n1, lo, hi, tile_in, beta);
      ctx->count_launch(3);
    }
  }
  ProfScope ps(ctx, KC_VEC, 40.0 * n);
  TvPostKernel<<<grid, 256, 0, ctx->stream>>>(seg, consts, beta, u, y_new, x);
  ctx->count_launch();
}

}  // namespace eps
