// 1-D total variation prox:  argmin_b 0.5*||b - v||^2 + lam * sum_k |b[k+1] - b[k]|.
// Reference: TotalVariation1DProx::ApplyVector (src/epsilon/prox/total_variation_1d.cc:12-24) calls glmgen's
// tf_dp (third_party/glmgen/c_lib/glmgen/src/tf/tf_dp.c:43-170), a sequential O(n) dynamic program.  The
// device path does not run that program: every length is solved by the parallel split-and-merge active-set
// solver of tv1d_par.cu, which reaches the same (unique) minimiser through the optimality conditions.
// This file holds the VectorProx wrapping around it (pre / post affine maps fused with the dual update).
// Compiled with -fmad=false.
#include "device.cuh"
#include "prox.cuh"

namespace eps {

namespace {

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
int Tv1dParallelSolve(Ctx* ctx, int n, double* v, double lam, double* beta, char* scratch,
                      int* h_flag, unsigned* host_state, const TvFuse* fuse);

// work layout (doubles): v[n] | beta[n] | scratch of the parallel solver (whose first part, the active set,
// persists between calls: warm start)
size_t Tv1dWorkDoubles(int64_t len) {
  return (size_t)(2 * len) + Tv1dParallelBytes(len) / 8 + 64;
}

void Tv1dProxStep(Ctx* ctx, const ElemSeg& seg, const double* consts, double* u, const double* y_old,
                  double* y_new, double* x, double* work, unsigned* tag_state) {
  const int n = seg.len;
  if (n <= 0) return;
  double* v = work;
  double* beta = v + n;
  char* par = reinterpret_cast<char*>(work + 2 * (size_t)n);
  par += (256 - (reinterpret_cast<uintptr_t>(par) & 255)) & 255;
  if (n == 1 || seg.lam == 0.0) {
    // no boundary / no penalty: the prox is the identity (tf_dp returns its input for lam = 0 as well)
    const int grid = GridFor(n, 256, 2, ctx->sm_count);
    ProfScope ps(ctx, KC_VEC, 72.0 * n);
    TvPreKernel<<<grid, 256, 0, ctx->stream>>>(seg, consts, u, y_old, v);
    TvPostKernel<<<grid, 256, 0, ctx->stream>>>(seg, consts, v, u, y_new, x);
    ctx->count_launch(2);
    return;
  }
  // the pre-map (w = u + y_old, v = B w + g) and the post-map (x = C (beta - g), y = A x, u -= y) run inside the
  // solver's first and last pass
  EPS_CHECK(!ctx->capturing, "the TV solver cannot be captured in a CUDA graph");
  const TvFuse fuse{seg, consts, u, y_old, y_new, x};
  const int rounds = Tv1dParallelSolve(ctx, n, v, seg.lam, beta, par, PinnedFlag(ctx), tag_state, &fuse);
  if (rounds < 0)
    throw Error(EPS_ERR_NUMERIC, Format("TOTAL_VARIATION_1D: the active-set rounds did not settle (n = %d, lambda = %g)",
                                        n, seg.lam));
}

}  // namespace eps
