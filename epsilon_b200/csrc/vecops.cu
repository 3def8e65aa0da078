// Elementwise vector kernels and deterministic dot products.
// Reference: the BlockVector algebra of src/epsilon/vector/block_vector.cc:9-93
// (+=, -=, *=, norm) -- here on flat device slabs.
#include "device.cuh"
#include "internal.h"

namespace eps {

namespace {

constexpr int kBlock = 256;
constexpr int kUnroll = 4;

template <typename F>
__global__ void __launch_bounds__(kBlock) ElemKernel(int64_t n, F f) {
  int64_t stride = (int64_t)gridDim.x * kBlock * kUnroll;
  for (int64_t base = (int64_t)blockIdx.x * kBlock * kUnroll + threadIdx.x; base < n; base += stride) {
#pragma unroll
    for (int k = 0; k < kUnroll; k++) {
      int64_t i = base + (int64_t)k * kBlock;
      if (i < n) f(i);
    }
  }
}

template <typename F>
void LaunchElem(Ctx* ctx, int64_t n, double bytes, F f) {
  if (n <= 0) return;
  ProfScope ps(ctx, KC_VEC, bytes);
  int grid = GridFor(n, kBlock, kUnroll, ctx->sm_count);
  ElemKernel<<<grid, kBlock, 0, ctx->stream>>>(n, f);
  ctx->count_launch();
}

__global__ void __launch_bounds__(kBlock) DotKernel(int64_t n, const double* __restrict__ x,
                                                    const double* __restrict__ y, double* partials,
                                                    unsigned* counter, double* out, int accumulate) {
  __shared__ double smem[32];
  double acc[1] = {0.0};
  int64_t stride = (int64_t)gridDim.x * kBlock * kUnroll;
  for (int64_t base = (int64_t)blockIdx.x * kBlock * kUnroll + threadIdx.x; base < n; base += stride) {
    double xv[kUnroll], yv[kUnroll];
#pragma unroll
    for (int k = 0; k < kUnroll; k++) {
      int64_t i = base + (int64_t)k * kBlock;
      xv[k] = i < n ? x[i] : 0.0;
      yv[k] = i < n ? y[i] : 0.0;
    }
#pragma unroll
    for (int k = 0; k < kUnroll; k++) acc[0] += xv[k] * yv[k];
  }
  GridSumFinish<1>(acc, smem, partials, counter, out, accumulate);
}

}  // namespace

void VecAxpby(Ctx* ctx, int64_t n, double a, const double* x, double b, const double* y, double* out) {
  LaunchElem(ctx, n, 24.0 * n, [=] __device__(int64_t i) { out[i] = a * x[i] + b * y[i]; });
}

void VecScale(Ctx* ctx, int64_t n, double a, const double* x, double* out) {
  LaunchElem(ctx, n, 16.0 * n, [=] __device__(int64_t i) { out[i] = a * x[i]; });
}

void VecMulAdd(Ctx* ctx, int64_t n, const double* d, const double* x, double b, const double* y, double* out) {
  if (y) {
    LaunchElem(ctx, n, 32.0 * n, [=] __device__(int64_t i) { out[i] = d[i] * x[i] + b * y[i]; });
  } else {
    LaunchElem(ctx, n, 24.0 * n, [=] __device__(int64_t i) { out[i] = d[i] * x[i]; });
  }
}

void VecFill(Ctx* ctx, int64_t n, double a, double* out) {
  LaunchElem(ctx, n, 8.0 * n, [=] __device__(int64_t i) { out[i] = a; });
}

void VecCopy(Ctx* ctx, int64_t n, const double* x, double* out) {
  if (n <= 0 || x == out) return;
  LaunchElem(ctx, n, 16.0 * n, [=] __device__(int64_t i) { out[i] = x[i]; });
}

void VecDot(Ctx* ctx, int64_t n, const double* x, const double* y, double* out_dev, bool accumulate) {
  ReduceScratch rs = GetReduceScratch(ctx);
  int grid = GridFor(n > 0 ? n : 1, kBlock, kUnroll, ctx->sm_count, 4);
  if (grid > 1024) grid = 1024;
  ProfScope ps(ctx, KC_REDUCE, 16.0 * n);
  DotKernel<<<grid, kBlock, 0, ctx->stream>>>(n, x, y, rs.partials, rs.counter, out_dev,
                                              accumulate ? 1 : 0);
  ctx->count_launch();
}

}  // namespace eps
