// Context, error plumbing and the profile-mode event bracket.
#include <cstdarg>
#include <mutex>

#include "internal.h"

namespace eps {

const char* const kKClassNames[KC_COUNT] = {
    "spmv_n", "spmv_t", "gemv_n", "gemv_t", "prox", "vec", "reduce",
    "kkt_dense", "tv1d", "allreduce", "other"};

static thread_local std::string g_last_error;

std::string Format(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  return buf;
}

void SetLastError(const std::string& msg) { g_last_error = msg; }
const std::string& LastError() { return g_last_error; }

Ctx::~Ctx() {
  if (comm) CommDestroy(this);
  for (int i = 0; i < 2; i++)
    if (prof_ev[i]) cudaEventDestroy(prof_ev[i]);
  split_scratch.release();
  seg_scratch.release();
  sell_scratch.release();
  red_partials.release();
  red_counter.release();
  if (h_flag) cudaFreeHost(h_flag);
  if (own_stream && stream) cudaStreamDestroy(stream);
}

ProfScope::ProfScope(Ctx* c, int k, double b) : ctx(c), kc(k), bytes(b) {
  if (!ctx->profile || ctx->capturing) return;
  if (!ctx->prof_ev[0]) {
    cudaEventCreate(&ctx->prof_ev[0]);
    cudaEventCreate(&ctx->prof_ev[1]);
  }
  cudaEventRecord(ctx->prof_ev[0], ctx->stream);
}

ProfScope::~ProfScope() {
  if (!ctx->profile || ctx->capturing) return;
  cudaEventRecord(ctx->prof_ev[1], ctx->stream);
  cudaEventSynchronize(ctx->prof_ev[1]);
  float ms = 0;
  cudaEventElapsedTime(&ms, ctx->prof_ev[0], ctx->prof_ev[1]);
  ctx->prof[kc].launches++;
  ctx->prof[kc].ms += ms;
  ctx->prof[kc].bytes += bytes;
}

ReduceScratch GetReduceScratch(Ctx* ctx) {
  if (ctx->red_partials.n == 0) {
    ctx->red_partials.alloc(ctx, 4096);
    ctx->red_counter.alloc(ctx, 16);
    ctx->red_counter.zero();
  }
  return ReduceScratch{ctx->red_partials.p, ctx->red_counter.p};
}

int* PinnedFlag(Ctx* ctx) {
  if (!ctx->h_flag) CUDA_CHECK(cudaMallocHost(&ctx->h_flag, 64));
  return ctx->h_flag;
}

double* GrowScratch(Ctx* ctx, DevBuf<double>* buf, size_t count) {
  if (buf->n < count) {
    EPS_CHECK(!ctx->capturing, "scratch growth during graph capture");
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    buf->alloc(ctx, count + count / 4);
    ctx->scratch_generation++;
  }
  return buf->p;
}

}  // namespace eps

using namespace eps;

extern "C" {

const char* eps_last_error(void) { return LastError().c_str(); }

int eps_ctx_create(int device, void* stream, eps_ctx** out) {
  try {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
      throw Error(EPS_ERR_NO_DEVICE,
                  Format("no CUDA device (%s); libepsgpu has no CPU fallback",
                         cudaGetErrorString(e)));
    EPS_CHECK(device >= 0 && device < count, "device %d of %d", device, count);
    CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
      throw Error(EPS_ERR_NO_DEVICE,
                  Format("device %d is sm_%d%d; this library is built for sm_100a only",
                         device, prop.major, prop.minor));
    auto* h = new eps_ctx();
    h->c.device = device;
    h->c.sm_count = prop.multiProcessorCount;
    if (stream) {
      h->c.stream = static_cast<cudaStream_t>(stream);
    } else {
      CUDA_CHECK(cudaStreamCreateWithFlags(&h->c.stream, cudaStreamNonBlocking));
      h->c.own_stream = true;
    }
    *out = h;
    return EPS_OK;
  } catch (const Error& e) {
    SetLastError(e.what());
    return e.code;
  }
}

void eps_ctx_destroy(eps_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->c.device);
  cudaStreamSynchronize(ctx->c.stream);
  delete ctx;
}

int64_t eps_ctx_kernel_launches(const eps_ctx* ctx) { return ctx->c.launches; }
int64_t eps_ctx_bytes_allocated(const eps_ctx* ctx) { return ctx->c.bytes_allocated; }

int eps_ctx_synchronize(eps_ctx* ctx) {
  cudaError_t e = cudaStreamSynchronize(ctx->c.stream);
  if (e != cudaSuccess) {
    SetLastError(cudaGetErrorString(e));
    return EPS_ERR_CUDA;
  }
  return EPS_OK;
}

}  // extern "C"
