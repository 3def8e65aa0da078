// Context, error plumbing and the profile-mode event bracket.
#include <cstdarg>
#include <mutex>

#include <thread>

#include <mutex>
#include <unordered_set>

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

namespace {
// Contexts that are still alive: buffers can outlive their context (a host language may finalise
// objects in any order at exit); they must then be freed without touching it.
std::mutex g_live_mu;
std::unordered_set<const Ctx*> g_live;
bool IsLive(const Ctx* c) {
  std::lock_guard<std::mutex> lk(g_live_mu);
  return g_live.count(c) != 0;
}
}  // namespace

bool CtxIsLive(const Ctx* c) { return IsLive(c); }

void RegisterCtx(const Ctx* c) {
  std::lock_guard<std::mutex> lk(g_live_mu);
  g_live.insert(c);
}

Ctx::~Ctx() {
  {
    std::lock_guard<std::mutex> lk(g_live_mu);
    g_live.erase(this);
  }
  if (comm) CommDestroy(this);
  for (int i = 0; i < 2; i++)
    if (prof_ev[i]) cudaEventDestroy(prof_ev[i]);
  split_scratch.release();
  seg_scratch.release();
  sell_scratch.release();
  red_partials.release();
  red_counter.release();
  if (h_flag) cudaFreeHost(h_flag);
  delete staging;
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

void* DevAlloc(Ctx* ctx, size_t bytes) {
  void* p = nullptr;
  if (ctx->pooled) CUDA_CHECK(cudaMallocAsync(&p, bytes, ctx->stream));
  else CUDA_CHECK(cudaMalloc(&p, bytes));
  ctx->bytes_allocated += (int64_t)bytes;
  return p;
}

void DevFree(Ctx* ctx, void* p, size_t bytes) {
  if (!p) return;
  if (ctx && !IsLive(ctx)) ctx = nullptr;  // context already destroyed: blocking free, no accounting
  if (ctx && ctx->pooled) {
    // stream-ordered: the memory returns to the pool after everything queued so far; a context
    // that is already gone (static buffers at process exit) falls back to the blocking free
    if (cudaFreeAsync(p, ctx->stream) != cudaSuccess) { cudaGetLastError(); cudaFree(p); }
  } else {
    cudaFree(p);
  }
  if (ctx) ctx->bytes_allocated -= (int64_t)bytes;
}

ReduceScratch GetReduceScratch(Ctx* ctx) {
  if (ctx->red_partials.n == 0) {
    ctx->red_partials.alloc(ctx, 4096);
    ctx->red_counter.alloc(ctx, 16);
    ctx->red_counter.zero();
  }
  return ReduceScratch{ctx->red_partials.p, ctx->red_counter.p};
}

// ---- staged upload ---------------------------------------------------------------------
// cudaMemcpyAsync from pageable memory runs at a few GB/s (one driver thread copies through
// a small bounce buffer).  The wire format hands us pageable bytes (Python strings in the
// reference, solvemodule.cc:58-72), so for large buffers kWorkers host threads copy
// interleaved chunks into their own pinned buffers and DMA them on their own streams.
namespace {
constexpr int kWorkers = 4;
constexpr int kSlots = 2;
constexpr size_t kChunk = (size_t)8 << 20;
constexpr size_t kStagedMin = (size_t)32 << 20;
}  // namespace

struct Ctx::Staging {
  char* buf[kWorkers][kSlots] = {};
  cudaStream_t stream[kWorkers] = {};
  cudaEvent_t ev[kWorkers][kSlots] = {};
  cudaEvent_t done[kWorkers] = {};
  cudaEvent_t start = nullptr;
  ~Staging() {
    for (int w = 0; w < kWorkers; w++) {
      for (int s = 0; s < kSlots; s++) {
        if (buf[w][s]) cudaFreeHost(buf[w][s]);
        if (ev[w][s]) cudaEventDestroy(ev[w][s]);
      }
      if (done[w]) cudaEventDestroy(done[w]);
      if (stream[w]) cudaStreamDestroy(stream[w]);
    }
    if (start) cudaEventDestroy(start);
  }
};

void UploadBytes(Ctx* ctx, void* dst, const void* src, size_t bytes) {
  if (bytes == 0) return;
  if (bytes < kStagedMin || ctx->capturing) {
    CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return;
  }
  {
    // page-locked source (cudaHostAlloc / cudaHostRegister by the caller): one DMA, no staging
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, src) == cudaSuccess && at.type == cudaMemoryTypeHost) {
      CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
      return;
    }
    cudaGetLastError();  // an unregistered pointer is not an error here
  }
  if (!ctx->staging) {
    auto* st = new Ctx::Staging();
    ctx->staging = st;
    for (int w = 0; w < kWorkers; w++) {
      CUDA_CHECK(cudaStreamCreateWithFlags(&st->stream[w], cudaStreamNonBlocking));
      CUDA_CHECK(cudaEventCreateWithFlags(&st->done[w], cudaEventDisableTiming));
      for (int s = 0; s < kSlots; s++) {
        CUDA_CHECK(cudaHostAlloc(&st->buf[w][s], kChunk, cudaHostAllocDefault));
        CUDA_CHECK(cudaEventCreateWithFlags(&st->ev[w][s], cudaEventDisableTiming));
      }
    }
    CUDA_CHECK(cudaEventCreateWithFlags(&st->start, cudaEventDisableTiming));
  }
  Ctx::Staging* st = ctx->staging;
  // the destination may still be in use by earlier work on the context stream
  CUDA_CHECK(cudaEventRecord(st->start, ctx->stream));
  const size_t nchunks = (bytes + kChunk - 1) / kChunk;
  const int device = ctx->device;
  cudaError_t errs[kWorkers];
  std::vector<std::thread> threads;
  for (int w = 0; w < kWorkers; w++) {
    errs[w] = cudaSuccess;
    threads.emplace_back([=, &errs]() {
      cudaError_t e = cudaSetDevice(device);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(st->stream[w], st->start, 0);
      int slot = 0;
      for (size_t c = w; c < nchunks && e == cudaSuccess; c += kWorkers, slot ^= 1) {
        const size_t off = c * kChunk;
        const size_t len = std::min(kChunk, bytes - off);
        e = cudaEventSynchronize(st->ev[w][slot]);  // the DMA that last used this slot is done
        if (e != cudaSuccess) break;
        memcpy(st->buf[w][slot], static_cast<const char*>(src) + off, len);
        e = cudaMemcpyAsync(static_cast<char*>(dst) + off, st->buf[w][slot], len, cudaMemcpyHostToDevice,
                            st->stream[w]);
        if (e == cudaSuccess) e = cudaEventRecord(st->ev[w][slot], st->stream[w]);
      }
      if (e == cudaSuccess) e = cudaEventRecord(st->done[w], st->stream[w]);
      errs[w] = e;
    });
  }
  for (auto& t : threads) t.join();
  for (int w = 0; w < kWorkers; w++) {
    CUDA_CHECK(errs[w]);
    CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, st->done[w], 0));
  }
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
    if (!getenv("EPS_NO_POOL")) {
      // keep freed blocks in the device's default pool: the second solve of a process (same shapes)
      // allocates in microseconds instead of paying cudaMalloc/cudaFree for every GB-sized buffer
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t keep = ~0ull;
        if (cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep) == cudaSuccess) h->c.pooled = true;
      }
      cudaGetLastError();
    }
    RegisterCtx(&h->c);
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
