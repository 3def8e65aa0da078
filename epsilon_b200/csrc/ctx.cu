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
  sell_scratch2.release();
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

namespace {
struct StepTrace {
  int from = -2;           // -2: not read yet, -1: off
  int sell_launches = 0;
  static constexpr int kSlots = 9, kStride = 256 * 8;
  DevBuf<unsigned long long> buf;
  int used = 0;
  bool dumped = false;
  struct Rec { int kind, ctas; char label[48]; } recs[kSlots];
};
StepTrace g_step_trace;
}  // namespace

static void StepTracePreset(Ctx* ctx, StepTrace& t) {
  // slot s is a stream pass unless s % 3 == 2 (pass, pass 2, tail, pass, ...): the stream layout wants its
  // "first warp done" word preset for atomicMin; nothing is written between the traced launches
  std::vector<unsigned long long> init((size_t)StepTrace::kSlots * StepTrace::kStride, 0ull);
  for (int s = 0; s < StepTrace::kSlots; s++)
    if (s % 3 != 2)
      for (size_t i = 0; i < StepTrace::kStride; i += 4) init[(size_t)s * StepTrace::kStride + i + 2] = ~0ull;
  t.buf.upload(init.data(), init.size());
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
}

unsigned long long* StepTraceSlot(Ctx* ctx, int kind, int ctas, const char* label) {
  StepTrace& t = g_step_trace;
  if (t.from == -2) t.from = getenv("EPS_STEP_TRACE") ? atoi(getenv("EPS_STEP_TRACE")) : -1;
  if (t.from < 0 || t.dumped || ctas > 256) return nullptr;
  if (kind == 0) t.sell_launches++;
  if (t.sell_launches < t.from) return nullptr;
  if (!t.buf.p) {
    t.buf.alloc(ctx, (size_t)StepTrace::kSlots * StepTrace::kStride);
    StepTracePreset(ctx, t);
    return nullptr;     // this launch ran behind a synchronisation: start with a later one
  }
  const bool pass1 = kind == 0 && strstr(label, "pre-phase") == nullptr;
  if (t.used == 0 && !pass1) return nullptr;            // the window opens at a first pass of the CG step
  if (t.used == StepTrace::kSlots) {
    t.dumped = true;
    std::vector<unsigned long long> h((size_t)StepTrace::kSlots * StepTrace::kStride);
    t.buf.download(h.data(), h.size());
    unsigned long long t0 = ~0ull;
    for (int s = 0; s < t.used; s++)
      for (int c = 0; c < t.recs[s].ctas; c++) {
        const unsigned long long v = h[(size_t)s * StepTrace::kStride + (size_t)c * (t.recs[s].kind ? 8 : 4)];
        if (v) t0 = std::min(t0, v);
      }
    std::string out = Format("[step trace] rank %d: us since the first stamp; per stamp min/median/max over the CTAs\n", ctx->rank);
    for (int s = 0; s < t.used; s++) {
      const int per = t.recs[s].kind ? 8 : 4;
      std::string line = Format("[step trace] rank %d %-28s", ctx->rank, t.recs[s].label);
      for (int k = 0; k < per; k++) {
        std::vector<double> v;
        for (int c = 0; c < t.recs[s].ctas; c++) {
          const unsigned long long x = h[(size_t)s * StepTrace::kStride + (size_t)c * per + k];
          if (x && x != ~0ull) v.push_back((double)(x - t0) * 1e-3);
        }
        if (v.empty()) { line += " | -"; continue; }
        std::sort(v.begin(), v.end());
        line += Format(" | %.1f/%.1f/%.1f", v.front(), v[v.size() / 2], v.back());
      }
      out += line + "\n";
    }
    fputs(out.c_str(), stderr);
    return nullptr;
  }
  const bool expect_tail = t.used % 3 == 2, expect_pass2 = t.used % 3 == 1;
  if ((kind == 1) != expect_tail || (kind == 0 && pass1 == expect_pass2)) {
    // not the pass / pass 2 / tail rhythm (a solve started in between): start over at the next first pass
    t.used = 0;
    StepTracePreset(ctx, t);
    return nullptr;
  }
  StepTrace::Rec& r = t.recs[t.used];
  r.kind = kind; r.ctas = ctas;
  snprintf(r.label, sizeof(r.label), "%s", label);
  return t.buf.p + (size_t)(t.used++) * StepTrace::kStride;
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
