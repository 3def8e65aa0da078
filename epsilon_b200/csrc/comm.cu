// Row-sharded multi-GPU: NCCL all-reduce of A'y and of scalar partial sums over
// NVLink 5 / NVSwitch.  The reference is single-process and single-threaded
// (tools/run_benchmarks.sh:15-17), so there is nothing it replaces; SURVEY section 8e.
// NCCL is bound at run time with dlopen so that a single-GPU process never needs it and
// so that the copy already loaded by torch.distributed (if any) is the one used.
#include <dlfcn.h>

#include "internal.h"

namespace eps {

namespace {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclFloat64 = 8 };  // ncclDataType_t
enum { ncclSum = 0 };      // ncclRedOp_t

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi& Api() {
  static NcclApi api;
  if (api.lib) return api;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) throw Error(EPS_ERR_NCCL, Format("cannot dlopen libnccl.so.2: %s", dlerror()));
  auto sym = [&](const char* s) {
    void* p = dlsym(api.lib, s);
    if (!p) throw Error(EPS_ERR_NCCL, Format("libnccl lacks %s", s));
    return p;
  };
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
  api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
  api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  return api;
}

#define NCCL_CHECK(expr)                                                                 \
  do {                                                                                   \
    ncclResult_t r_ = (expr);                                                            \
    if (r_ != ncclSuccess)                                                               \
      throw Error(EPS_ERR_NCCL, Format("%s: %s", #expr, Api().GetErrorString(r_)));      \
  } while (0)

}  // namespace

struct Comm {
  ncclComm_t comm = nullptr;
};

void CommUniqueId(void* out128) {
  ncclUniqueId id;
  NCCL_CHECK(Api().GetUniqueId(&id));
  memcpy(out128, &id, sizeof(id));
}

void CommInit(Ctx* ctx, int n_ranks, int rank, const void* unique_id) {
  EPS_CHECK(n_ranks >= 1 && rank >= 0 && rank < n_ranks, "rank %d of %d", rank, n_ranks);
  ctx->n_ranks = n_ranks;
  ctx->rank = rank;
  if (n_ranks == 1) return;
  CUDA_CHECK(cudaSetDevice(ctx->device));
  ncclUniqueId id;
  memcpy(&id, unique_id, sizeof(id));
  auto* c = new Comm();
  NCCL_CHECK(Api().CommInitRank(&c->comm, n_ranks, id, rank));
  ctx->comm = c;
  // map the peer buffers now (CUDA IPC opens cost ~0.1 s per peer): communicator set-up, not
  // solver set-up.  Grown on demand for longer vectors.
  const char* cap = getenv("EPS_PEER_CAPACITY");
  PeerEnsure(ctx, cap ? atoll(cap) : (int64_t)1 << 18);
}

void CommDestroy(Ctx* ctx) {
  if (ctx->peer) { PeerRelease(ctx->peer); delete ctx->peer; ctx->peer = nullptr; }
  if (!ctx->comm) return;
  if (ctx->comm->comm) Api().CommDestroy(ctx->comm->comm);
  delete ctx->comm;
  ctx->comm = nullptr;
}

void AllReduceSum(Ctx* ctx, double* buf, int64_t n) {
  if (ctx->n_ranks <= 1 || n <= 0) return;
  EPS_CHECK(ctx->comm, "communicator not initialised");
  ProfScope ps(ctx, KC_ALLREDUCE, 8.0 * n);
  NCCL_CHECK(Api().AllReduce(buf, buf, (size_t)n, ncclFloat64, ncclSum, ctx->comm->comm, ctx->stream));
  ctx->count_launch();
}

// ---- one-shot peer all-reduce: buffers mapped into every rank through CUDA IPC -------------
// Layout of each rank's block: [flags: 2*G u64 | counter at 2 KB | per-CTA flags 2*G*kPeerMaxCtas u64 at 4 KB,
// head padded to 64 KB][receive: 2*G*n f64][packet inboxes: 2*G*n + 2*n uint4].
static void PeerSetup(Ctx* ctx, int64_t n, PeerBuf* out) {
  out->ready = false;
  const int G = ctx->n_ranks;
  const size_t head = 65536;
  static_assert(4096 + 2 * kMaxRanks * kPeerMaxCtas * 8 <= 65536, "peer header");
  const size_t recv_bytes = sizeof(double) * 2 * (size_t)G * n;
  const size_t ll1_bytes = sizeof(uint4) * 2 * (size_t)G * n, ll2_bytes = sizeof(uint4) * 2 * (size_t)n;
  const bool use_ll = !getenv("EPS_PEER_LL") || atoi(getenv("EPS_PEER_LL")) != 0;
  const bool use_sn = use_ll && (!getenv("EPS_PEER_SENTINEL") || atoi(getenv("EPS_PEER_SENTINEL")) != 0);
  const size_t sn1_bytes = sizeof(double) * 2 * (size_t)G * n, sn2_bytes = sizeof(double) * 2 * (size_t)n;
  const size_t sn_off = head + recv_bytes + (use_ll ? ll1_bytes + ll2_bytes : 0);
  const size_t bytes = sn_off + (use_sn ? sn1_bytes + sn2_bytes : 0);
  CUDA_CHECK(cudaMalloc(&out->block, bytes));
  ctx->bytes_allocated += bytes;
  CUDA_CHECK(cudaMemsetAsync(out->block, 0, bytes, ctx->stream));
  if (use_sn) {
    // 64-bit pattern fill of the sentinel inboxes (cudaMemset only knows bytes / 32-bit words: two D2D copies
    // from a host vector would do as well; the driver's 32-bit 2D memset writes the two halves)
    std::vector<unsigned long long> fill((sn1_bytes + sn2_bytes) / 8, kPeerSentinel);
    CUDA_CHECK(cudaMemcpyAsync(static_cast<char*>(out->block) + sn_off, fill.data(), sn1_bytes + sn2_bytes,
                               cudaMemcpyHostToDevice, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  }
  cudaIpcMemHandle_t mine;
  int ok = cudaIpcGetMemHandle(&mine, out->block) == cudaSuccess ? 1 : 0;
  if (!ok) cudaGetLastError();
  // gather {ok, handle} of every rank
  struct Rec { int ok; int pad; cudaIpcMemHandle_t h; };
  Rec rec; memset(&rec, 0, sizeof(rec));
  rec.ok = ok; rec.h = mine;
  DevBuf<char> send(ctx, sizeof(Rec)), recv(ctx, sizeof(Rec) * G);
  CUDA_CHECK(cudaMemcpyAsync(send.p, &rec, sizeof(Rec), cudaMemcpyHostToDevice, ctx->stream));
  NCCL_CHECK(Api().AllGather(send.p, recv.p, sizeof(Rec), ncclInt8, ctx->comm->comm, ctx->stream));
  std::vector<Rec> all(G);
  CUDA_CHECK(cudaMemcpyAsync(all.data(), recv.p, sizeof(Rec) * G, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  int good = 1;
  for (int r = 0; r < G; r++) good &= all[r].ok;
  for (int r = 0; r < G && good; r++) {
    if (r == ctx->rank) { out->mapped[r] = nullptr; continue; }
    if (cudaIpcOpenMemHandle(&out->mapped[r], all[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      out->mapped[r] = nullptr;
      good = 0;
    }
  }
  // every rank must take the same decision
  DevBuf<double> agree(ctx, 1);
  double g = good ? 0.0 : 1.0;
  CUDA_CHECK(cudaMemcpyAsync(agree.p, &g, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  NCCL_CHECK(Api().AllReduce(agree.p, agree.p, 1, ncclFloat64, ncclSum, ctx->comm->comm, ctx->stream));
  CUDA_CHECK(cudaMemcpyAsync(&g, agree.p, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  if (g != 0.0) { PeerRelease(out); return; }
  out->x.G = G;
  out->x.rank = ctx->rank;
  out->x.n = n;
  for (int r = 0; r < G; r++) {
    char* base = static_cast<char*>(r == ctx->rank ? out->block : out->mapped[r]);
    out->x.flag[r] = reinterpret_cast<unsigned long long*>(base);
    out->x.cflag[r] = reinterpret_cast<unsigned long long*>(base + 4096);
    out->x.recv[r] = reinterpret_cast<double*>(base + head);
    out->x.ll1[r] = use_ll ? reinterpret_cast<uint4*>(base + head + recv_bytes) : nullptr;
    out->x.ll2[r] = use_ll ? reinterpret_cast<uint4*>(base + head + recv_bytes + ll1_bytes) : nullptr;
    out->x.sn1[r] = use_sn ? reinterpret_cast<double*>(base + sn_off) : nullptr;
    out->x.sn2[r] = use_sn ? reinterpret_cast<double*>(base + sn_off + sn1_bytes) : nullptr;
  }
  out->counter = reinterpret_cast<unsigned*>(static_cast<char*>(out->block) + 2048);
  out->seq = 0;
  out->ready = true;
}

PeerBuf* PeerEnsure(Ctx* ctx, int64_t n) {
  const int G = ctx->n_ranks;
  if (G <= 1 || G > kMaxRanks || !ctx->comm || getenv("EPS_NO_PEER")) return nullptr;
  if (ctx->peer && ctx->peer->ready && ctx->peer->x.n >= n) return ctx->peer;
  EPS_CHECK(!ctx->capturing, "peer buffer growth during graph capture");
  CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
  // Growth keeps the PeerBuf OBJECT: live solvers hold its address (KktSolver::peer) and read `x` afresh at every
  // launch, so they simply see the larger buffers.  Order: every rank closes its mappings of the old blocks, all
  // ranks meet (one tiny all-reduce), only then is the exported block freed.
  PeerBuf* pb = ctx->peer;
  if (pb) {
    for (int r = 0; r < kMaxRanks; r++)
      if (pb->mapped[r]) { cudaIpcCloseMemHandle(pb->mapped[r]); pb->mapped[r] = nullptr; }
    DevBuf<double> meet(ctx, 1);
    meet.zero();
    NCCL_CHECK(Api().AllReduce(meet.p, meet.p, 1, ncclFloat64, ncclSum, ctx->comm->comm, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    PeerRelease(pb);
    pb->x = PeerXchg();
  } else {
    pb = new PeerBuf();
    ctx->peer = pb;
  }
  PeerSetup(ctx, n, pb);
  return pb->ready ? pb : nullptr;
}

void PeerRelease(PeerBuf* pb) {
  for (int r = 0; r < kMaxRanks; r++)
    if (pb->mapped[r]) { cudaIpcCloseMemHandle(pb->mapped[r]); pb->mapped[r] = nullptr; }
  if (pb->block) { cudaFree(pb->block); pb->block = nullptr; }
  pb->ready = false;
}

}  // namespace eps

extern "C" {

int eps_nccl_unique_id(void* out_128_bytes) {
  try {
    eps::CommUniqueId(out_128_bytes);
    return EPS_OK;
  } catch (const eps::Error& e) {
    eps::SetLastError(e.what());
    return e.code;
  }
}

int eps_ctx_init_comm(eps_ctx* ctx, int n_ranks, int rank, const void* unique_id) {
  try {
    eps::CommInit(&ctx->c, n_ranks, rank, unique_id);
    return EPS_OK;
  } catch (const eps::Error& e) {
    eps::SetLastError(e.what());
    return e.code;
  }
}

}  // extern "C"
