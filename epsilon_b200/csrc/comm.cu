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
enum { ncclFloat64 = 8 };  // ncclDataType_t
enum { ncclSum = 0 };      // ncclRedOp_t

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
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
}

void CommDestroy(Ctx* ctx) {
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
