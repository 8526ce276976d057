// Data-parallel training over the GPUs of one box: replaces the reference's TrainNetworks / SyncNetworks
// (network_kernels.cu:295-484: one pthread per GPU, weights averaged THROUGH THE HOST every few iterations) with a
// gradient all-reduce over NVLink/NVSwitch:
//   * one process per GPU (torchrun), one NCCL communicator per process (unique id exchanged by the host plumbing),
//   * one ncclAllReduce(sum, fp32) per conv layer on its contiguous gradient bucket
//     [weight_updates | bias_updates | scale_updates], issued on a dedicated communication stream as soon as that
//     layer's weight-gradient kernel has been enqueued (event dependency), so it overlaps the remaining
//     dgrad / wgrad kernels of the backward pass,
//   * the update then runs on every GPU with batch = global batch (yc_network_update), which keeps replicas bit-
//     identical without ever exchanging weights.  Batchnorm statistics stay per GPU, as in the reference.
// NCCL is loaded with dlopen at first use (libnccl.so.2: the copy already mapped by the host process if there is one),
// so the library itself has no link-time dependency on it.
#include <dlfcn.h>
#include <string.h>

#include <vector>

#include "network_internal.hpp"

using namespace yc;

namespace {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclFloat = 7, ncclSum = 0 };  // nccl.h: ncclFloat32 = 7, ncclSum = 0

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
} g_nccl;

int load_nccl() {
  if (g_nccl.handle) return YC_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) {
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) {
    set_error("data-parallel training needs NCCL: dlopen(libnccl.so.2) failed: %s", dlerror());
    return YC_ERR_ARG;
  }
#define SYM(field, name)                                          \
  *(void**)(&g_nccl.field) = dlsym(h, name);                      \
  if (!g_nccl.field) {                                            \
    set_error("NCCL symbol %s not found", name);                  \
    return YC_ERR_ARG;                                            \
  }
  SYM(GetUniqueId, "ncclGetUniqueId")
  SYM(CommInitRank, "ncclCommInitRank")
  SYM(CommDestroy, "ncclCommDestroy")
  SYM(AllReduce, "ncclAllReduce")
  SYM(GroupStart, "ncclGroupStart")
  SYM(GroupEnd, "ncclGroupEnd")
  SYM(GetErrorString, "ncclGetErrorString")
  SYM(GetVersion, "ncclGetVersion")
#undef SYM
  g_nccl.handle = h;
  return YC_OK;
}

struct DpCtx {
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  cudaStream_t comm_stream = nullptr;
  std::vector<cudaEvent_t> ready;  // per layer: its weight gradient is complete on the compute stream
  cudaEvent_t done = nullptr;      // all all-reduces complete on the communication stream
  long long bytes_per_step = 0;
  int buckets = 0;
  bool overlap = true;   // all-reduce each bucket during backward; false: all buckets at once before the update
  bool pending = false;  // backward passes have run since the last reduction (overlap == false)
  // The buckets carry the momentum term divided by the world size (yc_network_update), which is only right if every
  // bucket is all-reduced exactly ONCE between two updates.  These two flags turn the sequences that would reduce a
  // bucket twice into errors instead of silently scaled gradients:
  bool reduced = false;  // the buckets hold all-reduced sums; no further backward pass may add to them before the update
};

#define NCCL_CALL(call)                                                                        \
  do {                                                                                         \
    ncclResult_t _r = (call);                                                                  \
    if (_r != 0) {                                                                             \
      set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(_r));      \
      return YC_ERR_CUDA;                                                                      \
    }                                                                                          \
  } while (0)

}  // namespace

// rank 0 creates the id (128 bytes); the host plumbing (torch.distributed / MPI / a file) hands it to every rank
extern "C" int yc_dp_unique_id(void* out128) {
  YC_REQUIRE(out128, "yc_dp_unique_id: null argument");
  int rc = load_nccl();
  if (rc) return rc;
  ncclUniqueId id;
  NCCL_CALL(g_nccl.GetUniqueId(&id));
  memcpy(out128, &id, sizeof(id));
  return YC_OK;
}

extern "C" int yc_dp_nccl_version(void) {
  if (load_nccl()) return -1;
  int v = 0;
  g_nccl.GetVersion(&v);
  return v;
}

// TrainNetworks' replica set-up (network_kernels.cu:405-484): joins this process' network to a world of `world`
// replicas.  Collective: every rank must call it.
extern "C" int yc_network_dp_init(yc_network* net, const void* unique_id128, int rank, int world) {
  YC_REQUIRE(net && unique_id128 && world >= 1 && rank >= 0 && rank < world, "yc_network_dp_init: bad arguments");
  YC_REQUIRE(net->flags & YC_NET_TRAIN, "yc_network_dp_init: network was not loaded with YC_NET_TRAIN");
  if (!net->finalized) {
    int rc = yc_network_finalize(net);
    if (rc) return rc;
  }
  int rc = load_nccl();
  if (rc) return rc;
  yc_dp_free(net);
  DpCtx* ctx = new DpCtx();
  ctx->rank = rank;
  ctx->world = world;
  ncclUniqueId id;
  memcpy(&id, unique_id128, sizeof(id));
  NCCL_CALL(g_nccl.CommInitRank(&ctx->comm, world, id, rank));
  NET_CUDA(cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking));
  ctx->ready.resize(net->layers.size(), nullptr);
  for (size_t i = 0; i < net->layers.size(); ++i)
    if (net->layers[i].type == L_CONV) {
      NET_CUDA(cudaEventCreateWithFlags(&ctx->ready[i], cudaEventDisableTiming));
      ctx->bytes_per_step += (long long)net->layers[i].bucket_floats * 4;
      ctx->buckets++;
    }
  NET_CUDA(cudaEventCreateWithFlags(&ctx->done, cudaEventDisableTiming));
  // with subdivisions > 1 the buckets accumulate over several backward passes: reduce once, before the update
  ctx->overlap = net->subdivisions == 1;
  net->dp = ctx;
  net->train_flags = world;  // yc_network_update divides by the GLOBAL batch
  return YC_OK;
}

int yc_dp_layer_ready(yc_network* net, int layer, cudaStream_t compute) {
  DpCtx* ctx = (DpCtx*)net->dp;
  if (!ctx || ctx->world <= 1) return YC_OK;
  if (ctx->reduced) {
    set_error("data parallelism: a backward pass after the gradients of this step were already all-reduced (a second "
              "yc_network_forward_backward before yc_network_update under per-layer overlap, or one after "
              "yc_network_dp_reduce): the buckets would be reduced twice.  With several backward passes per update load "
              "the network with subdivisions > 1 or call yc_network_dp_set_overlap(net, 0)");
    return YC_ERR_ARG;
  }
  if (!ctx->overlap) {
    ctx->pending = true;
    return YC_OK;
  }
  NLayer& L = net->layers[layer];
  NET_CUDA(cudaEventRecord(ctx->ready[layer], compute));
  NET_CUDA(cudaStreamWaitEvent(ctx->comm_stream, ctx->ready[layer], 0));
  NCCL_CALL(g_nccl.AllReduce(L.dw_packed, L.dw_packed, L.bucket_floats, ncclFloat, ncclSum, ctx->comm, ctx->comm_stream));
  return YC_OK;
}

int yc_dp_backward_done(yc_network* net, cudaStream_t compute) {
  DpCtx* ctx = (DpCtx*)net->dp;
  if (!ctx || ctx->world <= 1 || !ctx->overlap) return YC_OK;
  NET_CUDA(cudaEventRecord(ctx->done, ctx->comm_stream));
  NET_CUDA(cudaStreamWaitEvent(compute, ctx->done, 0));
  ctx->reduced = true;  // until yc_network_update
  return YC_OK;
}

// yc_network_update has consumed the reduced gradients: the next backward pass starts a new step
void yc_dp_update_done(yc_network* net) {
  DpCtx* ctx = (DpCtx*)net->dp;
  if (ctx) ctx->reduced = false;
}

// All buckets in one NCCL group, after everything enqueued on `compute`; `compute` then waits for the result.  Called by
// yc_network_update when the per-layer overlap is off (subdivisions > 1, or yc_network_dp_set_overlap(net, 0)).
int yc_dp_reduce_pending(yc_network* net, cudaStream_t compute) {
  DpCtx* ctx = (DpCtx*)net->dp;
  if (!ctx || ctx->world <= 1 || !ctx->pending) return YC_OK;
  NET_CUDA(cudaEventRecord(ctx->done, compute));
  NET_CUDA(cudaStreamWaitEvent(ctx->comm_stream, ctx->done, 0));
  NCCL_CALL(g_nccl.GroupStart());
  for (NLayer& L : net->layers)
    if (L.type == L_CONV)
      NCCL_CALL(g_nccl.AllReduce(L.dw_packed, L.dw_packed, L.bucket_floats, ncclFloat, ncclSum, ctx->comm, ctx->comm_stream));
  NCCL_CALL(g_nccl.GroupEnd());
  NET_CUDA(cudaEventRecord(ctx->done, ctx->comm_stream));
  NET_CUDA(cudaStreamWaitEvent(compute, ctx->done, 0));
  ctx->pending = false;
  ctx->reduced = true;  // until yc_network_update
  return YC_OK;
}

// on = 1 (default when subdivisions == 1): one all-reduce per layer bucket, overlapped with the backward pass;
// on = 0: gradients stay local until yc_network_dp_reduce / yc_network_update.
extern "C" int yc_network_dp_set_overlap(yc_network* net, int on) {
  YC_REQUIRE(net && net->dp, "yc_network_dp_set_overlap: network has no data-parallel context");
  DpCtx* ctx = (DpCtx*)net->dp;
  YC_REQUIRE(!on || net->subdivisions == 1, "yc_network_dp_set_overlap: per-layer overlap needs subdivisions=1");
  YC_REQUIRE(!ctx->pending, "yc_network_dp_set_overlap: un-reduced gradients pending");
  ctx->overlap = on != 0;
  return YC_OK;
}

// The caller throws this step's gradients away (timing loops that run backward passes without updates): the next
// backward pass may all-reduce again.  The buckets' CONTENTS are then meaningless until the caller zeroes them.
extern "C" int yc_network_dp_discard_step(yc_network* net) {
  YC_REQUIRE(net && net->dp, "yc_network_dp_discard_step: network has no data-parallel context");
  ((DpCtx*)net->dp)->reduced = false;
  return YC_OK;
}

// Reduce the accumulated local gradients now (no-op if nothing is pending).  Collective.
extern "C" int yc_network_dp_reduce(yc_network* net, void* stream) {
  YC_REQUIRE(net && net->dp, "yc_network_dp_reduce: network has no data-parallel context");
  return yc_dp_reduce_pending(net, stream ? (cudaStream_t)stream : net->stream);
}

void yc_dp_free(yc_network* net) {
  DpCtx* ctx = (DpCtx*)net->dp;
  if (!ctx) return;
  if (ctx->comm_stream) cudaStreamSynchronize(ctx->comm_stream);
  if (ctx->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(ctx->comm);
  for (cudaEvent_t e : ctx->ready)
    if (e) cudaEventDestroy(e);
  if (ctx->done) cudaEventDestroy(ctx->done);
  if (ctx->comm_stream) cudaStreamDestroy(ctx->comm_stream);
  delete ctx;
  net->dp = nullptr;
  net->train_flags = 0;
}

// what: 0 world, 1 rank, 2 buckets per step, 3 MiB all-reduced per step, 4 per-layer overlap on
extern "C" int yc_network_dp_info(const yc_network* net, int what) {
  if (!net || !net->dp) return what == 0 ? 1 : 0;
  const DpCtx* ctx = (const DpCtx*)net->dp;
  switch (what) {
    case 0: return ctx->world;
    case 1: return ctx->rank;
    case 2: return ctx->buckets;
    case 3: return (int)(ctx->bytes_per_step >> 20);
    case 4: return ctx->overlap ? 1 : 0;
  }
  return -1;
}
