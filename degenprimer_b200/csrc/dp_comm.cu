// Multi-GPU hit exchange: NCCL all-gather of the wire buffer, queued on the
// context's stream right behind the publish kernel (one exchange step
// at the end of the scan, KB-sized, latency bound).
//
// NCCL is bound at run time (dlopen of the libnccl.so.2 that torch ships and has
// already loaded in a torchrun process), so the library itself has no link-time
// dependency on it and single-GPU users never touch it.  Rank 0 makes the unique
// id, the host side (degenprimer_b200/distributed.py) broadcasts it with
// torch.distributed, every rank calls dp_comm_init.
#include <algorithm>
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include "dp_internal.cuh"

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { DP_NCCL_UINT8 = 1 };

struct dp_nccl_api {
  void* lib;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  const char* (*GetErrorString)(ncclResult_t);
};

static dp_nccl_api* nccl_api(dp_ctx* ctx) {
  static dp_nccl_api api;
  static bool tried = false, ok = false;
  if (!tried) {
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (api.lib) {
      api.GetUniqueId = (ncclResult_t(*)(ncclUniqueId*))dlsym(api.lib, "ncclGetUniqueId");
      api.CommInitRank = (ncclResult_t(*)(ncclComm_t*, int, ncclUniqueId, int))dlsym(api.lib, "ncclCommInitRank");
      api.AllGather = (ncclResult_t(*)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t))dlsym(api.lib, "ncclAllGather");
      api.CommDestroy = (ncclResult_t(*)(ncclComm_t))dlsym(api.lib, "ncclCommDestroy");
      api.GetErrorString = (const char* (*)(ncclResult_t))dlsym(api.lib, "ncclGetErrorString");
      ok = api.GetUniqueId && api.CommInitRank && api.AllGather && api.CommDestroy && api.GetErrorString;
    }
  }
  if (!ok) {
    dp_set_error(ctx, DP_ERR_STATE, "NCCL is not available: libnccl.so.2 could not be loaded (%s)",
                 api.lib ? "symbols missing" : "dlopen failed; import torch first or set LD_LIBRARY_PATH");
    return nullptr;
  }
  return &api;
}

#define DP_NCCL(ctx, api, call)                                                               \
  do {                                                                                        \
    ncclResult_t r__ = (call);                                                                \
    if (r__ != 0)                                                                             \
      return dp_set_error((ctx), DP_ERR_CUDA, "%s failed: %s", #call, (api)->GetErrorString(r__)); \
  } while (0)

void dp_comm_release(dp_ctx* ctx) {
  if (!ctx) return;
  if (ctx->comm_stream) cudaStreamSynchronize(ctx->comm_stream);
  if (ctx->comm) {
    dp_nccl_api* api = nccl_api(ctx);
    if (api) api->CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = nullptr;
  }
  for (int i = 0; i < 2; ++i) {
    if (ctx->d_wire_recv[i]) { cudaFree(ctx->d_wire_recv[i]); ctx->d_wire_recv[i] = nullptr; }
    if (ctx->h_wire_recv[i]) { cudaFreeHost(ctx->h_wire_recv[i]); ctx->h_wire_recv[i] = nullptr; }
    if (ctx->wire_slot[i]) { cudaFree(ctx->wire_slot[i]); ctx->wire_slot[i] = nullptr; }
    if (ctx->gather_done[i]) { cudaEventDestroy(ctx->gather_done[i]); ctx->gather_done[i] = nullptr; }
  }
  if (ctx->wire_ready) { cudaEventDestroy(ctx->wire_ready); ctx->wire_ready = nullptr; }
  if (ctx->comm_stream) { cudaStreamDestroy(ctx->comm_stream); ctx->comm_stream = nullptr; }
  if (ctx->wire_owned) { ctx->wire = nullptr; ctx->wire_bytes = 0; ctx->wire_owned = false; }
  ctx->comm_world = ctx->comm_rank = 0;
  ctx->comm_seq = 0;
}

int dp_comm_prepare(dp_ctx* ctx) {
  if (!ctx->comm) return DP_OK;
  const int slot = (int)(ctx->comm_seq & 1);
  ctx->wire = ctx->wire_slot[slot];
  // the gather that used this slot two calls ago must have read it before it is rewritten
  if (ctx->comm_seq >= 2) DP_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->gather_done[slot], 0));
  return DP_OK;
}

int dp_comm_enqueue(dp_ctx* ctx) {
  if (!ctx->comm) return DP_OK;
  dp_nccl_api* api = nccl_api(ctx);
  if (!api) return DP_ERR_STATE;
  const int slot = (int)(ctx->comm_seq & 1);
  DP_CUDA(ctx, cudaEventRecord(ctx->wire_ready, ctx->stream));
  DP_CUDA(ctx, cudaStreamWaitEvent(ctx->comm_stream, ctx->wire_ready, 0));
  DP_NCCL(ctx, api, api->AllGather(ctx->wire_slot[slot], ctx->d_wire_recv[slot], ctx->wire_bytes, DP_NCCL_UINT8,
                                   (ncclComm_t)ctx->comm, ctx->comm_stream));
  DP_CUDA(ctx, cudaMemcpyAsync(ctx->h_wire_recv[slot], ctx->d_wire_recv[slot], (size_t)ctx->comm_world * ctx->wire_bytes,
                               cudaMemcpyDeviceToHost, ctx->comm_stream));
  DP_CUDA(ctx, cudaEventRecord(ctx->gather_done[slot], ctx->comm_stream));
  ctx->comm_plan[slot] = ctx->last_plan;
  ++ctx->comm_seq;
  return DP_OK;
}

// slot of the call `age` calls ago (0 = the one that just returned); waits for its gather
static int comm_slot(dp_ctx* ctx, int age, int* slot) {
  if (!ctx->comm) return dp_set_error(ctx, DP_ERR_STATE, "no communicator: call dp_comm_init first");
  if (age < 0 || age > 1 || ctx->comm_seq < (uint64_t)age + 1)
    return dp_set_error(ctx, DP_ERR_ARG, "no gathered call of age %d (the exchange keeps the last two)", age);
  *slot = (int)((ctx->comm_seq - 1 - (uint64_t)age) & 1);
  DP_CUDA(ctx, cudaEventSynchronize(ctx->gather_done[*slot]));
  return DP_OK;
}

extern "C" {

int dp_comm_unique_id(dp_ctx* ctx, char id[128]) {
  if (!ctx || !id) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  dp_nccl_api* api = nccl_api(ctx);
  if (!api) return DP_ERR_STATE;
  ncclUniqueId u;
  DP_NCCL(ctx, api, api->GetUniqueId(&u));
  memcpy(id, u.internal, 128);
  return DP_OK;
}

int dp_comm_init(dp_ctx* ctx, const char id[128], int world, int rank, uint32_t max_hits) {
  if (!ctx || !id) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (world < 1 || rank < 0 || rank >= world || max_hits == 0)
    return dp_set_error(ctx, DP_ERR_ARG, "bad world/rank/max_hits");
  dp_nccl_api* api = nccl_api(ctx);
  if (!api) return DP_ERR_STATE;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  dp_comm_release(ctx);
  ncclUniqueId u;
  memcpy(u.internal, id, 128);
  ncclComm_t comm = nullptr;
  DP_NCCL(ctx, api, api->CommInitRank(&comm, world, u, rank));
  ctx->comm = comm;
  ctx->comm_world = world;
  ctx->comm_rank = rank;
  const uint64_t slot = 64 + (uint64_t)max_hits * sizeof(dp_raw_hit);
  int lo_prio = 0, hi_prio = 0;
  DP_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
  DP_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->comm_stream, cudaStreamNonBlocking, hi_prio));
  DP_CUDA(ctx, cudaEventCreateWithFlags(&ctx->wire_ready, cudaEventDisableTiming));
  for (int i = 0; i < 2; ++i) {
    DP_CUDA(ctx, cudaEventCreateWithFlags(&ctx->gather_done[i], cudaEventDisableTiming));
    DP_CUDA(ctx, cudaMalloc((void**)&ctx->wire_slot[i], slot));
    DP_CUDA(ctx, cudaMemset(ctx->wire_slot[i], 0, slot));
    DP_CUDA(ctx, cudaMalloc((void**)&ctx->d_wire_recv[i], (size_t)world * slot));
    DP_CUDA(ctx, cudaMallocHost((void**)&ctx->h_wire_recv[i], (size_t)world * slot));
    memset(ctx->h_wire_recv[i], 0, (size_t)world * slot);
  }
  ctx->wire = ctx->wire_slot[0];
  ctx->wire_owned = true;
  ctx->wire_bytes = slot;
  ctx->comm_seq = 0;
  return DP_OK;
}

int dp_comm_destroy(dp_ctx* ctx) {
  if (!ctx) return DP_ERR_ARG;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  dp_comm_release(ctx);
  return DP_OK;
}

int dp_comm_gathered(dp_ctx* ctx, int age, const void** host_ptr, uint64_t* slot_bytes, int* world) {
  if (!ctx || !host_ptr || !slot_bytes || !world) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  int slot = 0;
  int rc = comm_slot(ctx, age, &slot);
  if (rc) return rc;
  *host_ptr = ctx->h_wire_recv[slot];
  *slot_bytes = ctx->wire_bytes;
  *world = ctx->comm_world;
  return DP_OK;
}

int dp_comm_hits(dp_ctx* ctx, int age, dp_hit** hits_out, uint64_t* n_out) {
  if (!ctx || !hits_out || !n_out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *hits_out = nullptr;
  *n_out = 0;
  int s = 0;
  int rc = comm_slot(ctx, age, &s);
  if (rc) return rc;
  const char* recv = ctx->h_wire_recv[s];
  const dp_dg_plan& plan = ctx->comm_plan[s];
  const uint64_t slot = ctx->wire_bytes;
  const uint64_t cap = (slot - 64) / sizeof(dp_raw_hit);
  uint64_t total = 0;
  for (int r = 0; r < ctx->comm_world; ++r) {
    const uint64_t* h = (const uint64_t*)(recv + (size_t)r * slot);
    const uint64_t n = (uint32_t)h[0];
    if (n > cap) return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "rank %d had %llu hits, a slot holds %llu", r,
                                     (unsigned long long)n, (unsigned long long)cap);
    if (h[6] != 1) return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "rank %d scanned a template of %llu records; "
                                       "dp_comm_hits handles one segment per rank", r, (unsigned long long)h[6]);
    total += n;
  }
  dp_hit* out = (dp_hit*)dp_result_alloc(std::max<uint64_t>(1, total) * sizeof(dp_hit));
  if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu hits", (unsigned long long)total);
  uint64_t w = 0;
  for (int r = 0; r < ctx->comm_world; ++r) {
    const char* base = recv + (size_t)r * slot;
    const uint64_t* h = (const uint64_t*)base;
    const uint64_t n = (uint32_t)h[0], rec_len = h[5];
    const dp_raw_hit* raw = (const dp_raw_hit*)(base + 64);
    for (uint64_t i = 0; i < n; ++i, ++w) {
      const uint32_t pat = raw[i].pattern;
      const uint32_t L = plan.pat_len[pat];
      const uint32_t strand = plan.pat_strand[pat];
      out[w].pos = strand ? rec_len - L - raw[i].g : raw[i].g;     // record 0 starts at padded position 0
      out[w].record = (uint32_t)r;
      out[w].primer = plan.pat_primer[pat];
      out[w].strand = (uint8_t)strand;
      out[w].score3 = (uint8_t)raw[i].score3;
    }
  }
  std::sort(out, out + total, [](const dp_hit& a, const dp_hit& b) {
    if (a.primer != b.primer) return a.primer < b.primer;
    if (a.strand != b.strand) return a.strand < b.strand;
    if (a.record != b.record) return a.record < b.record;
    return a.pos < b.pos;
  });
  *hits_out = out;
  *n_out = total;
  return DP_OK;
}

}  // extern "C"
