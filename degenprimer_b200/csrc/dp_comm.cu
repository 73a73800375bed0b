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
  if (ctx->comm) {
    dp_nccl_api* api = nccl_api(ctx);
    if (api) api->CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = nullptr;
  }
  if (ctx->d_wire_recv) { cudaFree(ctx->d_wire_recv); ctx->d_wire_recv = nullptr; }
  if (ctx->h_wire_recv) { cudaFreeHost(ctx->h_wire_recv); ctx->h_wire_recv = nullptr; }
  if (ctx->wire_owned && ctx->wire) { cudaFree(ctx->wire); }
  if (ctx->wire_owned) { ctx->wire = nullptr; ctx->wire_bytes = 0; ctx->wire_owned = false; }
  ctx->comm_world = ctx->comm_rank = 0;
}

int dp_comm_enqueue(dp_ctx* ctx) {
  if (!ctx->comm) return DP_OK;
  dp_nccl_api* api = nccl_api(ctx);
  if (!api) return DP_ERR_STATE;
  DP_NCCL(ctx, api, api->AllGather(ctx->wire, ctx->d_wire_recv, ctx->wire_bytes, DP_NCCL_UINT8,
                                   (ncclComm_t)ctx->comm, ctx->stream));
  DP_CUDA(ctx, cudaMemcpyAsync(ctx->h_wire_recv, ctx->d_wire_recv, (size_t)ctx->comm_world * ctx->wire_bytes,
                               cudaMemcpyDeviceToHost, ctx->stream));
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
  DP_CUDA(ctx, cudaMalloc(&ctx->wire, slot));
  ctx->wire_owned = true;
  ctx->wire_bytes = slot;
  DP_CUDA(ctx, cudaMemset(ctx->wire, 0, slot));
  DP_CUDA(ctx, cudaMalloc((void**)&ctx->d_wire_recv, (size_t)world * slot));
  DP_CUDA(ctx, cudaMallocHost((void**)&ctx->h_wire_recv, (size_t)world * slot));
  memset(ctx->h_wire_recv, 0, (size_t)world * slot);
  return DP_OK;
}

int dp_comm_destroy(dp_ctx* ctx) {
  if (!ctx) return DP_ERR_ARG;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  dp_comm_release(ctx);
  return DP_OK;
}

int dp_comm_gathered(dp_ctx* ctx, const void** host_ptr, uint64_t* slot_bytes, int* world) {
  if (!ctx || !host_ptr || !slot_bytes || !world) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->comm) return dp_set_error(ctx, DP_ERR_STATE, "no communicator: call dp_comm_init first");
  *host_ptr = ctx->h_wire_recv;
  *slot_bytes = ctx->wire_bytes;
  *world = ctx->comm_world;
  return DP_OK;
}

int dp_comm_hits(dp_ctx* ctx, dp_hit** hits_out, uint64_t* n_out) {
  if (!ctx || !hits_out || !n_out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *hits_out = nullptr;
  *n_out = 0;
  if (!ctx->comm) return dp_set_error(ctx, DP_ERR_STATE, "no communicator: call dp_comm_init first");
  const uint64_t slot = ctx->wire_bytes;
  const uint64_t cap = (slot - 64) / sizeof(dp_raw_hit);
  uint64_t total = 0;
  for (int r = 0; r < ctx->comm_world; ++r) {
    const uint64_t* h = (const uint64_t*)(ctx->h_wire_recv + (size_t)r * slot);
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
    const char* base = ctx->h_wire_recv + (size_t)r * slot;
    const uint64_t* h = (const uint64_t*)base;
    const uint64_t n = (uint32_t)h[0], rec_len = h[5];
    const dp_raw_hit* raw = (const dp_raw_hit*)(base + 64);
    for (uint64_t i = 0; i < n; ++i, ++w) {
      const uint32_t pat = raw[i].pattern;
      const uint32_t L = ctx->last_plan.pat_len[pat];
      const uint32_t strand = ctx->last_plan.pat_strand[pat];
      out[w].pos = strand ? rec_len - L - raw[i].g : raw[i].g;     // record 0 starts at padded position 0
      out[w].record = (uint32_t)r;
      out[w].primer = ctx->last_plan.pat_primer[pat];
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
