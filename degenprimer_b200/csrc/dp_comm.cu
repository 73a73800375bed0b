// Multi-GPU hit exchange over peer memory (one exchange step at the end
// of the scan, KB-sized, latency bound).
//
// Every rank owns two receive buffers (world slots each: 64-byte header + raw hits)
// that its peers map through CUDA IPC.  The publish kernel of call c -- the last
// kernel of the call anyway -- stores this rank's slot into receive buffer c&1 of
// EVERY rank over NVLink and raises the slot's sequence word behind a system-scope
// fence.  There is no separate collective launch: the gather is the epilogue of the
// compute.  The same kernel first moves call c-1's slots (all ranks) from the local
// receive buffer to pinned host memory, which is also the flow control: a rank
// publishes call c only after it has seen every peer's call c-1, so peers are never
// more than one call apart and a buffer is never overwritten before it was drained.
//
// The host side (degenprimer_b200/distributed.py) moves the IPC handles between the
// ranks with torch.distributed; this file has no dependency on it or on NCCL.
#include <algorithm>
#include <vector>
#include <stdlib.h>
#include <string.h>

#include "dp_internal.cuh"

void dp_comm_release(dp_ctx* ctx) {
  if (!ctx->comm_world) return;
  cudaStreamSynchronize(ctx->stream);
  for (int b = 0; b < 2; ++b) {
    for (int r = 0; r < ctx->comm_world; ++r) {
      if (r != ctx->comm_rank && ctx->peer_mapped[b][r]) cudaIpcCloseMemHandle(ctx->peer_mapped[b][r]);
      ctx->peer_mapped[b][r] = nullptr;
    }
    if (ctx->d_recv[b]) cudaFree(ctx->d_recv[b]);
    if (ctx->d_peer[b]) cudaFree(ctx->d_peer[b]);
    if (ctx->h_gather[b]) cudaFreeHost(ctx->h_gather[b]);
    if (ctx->h_sorted[b]) cudaFreeHost(ctx->h_sorted[b]);
    ctx->h_sorted[b] = nullptr;
    ctx->d_recv[b] = nullptr;
    ctx->d_peer[b] = nullptr;
    ctx->h_gather[b] = nullptr;
  }
  if (ctx->comm_err) cudaFreeHost(ctx->comm_err);
  ctx->comm_err = nullptr;
  ctx->comm_world = ctx->comm_rank = 0;
  ctx->comm_slot_bytes = 0;
  ctx->comm_seq = 0;
  ctx->comm_layout.clear();
  ctx->comm_layout_base.clear();
}

// (length, strand, primer) of the patterns of `call`, for the kernel that converts its hits
static void fill_plan(dp_ctx* ctx, dp_exchange_args* X, uint64_t call) {
  const dp_dg_plan& plan = ctx->comm_plan[call & 3];
  memcpy(X->prev_len, plan.pat_len, DP_MAX_PATTERNS);
  memcpy(X->prev_strand, plan.pat_strand, DP_MAX_PATTERNS);
  memcpy(X->prev_primer, plan.pat_primer, DP_MAX_PATTERNS);
}

void dp_comm_fill(dp_ctx* ctx, dp_exchange_args* X) {
  X->world = 0;
  if (!ctx->comm_world || !ctx->d_peer[0]) return;
  const uint64_t c = ctx->comm_seq;
  X->world = ctx->comm_world;
  X->rank = ctx->comm_rank;
  X->seq = c;
  X->slot_bytes = ctx->comm_slot_bytes;
  X->slot_hits = (ctx->comm_slot_bytes - 64) / sizeof(dp_raw_hit);
  X->peers = ctx->d_peer[c & 1];
  X->recv_prev = ctx->d_recv[(c + 1) & 1];
  X->h_prev = ctx->h_gather[(c + 1) & 1];
  X->h_sorted = ctx->h_sorted[(c + 1) & 1];
  if (c) fill_plan(ctx, X, c - 1);
  X->err = ctx->comm_err;
  X->rec0_len = ctx->last_rec0_len;
  X->n_records = ctx->last_n_records;
  ctx->comm_plan[c & 3] = ctx->last_plan;
  ctx->comm_seq = c + 1;
}

// Host buffer that holds the call made `age` calls ago.  With c = the last call queued:
//   age 2  call c-2 was moved to the host by the publish kernel of call c-1, which has
//          completed (calls do not overlap): nothing to wait for, call c may be in flight;
//   age 1  call c-1 is moved by the publish kernel of call c: wait for the stream;
//   age 0  call c is still in the device receive buffer: one small kernel waits for the
//          slowest rank's slot and moves it.
static int comm_collect(dp_ctx* ctx, int age, int* buf, uint64_t* call_out) {
  if (!ctx->comm_world || !ctx->d_peer[0]) return dp_set_error(ctx, DP_ERR_STATE, "no exchange: call dp_comm_export and dp_comm_attach first");
  if (age < 0 || age > 2 || ctx->comm_seq < (uint64_t)age + 1)
    return dp_set_error(ctx, DP_ERR_ARG, "no gathered call of age %d (the exchange keeps the last three)", age);
  const uint64_t call = ctx->comm_seq - 1 - (uint64_t)age;
  const int b = (int)(call & 1);
  if (age == 0) {
    // the slots of the last call are still on the device (the next call would drain them)
    dp_exchange_args X;
    memset(&X, 0, sizeof X);
    X.world = ctx->comm_world;
    X.rank = ctx->comm_rank;
    X.seq = call;
    X.slot_bytes = ctx->comm_slot_bytes;
    X.slot_hits = (ctx->comm_slot_bytes - 64) / sizeof(dp_raw_hit);
    X.recv_prev = ctx->d_recv[b];
    X.h_prev = ctx->h_gather[b];
    X.h_sorted = ctx->h_sorted[b];
    fill_plan(ctx, &X, call);
    X.err = ctx->comm_err;
    dp_exchange_drain_launch(X, ctx->stream);
    DP_CUDA(ctx, cudaGetLastError());
  }
  if (age < 2) DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (*(volatile unsigned long long*)ctx->comm_err)
    return dp_set_error(ctx, DP_ERR_STATE, "exchange timed out waiting for a peer's call %llu: every rank must make the same calls",
                        *(volatile unsigned long long*)ctx->comm_err - 1);
  for (int r = 0; r < ctx->comm_world; ++r) {
    const uint64_t* h = (const uint64_t*)(ctx->h_gather[b] + (size_t)r * ctx->comm_slot_bytes);
    if (h[7] != call + 1)
      return dp_set_error(ctx, DP_ERR_STATE, "exchange: slot of rank %d holds call %lld, expected %llu", r,
                          (long long)h[7] - 1, (unsigned long long)call);
  }
  *buf = b;
  *call_out = call;
  return DP_OK;
}

extern "C" {

int dp_comm_export(dp_ctx* ctx, int world, int rank, uint32_t max_hits, char handles[128]) {
  if (!ctx || !handles) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (world < 1 || world > DP_MAX_RANKS || rank < 0 || rank >= world)
    return dp_set_error(ctx, DP_ERR_ARG, "world %d / rank %d out of range (at most %d ranks)", world, rank, DP_MAX_RANKS);
  if (sizeof(cudaIpcMemHandle_t) != 64) return dp_set_error(ctx, DP_ERR_STATE, "unexpected IPC handle size");
  // a slot never holds more than the first attempt of a call can produce (the hit workspace starts
  // at 2^20 records; a call that overflows it re-runs locally and flags its slot as truncated)
  if (max_hits > (1u << 20)) return dp_set_error(ctx, DP_ERR_ARG, "max_hits %u: a slot holds at most %u hits", max_hits, 1u << 20);
  dp_comm_release(ctx);
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint64_t slot = 64 + (uint64_t)std::max<uint32_t>(1, max_hits) * sizeof(dp_raw_hit);
  ctx->comm_slot_bytes = slot;
  ctx->comm_world = world;
  ctx->comm_rank = rank;
  ctx->comm_seq = 0;
  for (int b = 0; b < 2; ++b) {
    DP_CUDA(ctx, cudaMalloc(&ctx->d_recv[b], slot * world));
    DP_CUDA(ctx, cudaMemset(ctx->d_recv[b], 0, slot * world));
    DP_CUDA(ctx, cudaHostAlloc(&ctx->h_gather[b], slot * world, cudaHostAllocDefault));
    memset(ctx->h_gather[b], 0, slot * world);
    DP_CUDA(ctx, cudaHostAlloc(&ctx->h_sorted[b], 32 + DP_EXCHANGE_SORT_MAX * sizeof(dp_hit), cudaHostAllocDefault));
    memset(ctx->h_sorted[b], 0, 32);
    cudaIpcMemHandle_t h;
    DP_CUDA(ctx, cudaIpcGetMemHandle(&h, ctx->d_recv[b]));
    memcpy(handles + 64 * b, &h, 64);
  }
  DP_CUDA(ctx, cudaHostAlloc(&ctx->comm_err, sizeof(unsigned long long), cudaHostAllocDefault));
  *ctx->comm_err = 0ull;
  DP_CUDA(ctx, cudaDeviceSynchronize());
  return DP_OK;
}

int dp_comm_attach(dp_ctx* ctx, const char* all_handles) {
  if (!ctx || !all_handles) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->comm_world || !ctx->d_recv[0]) return dp_set_error(ctx, DP_ERR_STATE, "dp_comm_export first");
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  for (int b = 0; b < 2; ++b) {
    char* table[DP_MAX_RANKS];
    for (int r = 0; r < ctx->comm_world; ++r) {
      if (r == ctx->comm_rank) {
        ctx->peer_mapped[b][r] = ctx->d_recv[b];
      } else {
        cudaIpcMemHandle_t h;
        memcpy(&h, all_handles + (size_t)r * 128 + 64 * b, 64);
        DP_CUDA(ctx, cudaIpcOpenMemHandle(&ctx->peer_mapped[b][r], h, cudaIpcMemLazyEnablePeerAccess));
      }
      table[r] = (char*)ctx->peer_mapped[b][r];
    }
    DP_CUDA(ctx, cudaMalloc(&ctx->d_peer[b], sizeof(char*) * DP_MAX_RANKS));
    DP_CUDA(ctx, cudaMemcpy(ctx->d_peer[b], table, sizeof(char*) * ctx->comm_world, cudaMemcpyHostToDevice));
  }
  return DP_OK;
}

int dp_comm_destroy(dp_ctx* ctx) {
  if (!ctx) return DP_ERR_ARG;
  dp_comm_release(ctx);
  return DP_OK;
}

int dp_comm_gathered(dp_ctx* ctx, int age, const void** host_ptr, uint64_t* slot_bytes, int* world) {
  if (!ctx || !host_ptr || !slot_bytes || !world) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  int b = 0;
  uint64_t call = 0;
  int rc = comm_collect(ctx, age, &b, &call);
  if (rc) return rc;
  *host_ptr = ctx->h_gather[b];
  *slot_bytes = ctx->comm_slot_bytes;
  *world = ctx->comm_world;
  return DP_OK;
}

int dp_comm_set_layout(dp_ctx* ctx, const uint32_t* base, const uint64_t* local_start, const uint64_t* offset,
                       const uint32_t* record, const uint64_t* record_len) {
  if (!ctx) return DP_ERR_ARG;
  if (!ctx->comm_world) return dp_set_error(ctx, DP_ERR_STATE, "dp_comm_set_layout: dp_comm_export first");
  ctx->comm_layout.clear();
  ctx->comm_layout_base.clear();
  if (!base) return DP_OK;                       // back to one segment per rank
  if (!local_start || !offset || !record || !record_len) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  const int W = ctx->comm_world;
  for (int r = 0; r < W; ++r)
    if (base[r + 1] < base[r]) return dp_set_error(ctx, DP_ERR_ARG, "dp_comm_set_layout: base[] must not decrease");
  ctx->comm_layout_base.assign(base, base + W + 1);
  ctx->comm_layout.resize(base[W]);
  for (uint32_t e = 0; e < base[W]; ++e) {
    ctx->comm_layout[e].local_start = local_start[e];
    ctx->comm_layout[e].offset = offset[e];
    ctx->comm_layout[e].record = record[e];
    ctx->comm_layout[e].record_len = record_len[e];
  }
  return DP_OK;
}

int dp_comm_hits(dp_ctx* ctx, int age, dp_hit** hits_out, uint64_t* n_out) {
  if (!ctx || !hits_out || !n_out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *hits_out = nullptr;
  *n_out = 0;
  int s = 0;
  uint64_t call = 0;
  int rc = comm_collect(ctx, age, &s, &call);
  if (rc) return rc;
  const bool mapped = !ctx->comm_layout_base.empty();
  // the kernel that brought the slots to the host has converted and sorted them already
  // (one segment per rank, not too many hits): just take its array
  const uint64_t* sorted = ctx->h_sorted[s];
  const char* force = getenv("DEGENPRIMER_B200_EXCHANGE_SORT");      // "host": test hook for the path below
  if (!mapped && sorted[1] == 1 && sorted[2] == call + 1 && sorted[0] <= DP_EXCHANGE_SORT_MAX && !(force && force[0] == 'h')) {
    const uint64_t n = sorted[0];
    dp_hit* out = (dp_hit*)dp_result_alloc(std::max<uint64_t>(1, n) * sizeof(dp_hit));
    if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu hits", (unsigned long long)n);
    memcpy(out, sorted + 4, n * sizeof(dp_hit));
    *hits_out = out;
    *n_out = n;
    return DP_OK;
  }
  const char* recv = ctx->h_gather[s];
  const dp_dg_plan& plan = ctx->comm_plan[call & 3];
  const uint64_t slot = ctx->comm_slot_bytes;
  const uint64_t cap = (slot - 64) / sizeof(dp_raw_hit);
  uint64_t total = 0;
  for (int r = 0; r < ctx->comm_world; ++r) {
    const uint64_t* h = (const uint64_t*)(recv + (size_t)r * slot);
    const uint64_t n = (uint32_t)h[0];
    // the GPU wrote these lines over PCIe, so none is in a CPU cache: ask for all of them
    // now, the misses then overlap instead of being taken one by one in the loops below
    for (uint64_t off = 64; off < 64 + std::min(n, cap) * sizeof(dp_raw_hit); off += 64)
      __builtin_prefetch((const char*)h + off);
    if (n > cap || (h[DP_C_SATURATED] & 2ull))
      return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "rank %d had %llu hits, its slot holds fewer", r, (unsigned long long)n);
    if (!mapped && h[6] != 1) return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "rank %d scanned a template of %llu records; "
                                       "dp_comm_hits handles one segment per rank unless dp_comm_set_layout was called", r, (unsigned long long)h[6]);
    if (mapped && h[6] != ctx->comm_layout_base[r + 1] - ctx->comm_layout_base[r])
      return dp_set_error(ctx, DP_ERR_STATE, "rank %d scanned %llu records, the layout says %u", r, (unsigned long long)h[6],
                          ctx->comm_layout_base[r + 1] - ctx->comm_layout_base[r]);
    total += n;
  }
  dp_hit* out = (dp_hit*)dp_result_alloc(std::max<uint64_t>(1, total) * sizeof(dp_hit));
  if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu hits", (unsigned long long)total);
  // output order (primer, strand, record, pos): patterns are numbered in (primer, strand)
  // order, so a counting sort over patterns (x rank, when record = sender rank) leaves only
  // (record, pos) inside a bucket to sort
  const int W = mapped ? 1 : ctx->comm_world;
  static thread_local std::vector<uint64_t> bucket;
  bucket.assign((size_t)DP_MAX_PATTERNS * W + 1, 0);
  for (int r = 0; r < ctx->comm_world; ++r) {
    const char* base = recv + (size_t)r * slot;
    const uint64_t n = (uint32_t)((const uint64_t*)base)[0];
    const dp_raw_hit* raw = (const dp_raw_hit*)(base + 64);
    const uint32_t n_local = mapped ? ctx->comm_layout_base[r + 1] - ctx->comm_layout_base[r] : 1u;
    for (uint64_t i = 0; i < n; ++i) {
      if (raw[i].pattern >= DP_MAX_PATTERNS || raw[i].record >= n_local)
        return dp_result_free(out), dp_set_error(ctx, DP_ERR_STATE, "exchange: corrupt slot of rank %d", r);
      ++bucket[(size_t)raw[i].pattern * W + (mapped ? 0 : r) + 1];
    }
  }
  for (size_t i = 1; i < bucket.size(); ++i) bucket[i] += bucket[i - 1];
  for (int r = 0; r < ctx->comm_world; ++r) {
    const char* base = recv + (size_t)r * slot;
    const uint64_t* h = (const uint64_t*)base;
    const uint64_t n = (uint32_t)h[0], rec_len = h[5];
    const dp_raw_hit* raw = (const dp_raw_hit*)(base + 64);
    for (uint64_t i = 0; i < n; ++i) {
      const uint32_t pat = raw[i].pattern;
      const uint32_t L = plan.pat_len[pat];
      const uint32_t strand = plan.pat_strand[pat];
      dp_hit& o = out[bucket[(size_t)pat * W + (mapped ? 0 : r)]++];
      if (mapped) {
        // sender's local record -> the record of the whole genome it was cut from
        const dp_comm_piece& pc = ctx->comm_layout[ctx->comm_layout_base[r] + raw[i].record];
        const uint64_t fwd = pc.offset + (raw[i].g - pc.local_start);
        o.pos = strand ? pc.record_len - L - fwd : fwd;
        o.record = pc.record;
      } else {
        o.pos = strand ? rec_len - L - raw[i].g : raw[i].g;     // record 0 starts at padded position 0
        o.record = (uint32_t)r;
      }
      o.primer = plan.pat_primer[pat];
      o.strand = (uint8_t)strand;
      o.score3 = (uint8_t)raw[i].score3;
    }
  }
  // bucket[j] now is the END of bucket j
  uint64_t lo = 0;
  for (size_t j = 0; j + 1 < bucket.size(); ++j) {
    const uint64_t hi = bucket[j];
    if (hi - lo > 1)
      std::sort(out + lo, out + hi, [](const dp_hit& a, const dp_hit& b) {
        return a.record != b.record ? a.record < b.record : a.pos < b.pos;
      });
    lo = hi;
  }
  *hits_out = out;
  *n_out = total;
  return DP_OK;
}

}  // extern "C"
