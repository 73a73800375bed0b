// Context, error reporting, table upload.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "dp_internal.cuh"

static thread_local std::string g_create_error;

int dp_set_error(dp_ctx* ctx, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  else g_create_error = buf;
  return code;
}

// ---- host result blocks ---------------------------------------------------------
// Large result arrays are recycled: a fresh malloc of tens of MB is an mmap whose
// pages fault one by one while the results are copied in, which costs more than the
// kernels that produced them.  dp_free() parks big blocks here (a few, process wide)
// and the next call takes one that fits.
#include <mutex>
namespace {
struct result_block { void* p; size_t bytes; };
std::mutex g_blocks_mutex;
const int kParked = 16;                   // blocks kept for reuse: two per call in flight and then some
result_block g_blocks[kParked] = {};
const size_t kBigBlock = 1u << 20;
struct block_header { size_t bytes; size_t magic; };          // in front of every block
const size_t kMagic = 0x44504253554C5453ull;
}  // namespace

const size_t kMagicPinned = 0x44504250494E4E44ull;      // block came from cudaHostAlloc

void* dp_result_alloc(size_t bytes) {
  if (bytes == 0) bytes = 1;
  if (bytes >= kBigBlock) {
    {
      std::lock_guard<std::mutex> lock(g_blocks_mutex);
      int best = -1;
      for (int i = 0; i < kParked; ++i)
        if (g_blocks[i].p && g_blocks[i].bytes >= bytes && g_blocks[i].bytes <= 2 * bytes + kBigBlock &&
            (best < 0 || g_blocks[i].bytes < g_blocks[best].bytes))
          best = i;
      if (best >= 0) {
        void* p = g_blocks[best].p;
        g_blocks[best].p = nullptr;
        g_blocks[best].bytes = 0;
        return (char*)p + sizeof(block_header);
      }
    }
    // big blocks are page-locked: the device writes results straight into them (no staging copy);
    // they are recycled, so the cost of locking the pages is paid once
    void* raw = nullptr;
    const size_t want = bytes + bytes / 8;                  // a little slack: the next result of the loop fits too
    if (cudaHostAlloc(&raw, sizeof(block_header) + want, cudaHostAllocDefault) == cudaSuccess) {
      block_header* h = (block_header*)raw;
      h->bytes = want;
      h->magic = kMagicPinned;
      return (char*)h + sizeof(block_header);
    }
    cudaGetLastError();
  }
  block_header* h = (block_header*)malloc(sizeof(block_header) + bytes);
  if (!h) return nullptr;
  h->bytes = bytes;
  h->magic = kMagic;
  return (char*)h + sizeof(block_header);
}

bool dp_result_is_pinned(const void* p) {
  if (!p) return false;
  const block_header* h = (const block_header*)((const char*)p - sizeof(block_header));
  return h->magic == kMagicPinned;
}

static void block_release(block_header* h) {
  if (h->magic == kMagicPinned) cudaFreeHost(h);
  else free(h);
}

void dp_result_free(void* p) {
  if (!p) return;
  block_header* h = (block_header*)((char*)p - sizeof(block_header));
  if (h->magic != kMagic && h->magic != kMagicPinned) { free(p); return; }               // not ours: plain malloc
  if (h->bytes >= kBigBlock) {
    std::lock_guard<std::mutex> lock(g_blocks_mutex);
    int slot = -1;
    for (int i = 0; i < kParked; ++i) if (!g_blocks[i].p) { slot = i; break; }
    if (slot < 0) {                                            // evict the smallest parked block
      slot = 0;
      for (int i = 1; i < kParked; ++i) if (g_blocks[i].bytes < g_blocks[slot].bytes) slot = i;
      if (g_blocks[slot].bytes >= h->bytes) { block_release(h); return; }
      block_release((block_header*)g_blocks[slot].p);
    }
    g_blocks[slot].p = h;
    g_blocks[slot].bytes = h->bytes;
    return;
  }
  block_release(h);
}

cudaError_t dp_pool_alloc(dp_ctx* ctx, void** p, size_t bytes) {
  return cudaMallocFromPoolAsync(p, bytes ? bytes : 1, ctx->pool, ctx->stream);
}

void dp_pool_free(dp_ctx* ctx, void* p) {
  if (p) cudaFreeAsync(p, ctx->stream);
}

// Grow-only workspace.  Growing synchronises the stream (rare: first calls and overflows).
template <typename T>
static int ws_grow(dp_ctx* ctx, T** p, uint64_t* cap, uint64_t want, const char* what) {
  if (*cap >= want) return DP_OK;
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (*p) { cudaFree(*p); *p = nullptr; *cap = 0; }
  cudaError_t e = cudaMalloc((void**)p, want * sizeof(T));
  if (e != cudaSuccess) {
    cudaGetLastError();
    return dp_set_error(ctx, DP_ERR_NOMEM, "out of device memory for %llu %s", (unsigned long long)want, what);
  }
  *cap = want;
  return DP_OK;
}

int dp_ws_hits(dp_ctx* ctx, uint64_t cap) { return ws_grow(ctx, &ctx->d_hits, &ctx->hits_cap, cap, "hit records"); }
int dp_ws_dir(dp_ctx* ctx, uint64_t cap) { return ws_grow(ctx, &ctx->d_dir, &ctx->dir_cap, cap, "duplex directory entries"); }
int dp_ws_dimers(dp_ctx* ctx, uint64_t cap) { return ws_grow(ctx, &ctx->d_dimers, &ctx->dimers_cap, cap, "dimer records"); }
int dp_ws_codes(dp_ctx* ctx, uint64_t cap) { return ws_grow(ctx, &ctx->d_codes, &ctx->codes_cap, cap, "variant bytes"); }

int dp_ws_pinned(dp_ctx* ctx, size_t bytes) {
  if (ctx->h_res_bytes >= bytes) return DP_OK;
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->h_res) { cudaFreeHost(ctx->h_res); ctx->h_res = nullptr; ctx->h_res_bytes = 0; }
  size_t want = 1u << 20;
  while (want < bytes) want <<= 1;
  cudaError_t e = cudaMallocHost((void**)&ctx->h_res, want);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return dp_set_error(ctx, DP_ERR_NOMEM, "out of pinned host memory (%zu bytes)", want);
  }
  ctx->h_res_bytes = want;
  return DP_OK;
}

extern "C" {

int dp_abi_version(void) { return DP_ABI_VERSION; }

int dp_device_count(int* n) {
  if (!n) return DP_ERR_ARG;
  cudaError_t e = cudaGetDeviceCount(n);
  if (e != cudaSuccess) {
    *n = 0;
    g_create_error = cudaGetErrorString(e);
    return DP_ERR_CUDA;
  }
  return DP_OK;
}

const char* dp_create_error(void) { return g_create_error.c_str(); }

dp_ctx* dp_create(int device_ordinal) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    dp_set_error(nullptr, DP_ERR_CUDA, "no CUDA device: %s (this library has no CPU path)",
                 e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    return nullptr;
  }
  if (device_ordinal < 0 || device_ordinal >= n) {
    dp_set_error(nullptr, DP_ERR_ARG, "device ordinal %d out of range [0,%d)", device_ordinal, n);
    return nullptr;
  }
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device_ordinal)) != cudaSuccess) {
    dp_set_error(nullptr, DP_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    return nullptr;
  }
  if (prop.major != 10) {
    dp_set_error(nullptr, DP_ERR_CUDA,
                 "device %d is sm_%d%d; this library carries sm_100a code only",
                 device_ordinal, prop.major, prop.minor);
    return nullptr;
  }
  if ((e = cudaSetDevice(device_ordinal)) != cudaSuccess) {
    dp_set_error(nullptr, DP_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    return nullptr;
  }
  dp_ctx* ctx = new dp_ctx();
  ctx->device = device_ordinal;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->cc_major = prop.major;
  ctx->cc_minor = prop.minor;
  ctx->abort_flag = 0;
  ctx->tables_loaded = false;
  ctx->d_tables = nullptr;
  ctx->d_hp_tables = nullptr;
  ctx->hp_tables_loaded = false;
  ctx->pool = nullptr;
  ctx->d_hits = nullptr;  ctx->hits_cap = 0;
  ctx->d_dir = nullptr;   ctx->dir_cap = 0;
  ctx->d_dimers = nullptr; ctx->dimers_cap = 0;
  ctx->d_codes = nullptr; ctx->codes_cap = 0;
  ctx->d_counters = nullptr;
  ctx->counters_clean = false;
  ctx->h_res = nullptr;   ctx->h_res_bytes = 0;
  ctx->stage_bytes = 0;
  ctx->pending = nullptr;
  ctx->pending_free = nullptr;
  ctx->wire = nullptr;
  ctx->wire_bytes = 0;
  ctx->timing = true;
  ctx->comm_world = ctx->comm_rank = 0;
  ctx->comm_seq = 0;
  for (int i = 0; i < 2; ++i) {
    ctx->d_recv[i] = nullptr;
    ctx->d_peer[i] = nullptr;
    ctx->h_gather[i] = nullptr;
    ctx->h_sorted[i] = nullptr;
    for (int r = 0; r < DP_MAX_RANKS; ++r) ctx->peer_mapped[i][r] = nullptr;
  }
  ctx->comm_err = nullptr;
  ctx->comm_slot_bytes = 0;
  ctx->wire_owned = false;
  ctx->stream = ctx->copy_stream = nullptr;
  for (int i = 0; i < 2; ++i) {
    ctx->h_stage[i] = nullptr; ctx->d_stage[i] = nullptr;
    ctx->stage_ev[i] = nullptr; ctx->stage_copied[i] = nullptr;
  }
  cudaMemPoolProps pp;
  memset(&pp, 0, sizeof pp);
  pp.allocType = cudaMemAllocationTypePinned;
  pp.handleTypes = cudaMemHandleTypeNone;
  pp.location.type = cudaMemLocationTypeDevice;
  pp.location.id = device_ordinal;
  unsigned long long keep = ~0ull;    // never hand memory back between calls
  bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreate(&ctx->ev0) == cudaSuccess &&
            cudaEventCreate(&ctx->ev1) == cudaSuccess &&
            cudaEventCreate(&ctx->ev2) == cudaSuccess &&
            cudaEventCreate(&ctx->ev3) == cudaSuccess &&
            cudaEventCreate(&ctx->ev4) == cudaSuccess &&
            cudaEventCreate(&ctx->ev5) == cudaSuccess &&
            cudaMemPoolCreate(&ctx->pool, &pp) == cudaSuccess &&
            cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep) == cudaSuccess &&
            cudaMalloc(&ctx->d_counters, 16 * sizeof(unsigned long long)) == cudaSuccess &&
            cudaMalloc(&ctx->d_tables, sizeof(dp_nn_tables)) == cudaSuccess &&
            dp_ws_pinned(ctx, 1u << 20) == DP_OK;
  if (!ok) {
    dp_set_error(nullptr, DP_ERR_CUDA, "context resources: %s", cudaGetErrorString(cudaGetLastError()));
    delete ctx;
    return nullptr;
  }
  return ctx;
}

void dp_destroy(dp_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaStreamSynchronize(ctx->copy_stream);
  dp_comm_release(ctx);
  for (int i = 0; i < 2; ++i) {
    if (ctx->h_stage[i]) cudaFreeHost(ctx->h_stage[i]);
    if (ctx->d_stage[i]) cudaFree(ctx->d_stage[i]);
    if (ctx->stage_ev[i]) cudaEventDestroy(ctx->stage_ev[i]);
    if (ctx->stage_copied[i]) cudaEventDestroy(ctx->stage_copied[i]);
  }
  if (ctx->d_hits) cudaFree(ctx->d_hits);
  if (ctx->d_dir) cudaFree(ctx->d_dir);
  if (ctx->d_dimers) cudaFree(ctx->d_dimers);
  if (ctx->d_codes) cudaFree(ctx->d_codes);
  if (ctx->d_counters) cudaFree(ctx->d_counters);
  if (ctx->d_tables) cudaFree(ctx->d_tables);
  if (ctx->d_hp_tables) cudaFree(ctx->d_hp_tables);
  if (ctx->h_res) cudaFreeHost(ctx->h_res);
  if (ctx->pending && ctx->pending_free) ctx->pending_free(ctx->pending);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaEventDestroy(ctx->ev2);
  cudaEventDestroy(ctx->ev3);
  cudaEventDestroy(ctx->ev4);
  cudaEventDestroy(ctx->ev5);
  if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
  cudaStreamDestroy(ctx->copy_stream);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

void* dp_stream(dp_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int dp_synchronize(dp_ctx* ctx) {
  if (!ctx) return DP_ERR_ARG;
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return DP_OK;
}

const char* dp_last_error(const dp_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

void dp_free(void* p) { dp_result_free(p); }

int dp_abort_requested(dp_ctx* ctx, int flag) {
  if (!ctx) return 0;
  int prev = ctx->abort_flag;
  ctx->abort_flag = flag;
  return prev;
}

int dp_device_info(dp_ctx* ctx, char* name, int* sm_count, int* cc_major, int* cc_minor,
                   int* clock_khz, uint64_t* hbm_bytes) {
  if (!ctx) return DP_ERR_ARG;
  cudaDeviceProp prop;
  DP_CUDA(ctx, cudaGetDeviceProperties(&prop, ctx->device));
  if (name) { strncpy(name, prop.name, 255); name[255] = 0; }
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (clock_khz) {
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device);
    *clock_khz = khz;
  }
  if (hbm_bytes) *hbm_bytes = (uint64_t)prop.totalGlobalMem;
  return DP_OK;
}

int dp_tables_load(dp_ctx* ctx, const dp_nn_tables* tables) {
  if (!ctx || !tables) return DP_ERR_ARG;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  DP_CUDA(ctx, cudaMemcpyAsync(ctx->d_tables, tables, sizeof(dp_nn_tables), cudaMemcpyHostToDevice, ctx->stream));
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->tables_loaded = true;
  return DP_OK;
}

}  // extern "C"
