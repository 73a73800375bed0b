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
  ctx->d_hits = nullptr;
  ctx->hits_cap = 0;
  ctx->d_counters = nullptr;
  ctx->stage_bytes = 0;
  for (int i = 0; i < 2; ++i) { ctx->h_stage[i] = nullptr; ctx->d_stage[i] = nullptr; ctx->stage_ev[i] = nullptr; }
  bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreate(&ctx->ev0) == cudaSuccess &&
            cudaEventCreate(&ctx->ev1) == cudaSuccess &&
            cudaEventCreate(&ctx->ev2) == cudaSuccess &&
            cudaEventCreate(&ctx->ev3) == cudaSuccess &&
            cudaMalloc(&ctx->d_counters, 16 * sizeof(unsigned long long)) == cudaSuccess &&
            cudaMalloc(&ctx->d_tables, sizeof(dp_nn_tables)) == cudaSuccess;
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
  for (int i = 0; i < 2; ++i) {
    if (ctx->h_stage[i]) cudaFreeHost(ctx->h_stage[i]);
    if (ctx->d_stage[i]) cudaFree(ctx->d_stage[i]);
    if (ctx->stage_ev[i]) cudaEventDestroy(ctx->stage_ev[i]);
  }
  if (ctx->d_hits) cudaFree(ctx->d_hits);
  if (ctx->d_counters) cudaFree(ctx->d_counters);
  if (ctx->d_tables) cudaFree(ctx->d_tables);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaEventDestroy(ctx->ev2);
  cudaEventDestroy(ctx->ev3);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* dp_last_error(const dp_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

void dp_free(void* p) { free(p); }

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
