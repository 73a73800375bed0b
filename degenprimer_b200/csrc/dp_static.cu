// dp_variant_tm_static: per-variant Tm for callers that must not allocate.
//
// No counterpart in the reference.  Some callers cannot
// call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// dp_variant_tm stages the sequences and the results in pool allocations of the context; this variant
// stages them in ONE file-scope `static __device__` array of a fixed, modest size (DP_STATIC_TM_BYTES
// below: 296 KiB, held by every GPU that loads the library until the process exits) and calls no
// cudaMalloc / cudaFree / cudaMallocAsync / pool routine and no host-side allocating routine either.
// The arithmetic is dp_variant_tm_kernel's, statement for statement (TD_Functions.primer_template_Tr,
// DegenPrimer/TD_Functions.py:60-101), so the results are bit-identical.
//
// Every call on a device shares the array and finds the last call's contents in it: the status words
// are cleared by every call, the sequence and result areas are written before they are read.  Calls
// that overlap on one GPU (two contexts, two streams) RACE on it -- make them one after the other.
#include <string.h>

#include <mutex>

#include "dp_internal.cuh"

#define DP_STATIC_TM_MAX_SEQS 4096u                 // sequences per call
#define DP_STATIC_TM_MAX_SEQ_BYTES (256u << 10)     // n * len per call
#define DP_STATIC_OFF_TM 256u                       // doubles: one per sequence
#define DP_STATIC_OFF_SEQ (DP_STATIC_OFF_TM + DP_STATIC_TM_MAX_SEQS * 8u)
#define DP_STATIC_TM_BYTES (DP_STATIC_OFF_SEQ + DP_STATIC_TM_MAX_SEQ_BYTES)

// [0, 256): status (word 0 = 1 + index of the first byte that is not an unambiguous base, 0 = none);
// [256, 256 + 32 KiB): results; then the sequences as the caller gave them (ASCII).
// In HOST code this name is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress (static_scratch below), never from `&dp_static_scratch` or the bare name.
static __device__ __align__(256) unsigned char dp_static_scratch[DP_STATIC_TM_BYTES];

__device__ __forceinline__ double st_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double st_mul(double a, double b) { return __dmul_rn(a, b); }

// dp_variant_tm_kernel (dp_dg.cu) reading ASCII from the scratch instead of plane codes from a
// pool allocation; a foreign byte is reported through the status word instead of by the host.
__global__ void __launch_bounds__(128)
dp_variant_tm_static_kernel(const dp_nn_tables* __restrict__ Tp, unsigned char* __restrict__ scratch, uint32_t n,
                            uint32_t L, double conc, double cdeg, double DNA, double DMSO, double Na_eq_log) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const dp_nn_tables& T = *Tp;
  const unsigned char* s = scratch + DP_STATIC_OFF_SEQ + (uint64_t)i * L;
  double* tm = reinterpret_cast<double*>(scratch + DP_STATIC_OFF_TM);
  unsigned long long* status = reinterpret_cast<unsigned long long*>(scratch);
  for (uint32_t k = 0; k < L; ++k)
    if (dp_letter_of(s[k]) > 3u) {
      atomicMin(status + 1, (unsigned long long)i * L + k);      // word 1: smallest offending index (starts at ~0)
      status[0] = 1ull;
      tm[i] = 0.0;
      return;
    }
  const double P = conc, D = DNA;
  const double DUP = st_mul(cdeg, P < D ? P : D);
  const double K = __ddiv_rn(DUP, st_mul(st_add(P, -DUP), st_add(D, -DUP)));
  double dH = T.ini_dH, dS = T.ini_dS;
  const uint32_t first = dp_letter_of(s[0]), last = dp_letter_of(s[L - 1]);
  if (first == 0 || first == 2) { dH = st_add(dH, T.ter_dH); dS = st_add(dS, T.ter_dS); }
  if (last == 0 || last == 2) { dH = st_add(dH, T.ter_dH); dS = st_add(dS, T.ter_dS); }
  for (uint32_t k = 0; k + 1 < L; ++k) {
    const uint32_t x = dp_letter_of(s[k]), y = dp_letter_of(s[k + 1]);
    const uint32_t w = dp_letter_complement(y);           // perfect duplex
    dH = st_add(dH, T.stack_dH[x][y][w]);
    dS = st_add(dS, T.stack_dS[x][y][w]);
  }
  dS = st_add(dS, st_mul(st_mul(T.dS_Na, (double)L), Na_eq_log));
  const double denom = st_add(dS, -st_mul(T.R, log(K)));
  tm[i] = st_add(st_add(T.K0, __ddiv_rn(st_mul(dH, 1000.0), denom)), -st_mul(T.T_DMSO, DMSO));
}

// Device address of the array on `device` (current when this is called), looked up once per device.
static int static_scratch(dp_ctx* ctx, unsigned char** out) {
  static std::mutex mutex;
  static unsigned char* address[64] = {nullptr};
  if (ctx->device < 0 || ctx->device >= 64) return dp_set_error(ctx, DP_ERR_ARG, "device ordinal %d out of range", ctx->device);
  std::lock_guard<std::mutex> lock(mutex);
  if (!address[ctx->device]) {
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, dp_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return dp_set_error(ctx, DP_ERR_CUDA, "cudaGetSymbolAddress(dp_static_scratch) on device %d: %s", ctx->device,
                          cudaGetErrorString(e));
    }
    address[ctx->device] = p;
  }
  *out = address[ctx->device];
  return DP_OK;
}

extern "C" int dp_variant_tm_static_limits(uint32_t* max_seqs, uint64_t* max_seq_bytes, uint64_t* scratch_bytes) {
  if (max_seqs) *max_seqs = DP_STATIC_TM_MAX_SEQS;
  if (max_seq_bytes) *max_seq_bytes = DP_STATIC_TM_MAX_SEQ_BYTES;
  if (scratch_bytes) *scratch_bytes = DP_STATIC_TM_BYTES;
  return DP_OK;
}

extern "C" int dp_variant_tm_static(dp_ctx* ctx, uint32_t n, uint32_t len, const char* seqs, double concentration,
                                    double conversion_degree, const dp_conditions* cond, double* tm) {
  if (!ctx || !seqs || !cond || !tm) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_tables_load() has not been called");
  if (n == 0) return DP_OK;
  if (len == 0) return dp_set_error(ctx, DP_ERR_LENGTH, "empty sequence");
  if (!(conversion_degree > 0.0 && conversion_degree < 1.0))
    return dp_set_error(ctx, DP_ERR_ARG, "equilibrium ratio should be in the (0;1) interval");
  const uint64_t bytes = (uint64_t)n * len;
  if (n > DP_STATIC_TM_MAX_SEQS || bytes > DP_STATIC_TM_MAX_SEQ_BYTES)
    return dp_set_error(ctx, DP_ERR_UNSUPPORTED,
                        "dp_variant_tm_static: %u sequences of %u nt (%llu bytes) exceed the static scratch "
                        "(at most %u sequences and %u bytes per call); use dp_variant_tm",
                        n, len, (unsigned long long)bytes, DP_STATIC_TM_MAX_SEQS, DP_STATIC_TM_MAX_SEQ_BYTES);
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned char* scratch = nullptr;
  int rc = static_scratch(ctx, &scratch);
  if (rc) return rc;
  cudaStream_t st = ctx->stream;
  // the array keeps the last call's contents: the status words must start clean
  const unsigned long long clean[2] = {0ull, ~0ull};
  DP_CUDA(ctx, cudaMemcpyAsync(scratch, clean, sizeof clean, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(scratch + DP_STATIC_OFF_SEQ, seqs, bytes, cudaMemcpyHostToDevice, st));
  dp_variant_tm_static_kernel<<<(n + 127) / 128, 128, 0, st>>>(ctx->d_tables, scratch, n, len, concentration,
                                                               conversion_degree, cond->DNA, cond->DMSO, cond->Na_eq_log);
  DP_CUDA(ctx, cudaGetLastError());
  unsigned long long status[2] = {0ull, 0ull};
  DP_CUDA(ctx, cudaMemcpyAsync(status, scratch, sizeof status, cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaMemcpyAsync(tm, scratch + DP_STATIC_OFF_TM, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaStreamSynchronize(st));
  if (status[0]) {
    const unsigned long long at = status[1];
    const char c = at < bytes ? seqs[at] : '?';
    return dp_set_error(ctx, DP_ERR_ALPHABET, "byte %llu ('%c') is not an unambiguous base", at,
                        (c >= 32 && c < 127) ? c : '?');
  }
  return DP_OK;
}
