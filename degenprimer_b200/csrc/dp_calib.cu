// Integer-pipe calibration micro-kernels: the denominators of the integer
// roofline (MEASURED_PEAKS.json has no integer figure, so LOP3 /
// POPC / LDS issue rates are measured on the box, at the clock actually held).
#include "dp_internal.cuh"

#define CAL_ITERS 4096
#define CAL_ILP 8

// 8 independent LOP3 chains per thread, 3-input xor-and mix that cannot fold
__global__ void __launch_bounds__(256) dp_cal_lop3(uint32_t* out, uint32_t seed) {
  uint32_t x[CAL_ILP];
  const uint32_t k1 = out[threadIdx.x] | seed;      // opaque to the compiler
#pragma unroll
  for (int i = 0; i < CAL_ILP; ++i) x[i] = seed + i * 977u + threadIdx.x;
#pragma unroll 1
  for (int it = 0; it < CAL_ITERS / 8; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
#pragma unroll
      for (int i = 0; i < CAL_ILP; ++i) x[i] = (x[i] & x[(i + 1) % CAL_ILP]) ^ k1;   // one LOP3, inputs keep changing
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < CAL_ILP; ++i) s ^= x[i];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) dp_cal_popc(uint32_t* out, uint32_t seed) {
  uint32_t x[CAL_ILP];
#pragma unroll
  for (int i = 0; i < CAL_ILP; ++i) x[i] = seed + i * 977u + threadIdx.x * 7919u;
#pragma unroll 1
  for (int it = 0; it < CAL_ITERS / 8; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
#pragma unroll
      for (int i = 0; i < CAL_ILP; ++i) x[i] = __popc(x[i]) ^ (x[i] << 3);   // POPC + one LOP3/SHF-class op
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < CAL_ILP; ++i) s ^= x[i];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// conflict-free LDS.32 stream: every lane reads its own bank
__global__ void __launch_bounds__(256) dp_cal_lds(uint32_t* out, uint32_t seed) {
  __shared__ uint32_t sm[32 * 256];
  for (int i = threadIdx.x; i < 32 * 256; i += 256) sm[i] = seed ^ i;
  __syncthreads();
  uint32_t acc[CAL_ILP];
#pragma unroll
  for (int i = 0; i < CAL_ILP; ++i) acc[i] = 0;
  const volatile uint32_t* col = sm + threadIdx.x;
#pragma unroll 1
  for (int it = 0; it < CAL_ITERS / 32; ++it) {
#pragma unroll
    for (int u = 0; u < 32; ++u) acc[u % CAL_ILP] ^= col[u * 256];
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < CAL_ILP; ++i) s ^= acc[i];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// which: 0 = LOP3, 1 = POPC, 2 = LDS.32.  Returns lane-operations per second
// over the whole chip (warp instructions x 32), best of `reps` timed launches.
extern "C" int dp_calibrate(dp_ctx* ctx, int which, int reps, double* lane_ops_per_s, double* ms_best) {
  if (!ctx || !lane_ops_per_s) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  uint32_t* d_out = nullptr;
  const int blocks = ctx->sm_count * 8, threads = 256;
  DP_CUDA(ctx, cudaMalloc(&d_out, (size_t)blocks * threads * sizeof(uint32_t)));
  double ops;
  if (which == 0) ops = (double)CAL_ITERS * CAL_ILP;
  else if (which == 1) ops = (double)CAL_ITERS * CAL_ILP;
  else if (which == 2) ops = (double)CAL_ITERS;
  else { cudaFree(d_out); return dp_set_error(ctx, DP_ERR_ARG, "which must be 0..2"); }
  ops *= (double)blocks * threads;
  float best = 1e30f;
  if (reps < 1) reps = 1;
  for (int r = 0; r < reps + 2; ++r) {
    cudaEventRecord(ctx->ev0, ctx->stream);
    if (which == 0) dp_cal_lop3<<<blocks, threads, 0, ctx->stream>>>(d_out, 12345u + r);
    else if (which == 1) dp_cal_popc<<<blocks, threads, 0, ctx->stream>>>(d_out, 12345u + r);
    else dp_cal_lds<<<blocks, threads, 0, ctx->stream>>>(d_out, 12345u + r);
    cudaEventRecord(ctx->ev1, ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { cudaFree(d_out); return dp_set_error(ctx, DP_ERR_CUDA, "calibration: %s", cudaGetErrorString(e)); }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    if (r >= 2 && ms < best) best = ms;
  }
  cudaFree(d_out);
  *lane_ops_per_s = ops / (best * 1e-3);
  if (ms_best) *ms_best = best;
  return DP_OK;
}
