// Product pairing on the device: the fwd/rev windowed join of
// MixtureFactory.create_PCR_mixture, DegenPrimer/PCR_Mixture.py:216-236.
//
// The reference walks the position-sorted forward sites and, for each, the reverse sites from a
// moving start index:
//     rev_start = 0
//     for fwd_pi, fwd_pos in good_fwd:                      start = fwd_pos + 1
//         for rev_pi, rev_pos in good_rev[rev_start:]:      end   = rev_pos - 1
//             if start >= end: rev_start = rev_pi; continue          # rev_pi indexes the SLICE
//             if not min <= end-start+1 <= max: break                # too short breaks as well
//             add_product(start, end, ...)
// Both lists are sorted, so for forward site f with slice start r:
//   c_f = #{rev_pos <= fwd_pos_f + 2}            the sites the `continue` branch skips are [r, c_f)
//   j0  = max(r, c_f)                            first candidate (absolute index)
//   r'  = c_f - r - 1  if c_f > r  else  r       the RELATIVE index of the last skipped site
//   products of f = [j0, u_f) if min <= len(j0) <= max else none,  u_f = #{rev_pos <= fwd_pos_f + 1 + max}
// (lengths grow with j, so the first candidate decides whether the loop breaks at once).
// c_f and u_f are binary searches (one thread per forward site); only the scalar recurrence r -> r'
// is sequential: one thread walks it with the c values staged through shared memory (~10 clocks per
// site); counts, their prefix sum and the pair list are parallel again.
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "dp_internal.cuh"

#define DP_MIX_THREADS 256
#define DP_MIX_CHUNK 4096

__device__ __forceinline__ uint32_t mix_upper_bound(const int64_t* __restrict__ a, uint32_t n, int64_t x) {
  uint32_t lo = 0, hi = n;                      // first index with a[i] > x
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (a[mid] <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(DP_MIX_THREADS)
dp_mix_bounds_kernel(const int64_t* __restrict__ fwd, uint32_t n_fwd, const int64_t* __restrict__ rev, uint32_t n_rev,
                     int64_t max_amplicon, uint32_t* __restrict__ c, uint32_t* __restrict__ u) {
  const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_fwd) return;
  c[f] = mix_upper_bound(rev, n_rev, fwd[f] + 2);
  u[f] = mix_upper_bound(rev, n_rev, fwd[f] + 1 + max_amplicon);
}

// the sequential part: rs[f] = slice start the reference uses for forward site f
__global__ void __launch_bounds__(DP_MIX_THREADS)
dp_mix_recurrence_kernel(const uint32_t* __restrict__ c, uint32_t n_fwd, uint32_t* __restrict__ rs) {
  __shared__ uint32_t sc[DP_MIX_CHUNK];
  __shared__ uint32_t sr[DP_MIX_CHUNK];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0u;
  for (uint32_t base = 0; base < n_fwd; base += DP_MIX_CHUNK) {
    const uint32_t m = min((uint32_t)DP_MIX_CHUNK, n_fwd - base);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < m; i += blockDim.x) sc[i] = c[base + i];
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t r = carry;
      for (uint32_t i = 0; i < m; ++i) {
        sr[i] = r;
        const uint32_t cf = sc[i];
        r = cf > r ? cf - r - 1u : r;
      }
      carry = r;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < m; i += blockDim.x) rs[base + i] = sr[i];
  }
}

__global__ void __launch_bounds__(DP_MIX_THREADS)
dp_mix_count_kernel(const int64_t* __restrict__ fwd, uint32_t n_fwd, const int64_t* __restrict__ rev, uint32_t n_rev,
                    int64_t min_amplicon, int64_t max_amplicon, const uint32_t* __restrict__ c,
                    const uint32_t* __restrict__ u, const uint32_t* __restrict__ rs, uint32_t* __restrict__ j0_out,
                    uint32_t* __restrict__ count) {
  const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_fwd) return;
  const uint32_t j0 = max(rs[f], c[f]);
  uint32_t n = 0;
  if (j0 < n_rev) {
    const int64_t len = rev[j0] - fwd[f] - 1;           // end - start + 1
    if (len >= min_amplicon && len <= max_amplicon) n = u[f] - j0;
  }
  j0_out[f] = j0;
  count[f] = n;
}

// exclusive prefix sum of count[] by one block (n is the number of forward sites: small)
__global__ void __launch_bounds__(1024)
dp_mix_scan_kernel(const uint32_t* __restrict__ count, uint32_t n, unsigned long long* __restrict__ offset,
                   unsigned long long* __restrict__ total) {
  __shared__ unsigned long long part[1024];
  const uint32_t per = (n + 1023u) / 1024u;
  const uint32_t a = min(n, threadIdx.x * per), b = min(n, a + per);
  unsigned long long s = 0;
  for (uint32_t i = a; i < b; ++i) s += count[i];
  part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long run = 0;
    for (int i = 0; i < 1024; ++i) { const unsigned long long x = part[i]; part[i] = run; run += x; }
    *total = run;
  }
  __syncthreads();
  unsigned long long run = part[threadIdx.x];
  for (uint32_t i = a; i < b; ++i) { offset[i] = run; run += count[i]; }
}

__global__ void __launch_bounds__(DP_MIX_THREADS)
dp_mix_emit_kernel(uint32_t n_fwd, const uint32_t* __restrict__ j0, const uint32_t* __restrict__ count,
                   const unsigned long long* __restrict__ offset, dp_pair* __restrict__ pairs) {
  // one warp per forward site, lanes over its products
  const uint32_t f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31u;
  if (f >= n_fwd) return;
  const uint32_t n = count[f], first = j0[f];
  dp_pair* dst = pairs + offset[f];
  for (uint32_t k = lane; k < n; k += 32u) {
    dp_pair p;
    p.fwd = f;
    p.rev = first + k;
    dst[k] = p;
  }
}

extern "C" int dp_pair_products(dp_ctx* ctx, uint32_t n_fwd, const int64_t* fwd_pos, uint32_t n_rev,
                                const int64_t* rev_pos, int64_t min_amplicon, int64_t max_amplicon,
                                dp_pair** pairs_out, uint64_t* n_pairs_out, uint32_t* rev_start_out) {
  if (!ctx || !pairs_out || !n_pairs_out || (n_fwd && !fwd_pos) || (n_rev && !rev_pos))
    return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *pairs_out = nullptr;
  *n_pairs_out = 0;
  for (uint32_t i = 1; i < n_fwd; ++i)
    if (fwd_pos[i] < fwd_pos[i - 1]) return dp_set_error(ctx, DP_ERR_ARG, "forward positions are not sorted (index %u)", i);
  for (uint32_t i = 1; i < n_rev; ++i)
    if (rev_pos[i] < rev_pos[i - 1]) return dp_set_error(ctx, DP_ERR_ARG, "reverse positions are not sorted (index %u)", i);
  if (n_fwd == 0 || n_rev == 0) {
    *pairs_out = (dp_pair*)dp_result_alloc(sizeof(dp_pair));
    return *pairs_out ? DP_OK : dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory");
  }
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int64_t *d_fwd = nullptr, *d_rev = nullptr;
  uint32_t *d_c = nullptr, *d_u = nullptr, *d_rs = nullptr, *d_j0 = nullptr, *d_cnt = nullptr;
  unsigned long long *d_off = nullptr, *d_total = nullptr;
  dp_pair* d_pairs = nullptr;
  struct cleanup {
    dp_ctx* ctx;
    void** p[10];
    ~cleanup() { for (void** q : p) if (q && *q) dp_pool_free(ctx, *q); }
  } guard{ctx, {(void**)&d_fwd, (void**)&d_rev, (void**)&d_c, (void**)&d_u, (void**)&d_rs, (void**)&d_j0,
                (void**)&d_cnt, (void**)&d_off, (void**)&d_total, (void**)&d_pairs}};
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_fwd, n_fwd * sizeof(int64_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_rev, n_rev * sizeof(int64_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_c, n_fwd * sizeof(uint32_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_u, n_fwd * sizeof(uint32_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_rs, n_fwd * sizeof(uint32_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_j0, n_fwd * sizeof(uint32_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_cnt, n_fwd * sizeof(uint32_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_off, n_fwd * sizeof(unsigned long long)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_total, sizeof(unsigned long long)));
  DP_CUDA(ctx, cudaMemcpyAsync(d_fwd, fwd_pos, n_fwd * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_rev, rev_pos, n_rev * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  const unsigned blocks = (n_fwd + DP_MIX_THREADS - 1) / DP_MIX_THREADS;
  dp_mix_bounds_kernel<<<blocks, DP_MIX_THREADS, 0, st>>>(d_fwd, n_fwd, d_rev, n_rev, max_amplicon, d_c, d_u);
  dp_mix_recurrence_kernel<<<1, DP_MIX_THREADS, 0, st>>>(d_c, n_fwd, d_rs);
  dp_mix_count_kernel<<<blocks, DP_MIX_THREADS, 0, st>>>(d_fwd, n_fwd, d_rev, n_rev, min_amplicon, max_amplicon, d_c, d_u,
                                                         d_rs, d_j0, d_cnt);
  dp_mix_scan_kernel<<<1, 1024, 0, st>>>(d_cnt, n_fwd, d_off, d_total);
  DP_CUDA(ctx, cudaGetLastError());
  unsigned long long total = 0;
  DP_CUDA(ctx, cudaMemcpyAsync(&total, d_total, sizeof total, cudaMemcpyDeviceToHost, st));
  if (rev_start_out)
    DP_CUDA(ctx, cudaMemcpyAsync(rev_start_out, d_rs, n_fwd * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaStreamSynchronize(st));
  dp_pair* out = (dp_pair*)dp_result_alloc(std::max<unsigned long long>(1, total) * sizeof(dp_pair));
  if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu products", total);
  if (total) {
    cudaError_t e = dp_pool_alloc(ctx, (void**)&d_pairs, total * sizeof(dp_pair));
    if (e == cudaSuccess) {
      const unsigned long long threads = (unsigned long long)n_fwd * 32ull;
      dp_mix_emit_kernel<<<(unsigned)((threads + DP_MIX_THREADS - 1) / DP_MIX_THREADS), DP_MIX_THREADS, 0, st>>>(
          n_fwd, d_j0, d_cnt, d_off, d_pairs);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_pairs, total * sizeof(dp_pair), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      dp_result_free(out);
      return dp_set_error(ctx, DP_ERR_CUDA, "dp_pair_products: %s", cudaGetErrorString(e));
    }
  }
  *pairs_out = out;
  *n_pairs_out = total;
  return DP_OK;
}
