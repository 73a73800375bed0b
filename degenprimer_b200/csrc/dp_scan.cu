// Annealing-site scan kernels + the host pipeline of dp_scan / dp_find / dp_duplex_dg.
//
// Replaces SearchEngine._find / _find_mp / _find_in_slice / _find_in_chunk
// (DegenPrimer/SearchEngine.py:190-211,257-359).  The reference correlates
// root-of-unity maps with FFTs; the value it thresholds is exactly
//   score3(q) = 3*#full + 2*#partial + #nonACGT   (thirds)
// and a window is a hit iff score3 >= 3*max(1, L-k) - 1.  Both strands are
// scanned from the forward planes: the reverse strand uses reverse-complemented
// masks and reports pos_rev = T - L - pos_fwd.
//
// Pipeline of one call (no host round trip between the stages):
//   memset counters -> bit-sliced prefilter kernel -> run-boundary kernel
//   -> per-window kernel for patterns the prefilter cannot take
//   [-> duplex kernel over (hit, variant) pairs, sizes read on the device]
//   -> counters + a small speculative prefix of the results to pinned memory
//   -> ONE stream synchronise (a second copy only when the results are large).
#include <algorithm>
#include <memory>
#include <stdlib.h>
#include <string.h>

#include "dp_scan.cuh"
#include "dp_scan_bitslice.cuh"
#include "dp_jit.cuh"

// per-kernel timing of a call (six event records + four elapsed-time queries, ~15 us of
// host time) can be switched off per context: dp_set_timing
#define DP_EVENT(ctx, ev, st)                                         \
  do {                                                                \
    if ((ctx)->timing) DP_CUDA(ctx, cudaEventRecord((ctx)->ev, st));  \
  } while (0)
static void drop_resident(dp_ctx* ctx);      // the hits of the last dp_find leave the device (dp_find_again)

static inline void dp_elapsed(dp_ctx* ctx, float* ms, cudaEvent_t a, cudaEvent_t b) {
  *ms = 0.f;
  if (ctx->timing) cudaEventElapsedTime(ms, a, b);
}

// ---- kernel 1: every window, XOR/AND-style mask select + popcount ---------------
// One thread owns one plane word = 32 consecutive window starts; windows that
// touch non-ACGT bases are scored here too (1/3 per unknown base).
template <int NW, int NP>
__global__ void __launch_bounds__(256)
dp_scan_popc_kernel(const __grid_constant__ dp_scan_args A, const __grid_constant__ dp_pattern_set<NP> PS,
                    uint64_t pmask0, uint64_t pmask1) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t w = A.word_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < A.word_end; w += stride) {
    uint32_t H[NW + 1], Lw[NW + 1], E[NW + 1];
#pragma unroll
    for (int k = 0; k <= NW; ++k) {
      H[k] = __ldg(A.hi + w + k);
      Lw[k] = __ldg(A.lo + w + k);
      E[k] = __ldg(A.exc + w + k);
    }
    for (int b = 0; b < 32; ++b) {
      uint32_t whi[NW], wlo[NW], wex[NW];
#pragma unroll
      for (int k = 0; k < NW; ++k) {
        whi[k] = __funnelshift_r(H[k], H[k + 1], b);
        wlo[k] = __funnelshift_r(Lw[k], Lw[k + 1], b);
        wex[k] = __funnelshift_r(E[k], E[k + 1], b);
      }
      const uint64_t g = w * 32 + b;
      for (uint32_t p = 0; p < A.n_patterns; ++p) {
        uint64_t bit = p < 64 ? (pmask0 >> p) & 1ull : (pmask1 >> (p - 64)) & 1ull;
        if (!bit) continue;
        const dp_pattern& P = PS.pat[p];
        uint32_t s = dp_score3<NW>(P, whi, wlo, wex);
        if (s >= P.thr3) dp_emit_hit(A, g, p, P.len, s, P.n_var);
      }
    }
  }
}

// ---- kernel 3: windows that touch a non-ACGT base, for the prefiltered patterns ---
// The bit-sliced kernel skips every window that contains an unknown base (1/3 of
// a match is not a hard mismatch).  Those windows are enumerated from the sorted
// exception-run table instead of re-reading the template: a window belongs to the
// run that holds its FIRST unknown base, i.e. for run [start, end) with the
// previous run ending at prev_end
//   starts in [max(prev_end, start-(L-1)), start)  reach into the run from the left,
//   starts in [max(start, end-e_max), end)         begin inside the run; a window
//     with more than e_max = (3L - thr3)/2 unknown bases cannot reach thr3.
// One thread per (run, candidate start); O(runs * L) work instead of O(T).
template <int NP>
__global__ void __launch_bounds__(128)
dp_scan_runs_kernel(const __grid_constant__ dp_scan_args A, const __grid_constant__ dp_pattern_set<NP> PS,
                    uint64_t pmask0, uint64_t pmask1, uint32_t left /* Lmax-1 */, uint32_t emax) {
  const uint32_t span = left + emax;
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t r = idx / span;
  const uint32_t j = (uint32_t)(idx % span);
  if (r >= A.n_runs) return;
  const uint64_t start = A.runs[r].start, end = A.runs[r].end;
  const uint64_t prev_end = r ? A.runs[r - 1].end : 0;
  uint64_t g;
  if (j < left) {
    const uint64_t d = left - j;                 // 1 .. Lmax-1 bases in front of the run
    if (start < d) return;
    g = start - d;
    if (g < prev_end) return;                    // an earlier run owns this window
  } else {
    const uint64_t d = emax - (j - left);        // 1 .. emax bases before the run's end
    if (end - start < d) return;
    g = end - d;
  }
  for (uint32_t p = 0; p < A.n_patterns; ++p) {
    uint64_t bit = p < 64 ? (pmask0 >> p) & 1ull : (pmask1 >> (p - 64)) & 1ull;
    if (!bit) continue;
    const dp_pattern& P = PS.pat[p];
    if (g < start && g + P.len <= start) continue;          // does not reach the run
    if (g >= start && 2ull * min((unsigned long long)(end - g), (unsigned long long)P.len) > 3ull * P.len - P.thr3) continue;   // too many unknown bases
    dp_verify_window(A, P, g, p, true);
  }
}

// One block of 1024 threads (a 128-thread block that fits next to a resident scan block was tried: no gain, and
// the exchange at 8 ranks is latency bound on this kernel).
#ifndef DP_PUBLISH_THREADS
#define DP_PUBLISH_THREADS 1024
#define DP_PUBLISH_MIN_BLOCKS 1
#endif

// Block-wide copy of 64-bit words with eight independent loads in flight per thread: the copies of the publish
// kernel are latency bound (L2 / NVLink / PCIe round trips), not bandwidth bound.  `src` may be memory a peer GPU
// has just written: it is read with volatile loads.
__device__ __forceinline__ void dp_block_copy(uint64_t* __restrict__ dst, const uint64_t* src, uint64_t n) {
  const volatile uint64_t* s = (const volatile uint64_t*)src;
  const uint64_t step = blockDim.x;
  uint64_t i = threadIdx.x;
  for (; i + 7 * step < n; i += 8 * step) {
    uint64_t v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = s[i + k * step];
#pragma unroll
    for (int k = 0; k < 8; ++k) dst[i + k * step] = v[k];
  }
  for (; i < n; i += step) dst[i] = s[i];
}

// ---- kernel 4: publish -------------------------------------------------------------
// Last kernel of a call: one block writes the counters and the first few result
// records straight into pinned host memory (zero-copy over PCIe: a few KB) and
// leaves the counters zeroed for the next call.  Replaces four small D2H copies
// and a memset; the host only synchronises the stream and reads.
struct dp_publish_args {
  unsigned long long* counters;
  const uint64_t* hits;        // dp_raw_hit = 3 words
  const uint64_t* dir;
  const uint64_t* dimers;      // dp_raw_dimer = 2 words
  uint64_t hits_cap, dir_cap, dimers_cap;
  uint64_t* h_counters;
  uint64_t* h_hits;
  uint64_t* h_dir;
  uint64_t* h_dimers;
  uint32_t spec_hits, spec_pairs, spec_dimers;
  int with_dg;
  uint64_t* wire;              // optional device mirror {counters[8], raw hits...} for the multi-GPU gather
  uint64_t wire_hits;          // raw hit records the mirror can hold
  uint64_t rec0_len, n_records;   // sender's template, for receivers: header[5], header[6]
  dp_exchange_args X;          // multi-GPU exchange over peer memory (world == 0: none)
};

// Exchange, flow control (call c >= 1): every rank's slot of call c-1 must have landed in
// this rank's receive buffer (c-1)&1 -- header word 7 holds c once the slot is complete.
// Peers run at most one call apart, so the poll normally succeeds at once; it gives up
// after ~2 s and raises the error flag.  Returns false on a timeout (uniform over the block).
__device__ bool dp_exchange_wait(const dp_exchange_args& X, uint64_t call) {
  __shared__ int ok;
  const uint64_t want = call + 1;
  if (threadIdx.x == 0) ok = 1;
  __syncthreads();
  if ((int)threadIdx.x < X.world) {
    const volatile uint64_t* flag = (const volatile uint64_t*)(X.recv_prev + threadIdx.x * X.slot_bytes) + 7;
    unsigned long long t0 = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (*flag != want) {
      __nanosleep(256);
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 2000000000ull) { ok = 0; break; }
    }
    __threadfence();               // the slot's words are read after its flag
  }
  __syncthreads();
  if (!ok && threadIdx.x == 0) *X.err = want;
  return ok != 0;
}

// The slots of a call that has landed: local receive buffer -> pinned host memory, raw, and
// (one segment per rank, at most DP_EXCHANGE_SORT_MAX hits in all) also as the final result:
// dp_hit records in reference coordinates, sorted by (primer, strand, record = sender, pos).
// Every rank needs the hits of all ranks, so this conversion is N times the work of a rank's
// own hits; done here it costs a few microseconds of one block instead of host time per call.
__device__ void dp_exchange_to_host(const dp_exchange_args& X, uint64_t call) {
  __shared__ uint64_t keys[DP_EXCHANGE_SORT_MAX];
  __shared__ uint32_t first[DP_MAX_RANKS + 1];
  __shared__ int sortable;
  for (int r = 0; r < X.world; ++r) {
    const volatile uint64_t* src = (const volatile uint64_t*)(X.recv_prev + r * X.slot_bytes);
    uint64_t* dst = (uint64_t*)(X.h_prev + r * X.slot_bytes);
    const uint64_t n = min((uint64_t)(uint32_t)src[DP_C_HITS], X.slot_hits);
    dp_block_copy(dst, (const uint64_t*)src, 8 + 3 * n);
  }
  if (threadIdx.x == 0) {
    uint32_t total = 0;
    int ok = 1;
    for (int r = 0; r < X.world; ++r) {
      const volatile uint64_t* h = (const volatile uint64_t*)(X.recv_prev + r * X.slot_bytes);
      const uint64_t n = (uint32_t)h[DP_C_HITS];
      if (n > X.slot_hits || h[6] != 1 || (h[DP_C_SATURATED] & 2ull)) ok = 0;
      first[r] = total;
      total += (uint32_t)min(n, X.slot_hits);
    }
    first[X.world] = total;
    sortable = ok && total <= DP_EXCHANGE_SORT_MAX;
  }
  __syncthreads();
  if (!sortable) {
    if (threadIdx.x == 0) { X.h_sorted[1] = 0; X.h_sorted[2] = call + 1; }
    return;
  }
  const uint32_t total = first[X.world];
  // key, position and the packed rest of record i of the landed call
  auto record = [&](uint32_t i, uint64_t& key, uint64_t& pos, uint32_t& meta) {
    int r = 0;
    while (i >= first[r + 1]) ++r;
    const volatile uint64_t* h = (const volatile uint64_t*)(X.recv_prev + r * X.slot_bytes);
    const uint64_t g = h[8 + 3 * (i - first[r])], w1 = h[8 + 3 * (i - first[r]) + 1];
    const uint32_t pat = (uint32_t)(w1 >> 32) & 0xFFFFu, score3 = (uint32_t)(w1 >> 48);
    const uint32_t strand = X.prev_strand[pat & (DP_MAX_PATTERNS - 1)];
    pos = strand ? h[5] - X.prev_len[pat & (DP_MAX_PATTERNS - 1)] - g : g;   // record 0 starts at padded position 0
    key = ((uint64_t)pat << 56) | ((uint64_t)r << 48) | (pos & 0xFFFFFFFFFFFFull);
    meta = (uint32_t)r | ((uint32_t)X.prev_primer[pat & (DP_MAX_PATTERNS - 1)] << 8) | (strand << 16) | ((score3 & 0xFFu) << 17);
  };
  for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) {
    uint64_t key, pos;
    uint32_t meta;
    record(i, key, pos, meta);
    keys[i] = key;
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) {
    uint64_t key, pos;
    uint32_t meta;
    record(i, key, pos, meta);
    uint32_t rank = 0;
    for (uint32_t j = 0; j < total; ++j) rank += keys[j] < key;       // keys are distinct
    // dp_hit {pos u64; record u32; primer u16; strand u8; score3 u8}
    X.h_sorted[4 + 2 * rank] = pos;
    X.h_sorted[4 + 2 * rank + 1] = (uint64_t)(meta & 0xFFu) | ((uint64_t)((meta >> 8) & 0xFFu) << 32) |
                                   ((uint64_t)((meta >> 16) & 1u) << 48) | ((uint64_t)(meta >> 17) << 56);
  }
  if (threadIdx.x == 0) { X.h_sorted[0] = total; X.h_sorted[1] = 1; X.h_sorted[2] = call + 1; }
}

// standalone drain: brings the slots of the LAST call to the host (dp_comm_hits with age 0)
__global__ void __launch_bounds__(DP_PUBLISH_THREADS, DP_PUBLISH_MIN_BLOCKS)
dp_exchange_drain_kernel(const __grid_constant__ dp_exchange_args X) {
  if (dp_exchange_wait(X, X.seq)) dp_exchange_to_host(X, X.seq);
}

void dp_exchange_drain_launch(const dp_exchange_args& X, cudaStream_t st) {
  dp_exchange_drain_kernel<<<1, DP_PUBLISH_THREADS, 0, st>>>(X);
}

__global__ void __launch_bounds__(DP_PUBLISH_THREADS, DP_PUBLISH_MIN_BLOCKS)
dp_publish_kernel(const __grid_constant__ dp_publish_args P) {
  __shared__ unsigned long long c[8];
  if (threadIdx.x < 8) c[threadIdx.x] = P.counters[threadIdx.x];
  __syncthreads();
  // multi-GPU exchange: the slot's stores travel over NVLink while the host copies below go
  // out over PCIe; the sequence words that publish the slot are stored at the very end
  bool landed = false, sent = false;
  if (P.X.world) {
    // flow control: last call's slots of all ranks are here.  That also proves that every
    // peer has consumed the buffer this call is about to overwrite: a peer raises the
    // sequence word of its call c-1 only after it has moved call c-2 out of that buffer
    // (end of this kernel).
    landed = P.X.seq ? dp_exchange_wait(P.X, P.X.seq - 1) : false;
    sent = landed || !P.X.seq;
    if (sent) {
      // this rank's slot of call c goes into every rank's receive buffer c&1
      const uint64_t n_wire = min(min((uint64_t)(uint32_t)c[DP_C_HITS], P.hits_cap), P.X.slot_hits);
      for (int r = 0; r < P.X.world; ++r) {
        uint64_t* dst = (uint64_t*)(P.X.peers[r] + (uint64_t)P.X.rank * P.X.slot_bytes);
        dp_block_copy(dst + 8, P.hits, n_wire * 3);
        // word 4 bit 1: the slot carries fewer records than word 0 says (receivers fall back)
        if (threadIdx.x < 7)
          dst[threadIdx.x] = threadIdx.x == 5 ? P.X.rec0_len : threadIdx.x == 6 ? P.X.n_records :
                             threadIdx.x == DP_C_SATURATED ? (c[DP_C_SATURATED] & 1ull) | ((uint64_t)(uint32_t)c[DP_C_HITS] > n_wire ? 2ull : 0ull)
                                                           : c[threadIdx.x];
      }
    }
  }
  const uint64_t n_hits = min(min((uint64_t)(uint32_t)c[DP_C_HITS], P.hits_cap), (uint64_t)P.spec_hits);
  dp_block_copy(P.h_hits, P.hits, n_hits * 3);
  if (P.with_dg) {
    const uint64_t n_pairs = min(min((uint64_t)(c[DP_C_HITS] >> 32), P.dir_cap), (uint64_t)P.spec_pairs);
    const uint64_t n_dim = min(min((uint64_t)c[DP_C_DIMERS], P.dimers_cap), (uint64_t)P.spec_dimers);
    dp_block_copy(P.h_dir, P.dir, n_pairs);
    dp_block_copy(P.h_dimers, P.dimers, n_dim * 2);
  }
  if (P.wire) {
    const uint64_t n_wire = min(min((uint64_t)(uint32_t)c[DP_C_HITS], P.hits_cap), P.wire_hits);
    dp_block_copy(P.wire + 8, P.hits, n_wire * 3);
    if (threadIdx.x < 8)
      P.wire[threadIdx.x] = threadIdx.x == 5 ? P.rec0_len : threadIdx.x == 6 ? P.n_records : c[threadIdx.x];
  }
  if (landed) dp_exchange_to_host(P.X, P.X.seq - 1);      // last call's slots of all ranks -> pinned host memory
  if (threadIdx.x < 8) {
    P.h_counters[threadIdx.x] = c[threadIdx.x];
    P.counters[threadIdx.x] = 0ull;
  }
  if (sent) {
    // header word 7 (= c+1) publishes the slot to a peer.  It must come after this rank's
    // reads of receive buffer (c-1)&1 above: the peer's next call overwrites that buffer as
    // soon as it sees the word.  The fence orders those reads and the slot's stores before it.
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < P.X.world) {
      volatile uint64_t* flag = (volatile uint64_t*)(P.X.peers[threadIdx.x] + (uint64_t)P.X.rank * P.X.slot_bytes) + 7;
      *flag = P.X.seq + 1;
    }
  }
}

// ---- host: patterns ---------------------------------------------------------------
// weights in thirds of letter `l` against plane code c (A C T G): SURVEY §8a A2
static void letter_weights(uint32_t l, uint32_t full[4], uint32_t half[4]) {
  // order of plane codes: A=0 C=1 T=2 G=3
  static const unsigned char W[15][4] = {
      /*A*/ {3, 0, 0, 0}, /*C*/ {0, 3, 0, 0}, /*T*/ {0, 0, 3, 0}, /*G*/ {0, 0, 0, 3},
      /*R*/ {3, 0, 0, 3}, /*Y*/ {0, 3, 3, 0}, /*S*/ {0, 2, 0, 2}, /*W*/ {2, 0, 2, 0},
      /*K*/ {0, 0, 3, 3}, /*M*/ {3, 3, 0, 0}, /*B*/ {0, 2, 3, 2}, /*D*/ {2, 0, 2, 3},
      /*H*/ {2, 3, 2, 0}, /*V*/ {3, 2, 0, 2}, /*N*/ {2, 2, 2, 2}};
  for (int c = 0; c < 4; ++c) {
    full[c] = W[l][c] == 3;
    half[c] = W[l][c] == 2;
  }
}

static const int COMP_CODE[4] = {2, 3, 0, 1};   // A<->T, C<->G in plane codes

// Builds the (primer, strand) patterns of one call.  Returns DP_OK or an error.
static int build_patterns(dp_ctx* ctx, uint32_t n_primers, const char* const* masters,
                          const int32_t* kmax, int strand_mask, const uint32_t* var_count,
                          std::vector<dp_pattern>& pats,
                          std::vector<std::vector<uint8_t>>& letters /* per pattern, window order */) {
  pats.clear();
  letters.clear();
  for (uint32_t i = 0; i < n_primers; ++i) {
    const char* m = masters[i];
    size_t L = m ? strlen(m) : 0;
    if (L == 0) return dp_set_error(ctx, DP_ERR_LENGTH, "primer %u is empty", i);
    if (L > DP_MAX_PRIMER_LEN)
      return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "primer %u is %zu nt; at most %d supported", i, L, DP_MAX_PRIMER_LEN);
    std::vector<uint8_t> lt(L);
    for (size_t j = 0; j < L; ++j) {
      uint32_t l = dp_letter_of((unsigned char)m[j]);
      if (l == DP_NOT_A_LETTER)
        return dp_set_error(ctx, DP_ERR_ALPHABET, "primer %u position %zu: '%c' is not an IUPAC DNA letter", i, j, m[j]);
      lt[j] = (uint8_t)l;
    }
    long long k = kmax ? kmax[i] : 0;
    long long need = (long long)L - k;
    if (need < 1) need = 1;
    for (int strand = 0; strand < 2; ++strand) {
      if (!(strand_mask & (1 << strand))) continue;
      dp_pattern P;
      memset(&P, 0, sizeof P);
      P.len = (uint16_t)L;
      P.thr3 = (uint16_t)(3 * need - 1);
      P.primer = (uint16_t)i;
      P.strand = (uint8_t)strand;
      P.n_var = var_count ? var_count[i] : 1u;
      std::vector<uint8_t> wl(L);
      for (size_t j = 0; j < L; ++j) {
        // window position j on the forward planes
        uint32_t l = strand ? lt[L - 1 - j] : lt[j];
        uint32_t full[4], half[4];
        letter_weights(l, full, half);
        for (int c = 0; c < 4; ++c) {
          int src = strand ? COMP_CODE[c] : c;   // template base c pairs with letter via its complement
          if (full[src]) P.f[c][j >> 5] |= 1u << (j & 31);
          if (half[src]) { P.h[c][j >> 5] |= 1u << (j & 31); P.has_half = 1; }
        }
        P.lmask[j >> 5] |= 1u << (j & 31);
        wl[j] = (uint8_t)(strand ? dp_letter_complement(l) : l);
      }
      pats.push_back(P);
      letters.push_back(wl);
    }
  }
  if (pats.size() > DP_MAX_PATTERNS)
    return dp_set_error(ctx, DP_ERR_ARG, "too many primers in one call (%u > %d)", n_primers, DP_MAX_PRIMERS);
  return DP_OK;
}

// ---- host: one scan job -------------------------------------------------------------
// the part of a job that only depends on the primer arguments: shared between the cache and
// the calls that use it (copying a job is then a few words, not vectors and source texts)
struct dp_job_core {
  std::vector<dp_pattern> pats;
  std::vector<std::vector<uint8_t>> letters;
  dp_bitslice_plan plan;
  std::vector<dp_jit_unit> jit;     // the same plan as run-time specialised kernels (dp_jit.cu)
};

struct dp_scan_job {
  std::shared_ptr<dp_job_core> core;
  bool jit_ready;                   // every unit has a compiled kernel
  int used_algo;
  int NW;
  uint32_t max_len;
  dp_scan_args A;
  uint64_t pmask_all[2];
};

// Finds (compile = false) or makes (compile = true) the specialised kernels of a job.
static int jit_attach(dp_ctx* ctx, dp_scan_job& J, bool compile) {
  J.jit_ready = !J.core->jit.empty();
  for (dp_jit_unit& u : J.core->jit) {
    if (!u.kernel || !u.kernel->fn) {
      if (compile) {
        int rc = dp_jit_compile(ctx, u.source, u.threads, &u.kernel);
        if (rc) {
          J.jit_ready = false;
          return rc;
        }
      } else {
        u.kernel = dp_jit_lookup(u.source);
      }
    }
    if (!u.kernel || !u.kernel->fn) J.jit_ready = false;
  }
  return DP_OK;
}

// patterns, prefilter plan and kernels of a call's primer arguments
static int job_build(dp_ctx* ctx, uint32_t n_primers, const char* const* masters, const int32_t* max_mismatches,
                     int strand_mask, int algo, const uint32_t* var_count, dp_scan_job& J,
                     bool compile = false) {
  J.core = std::make_shared<dp_job_core>();
  int rc = build_patterns(ctx, n_primers, masters, max_mismatches, strand_mask, var_count, J.core->pats, J.core->letters);
  if (rc) return rc;
  J.max_len = 0;
  for (auto& p : J.core->pats) J.max_len = std::max<uint32_t>(J.max_len, p.len);
  const uint32_t n_pat = (uint32_t)J.core->pats.size();
  J.pmask_all[0] = J.pmask_all[1] = 0;
  for (uint32_t p = 0; p < n_pat; ++p) J.pmask_all[p >> 6] |= 1ull << (p & 63);
  J.used_algo = DP_ALGO_POPC;
  J.core->jit.clear();
  J.jit_ready = false;
  if (algo != DP_ALGO_POPC) {
    if (dp_bitslice_make_plan(J.core->pats, J.core->letters, J.core->plan)) J.used_algo = DP_ALGO_BITSLICE;
  }
  if (J.used_algo == DP_ALGO_BITSLICE && algo != DP_ALGO_BITSLICE) {
    std::vector<dp_jit_entry> entries;
    for (const dp_bitslice_group& g : J.core->plan.groups)
      for (const dp_bs_pattern& bp : g.pats) entries.push_back(dp_jit_entry{bp.pattern, bp.anchor, g.K});
    std::sort(entries.begin(), entries.end(), [](const dp_jit_entry& a, const dp_jit_entry& b) { return a.pattern < b.pattern; });
    {
      dp_jit_make_units(J.core->pats, J.core->letters, entries, J.core->jit);
      rc = jit_attach(ctx, J, compile || algo == DP_ALGO_JIT);
      if (rc) return rc;
    }
  }
  J.NW = J.max_len > 32 ? 2 : 1;
  memset(&J.A, 0, sizeof J.A);
  return DP_OK;
}

// The primer side of a job (patterns, prefilter plan) depends only on the call's
// primer arguments: the last one is kept per thread, because the reference calls
// find() once per template with the same primers.
static int scan_prepare(dp_ctx* ctx, const dp_template* t, uint32_t n_primers, const char* const* masters,
                        const int32_t* max_mismatches, uint64_t seg_begin, uint64_t seg_end, int strand_mask,
                        int algo, const uint32_t* var_count, dp_scan_job& J) {
  if (n_primers == 0 || n_primers > DP_MAX_PRIMERS)
    return dp_set_error(ctx, DP_ERR_ARG, "n_primers must be in [1,%d]", DP_MAX_PRIMERS);
  if (!(strand_mask & DP_STRAND_BOTH)) return dp_set_error(ctx, DP_ERR_ARG, "empty strand mask");
  if (algo != DP_ALGO_AUTO && algo != DP_ALGO_BITSLICE && algo != DP_ALGO_POPC && algo != DP_ALGO_JIT)
    return dp_set_error(ctx, DP_ERR_ARG, "unknown algo %d", algo);
  std::string key;
  key.reserve(64 * n_primers);
  for (uint32_t i = 0; i < n_primers; ++i) {
    key += masters[i] ? masters[i] : "";
    key += '/';
    key += std::to_string(max_mismatches ? max_mismatches[i] : 0);
    key += '/';
    key += std::to_string(var_count ? (long long)var_count[i] : -1ll);
    key += ';';
  }
  key += std::to_string(strand_mask) + ',' + std::to_string(algo);
  // the last few jobs of this thread (the reference alternates between the primers of a pair)
  struct cache_entry { std::string key; dp_scan_job job; uint64_t used; };
  static thread_local std::vector<cache_entry> cache;
  static thread_local uint64_t cache_clock = 0;
  cache_entry* hit = nullptr;
  for (cache_entry& e : cache)
    if (e.key == key) hit = &e;
  if (hit) {
    if (!hit->job.jit_ready && !hit->job.core->jit.empty()) jit_attach(ctx, hit->job, false);   // prepared since?
    hit->used = ++cache_clock;
    J = hit->job;
  } else {
    int rc = job_build(ctx, n_primers, masters, max_mismatches, strand_mask, algo, var_count, J);
    if (rc) return rc;
    if (cache.size() < 8) {
      cache.push_back(cache_entry());
      hit = &cache.back();
    } else {
      hit = &cache[0];
      for (cache_entry& e : cache)
        if (e.used < hit->used) hit = &e;
    }
    hit->key = key;
    hit->job = J;
    hit->used = ++cache_clock;
  }
  if (t->records.size() == 1 && t->records[0].len < J.max_len && !ctx->comm_world)
    return dp_set_error(ctx, DP_ERR_LENGTH, "template (%llu) is shorter than the primer (%u)",
                        (unsigned long long)t->records[0].len, J.max_len);
  dp_scan_args& A = J.A;
  A.lo = t->d_lo; A.hi = t->d_hi; A.exc = t->d_exc;
  A.rec_start = t->d_rec_start; A.rec_len = t->d_rec_len;
  A.n_records = (uint32_t)t->records.size();
  A.n_patterns = (uint32_t)J.core->pats.size();
  A.n_words = t->n_words;
  if (seg_end > t->n_words * 32) seg_end = t->n_words * 32;
  if (seg_begin > seg_end) seg_begin = seg_end;
  A.seg_begin = seg_begin; A.seg_end = seg_end;
  A.word_begin = seg_begin / 32;
  A.word_end = (seg_end + 31) / 32;
  A.counters = ctx->d_counters;
  A.runs = t->d_runs;
  A.n_runs = t->runs.size();
  ctx->last_rec0_len = t->records.empty() ? 0 : t->records[0].len;
  ctx->last_n_records = t->records.size();
  return DP_OK;
}

template <int NP>
static void fill_set(dp_pattern_set<NP>& PS, const std::vector<dp_pattern>& pats) {
  memset(&PS, 0, sizeof PS);
  memcpy(PS.pat, pats.data(), pats.size() * sizeof(dp_pattern));
}

template <int NP>
static int launch_popc(dp_ctx* ctx, const dp_scan_job& J, const uint64_t mask[2]) {
  static thread_local dp_pattern_set<NP> PS;     // up to 12 KB: keep it off the stack
  fill_set(PS, J.core->pats);
  const uint64_t n_words = J.A.word_end - J.A.word_begin;
  unsigned blocks = (unsigned)std::min<uint64_t>((n_words + 255) / 256, (uint64_t)ctx->sm_count * 8);
  if (J.NW == 1) dp_scan_popc_kernel<1, NP><<<blocks, 256, 0, ctx->stream>>>(J.A, PS, mask[0], mask[1]);
  else dp_scan_popc_kernel<2, NP><<<blocks, 256, 0, ctx->stream>>>(J.A, PS, mask[0], mask[1]);
  DP_CUDA(ctx, cudaGetLastError());
  return DP_OK;
}

template <int NP>
static int launch_runs(dp_ctx* ctx, const dp_scan_job& J, const uint64_t mask[2], uint32_t left, uint32_t emax) {
  static thread_local dp_pattern_set<NP> PS;
  fill_set(PS, J.core->pats);
  const uint64_t n = J.A.n_runs * (uint64_t)(left + emax);
  dp_scan_runs_kernel<NP><<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(J.A, PS, mask[0], mask[1], left, emax);
  DP_CUDA(ctx, cudaGetLastError());
  return DP_OK;
}

// Resets the counters and queues every scan kernel of the job.  Events: ev1 before the
// first kernel, ev3 after the dominant one, ev2 after the last.
static int scan_enqueue(dp_ctx* ctx, dp_scan_job& J, uint32_t* launches) {
  dp_scan_args& A = J.A;
  A.hits = ctx->d_hits;
  A.hits_cap = ctx->hits_cap;
  for (size_t p = 0; p < J.core->pats.size(); ++p) {
    ctx->last_plan.pat_len[p] = (uint8_t)J.core->pats[p].len;
    ctx->last_plan.pat_strand[p] = J.core->pats[p].strand;
    ctx->last_plan.pat_primer[p] = (uint8_t)J.core->pats[p].primer;
  }
  if (!ctx->counters_clean)
    DP_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
  ctx->counters_clean = false;
  DP_EVENT(ctx, ev1, ctx->stream);
  const bool small = J.core->pats.size() <= DP_NP_SMALL;
  const uint64_t n_scan_words = A.word_end - A.word_begin;
  int rc = DP_OK;
  if (n_scan_words) {
    if (J.used_algo == DP_ALGO_BITSLICE) {
      if (J.jit_ready) {
        for (const dp_jit_unit& u : J.core->jit) {
          rc = dp_jit_launch(ctx, A, J.core->pats, u);
          if (rc) return rc;
          ++*launches;
        }
      } else {
        rc = dp_bitslice_launch(ctx, A, J.core->pats, J.core->plan, launches);
        if (rc) return rc;
      }
      DP_EVENT(ctx, ev3, ctx->stream);
      if (A.n_runs && !J.jit_ready) {               // (the specialised kernels do this at their end themselves)
        // windows touching non-ACGT bases, for the prefiltered patterns
        uint32_t left = 0, emax = 0;
        for (size_t p = 0; p < J.core->pats.size(); ++p) {
          if (!((J.core->plan.pmask[p >> 6] >> (p & 63)) & 1ull)) continue;
          left = std::max<uint32_t>(left, J.core->pats[p].len - 1u);
          emax = std::max<uint32_t>(emax, (3u * J.core->pats[p].len - J.core->pats[p].thr3) / 2u);
        }
        emax = std::max<uint32_t>(1u, std::min<uint32_t>(emax, J.max_len));
        rc = small ? launch_runs<DP_NP_SMALL>(ctx, J, J.core->plan.pmask, left, emax)
                   : launch_runs<DP_MAX_PATTERNS>(ctx, J, J.core->plan.pmask, left, emax);
        if (rc) return rc;
        ++*launches;
      }
      uint64_t rest[2] = {J.pmask_all[0] & ~J.core->plan.pmask[0], J.pmask_all[1] & ~J.core->plan.pmask[1]};
      if (rest[0] | rest[1]) {
        rc = small ? launch_popc<DP_NP_SMALL>(ctx, J, rest) : launch_popc<DP_MAX_PATTERNS>(ctx, J, rest);
        if (rc) return rc;
        ++*launches;
      }
    } else {
      rc = small ? launch_popc<DP_NP_SMALL>(ctx, J, J.pmask_all) : launch_popc<DP_MAX_PATTERNS>(ctx, J, J.pmask_all);
      if (rc) return rc;
      ++*launches;
      DP_EVENT(ctx, ev3, ctx->stream);
    }
  } else {
    DP_EVENT(ctx, ev3, ctx->stream);
  }
  DP_EVENT(ctx, ev2, ctx->stream);
  return DP_OK;
}

// ---- host: results ---------------------------------------------------------------------
// Speculative prefix copied together with the counters, so that small results
// (the common case) cost a single synchronise.
#define DP_SPEC_HITS 1024u
#define DP_SPEC_PAIRS 4096u
#define DP_SPEC_DIMERS 2048u
#define DP_RES_OFF_HITS 128u
#define DP_RES_OFF_DIR (DP_RES_OFF_HITS + DP_SPEC_HITS * (unsigned)sizeof(dp_raw_hit))
#define DP_RES_OFF_DIMERS (DP_RES_OFF_DIR + DP_SPEC_PAIRS * (unsigned)sizeof(uint64_t))
#define DP_RES_SPEC_END (DP_RES_OFF_DIMERS + DP_SPEC_DIMERS * (unsigned)sizeof(dp_raw_dimer))

struct dp_sizes {
  uint64_t n_hits, n_pairs, n_dimers, tried, cand;
  bool saturated;
};

static int publish_enqueue(dp_ctx* ctx, bool with_dg, bool exchange = false) {
  cudaStream_t st = ctx->stream;
  dp_publish_args P;
  P.X.world = 0;
  if (exchange) dp_comm_fill(ctx, &P.X);
  P.counters = ctx->d_counters;
  P.hits = (const uint64_t*)ctx->d_hits;
  P.dir = ctx->d_dir;
  P.dimers = (const uint64_t*)ctx->d_dimers;
  P.hits_cap = ctx->hits_cap;
  P.dir_cap = ctx->dir_cap;
  P.dimers_cap = ctx->dimers_cap;
  P.h_counters = (uint64_t*)ctx->h_res;
  P.h_hits = (uint64_t*)(ctx->h_res + DP_RES_OFF_HITS);
  P.h_dir = (uint64_t*)(ctx->h_res + DP_RES_OFF_DIR);
  P.h_dimers = (uint64_t*)(ctx->h_res + DP_RES_OFF_DIMERS);
  P.spec_hits = DP_SPEC_HITS;
  P.spec_pairs = DP_SPEC_PAIRS;
  P.spec_dimers = DP_SPEC_DIMERS;
  P.with_dg = with_dg ? 1 : 0;
  P.wire = (uint64_t*)ctx->wire;
  P.wire_hits = ctx->wire_bytes >= 64 ? (ctx->wire_bytes - 64) / sizeof(dp_raw_hit) : 0;
  if (ctx->wire_bytes < 64) P.wire = nullptr;
  P.rec0_len = ctx->last_rec0_len;
  P.n_records = ctx->last_n_records;
  dp_publish_kernel<<<1, DP_PUBLISH_THREADS, 0, st>>>(P);
  DP_CUDA(ctx, cudaGetLastError());
  DP_EVENT(ctx, ev5, st);
  return DP_OK;
}

static int read_sizes(dp_ctx* ctx, dp_sizes& S) {
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->counters_clean = true;
  const unsigned long long* c = (const unsigned long long*)ctx->h_res;
  S.n_hits = (uint32_t)c[DP_C_HITS];
  S.n_pairs = c[DP_C_HITS] >> 32;
  S.cand = c[DP_C_CAND];
  S.n_dimers = c[DP_C_DIMERS];
  S.tried = c[DP_C_TRIED];
  S.saturated = c[DP_C_SATURATED] != 0;
  return DP_OK;
}

static int fetch_sizes(dp_ctx* ctx, bool with_dg, dp_sizes& S) {
  int rc = publish_enqueue(ctx, with_dg);
  if (rc) return rc;
  return read_sizes(ctx, S);
}

// Pointers to the complete raw results on the host (inside ctx->h_res, or inside
// `big` when they do not fit the pinned buffer).
struct dp_host_results {
  const dp_raw_hit* hits;
  const uint64_t* dir;
  const dp_raw_dimer* dimers;
  std::vector<char> big;
};

static bool results_are_large(bool with_dg, const dp_sizes& S) {
  return with_dg && (S.n_pairs > DP_SPEC_PAIRS || S.n_dimers > DP_SPEC_DIMERS);
}

static int fetch_all(dp_ctx* ctx, bool with_dg, const dp_sizes& S, dp_host_results& R) {
  const bool spec_ok = S.n_hits <= DP_SPEC_HITS && (!with_dg || (S.n_pairs <= DP_SPEC_PAIRS && S.n_dimers <= DP_SPEC_DIMERS));
  if (spec_ok) {
    R.hits = (const dp_raw_hit*)(ctx->h_res + DP_RES_OFF_HITS);
    R.dir = (const uint64_t*)(ctx->h_res + DP_RES_OFF_DIR);
    R.dimers = (const dp_raw_dimer*)(ctx->h_res + DP_RES_OFF_DIMERS);
    // written by the GPU over PCIe: no line is in a CPU cache, let the misses overlap
    for (size_t off = 0; off < S.n_hits * sizeof(dp_raw_hit); off += 64) __builtin_prefetch((const char*)R.hits + off);
    if (with_dg) {
      for (size_t off = 0; off < S.n_pairs * sizeof(uint64_t); off += 64) __builtin_prefetch((const char*)R.dir + off);
      for (size_t off = 0; off < S.n_dimers * sizeof(dp_raw_dimer); off += 64) __builtin_prefetch((const char*)R.dimers + off);
    }
    return DP_OK;
  }
  const size_t b_hits = S.n_hits * sizeof(dp_raw_hit);
  const size_t b_dir = with_dg ? S.n_pairs * sizeof(uint64_t) : 0;
  const size_t b_dim = with_dg ? S.n_dimers * sizeof(dp_raw_dimer) : 0;
  const size_t total = b_hits + b_dir + b_dim;
  char* base;
  if (total <= (256u << 20)) {
    int rc = dp_ws_pinned(ctx, total);
    if (rc) return rc;
    base = ctx->h_res;
  } else {
    R.big.resize(total);
    base = R.big.data();
  }
  cudaStream_t st = ctx->stream;
  if (b_hits) DP_CUDA(ctx, cudaMemcpyAsync(base, ctx->d_hits, b_hits, cudaMemcpyDeviceToHost, st));
  if (b_dir) DP_CUDA(ctx, cudaMemcpyAsync(base + b_hits, ctx->d_dir, b_dir, cudaMemcpyDeviceToHost, st));
  if (b_dim) DP_CUDA(ctx, cudaMemcpyAsync(base + b_hits + b_dir, ctx->d_dimers, b_dim, cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaStreamSynchronize(st));
  R.hits = (const dp_raw_hit*)base;
  R.dir = (const uint64_t*)(base + b_hits);
  R.dimers = (const dp_raw_dimer*)(base + b_hits + b_dir);
  return DP_OK;
}

// raw hits -> dp_hit in reference coordinates, sorted by (primer, strand, record, pos).
// order[i] = raw slot of output hit i.
static int hits_to_abi(dp_ctx* ctx, const dp_template* t, const dp_scan_job& J, const dp_raw_hit* raw,
                       uint64_t n, dp_hit** out_p, std::vector<uint32_t>* order_p) {
  dp_hit* out = (dp_hit*)dp_result_alloc(std::max<uint64_t>(1, n) * sizeof(dp_hit));
  if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu hits", (unsigned long long)n);
  std::vector<dp_hit> tmp(n);
  for (uint64_t i = 0; i < n; ++i) {
    const dp_raw_hit& h = raw[i];
    const dp_pattern& P = J.core->pats[h.pattern];
    const dp_record& rec = t->records[h.record];
    uint64_t off = h.g - rec.start;
    tmp[i].pos = P.strand ? rec.len - P.len - off : off;
    tmp[i].record = h.record;
    tmp[i].primer = P.primer;
    tmp[i].strand = P.strand;
    tmp[i].score3 = (uint8_t)h.score3;
  }
  std::vector<uint32_t> order(n);
  bool packable = true;                // (primer, strand, record, pos) fits one 64-bit key
  for (uint64_t i = 0; i < n && packable; ++i)
    packable = tmp[i].primer < 64 && tmp[i].record < (1u << 17) && tmp[i].pos < (1ull << 39);
  if (packable) {
    struct keyed { uint64_t key; uint32_t idx; };
    std::vector<keyed> ks(n);
    for (uint64_t i = 0; i < n; ++i) {
      ks[i].key = ((uint64_t)tmp[i].primer << 57) | ((uint64_t)tmp[i].strand << 56) | ((uint64_t)tmp[i].record << 39) | tmp[i].pos;
      ks[i].idx = (uint32_t)i;
    }
    std::sort(ks.begin(), ks.end(), [](const keyed& a, const keyed& b) { return a.key < b.key; });
    for (uint64_t i = 0; i < n; ++i) order[i] = ks[i].idx;
  } else {
    for (uint64_t i = 0; i < n; ++i) order[i] = (uint32_t)i;
    std::sort(order.begin(), order.end(), [&tmp](uint32_t x, uint32_t y) {
      const dp_hit& a = tmp[x];
      const dp_hit& b = tmp[y];
      if (a.primer != b.primer) return a.primer < b.primer;
      if (a.strand != b.strand) return a.strand < b.strand;
      if (a.record != b.record) return a.record < b.record;
      return a.pos < b.pos;
    });
  }
  for (uint64_t i = 0; i < n; ++i) out[i] = tmp[order[i]];
  *out_p = out;
  if (order_p) order_p->swap(order);
  return DP_OK;
}

// dimers grouped by (hit in `order`, variant); inside a group the device already
// sorted by dG (stable).  var_of_hit(i) = number of variants of output hit i.
template <typename VarCount>
static int dimers_to_abi(dp_ctx* ctx, const dp_host_results& R, const dp_sizes& S, const std::vector<uint32_t>& order,
                         VarCount var_of_hit, dp_dimer** out_p, uint64_t* n_out_p) {
  dp_dimer* out = (dp_dimer*)dp_result_alloc(std::max<uint64_t>(1, S.n_dimers) * sizeof(dp_dimer));
  if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu dimers", (unsigned long long)S.n_dimers);
  uint64_t w = 0;
  for (size_t i = 0; i < order.size(); ++i) {
    const dp_raw_hit& h = R.hits[order[i]];
    const uint32_t V = var_of_hit(i);
    for (uint32_t v = 0; v < V; ++v) {
      const uint64_t e = R.dir[(uint64_t)h.pair_base + v];
      const uint64_t off = e >> 16;
      const uint32_t cnt = (uint32_t)(e & 0xFFFFu);
      if (off + cnt > S.n_dimers || w + cnt > S.n_dimers) {
        dp_result_free(out);
        return dp_set_error(ctx, DP_ERR_STATE, "internal: duplex directory is inconsistent");
      }
      for (uint32_t k = 0; k < cnt; ++k) {
        out[w].hit = (uint32_t)i;
        out[w].variant = v;
        out[w].match_mask = R.dimers[off + k].match_mask;
        out[w].dG = R.dimers[off + k].dG;
        ++w;
      }
    }
  }
  *out_p = out;
  *n_out_p = w;
  return DP_OK;
}

// Large results: the duplex kernel left the number of dimers of every hit in the raw hit
// records; the host turns them into output offsets in sorted-hit order, the device writes final
// dp_dimer records, ONE copy brings them back -- straight into the (page-locked, recycled) block
// that is handed to the caller.  order[i] = raw slot of output hit i; raw = the raw hits on the
// host (null: fetched here).
static int dimers_large(dp_ctx* ctx, const dp_dg_plan& plan, const std::vector<uint32_t>& order, const dp_sizes& S,
                        const dp_raw_hit* raw, dp_dimer** out_p, uint64_t* n_out_p) {
  const uint64_t n = order.size();
  cudaStream_t st = ctx->stream;
  std::vector<dp_raw_hit> fetched;
  if (!raw) {
    fetched.resize(n);
    DP_CUDA(ctx, cudaMemcpyAsync(fetched.data(), ctx->d_hits, n * sizeof(dp_raw_hit), cudaMemcpyDeviceToHost, st));
    DP_CUDA(ctx, cudaStreamSynchronize(st));
    raw = fetched.data();
  }
  // `raw` usually points into the pinned staging area that is reused (and may be re-allocated)
  // below: read it first
  std::vector<dp_slot_info_host> info(n);
  uint64_t run = 0;
  for (uint64_t i = 0; i < n; ++i) {
    dp_slot_info_host& x = info[order[i]];
    x.count = raw[order[i]].reserved;
    x.final_off = run;
    x.rank = (uint32_t)i;
    run += x.count;
  }
  if (run != S.n_dimers)
    return dp_set_error(ctx, DP_ERR_STATE, "internal: per-hit dimer counts (%llu) disagree with the total (%llu)",
                        (unsigned long long)run, (unsigned long long)S.n_dimers);
  int rc = dp_ws_pinned(ctx, n * sizeof(dp_slot_info_host));
  if (rc) return rc;
  dp_slot_info_host* h = (dp_slot_info_host*)ctx->h_res;
  memcpy(h, info.data(), n * sizeof(dp_slot_info_host));
  void* d_info = nullptr;
  dp_dimer* d_final = nullptr;
  DP_CUDA(ctx, dp_pool_alloc(ctx, &d_info, std::max<uint64_t>(1, n) * sizeof(dp_slot_info_host)));
  cudaError_t e = cudaMemcpyAsync(d_info, h, n * sizeof(dp_slot_info_host), cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = dp_pool_alloc(ctx, (void**)&d_final, std::max<uint64_t>(1, run) * sizeof(dp_dimer));
  dp_dimer* out = nullptr;
  if (e == cudaSuccess) {
    rc = dp_dg_gather(ctx, plan, n, d_info, d_final);
    if (rc == DP_OK) {
      out = (dp_dimer*)dp_result_alloc(std::max<uint64_t>(1, run) * sizeof(dp_dimer));
      if (!out) rc = dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu dimers", (unsigned long long)run);
    }
    const size_t bytes = run * sizeof(dp_dimer);
    if (rc == DP_OK && run) {
      // the caller's block is page-locked when it is big (dp_result_alloc): DMA straight into it
      e = cudaMemcpyAsync(out, d_final, bytes, cudaMemcpyDeviceToHost, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    }
  }
  dp_pool_free(ctx, d_info);
  dp_pool_free(ctx, d_final);
  if (rc) { dp_result_free(out); return rc; }
  if (e != cudaSuccess) { dp_result_free(out); return dp_set_error(ctx, DP_ERR_CUDA, "large-result gather: %s", cudaGetErrorString(e)); }
  *out_p = out;
  *n_out_p = run;
  return DP_OK;
}

static uint64_t count_windows(const dp_template* t, const dp_scan_job& J) {
  uint64_t windows = 0;
  for (auto& p : J.core->pats)
    for (auto& rec : t->records)
      if (rec.len >= p.len) windows += rec.len - p.len + 1;
  return windows;
}

// variants -> plane codes on the device (skipped when the same blob is already there)
static int upload_codes(dp_ctx* ctx, const char* variants, uint64_t n_bytes) {
  std::vector<uint8_t>& last = ctx->codes_shadow;     // what ctx->d_codes holds
  static thread_local std::vector<uint8_t> codes;
  codes.resize(n_bytes);
  for (uint64_t i = 0; i < n_bytes; ++i) {
    uint32_t l = dp_letter_of((unsigned char)variants[i]);
    if (l > 3)
      return dp_set_error(ctx, DP_ERR_ALPHABET, "byte %llu ('%c') of the variants is not an unambiguous base",
                          (unsigned long long)i, variants[i]);
    codes[i] = (uint8_t)l;
  }
  if (last == codes && ctx->codes_cap >= n_bytes && !last.empty()) return DP_OK;
  drop_resident(ctx);
  int rc = dp_ws_codes(ctx, std::max<uint64_t>(n_bytes, 4096));
  if (rc) return rc;
  rc = dp_ws_pinned(ctx, DP_RES_SPEC_END + n_bytes);
  if (rc) return rc;
  // staged through pinned memory behind the result area; the stream is idle between calls
  memcpy(ctx->h_res + DP_RES_SPEC_END, codes.data(), n_bytes);
  DP_CUDA(ctx, cudaMemcpyAsync(ctx->d_codes, ctx->h_res + DP_RES_SPEC_END, n_bytes, cudaMemcpyHostToDevice, ctx->stream));
  last = codes;
  return DP_OK;
}

static int fill_dg_launch(dp_ctx* ctx, const dp_conditions* cond, dp_dg_launch& L) {
  memset(&L.plan, 0, sizeof L.plan);
  L.PCR_T = cond->PCR_T;
  L.Na_eq_log = cond->Na_eq_log;
  L.max_dimer_dG = cond->max_dimer_dG;
  L.max_dG_for_K = cond->max_dG_for_K;
  return DP_OK;
}

static void fill_scan_stats(dp_ctx* ctx, const dp_template* t, const dp_scan_job& J, const dp_sizes& S,
                            float ms_scan, float ms_main, uint32_t launches, dp_scan_stats* stats) {
  if (!stats) return;
  stats->windows = count_windows(t, J);
  stats->candidates = S.cand;
  stats->hits = S.n_hits;
  stats->ms_scan = ms_scan;
  stats->ms_main = ms_main;
  float ms_total = 0.f;
  dp_elapsed(ctx, &ms_total, ctx->ev0, ctx->ev5);
  stats->ms_total = ms_total;
  stats->launches = launches;
  stats->algo = (uint32_t)(J.used_algo == DP_ALGO_BITSLICE && J.jit_ready ? DP_ALGO_JIT : J.used_algo);
}

// scan [+ duplex] with overflow handling, split so that a caller can queue its own
// work (the multi-GPU hit gather) between the enqueue and the wait.
struct dp_pipeline_state {
  bool need_scan;
  int attempt;
  float ms_scan, ms_main, ms_dg;
  uint32_t scan_launches, dg_launches;
  dp_sizes S;                       // sizes of the previous attempt (duplex-only re-runs)
  bool overflowed;                  // the first attempt did not fit: results were re-made after the wait
  bool no_exchange;                 // dp_find_again: a duplex-only call is not part of the multi-GPU exchange
};

static int pipeline_reserve(dp_ctx* ctx, bool with_dg) {
  int rc = dp_ws_hits(ctx, std::max<uint64_t>(ctx->hits_cap, 1u << 20));
  if (rc) return rc;
  if (with_dg) {
    rc = dp_ws_dir(ctx, std::max<uint64_t>(ctx->dir_cap, 1u << 22));
    if (rc) return rc;
    rc = dp_ws_dimers(ctx, std::max<uint64_t>(ctx->dimers_cap, 1u << 22));
    if (rc) return rc;
  }
  return DP_OK;
}

// queues one attempt: scan kernels (or the restored counters), duplex kernel, publish
static int pipeline_enqueue(dp_ctx* ctx, const dp_template* t, dp_scan_job& J, const dp_dg_launch* dg,
                            dp_pipeline_state& P) {
  int rc;
  if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  if (P.need_scan) {
    rc = scan_enqueue(ctx, J, &P.scan_launches);
    if (rc) return rc;
  } else {
    // duplex stage again on the hits that are still on the device: the publish kernel
    // zeroed the counters, put the hit / pair counts back
    unsigned long long* hc = (unsigned long long*)ctx->h_res + 8;
    for (int i = 0; i < 8; ++i) hc[i] = 0;
    hc[DP_C_HITS] = P.S.n_hits | (P.S.n_pairs << 32);
    hc[DP_C_CAND] = P.S.cand;
    DP_CUDA(ctx, cudaMemcpyAsync(ctx->d_counters, hc, 8 * sizeof(unsigned long long), cudaMemcpyHostToDevice, ctx->stream));
    ctx->counters_clean = false;
    DP_EVENT(ctx, ev2, ctx->stream);
  }
  if (dg) {
    rc = dp_dg_enqueue(ctx, t, *dg, 0, &P.dg_launches, /*hits_fresh=*/P.need_scan);
    if (rc) return rc;
  }
  DP_EVENT(ctx, ev4, ctx->stream);
  // multi-GPU: the first attempt of a call carries the exchange (a re-run after an overflow
  // must not: every rank exchanges exactly once per call)
  return publish_enqueue(ctx, dg != nullptr, !P.overflowed && !P.no_exchange);
}

// waits for the queued attempt; grows buffers and repeats while something overflowed
static int pipeline_wait(dp_ctx* ctx, const dp_template* t, dp_scan_job& J, const dp_dg_launch* dg,
                         dp_pipeline_state& P, dp_sizes& S, dp_host_results& R) {
  int rc;
  for (;;) {
    rc = read_sizes(ctx, S);
    if (rc) return rc;
    float ms = 0.f;
    if (P.need_scan) {
      dp_elapsed(ctx, &ms, ctx->ev1, ctx->ev2); P.ms_scan += ms;
      dp_elapsed(ctx, &P.ms_main, ctx->ev1, ctx->ev3);
    }
    if (dg) { dp_elapsed(ctx, &ms, ctx->ev2, ctx->ev4); P.ms_dg += ms; }
    if (S.saturated)
      return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "more than 4e9 hits or duplexes in one call: lower max_mismatches or split the template");
    bool again = false;
    if (P.need_scan && S.n_hits > ctx->hits_cap) {
      rc = dp_ws_hits(ctx, S.n_hits + (S.n_hits >> 3) + 1024);
      if (rc) return rc;
      again = true;
    } else {
      P.need_scan = false;
      if (dg && (S.n_pairs > ctx->dir_cap || S.n_dimers > ctx->dimers_cap)) {
        // (while the directory overflowed the dimer count is only a lower bound; the loop repeats)
        if (S.n_pairs > ctx->dir_cap) { rc = dp_ws_dir(ctx, S.n_pairs + (S.n_pairs >> 3) + 1024); if (rc) return rc; }
        if (S.n_dimers > ctx->dimers_cap) { rc = dp_ws_dimers(ctx, S.n_dimers + (S.n_dimers >> 2) + 1024); if (rc) return rc; }
        again = true;
      }
    }
    if (!again) break;
    if (++P.attempt == 5) return dp_set_error(ctx, DP_ERR_STATE, "internal: result buffers keep overflowing");
    P.overflowed = true;
    P.S = S;
    rc = pipeline_enqueue(ctx, t, J, dg, P);
    if (rc) return rc;
  }
  // large duplex results are laid out by the device (dimers_large): only the hits come back here
  return fetch_all(ctx, dg != nullptr && !results_are_large(true, S), S, R);
}

static void pipeline_init(dp_pipeline_state& P) {
  memset(&P, 0, sizeof P);
  P.need_scan = true;
}

extern "C" int dp_scan(dp_ctx* ctx, const dp_template* t, uint32_t n_primers,
                       const char* const* masters, const int32_t* max_mismatches,
                       uint64_t seg_begin, uint64_t seg_end, int strand_mask, int algo,
                       dp_hit** hits_out, uint64_t* n_hits_out, dp_scan_stats* stats) {
  if (!ctx || !t || !masters || !hits_out || !n_hits_out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *hits_out = nullptr;
  *n_hits_out = 0;
  if (stats) memset(stats, 0, sizeof *stats);
  if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  drop_resident(ctx);
  dp_scan_job J;
  int rc = scan_prepare(ctx, t, n_primers, masters, max_mismatches, seg_begin, seg_end, strand_mask, algo, nullptr, J);
  if (rc) return rc;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  DP_EVENT(ctx, ev0, ctx->stream);
  dp_sizes S;
  dp_host_results R;
  dp_pipeline_state P;
  pipeline_init(P);
  rc = pipeline_reserve(ctx, false);
  if (rc) return rc;
  rc = pipeline_enqueue(ctx, t, J, nullptr, P);
  if (rc) return rc;
  rc = pipeline_wait(ctx, t, J, nullptr, P, S, R);
  if (rc) return rc;
  rc = hits_to_abi(ctx, t, J, R.hits, S.n_hits, hits_out, nullptr);
  if (rc) return rc;
  *n_hits_out = S.n_hits;
  fill_scan_stats(ctx, t, J, S, P.ms_scan, P.ms_main, P.scan_launches, stats);
  return DP_OK;
}

// A fused call between its two halves (one per context; not re-entrant).
struct dp_find_pending {
  bool active = false;
  const dp_template* t = nullptr;
  uint64_t t_uid = 0;                 // dp_find_again: the template must still be alive
  dp_scan_job J;
  dp_dg_launch DL;
  dp_pipeline_state P;
  std::vector<uint32_t> var_count;
  // after a successful dp_find_end the hits are still on the device (dp_find_again)
  bool resident = false;
  dp_sizes S;                         // sizes of the finished call
  std::vector<uint32_t> order;        // raw slot of output hit i
  std::vector<uint16_t> primer_of;    // primer of output hit i
};

// any other call on the context overwrites the hit workspace
static void drop_resident(dp_ctx* ctx) {
  if (ctx->pending) ((dp_find_pending*)ctx->pending)->resident = false;
}

static void pending_free(void* p) { delete (dp_find_pending*)p; }

extern "C" int dp_find_begin(dp_ctx* ctx, const dp_template* t, uint32_t n_primers,
                             const char* const* masters, const int32_t* max_mismatches,
                             uint64_t seg_begin, uint64_t seg_end, int strand_mask, int algo,
                             const uint32_t* var_count, const uint64_t* var_offset, const char* variants,
                             const dp_conditions* cond) {
  if (!ctx || !t || !masters || !var_count || !var_offset || !variants || !cond)
    return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_tables_load() has not been called");
  if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  if (!ctx->pending) { ctx->pending = new dp_find_pending(); ctx->pending_free = pending_free; }
  dp_find_pending& F = *(dp_find_pending*)ctx->pending;
  if (F.active) return dp_set_error(ctx, DP_ERR_STATE, "dp_find_begin: the previous call was not finished with dp_find_end");
  F.resident = false;
  int rc = scan_prepare(ctx, t, n_primers, masters, max_mismatches, seg_begin, seg_end, strand_mask, algo, var_count, F.J);
  if (rc) return rc;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  DP_EVENT(ctx, ev0, ctx->stream);
  // variants of every primer, as plane codes
  uint64_t var_bytes = 0;
  for (uint32_t p = 0; p < n_primers; ++p) {
    uint64_t L = strlen(masters[p]);
    var_bytes = std::max<uint64_t>(var_bytes, var_offset[p] + (uint64_t)var_count[p] * L);
    if (var_offset[p] + (uint64_t)var_count[p] * L > 0xFFFFFFFFull)
      return dp_set_error(ctx, DP_ERR_ARG, "variant blob larger than 4 GB");
  }
  rc = upload_codes(ctx, variants, var_bytes);
  if (rc) return rc;
  fill_dg_launch(ctx, cond, F.DL);
  for (size_t p = 0; p < F.J.core->pats.size(); ++p) {
    F.DL.plan.pat_len[p] = (uint8_t)F.J.core->pats[p].len;
    F.DL.plan.pat_strand[p] = F.J.core->pats[p].strand;
    F.DL.plan.pat_primer[p] = (uint8_t)F.J.core->pats[p].primer;
  }
  for (uint32_t p = 0; p < n_primers; ++p) {
    F.DL.plan.var_count[p] = var_count[p];
    F.DL.plan.var_off[p] = (uint32_t)var_offset[p];
  }
  F.var_count.assign(var_count, var_count + n_primers);
  F.t = t;
  F.t_uid = t->uid;
  pipeline_init(F.P);
  rc = pipeline_reserve(ctx, true);
  if (rc) return rc;
  rc = pipeline_enqueue(ctx, t, F.J, &F.DL, F.P);
  if (rc) return rc;
  F.active = true;
  return DP_OK;
}

extern "C" int dp_find_end(dp_ctx* ctx, dp_hit** hits_out, uint64_t* n_hits_out,
                           dp_dimer** dimers_out, uint64_t* n_dimers_out,
                           dp_scan_stats* stats, dp_dg_stats* dg_stats) {
  if (!ctx || !hits_out || !n_hits_out || !dimers_out || !n_dimers_out)
    return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *hits_out = nullptr; *n_hits_out = 0;
  *dimers_out = nullptr; *n_dimers_out = 0;
  if (stats) memset(stats, 0, sizeof *stats);
  if (dg_stats) memset(dg_stats, 0, sizeof *dg_stats);
  dp_find_pending* Fp = (dp_find_pending*)ctx->pending;
  if (!Fp || !Fp->active) return dp_set_error(ctx, DP_ERR_STATE, "dp_find_end without dp_find_begin");
  dp_find_pending& F = *Fp;
  F.active = false;
  dp_sizes S;
  dp_host_results R;
  int rc = pipeline_wait(ctx, F.t, F.J, &F.DL, F.P, S, R);
  if (rc) return rc;
  std::vector<uint32_t> order;
  rc = hits_to_abi(ctx, F.t, F.J, R.hits, S.n_hits, hits_out, &order);
  if (rc) return rc;
  *n_hits_out = S.n_hits;
  const dp_hit* hits = *hits_out;
  const std::vector<uint32_t>& vc = F.var_count;
  if (results_are_large(true, S)) {
    rc = dimers_large(ctx, F.DL.plan, order, S, R.hits, dimers_out, n_dimers_out);
    ++F.P.dg_launches;                                     // the gather kernel
  }
  else
    rc = dimers_to_abi(ctx, R, S, order, [&](size_t i) { return vc[hits[i].primer]; }, dimers_out, n_dimers_out);
  if (rc) { dp_result_free(*hits_out); *hits_out = nullptr; *n_hits_out = 0; return rc; }
  F.S = S;
  F.order.swap(order);
  F.primer_of.resize(S.n_hits);
  for (uint64_t i = 0; i < S.n_hits; ++i) F.primer_of[i] = hits[i].primer;
  F.resident = true;
  fill_scan_stats(ctx, F.t, F.J, S, F.P.ms_scan, F.P.ms_main, F.P.scan_launches, stats);
  if (stats) stats->reserved = F.P.overflowed ? 1.f : 0.f;     // the wire mirror of the first attempt is stale
  if (dg_stats) {
    dg_stats->duplexes = S.n_pairs;
    dg_stats->dimers_tried = S.tried;
    dg_stats->dimers_kept = *n_dimers_out;
    dg_stats->ms_kernel = F.P.ms_dg;
    dp_elapsed(ctx, &dg_stats->ms_total, ctx->ev2, ctx->ev5);
    dg_stats->launches = F.P.dg_launches;
  }
  return DP_OK;
}

// The duplex stage of the last dp_find of this context again, under other conditions: the
// hits are still on the device, nothing is scanned or uploaded (the optimiser's inner loop,
// DegenPrimer/PCR_Optimizer.py:162-214 re-evaluates every duplex for each trial point).
extern "C" int dp_find_again(dp_ctx* ctx, const dp_conditions* cond, dp_dimer** dimers_out, uint64_t* n_dimers_out,
                             dp_dg_stats* dg_stats) {
  if (!ctx || !cond || !dimers_out || !n_dimers_out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *dimers_out = nullptr; *n_dimers_out = 0;
  if (dg_stats) memset(dg_stats, 0, sizeof *dg_stats);
  dp_find_pending* Fp = (dp_find_pending*)ctx->pending;
  if (!Fp || Fp->active || !Fp->resident)
    return dp_set_error(ctx, DP_ERR_STATE, "dp_find_again: the hits of the last dp_find are not on the device any more "
                                           "(another call used the context, or there was none)");
  if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  dp_find_pending& F = *Fp;
  if (!dp_template_alive(F.t_uid)) {
    F.resident = false;
    return dp_set_error(ctx, DP_ERR_STATE, "dp_find_again: the template of the last dp_find has been freed");
  }
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  F.DL.PCR_T = cond->PCR_T;
  F.DL.Na_eq_log = cond->Na_eq_log;
  F.DL.max_dimer_dG = cond->max_dimer_dG;
  F.DL.max_dG_for_K = cond->max_dG_for_K;
  pipeline_init(F.P);
  F.P.need_scan = false;
  F.P.no_exchange = true;
  F.P.S = F.S;
  F.resident = false;                   // until this call has succeeded
  DP_EVENT(ctx, ev0, ctx->stream);
  int rc = pipeline_enqueue(ctx, F.t, F.J, &F.DL, F.P);
  if (rc) return rc;
  dp_sizes S;
  dp_host_results R;
  rc = pipeline_wait(ctx, F.t, F.J, &F.DL, F.P, S, R);
  if (rc) return rc;
  if (S.n_hits != F.S.n_hits || S.n_pairs != F.S.n_pairs)
    return dp_set_error(ctx, DP_ERR_STATE, "internal: resident hits changed between dp_find and dp_find_again");
  const std::vector<uint16_t>& primer_of = F.primer_of;
  const std::vector<uint32_t>& vc = F.var_count;
  if (results_are_large(true, S))
    rc = dimers_large(ctx, F.DL.plan, F.order, S, R.hits, dimers_out, n_dimers_out);
  else
    rc = dimers_to_abi(ctx, R, S, F.order, [&](size_t i) { return vc[primer_of[i]]; }, dimers_out, n_dimers_out);
  if (rc) return rc;
  F.S = S;
  F.resident = true;
  if (dg_stats) {
    dg_stats->duplexes = S.n_pairs;
    dg_stats->dimers_tried = S.tried;
    dg_stats->dimers_kept = *n_dimers_out;
    dg_stats->ms_kernel = F.P.ms_dg;
    dp_elapsed(ctx, &dg_stats->ms_total, ctx->ev2, ctx->ev5);
    dg_stats->launches = F.P.dg_launches;
  }
  return DP_OK;
}

extern "C" int dp_find(dp_ctx* ctx, const dp_template* t, uint32_t n_primers,
                       const char* const* masters, const int32_t* max_mismatches,
                       uint64_t seg_begin, uint64_t seg_end, int strand_mask, int algo,
                       const uint32_t* var_count, const uint64_t* var_offset, const char* variants,
                       const dp_conditions* cond, dp_hit** hits_out, uint64_t* n_hits_out,
                       dp_dimer** dimers_out, uint64_t* n_dimers_out,
                       dp_scan_stats* stats, dp_dg_stats* dg_stats) {
  if (!hits_out || !n_hits_out || !dimers_out || !n_dimers_out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *hits_out = nullptr; *n_hits_out = 0;
  *dimers_out = nullptr; *n_dimers_out = 0;
  if (stats) memset(stats, 0, sizeof *stats);
  if (dg_stats) memset(dg_stats, 0, sizeof *dg_stats);
  int rc = dp_find_begin(ctx, t, n_primers, masters, max_mismatches, seg_begin, seg_end, strand_mask, algo,
                         var_count, var_offset, variants, cond);
  if (rc) return rc;
  return dp_find_end(ctx, hits_out, n_hits_out, dimers_out, n_dimers_out, stats, dg_stats);
}

extern "C" int dp_set_timing(dp_ctx* ctx, int on) {
  if (!ctx) return DP_ERR_ARG;
  ctx->timing = on != 0;
  return DP_OK;
}

extern "C" int dp_scan_prepare_kernels(dp_ctx* ctx, uint32_t n_primers, const char* const* masters,
                                       const int32_t* max_mismatches, int strand_mask) {
  if (!ctx || !masters) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (n_primers == 0 || n_primers > DP_MAX_PRIMERS)
    return dp_set_error(ctx, DP_ERR_ARG, "n_primers must be in [1,%d]", DP_MAX_PRIMERS);
  if (!(strand_mask & DP_STRAND_BOTH)) return dp_set_error(ctx, DP_ERR_ARG, "empty strand mask");
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  dp_scan_job J;
  return job_build(ctx, n_primers, masters, max_mismatches, strand_mask, DP_ALGO_AUTO, nullptr, J, true);
}

// Build check without a device: generates the specialised kernel source of a pattern set,
// compiles it with NVRTC and (optionally) writes source / cubin of the first unit to files.
// Returns the number of launch units, or a negative DP_ERR code (message via dp_last_error(NULL)).
extern "C" int dp_jit_selftest(uint32_t n_primers, const char* const* masters, const int32_t* max_mismatches,
                               int strand_mask, const char* source_path, const char* cubin_path) {
  dp_scan_job J;
  J.core = std::make_shared<dp_job_core>();
  int rc = build_patterns(nullptr, n_primers, masters, max_mismatches, strand_mask, nullptr, J.core->pats, J.core->letters);
  if (rc) return rc;
  if (!dp_bitslice_make_plan(J.core->pats, J.core->letters, J.core->plan)) return 0;
  std::vector<dp_jit_entry> entries;
  for (const dp_bitslice_group& g : J.core->plan.groups)
    for (const dp_bs_pattern& bp : g.pats) entries.push_back(dp_jit_entry{bp.pattern, bp.anchor, g.K});
  std::sort(entries.begin(), entries.end(), [](const dp_jit_entry& a, const dp_jit_entry& b) { return a.pattern < b.pattern; });
  dp_jit_make_units(J.core->pats, J.core->letters, entries, J.core->jit);
  for (size_t i = 0; i < J.core->jit.size(); ++i) {
    std::string log;
    std::string cubin = dp_jit_compile_only(J.core->jit[i].source, log);
    if (i == 0 && source_path) {
      FILE* f = fopen(source_path, "wb");
      if (f) { fwrite(J.core->jit[i].source.data(), 1, J.core->jit[i].source.size(), f); fclose(f); }
    }
    if (cubin.empty()) return dp_set_error(nullptr, DP_ERR_STATE, "%s", log.c_str());
    if (i == 0 && cubin_path) {
      FILE* f = fopen(cubin_path, "wb");
      if (f) { fwrite(cubin.data(), 1, cubin.size(), f); fclose(f); }
    }
  }
  return (int)J.core->jit.size();
}

// The prefilter plan of a primer set as text (one "//PLAN" line per pattern: K, groups needed,
// estimated alive probability, the window positions of every group).  No device needed.
extern "C" int dp_jit_describe(uint32_t n_primers, const char* const* masters, const int32_t* max_mismatches,
                               int strand_mask, char* out, uint64_t capacity) {
  if (!masters || !out || !capacity) return DP_ERR_ARG;
  out[0] = 0;
  dp_scan_job J;
  J.core = std::make_shared<dp_job_core>();
  int rc = build_patterns(nullptr, n_primers, masters, max_mismatches, strand_mask, nullptr, J.core->pats, J.core->letters);
  if (rc) return rc;
  if (!dp_bitslice_make_plan(J.core->pats, J.core->letters, J.core->plan)) return 0;
  std::vector<dp_jit_entry> entries;
  for (const dp_bitslice_group& g : J.core->plan.groups)
    for (const dp_bs_pattern& bp : g.pats) entries.push_back(dp_jit_entry{bp.pattern, bp.anchor, g.K});
  std::sort(entries.begin(), entries.end(), [](const dp_jit_entry& a, const dp_jit_entry& b) { return a.pattern < b.pattern; });
  dp_jit_make_units(J.core->pats, J.core->letters, entries, J.core->jit);
  std::string text;
  for (const dp_jit_unit& u : J.core->jit) {
    size_t at = 0;
    while (at < u.source.size()) {
      size_t eol = u.source.find('\n', at);
      if (eol == std::string::npos) eol = u.source.size();
      if (u.source.compare(at, 6, "//PLAN") == 0) text += u.source.substr(at, eol - at) + "\n";
      else break;
      at = eol + 1;
    }
  }
  if (text.size() + 1 > capacity) return dp_set_error(nullptr, DP_ERR_ARG, "dp_jit_describe: %zu bytes needed", text.size() + 1);
  memcpy(out, text.c_str(), text.size() + 1);
  return (int)J.core->jit.size();
}

extern "C" int dp_set_wire(dp_ctx* ctx, void* device_ptr, uint64_t bytes) {
  if (!ctx) return DP_ERR_ARG;
  if (device_ptr && bytes < 64 + sizeof(dp_raw_hit))
    return dp_set_error(ctx, DP_ERR_ARG, "wire buffer must hold the 64-byte header and at least one 24-byte record");
  if (((uintptr_t)device_ptr & 7u) != 0) return dp_set_error(ctx, DP_ERR_ARG, "wire buffer must be 8-byte aligned");
  ctx->wire = device_ptr;
  ctx->wire_bytes = device_ptr ? bytes : 0;
  return DP_OK;
}

// Duplex stage alone, on hits handed back by the caller (the reference's optimiser
// re-runs it per step under new conditions, DegenPrimer/PCR_Optimizer.py:162-169).
extern "C" int dp_duplex_dg(dp_ctx* ctx, const dp_template* t, const dp_hit* hits, uint64_t n_hits,
                            uint32_t n_primers, const uint8_t* primer_len, const uint32_t* var_count,
                            const uint64_t* var_offset, const char* variants, const dp_conditions* cond,
                            dp_dimer** dimers_out, uint64_t* n_dimers_out, dp_dg_stats* stats) {
  if (!ctx || !t || !cond || !dimers_out || !n_dimers_out || (n_hits && !hits) ||
      (n_primers && (!primer_len || !var_count || !var_offset || !variants)))
    return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *dimers_out = nullptr;
  *n_dimers_out = 0;
  if (stats) memset(stats, 0, sizeof *stats);
  if (!ctx->tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_tables_load() has not been called");
  if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  if (n_hits == 0) { *dimers_out = (dp_dimer*)dp_result_alloc(sizeof(dp_dimer)); return DP_OK; }
  drop_resident(ctx);
  if (n_hits > 0xF0000000ull) return dp_set_error(ctx, DP_ERR_ARG, "more than 4e9 hits in one call");
  if (n_primers > DP_MAX_PRIMERS) return dp_set_error(ctx, DP_ERR_ARG, "n_primers must be in [1,%d]", DP_MAX_PRIMERS);
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  DP_EVENT(ctx, ev0, ctx->stream);
  uint64_t var_bytes = 0;
  for (uint32_t p = 0; p < n_primers; ++p) {
    if (primer_len[p] == 0 || primer_len[p] > DP_MAX_PRIMER_LEN)
      return dp_set_error(ctx, DP_ERR_ARG, "primer %u: length outside [1,%d]", p, DP_MAX_PRIMER_LEN);
    var_bytes = std::max<uint64_t>(var_bytes, var_offset[p] + (uint64_t)var_count[p] * primer_len[p]);
  }
  if (var_bytes > 0xFFFFFFFFull) return dp_set_error(ctx, DP_ERR_ARG, "variant blob larger than 4 GB");
  int rc = upload_codes(ctx, variants, var_bytes);
  if (rc) return rc;
  dp_dg_launch DL;
  fill_dg_launch(ctx, cond, DL);
  for (uint32_t p = 0; p < n_primers; ++p) {
    for (int s = 0; s < 2; ++s) {
      DL.plan.pat_len[2 * p + s] = primer_len[p];
      DL.plan.pat_strand[2 * p + s] = (uint8_t)s;
      DL.plan.pat_primer[2 * p + s] = (uint8_t)p;
    }
    DL.plan.var_count[p] = var_count[p];
    DL.plan.var_off[p] = (uint32_t)var_offset[p];
  }
  // raw hits in the caller's order
  rc = dp_ws_hits(ctx, std::max<uint64_t>(std::max<uint64_t>(ctx->hits_cap, 1u << 20), n_hits));
  if (rc) return rc;
  rc = dp_ws_pinned(ctx, DP_RES_SPEC_END + var_bytes + n_hits * sizeof(dp_raw_hit) + 64);
  if (rc) return rc;
  dp_raw_hit* raw = (dp_raw_hit*)(ctx->h_res + ((DP_RES_SPEC_END + var_bytes + 63) & ~(size_t)63));
  uint64_t pairs = 0;
  for (uint64_t h = 0; h < n_hits; ++h) {
    const dp_hit& x = hits[h];
    if (x.primer >= n_primers) return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: primer %u out of range", (unsigned long long)h, x.primer);
    if (x.record >= t->records.size()) return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: record out of range", (unsigned long long)h);
    const dp_record& rec = t->records[x.record];
    const uint32_t L = primer_len[x.primer];
    if (x.pos + L > rec.len) return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: window outside its record", (unsigned long long)h);
    raw[h].g = rec.start + (x.strand ? rec.len - L - x.pos : x.pos);
    raw[h].record = x.record;
    raw[h].pattern = (uint16_t)(2 * x.primer + (x.strand ? 1 : 0));
    raw[h].score3 = x.score3;
    raw[h].pair_base = (uint32_t)pairs;
    raw[h].reserved = 0;
    pairs += var_count[x.primer];
    if (pairs > 0xF0000000ull) return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "more than 4e9 (hit, variant) duplexes in one call");
  }
  rc = dp_ws_dir(ctx, std::max<uint64_t>(std::max<uint64_t>(ctx->dir_cap, 1u << 22), pairs));
  if (rc) return rc;
  rc = dp_ws_dimers(ctx, std::max<uint64_t>(ctx->dimers_cap, 1u << 22));
  if (rc) return rc;
  cudaStream_t st = ctx->stream;
  unsigned long long* hc = (unsigned long long*)ctx->h_res + 8;     // behind the counter read-back area
  for (int i = 0; i < 8; ++i) hc[i] = 0;
  hc[DP_C_HITS] = n_hits | (pairs << 32);
  DP_CUDA(ctx, cudaMemcpyAsync(ctx->d_hits, raw, n_hits * sizeof(dp_raw_hit), cudaMemcpyHostToDevice, st));
  dp_sizes S;
  dp_host_results R;
  float ms_dg = 0.f;
  uint32_t launches = 0;
  for (int attempt = 0;; ++attempt) {
    if (attempt == 4) return dp_set_error(ctx, DP_ERR_STATE, "internal: dimer buffer keeps overflowing");
    DP_CUDA(ctx, cudaMemcpyAsync(ctx->d_counters, hc, 8 * sizeof(unsigned long long), cudaMemcpyHostToDevice, st));
    ctx->counters_clean = false;
    DP_EVENT(ctx, ev2, st);
    rc = dp_dg_enqueue(ctx, t, DL, pairs, &launches, /*hits_fresh=*/true);      // uploaded with reserved = 0
    if (rc) return rc;
    DP_EVENT(ctx, ev4, st);
    // only directory + dimers are wanted back; the hit prefix copy is harmless
    rc = fetch_sizes(ctx, true, S);
    if (rc) return rc;
    float ms = 0.f;
    dp_elapsed(ctx, &ms, ctx->ev2, ctx->ev4);
    ms_dg += ms;
    if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
    if (S.n_dimers > ctx->dimers_cap) {
      rc = dp_ws_dimers(ctx, S.n_dimers + (S.n_dimers >> 3) + 1024);
      if (rc) return rc;
      continue;
    }
    break;
  }
  std::vector<uint32_t> order(n_hits);
  for (uint64_t i = 0; i < n_hits; ++i) order[i] = (uint32_t)i;
  if (results_are_large(true, S)) {
    rc = dimers_large(ctx, DL.plan, order, S, nullptr, dimers_out, n_dimers_out);
  } else {
    // small: directory + dimers already sit in the published prefix; the raw hits are
    // still in the pinned staging area behind it
    S.n_hits = 0;
    rc = fetch_all(ctx, true, S, R);
    if (rc) return rc;
    R.hits = raw;
    rc = dimers_to_abi(ctx, R, S, order, [&](size_t i) { return var_count[hits[i].primer]; }, dimers_out, n_dimers_out);
  }
  if (rc) return rc;
  if (stats) {
    stats->duplexes = pairs;
    stats->dimers_tried = S.tried;
    stats->dimers_kept = *n_dimers_out;
    stats->ms_kernel = ms_dg;
    dp_elapsed(ctx, &stats->ms_total, ctx->ev0, ctx->ev5);
    stats->launches = launches;
  }
  return DP_OK;
}
