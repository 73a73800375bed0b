// Device-side structures and helpers of the scan kernels.  This header is compiled
// twice: by nvcc into the library, and by NVRTC at run time as the header of the
// specialised scan kernels (dp_jit.cu embeds its text), so it must stay free of host
// headers.
#pragma once
#ifdef __CUDACC_RTC__
typedef unsigned char uint8_t;
typedef unsigned short uint16_t;
typedef unsigned int uint32_t;
typedef int int32_t;
typedef unsigned long long uint64_t;
#else
#include <stdint.h>
#endif

// One maximal run of a single non-ACGT letter, in padded-concatenation coords.
struct dp_exc_run {
  uint64_t start;
  uint64_t end;      // exclusive
  uint32_t letter;   // letter index 4..14, or DP_NOT_A_LETTER for record padding
  uint32_t pad;
};

// raw hit as written by the scan kernels (padded-concatenation coordinates).
// pair_base = number of (hit, variant) pairs of all earlier slots: slot and
// pair_base come from ONE packed 64-bit atomic, so pair_base is monotone in slot.
struct dp_raw_hit {
  uint64_t g;          // forward window start in the padded concatenation
  uint32_t record;
  uint16_t pattern;
  uint16_t score3;
  uint32_t pair_base;
  uint32_t reserved;
};

// counters[] slots (device, 16 x u64)
enum {
  DP_C_HITS = 0,      // low 32 bits: hits wanted, high 32 bits: (hit, variant) pairs
  DP_C_CAND = 1,      // windows that reached the exact verify
  DP_C_DIMERS = 2,    // dimer slots wanted
  DP_C_TRIED = 3,     // dimers whose dG was computed
  DP_C_SATURATED = 4, // the packed hit counter came close to 2^32
  DP_C_CHUNK0 = 5,    // 5..7: next chunk of words of launch unit 0..2 of the specialised scan kernel
  DP_C_PACK0 = 8,     // pack kernel: boundary entries wanted
  DP_C_PACK1 = 9      // pack kernel: first foreign byte
};

// One (primer, strand) pattern.  Masks are indexed by plane code (A C T G) and
// cover primer positions 0..63 (two words, bit j = position j of the WINDOW on
// the forward strand; reverse-strand patterns are stored reverse-complemented).
//   f: base scores 3/3 at this position,  h: base scores 2/3   
struct dp_pattern {
  uint32_t f[4][2];
  uint32_t h[4][2];
  uint32_t lmask[2];
  uint16_t len;
  uint16_t thr3;       // hit iff score3 >= thr3  (3*max(1,L-k)-1)
  uint16_t primer;
  uint8_t strand;
  uint8_t has_half;
  uint32_t n_var;      // unambiguous variants of the primer (duplex pairs per hit)
  uint32_t reserved;
};

// The patterns of a call travel as a __grid_constant__ kernel parameter (constant
// bank 0, uniform LDC access, no symbol upload, nothing shared between contexts).
// Two capacities so that the common 1-4 primer call ships < 1 KB per launch.
#define DP_NP_SMALL 8
template <int NP>
struct dp_pattern_set {
  dp_pattern pat[NP];
};

struct dp_scan_args {
  const uint32_t* lo;
  const uint32_t* hi;
  const uint32_t* exc;
  const uint64_t* rec_start;
  const uint64_t* rec_len;
  uint32_t n_records;
  uint32_t n_patterns;
  uint64_t n_words;       // words holding record data
  uint64_t word_begin;    // first word scanned
  uint64_t word_end;
  uint64_t seg_begin;     // window starts allowed in [seg_begin, seg_end)
  uint64_t seg_end;
  dp_raw_hit* hits;
  uint64_t hits_cap;
  unsigned long long* counters;   // DP_C_* slots
  const dp_exc_run* runs;         // exception runs (sorted) for the run-boundary pass
  uint64_t n_runs;
};

__device__ __forceinline__ uint32_t dp_mux4(uint32_t hi, uint32_t lo, uint32_t a, uint32_t c,
                                            uint32_t t, uint32_t g) {
  uint32_t u0 = (lo & c) | (~lo & a);    // hi = 0: A or C
  uint32_t u1 = (lo & g) | (~lo & t);    // hi = 1: T or G
  return (hi & u1) | (~hi & u0);
}

// exact thirds score of one window against one pattern; words 0/1 hold window
// positions 0..31 / 32..63
template <int NW>
__device__ __forceinline__ uint32_t dp_score3(const dp_pattern& p, const uint32_t* whi,
                                              const uint32_t* wlo, const uint32_t* wex) {
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < NW; ++k) {
    uint32_t ok = ~wex[k];
    uint32_t mf = dp_mux4(whi[k], wlo[k], p.f[0][k], p.f[1][k], p.f[2][k], p.f[3][k]) & ok;
    s += 3u * __popc(mf);
    if (p.has_half) {
      uint32_t mh = dp_mux4(whi[k], wlo[k], p.h[0][k], p.h[1][k], p.h[2][k], p.h[3][k]) & ok;
      s += 2u * __popc(mh);
    }
    s += __popc(wex[k] & p.lmask[k]);
  }
  return s;
}

// record that contains padded position g (records are word aligned, sorted)
__device__ __forceinline__ uint32_t dp_find_record(const uint64_t* rec_start, uint32_t n, uint64_t g) {
  uint32_t a = 0, b = n;
  while (a < b) {
    uint32_t m = (a + b) >> 1;
    if (rec_start[m] <= g) a = m + 1; else b = m;
  }
  return a ? a - 1 : 0;
}

__device__ __forceinline__ void dp_emit_hit(const dp_scan_args& A, uint64_t g, uint32_t pat,
                                            uint32_t len, uint32_t score, uint32_t n_var) {
  if (g < A.seg_begin || g >= A.seg_end) return;
  uint32_t r = dp_find_record(A.rec_start, A.n_records, g);
  uint64_t off = g - A.rec_start[r];
  if (off + len > A.rec_len[r]) return;          // window runs past its record
  // one packed atomic hands out the slot (low half) and the first (hit, variant)
  // pair index (high half) together, so pair_base is monotone in slot
  const unsigned long long seen = *(volatile unsigned long long*)&A.counters[DP_C_HITS];
  if ((uint32_t)seen > 0xF0000000u || (seen >> 32) > 0xF0000000ull) {
    A.counters[DP_C_SATURATED] = 1ull;
    return;
  }
  const unsigned long long old = atomicAdd(&A.counters[DP_C_HITS], 1ull | ((unsigned long long)n_var << 32));
  const uint32_t slot = (uint32_t)old;
  if (slot < A.hits_cap) {
    dp_raw_hit h;
    h.g = g; h.record = r; h.pattern = (uint16_t)pat; h.score3 = (uint16_t)score;
    h.pair_base = (uint32_t)(old >> 32); h.reserved = 0;
    A.hits[slot] = h;
  }
}

// exact thirds score of window g against pattern P; with_exc = false leaves windows that touch
// a non-ACGT base to the run-boundary pass.  True iff the window reaches the threshold.
__device__ __forceinline__ bool dp_score_window(const dp_scan_args& A, const dp_pattern& P, uint64_t g,
                                                bool with_exc, uint32_t* score) {
  const uint64_t wi = g >> 5;
  const int sh = (int)(g & 31);
  uint32_t whi[2], wlo[2], wex[2];
  {
    const uint32_t a0 = __ldg(A.hi + wi), a1 = __ldg(A.hi + wi + 1), a2 = __ldg(A.hi + wi + 2);
    whi[0] = __funnelshift_r(a0, a1, sh); whi[1] = __funnelshift_r(a1, a2, sh);
    const uint32_t b0 = __ldg(A.lo + wi), b1 = __ldg(A.lo + wi + 1), b2 = __ldg(A.lo + wi + 2);
    wlo[0] = __funnelshift_r(b0, b1, sh); wlo[1] = __funnelshift_r(b1, b2, sh);
    const uint32_t e0 = __ldg(A.exc + wi), e1 = __ldg(A.exc + wi + 1), e2 = __ldg(A.exc + wi + 2);
    wex[0] = __funnelshift_r(e0, e1, sh); wex[1] = __funnelshift_r(e1, e2, sh);
  }
  if (!with_exc && ((wex[0] & P.lmask[0]) | (wex[1] & P.lmask[1]))) return false;
  const uint32_t s = dp_score3<2>(P, whi, wlo, wex);
  *score = s;
  return s >= P.thr3;
}

__device__ __forceinline__ void dp_verify_window(const dp_scan_args& A, const dp_pattern& P, uint64_t g,
                                                 uint32_t pat, bool with_exc) {
  uint32_t s;
  if (dp_score_window(A, P, g, with_exc, &s)) dp_emit_hit(A, g, pat, P.len, s, P.n_var);
}

// ---- windows that touch non-ACGT bases, for a set of prefiltered patterns --------------------
// A window belongs to the run that holds its FIRST unknown base: for run [start, end) with the previous run
// ending at prev_end, starts in [max(prev_end, start-(L-1)), start) reach into the run from the left, starts in
// [max(start, end-emax), end) begin inside it (a window with more than emax = (3L - thr3)/2 unknown bases cannot
// reach thr3).  Items (run, candidate start) are dealt to threads `first, first + stride, ...`.
__device__ __forceinline__ void dp_scan_runs_device(const dp_scan_args& A, const dp_pattern* pats, const uint8_t* pattern_of,
                                                    uint32_t n_pats, uint64_t first, uint64_t stride) {
  if (!A.n_runs) return;
  uint32_t left = 0, emax = 0;
  for (uint32_t p = 0; p < n_pats; ++p) {
    left = max(left, (uint32_t)pats[p].len - 1u);
    emax = max(emax, (3u * pats[p].len - pats[p].thr3) / 2u);
  }
  emax = max(1u, min(emax, left + 1u));
  const uint32_t span = left + emax;
  const uint64_t n_items = A.n_runs * (uint64_t)span;
  for (uint64_t idx = first; idx < n_items; idx += stride) {
    const uint64_t r = idx / span;
    const uint32_t j = (uint32_t)(idx % span);
    const uint64_t start = A.runs[r].start, end = A.runs[r].end;
    const uint64_t prev_end = r ? A.runs[r - 1].end : 0;
    uint64_t g;
    if (j < left) {
      const uint64_t d = left - j;                 // 1 .. Lmax-1 bases in front of the run
      if (start < d) continue;
      g = start - d;
      if (g < prev_end) continue;                  // an earlier run owns this window
    } else {
      const uint64_t d = emax - (j - left);        // 1 .. emax bases before the run's end
      if (end - start < d) continue;
      g = end - d;
    }
    for (uint32_t p = 0; p < n_pats; ++p) {
      const dp_pattern& P = pats[p];
      if (g < start && g + P.len <= start) continue;          // does not reach the run
      if (g >= start && 2ull * min((unsigned long long)(end - g), (unsigned long long)P.len) > 3ull * P.len - P.thr3) continue;
      dp_verify_window(A, P, g, pattern_of[p], true);
    }
  }
}

// ---- deferred exact verify of the run-time specialised kernels ------------------------------
// The prefilter leaves ~1e-4 of the windows alive.  Verifying them where they are found costs a
// whole warp ~250 issue slots per candidate (31 lanes idle).  Instead every warp parks its
// candidates in a small shared-memory queue and verifies 32 of them at once, one per lane; the
// hits of such a batch take ONE atomic (slot and pair index handed out by ballot / prefix sum).
#define DP_Q_CAP 64
#define DP_JIT_CHUNK 16       // words per thread and chunk of the specialised scan kernels ...
#define DP_JIT_CHUNK_TAIL 4   // ... and in the last sixteenth of a launch

// every lane of a converged warp calls this; lanes with `hit` append a record
__device__ __forceinline__ void dp_emit_hits_warp(const dp_scan_args& A, bool hit, uint64_t g, uint32_t r,
                                                  uint32_t pat, uint32_t score, uint32_t n_var) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, hit);
  if (!m) return;
  const uint32_t v = hit ? n_var : 0u;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t x = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= (uint32_t)o) incl += x;
  }
  const uint32_t total_var = __shfl_sync(0xFFFFFFFFu, incl, 31);
  const int leader = __ffs(m) - 1;
  unsigned long long old = 0;
  if ((int)lane == leader) {
    const unsigned long long seen = *(volatile unsigned long long*)&A.counters[DP_C_HITS];
    if ((uint32_t)seen > 0xF0000000u || (seen >> 32) > 0xF0000000ull) {
      A.counters[DP_C_SATURATED] = 1ull;
      old = ~0ull;
    } else {
      old = atomicAdd(&A.counters[DP_C_HITS], (unsigned long long)__popc(m) | ((unsigned long long)total_var << 32));
    }
  }
  old = __shfl_sync(0xFFFFFFFFu, old, leader);
  if (old == ~0ull || !hit) return;
  // slot and pair_base both grow with the lane number: pair_base stays monotone in slot
  const uint32_t slot = (uint32_t)old + __popc(m & ((1u << lane) - 1u));
  if (slot < A.hits_cap) {
    dp_raw_hit h;
    h.g = g; h.record = r; h.pattern = (uint16_t)pat; h.score3 = (uint16_t)score;
    h.pair_base = (uint32_t)(old >> 32) + (incl - v); h.reserved = 0;
    A.hits[slot] = h;
  }
}

// one queued candidate per lane (entry = g << 8 | slot in `pats`); `valid` lanes verify theirs
static __device__ __noinline__ void dp_verify_warp(const dp_scan_args& A, const dp_pattern* pats, const uint8_t* pattern_of,
                                            unsigned long long entry, bool valid) {
  bool hit = false;
  uint64_t g = 0;
  uint32_t r = 0, slot = 0, score = 0, n_var = 0;
  if (valid) {
    g = entry >> 8;
    slot = (uint32_t)(entry & 0xFFull);
    const dp_pattern& P = pats[slot];
    if (dp_score_window(A, P, g, false, &score)) {
      r = dp_find_record(A.rec_start, A.n_records, g);
      hit = g - A.rec_start[r] + P.len <= A.rec_len[r];        // a window must not run past its record
      n_var = P.n_var;
    }
  }
  dp_emit_hits_warp(A, hit, g, r, pattern_of[slot], score, n_var);
}

// Appends the alive windows of word w (bit b = window 32w+b) of pattern `slot` to the warp's
// queue q (n entries held, n < 32 on entry) and verifies a batch when 32 are there.  Called by
// all lanes of the warp together; returns the new n.
static __device__ __noinline__ uint32_t dp_queue_push(const dp_scan_args& A, const dp_pattern* pats, const uint8_t* pattern_of,
                                               unsigned long long* q, uint32_t n, uint32_t alive, uint64_t w,
                                               uint32_t slot, unsigned long long& n_cand) {
  const uint32_t lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
  while (__any_sync(0xFFFFFFFFu, alive != 0u)) {
    bool mine = false;
    unsigned long long e = 0ull;
    if (alive) {
      const int b = __ffs(alive) - 1;
      alive &= alive - 1u;
      ++n_cand;
      const uint64_t g = w * 32 + b;
      if (g >= A.seg_begin && g < A.seg_end) { mine = true; e = (g << 8) | slot; }
    }
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, mine);
    if (mine) q[n + __popc(m & below)] = e;
    n += __popc(m);
    __syncwarp();
    if (n >= 32u) {
      dp_verify_warp(A, pats, pattern_of, q[lane], true);
      __syncwarp();
      const unsigned long long moved = lane + 32u < n ? q[lane + 32u] : 0ull;
      __syncwarp();
      if (lane + 32u < n) q[lane] = moved;
      n -= 32u;
      __syncwarp();
    }
  }
  return n;
}

// ---- bit-sliced population count of N one-bit planes ----------------------------
template <int N>
struct dp_level {
  // Reduces N same-weight words to one; writes floor(N/2) next-weight carries.
  __device__ __forceinline__ static uint32_t run(const uint32_t* in, uint32_t* c) {
    if constexpr (N == 1) {
      return in[0];
    } else if constexpr (N == 2) {
      c[0] = in[0] & in[1];
      return in[0] ^ in[1];
    } else {
      constexpr int G = N / 3, R = N % 3;
      uint32_t s[G + R];
#pragma unroll
      for (int g = 0; g < G; ++g) {
        uint32_t a = in[3 * g], b = in[3 * g + 1], d = in[3 * g + 2];
        s[g] = a ^ b ^ d;
        c[g] = (a & b) | (d & (a | b));
      }
#pragma unroll
      for (int r = 0; r < R; ++r) s[G + r] = in[3 * G + r];
      return dp_level<G + R>::run(s, c + G);
    }
  }
};

// bits[] receives the binary count of the N planes; returns number of bits
template <int N, int B>
struct dp_count {
  __device__ __forceinline__ static void run(const uint32_t* in, uint32_t* bits) {
    if constexpr (N == 0) {
      return;
    } else {
      uint32_t c[N / 2 + 1];
      bits[B] = dp_level<N>::run(in, c);
      dp_count<N / 2, B + 1>::run(c, bits);
    }
  }
};

__host__ __device__ constexpr int dp_bits_for(int n) { return n == 0 ? 0 : 1 + dp_bits_for(n / 2); }

// word of windows whose plane count is >= THR
template <int N, int THR>
__device__ __forceinline__ uint32_t dp_count_ge(const uint32_t* in) {
  constexpr int NB = dp_bits_for(N);
  uint32_t bits[NB];
  dp_count<N, 0>::run(in, bits);
  uint32_t ge = 0u, eq = ~0u;
#pragma unroll
  for (int b = NB - 1; b >= 0; --b) {
    if ((THR >> b) & 1) {
      eq &= bits[b];
    } else {
      ge |= eq & bits[b];
      eq &= ~bits[b];
    }
  }
  return ge | eq;
}

