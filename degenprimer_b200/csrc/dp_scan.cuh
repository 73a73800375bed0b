// Scan: shared device structures (patterns in constant memory, exact verify).
#pragma once
#include "dp_internal.cuh"

#define DP_MAX_PATTERNS (2 * DP_MAX_PRIMERS)

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
};

// raw hit as written by the kernels (padded-concatenation coordinates)
struct dp_raw_hit {
  uint64_t g;          // forward window start in the padded concatenation
  uint32_t record;
  uint16_t pattern;
  uint16_t score3;
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
  unsigned long long* counters;   // [0] hits wanted, [1] candidates verified
};

// defined once: this header is included by dp_scan.cu only
__constant__ dp_pattern c_patterns[DP_MAX_PATTERNS];

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
                                            uint32_t len, uint32_t score) {
  if (g < A.seg_begin || g >= A.seg_end) return;
  uint32_t r = dp_find_record(A.rec_start, A.n_records, g);
  uint64_t off = g - A.rec_start[r];
  if (off + len > A.rec_len[r]) return;          // window runs past its record
  unsigned long long slot = atomicAdd(&A.counters[0], 1ull);
  if (slot < A.hits_cap) {
    dp_raw_hit h;
    h.g = g; h.record = r; h.pattern = (uint16_t)pat; h.score3 = (uint16_t)score;
    A.hits[slot] = h;
  }
}
