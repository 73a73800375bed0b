// Duplex thermodynamics: dp_duplex_dg, dp_variant_tm, dp_dimer_dg.
//
// Replaces, for every (site x unambiguous variant):
//   SearchEngine._compile_duplexes_for_position  DegenPrimer/SearchEngine.py:106-118
//   Duplex.__init__ / _recalculate               DegenPrimer/SecStructures.py:231-252,294-324
//   Dimer.from_sequences / alternatives          DegenPrimer/SecStructures.py:126-152,173-191
//   dimer_dG_corrected                           DegenPrimer/TD_Functions.py:124-206
// fp64 throughout; every add/mul goes through __dadd_rn/__dmul_rn so nothing is
// contracted into an FMA and the summation order is the reference's.
#include <algorithm>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "dp_internal.cuh"

#define DP_WIN 64     // bytes per stored window

struct dp_dg_args {
  const dp_nn_tables* T;
  double PCR_T, Na_eq_log, max_dimer_dG, max_dG_for_K;
};

__device__ __forceinline__ double f64_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double f64_mul(double a, double b) { return __dmul_rn(a, b); }

// ---- generic dimer dG -------------------------------------------------------------
// s: seq1 letters (5'->3'), s_len; r: seq2 reversed = 3'->5' letters, r_len;
// mask: matched seq1 indices; off: Dimer.offset.  Returns NaN where the
// reference would raise (a table key that does not exist).
// The two strands are anything indexable: byte arrays, or dp_seq_planes (a variant kept as two
// 64-bit planes in registers).
struct dp_seq_planes {
  uint64_t hi, lo;
  __device__ __forceinline__ uint32_t operator[](int i) const {
    return (uint32_t)((hi >> i) & 1ull) * 2u + (uint32_t)((lo >> i) & 1ull);
  }
};

template <typename SeqS, typename SeqR>
__device__ double dp_dimer_dg_device(const dp_dg_args& A, const SeqS& s, int s_len,
                                     const SeqR& r, int r_len, uint64_t mask, int off) {
  const dp_nn_tables& T = *A.T;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  if (!mask) return nan;
  const int f0 = __ffsll((long long)mask) - 1;
  const int f1 = 63 - __clzll((long long)mask);
  double dG = 0.0, dH = T.ini_dH, dS = T.ini_dS;
  bool bad = false;
  // left end — TD_Functions.py:140-158
  if (f0 == 0 && off < 0) {
    int r0 = f0 - off;
    uint32_t x = r[r0], y = r[r0 - 1], z = s[0];
    if (x > 3 || y > 3 || z > 3) bad = true;
    else { dH = f64_add(dH, T.dangling_dH[1][x][y][z]); dS = f64_add(dS, T.dangling_dS[1][x][y][z]); }
  } else if (f0 == 0 && off == 0) {
    if (s[0] == 0 || s[0] == 2) { dH = f64_add(dH, T.ter_dH); dS = f64_add(dS, T.ter_dS); }
  } else if (f0 == off) {
    uint32_t x = s[f0 - 1], y = s[f0], z = r[0];
    if (x > 3 || y > 3 || z > 3) bad = true;
    else { dH = f64_add(dH, T.dangling_dH[0][x][y][z]); dS = f64_add(dS, T.dangling_dS[0][x][y][z]); }
  } else if (f0 > off) {
    dG = f64_add(dG, T.terminal_mismatch);
  }
  // right end — TD_Functions.py:160-178
  const int r1 = f1 - off;
  if (f1 == s_len - 1 && r1 < r_len - 1) {
    uint32_t x = r[r1 + 1], y = r[r1], z = s[s_len - 1];
    if (x > 3 || y > 3 || z > 3) bad = true;
    else { dH = f64_add(dH, T.dangling_dH[0][x][y][z]); dS = f64_add(dS, T.dangling_dS[0][x][y][z]); }
  } else if (r1 == r_len - 1 && f1 < s_len - 1) {
    uint32_t x = s[f1], y = s[f1 + 1], z = r[r_len - 1];
    if (x > 3 || y > 3 || z > 3) bad = true;
    else { dH = f64_add(dH, T.dangling_dH[1][x][y][z]); dS = f64_add(dS, T.dangling_dS[1][x][y][z]); }
  } else if (f1 < s_len - 1 && r1 < r_len - 1) {
    dG = f64_add(dG, T.terminal_mismatch);
  } else if (f1 == s_len - 1 && r1 == r_len - 1) {
    if (s[s_len - 1] == 0 || s[s_len - 1] == 2) { dH = f64_add(dH, T.ter_dH); dS = f64_add(dS, T.ter_dS); }
  }
  // stacks, |x|, |xx| and internal loops — TD_Functions.py:180-202
  uint64_t m = mask & (mask - 1);     // matches after the first
  int a = f0;
  while (m) {
    const int b = __ffsll((long long)m) - 1;
    m &= m - 1;
    const int gap = b - a;
    if (gap < 4) {
      uint32_t x = s[a], y = s[a + 1], z = r[a - off], w = r[a - off + 1];
      if (x > 3 || y > 3 || w > 14 || z != dp_letter_complement(x)) bad = true;
      else { dH = f64_add(dH, T.stack_dH[x][y][w]); dS = f64_add(dS, T.stack_dS[x][y][w]); }
      if (gap > 1) {
        uint32_t x2 = s[b - 1], y2 = s[b], z2 = r[b - off], w2 = r[b - off - 1];
        if (x2 > 3 || y2 > 3 || w2 > 14 || z2 != dp_letter_complement(y2)) bad = true;
        else { dH = f64_add(dH, T.rstack_dH[x2][y2][w2]); dS = f64_add(dS, T.rstack_dS[x2][y2][w2]); }
      }
    } else {
      dS = f64_add(dS, T.loop_dS[gap - 1]);
    }
    a = b;
  }
  // salt and temperature — TD_Functions.py:204-205
  dS = f64_add(dS, f64_mul(f64_mul(T.dS_Na, (double)(f1 - f0)), A.Na_eq_log));
  const double tds = __ddiv_rn(f64_mul(f64_add(A.PCR_T, -T.K0), dS), 1000.0);
  const double res = f64_add(f64_add(dG, dH), -tds);
  return bad ? nan : res;
}

// ---- window letters ----------------------------------------------------------------
// letter index of padded position g (plane code, or the run's IUPAC letter)
__device__ __forceinline__ uint32_t dp_letter_at(const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi,
                                                 const uint32_t* __restrict__ exc, const dp_exc_run* __restrict__ runs,
                                                 uint64_t n_runs, uint64_t g) {
  const uint64_t wi = g >> 5;
  const uint32_t b = (uint32_t)(g & 31);
  if ((__ldg(exc + wi) >> b) & 1u) {
    uint64_t lo_i = 0, hi_i = n_runs;
    while (lo_i < hi_i) {
      uint64_t mid = (lo_i + hi_i) >> 1;
      if (runs[mid].start <= g) lo_i = mid + 1; else hi_i = mid;
    }
    if (lo_i) { const dp_exc_run& rr = runs[lo_i - 1]; if (g < rr.end) return rr.letter; }
    return 14u;
  }
  return ((__ldg(hi + wi) >> b) & 1u) * 2u + ((__ldg(lo + wi) >> b) & 1u);
}

// win[h*64 + i] = letter index of window position i as the reference's
// strand string shows it (reverse hits: reverse complement of the forward bases)
__global__ void dp_dg_windows_kernel(const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi,
                                     const uint32_t* __restrict__ exc, const dp_exc_run* __restrict__ runs,
                                     uint64_t n_runs, const uint64_t* __restrict__ hit_g,
                                     const uint8_t* __restrict__ hit_len, const uint8_t* __restrict__ hit_strand,
                                     uint64_t n_hits, uint8_t* __restrict__ win) {
  uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t h = idx / DP_WIN;
  int i = (int)(idx % DP_WIN);
  if (h >= n_hits) return;
  int L = hit_len[h];
  uint8_t out = DP_NOT_A_LETTER;
  if (i < L) {
    int strand = hit_strand[h];
    uint64_t g = hit_g[h] + (strand ? (uint64_t)(L - 1 - i) : (uint64_t)i);
    uint32_t letter = dp_letter_at(lo, hi, exc, runs, n_runs, g);
    out = (uint8_t)(strand ? dp_letter_complement(letter) : letter);
  }
  win[idx] = out;
}

// ---- per (hit, variant) duplex: the whole Duplex.__init__ / _recalculate in one pass ----
// Grid-stride over the (hit, variant) pairs of the raw hits in `hits`; the pair
// and hit counts are read from the device counters, so the kernel can be queued
// right behind the scan without a host round trip.  Per pair:
//   window letters from the planes -> match mask -> [full, alternatives...] in the
//   reference's order -> dG of each -> filter -> stable order by dG -> one atomic
//   reservation of the output range -> directory entry offset << 16 | count.
struct dp_duplex_args {
  dp_dg_args dg;
  const uint32_t* lo;
  const uint32_t* hi;
  const uint32_t* exc;
  const dp_exc_run* runs;
  uint64_t n_runs;
  dp_raw_hit* hits;             // .reserved receives the number of dimers of the hit
  uint64_t hits_cap;
  const uint8_t* codes;
  uint64_t* dir;
  uint64_t dir_cap;
  dp_raw_dimer* out;
  uint64_t out_cap;
  unsigned long long* counters;
};

// Geometry (measured on B200, tools/dg_variants.sh / n8_variants.sh): 4 warps per block, 16-entry columns, 4 blocks
// per SM.  Two-warp blocks (5K registers, 23 KB shared memory) fit on an SM NEXT TO the resident block of another
// call's scan kernel, but the overlap bought nothing measurable while the smaller columns overflow more often:
// config 4 0.228 vs 0.173 ms, and 0.152 vs 0.075 ms on a 1/8 shard.
#ifndef DP_DUPLEX_WARPS
#define DP_DUPLEX_WARPS 4
#endif
#define DP_DUPLEX_THREADS (32 * DP_DUPLEX_WARPS)
#ifndef DP_LANE_LIST
#define DP_LANE_LIST 16        // dimers one lane's column of shared memory holds (more: a slow path, same results)
#endif
#ifndef DP_DUPLEX_BLOCKS_PER_SM
#define DP_DUPLEX_BLOCKS_PER_SM 4
#endif
#define DP_DUPLEX_TABLE_BYTES ((sizeof(dp_nn_tables) + 15) / 16 * 16)
#define DP_DUPLEX_SMEM (DP_DUPLEX_TABLE_BYTES + DP_DUPLEX_WARPS * sizeof(dp_duplex_scratch))

// per-warp scratch: the window's complement strand and every lane's list of surviving dimers
// ([entry][lane]: a lane's accesses stay in its own bank)
struct dp_duplex_scratch {
  uint8_t r[DP_WIN];
  double kdg[DP_LANE_LIST * 32];
  uint64_t kmask[DP_LANE_LIST * 32];
  uint16_t kd[DP_LANE_LIST * 32];        // construction index of the entry (ties keep that order)
  uint16_t kn[32];                       // survivors of every lane in this pass
  uint8_t idx[DP_WIN * 32];              // [position][lane]: NN-table index of the stack (s[i] s[i+1] / r[i+1])
};

// dimer_dG_corrected (TD_Functions.py:124-206) for a primer-template Duplex: both strands have
// the primer's length and offset 0, so no dangling end is reachable and the end terms reduce to
// "terminal AT" or "terminal mismatch"; the stack of two adjacent positions is looked up through
// the lane's precomputed table index.  Same additions in the same order as dp_dimer_dg_device.
__device__ __forceinline__ double dp_duplex_dimer_dg(const dp_dg_args& A, const dp_seq_planes& s, const uint8_t* r,
                                                     const uint8_t* idx /* stride 32 */, int L, uint64_t mask) {
  const dp_nn_tables& T = *A.T;
  const double* sdH = &T.stack_dH[0][0][0];
  const double* sdS = &T.stack_dS[0][0][0];
  const int f0 = __ffsll((long long)mask) - 1;
  const int f1 = 63 - __clzll((long long)mask);
  double dG = 0.0, dH = T.ini_dH, dS = T.ini_dS;
  if (f0 == 0) {                                         // left end, TD_Functions.py:140-158
    const uint32_t x = s[0];
    if (x == 0 || x == 2) { dH = f64_add(dH, T.ter_dH); dS = f64_add(dS, T.ter_dS); }
  } else {
    dG = f64_add(dG, T.terminal_mismatch);
  }
  if (f1 == L - 1) {                                     // right end, :160-178
    const uint32_t x = s[L - 1];
    if (x == 0 || x == 2) { dH = f64_add(dH, T.ter_dH); dS = f64_add(dS, T.ter_dS); }
  } else {
    dG = f64_add(dG, T.terminal_mismatch);
  }
  uint64_t m = mask & (mask - 1);                        // stacks, |x|, |xx|, loops, :180-202
  int a = f0;
  while (m) {
    const int b = __ffsll((long long)m) - 1;
    m &= m - 1;
    const int gap = b - a;
    if (gap < 4) {
      const uint32_t i = idx[32 * a];
      dH = f64_add(dH, sdH[i]);
      dS = f64_add(dS, sdS[i]);
      if (gap > 1) {
        const uint32_t x2 = s[b - 1], y2 = s[b], w2 = r[b - 1];
        dH = f64_add(dH, T.rstack_dH[x2][y2][w2]);
        dS = f64_add(dS, T.rstack_dS[x2][y2][w2]);
      }
    } else {
      dS = f64_add(dS, T.loop_dS[gap - 1]);
    }
    a = b;
  }
  dS = f64_add(dS, f64_mul(f64_mul(T.dS_Na, (double)(f1 - f0)), A.Na_eq_log));      // :204-205
  const double tds = __ddiv_rn(f64_mul(f64_add(A.PCR_T, -T.K0), dS), 1000.0);
  return f64_add(f64_add(dG, dH), -tds);
}

// d-th dimer of [full, alternatives...] in the reference's construction order
// (Dimer.alternatives, SecStructures.py:126-152): M split into n maximal runs of
// consecutive indices; d = 0 is the full dimer; then for each left cut l = 0..n-2:
// runs [l, r) for r = n-1 down to l+1, followed by runs [l+1, n).
__device__ __forceinline__ uint64_t dp_nth_dimer(uint64_t M, int n, uint32_t d, const uint8_t* run_start,
                                                 const uint8_t* run_end) {
  if (d == 0) return M;
  int l = 0;
  uint32_t rem = d - 1;
  while (rem >= (uint32_t)(n - l)) { rem -= (uint32_t)(n - l); ++l; }
  int a, b;
  if (rem < (uint32_t)(n - 1 - l)) { a = l; b = n - 2 - (int)rem; }
  else { a = l + 1; b = n - 1; }
  const uint64_t lo_m = ~0ull << run_start[a];
  const uint64_t hi_m = (2ull << run_end[b]) - 1ull;     // run_end = 63 wraps to all ones
  return M & lo_m & hi_m;
}

// Calls f(sub) for every dimer of [full, alternatives...] of match mask M, in that construction
// order, walking the run boundaries with bit tricks (no run table).
template <typename F>
__device__ __forceinline__ void dp_for_each_dimer(uint64_t M, F&& f) {
  f(M);
  const uint64_t S = M & ~(M << 1);                      // first index of every run
  const uint64_t E = M & ~(M >> 1);                      // last index of every run
  const int n = __popcll(S);
  uint64_t Sl = S;
  for (int l = 0; l + 1 < n; ++l) {
    const int start_l = __ffsll((long long)Sl) - 1;
    Sl &= Sl - 1ull;
    const int start_l1 = __ffsll((long long)Sl) - 1;
    const uint64_t lo_m = ~0ull << start_l;
    uint64_t Er = E & ~(1ull << (63 - __clzll((long long)E)));      // ends of runs 0..n-2
    for (int r = n - 1; r > l; --r) {                     // runs [l, r): ends with run r-1
      const int e = 63 - __clzll((long long)Er);
      Er &= ~(1ull << e);
      f(M & lo_m & ((2ull << e) - 1ull));
    }
    f(M & (~0ull << start_l1));                           // runs [l+1, n)
  }
}

// lanes that share one variant of a primer with V variants, and the passes a hit of it needs
__host__ __device__ __forceinline__ uint32_t dp_lanes_per_variant(uint32_t V) {
  uint32_t J = 1u;                                       // 32, 16, 8, 4, 2, 1 for V = 1, 2, 3-4, 5-8, 9-16, more
  while (J < 32u && J * 2u * V <= 32u) J *= 2u;
  return J;
}
__host__ __device__ __forceinline__ uint32_t dp_passes(uint32_t V) {
  const uint32_t per_pass = 32u / dp_lanes_per_variant(V);
  return V ? (V + per_pass - 1u) / per_pass : 0u;
}

// dimers per hit are summed over the hit's passes: start from zero (the hits may be resident ones
// that have been through the duplex stage before, dp_find_again)
__global__ void __launch_bounds__(128)
dp_duplex_clear_kernel(dp_raw_hit* __restrict__ hits, uint64_t hits_cap, const unsigned long long* __restrict__ counters) {
  const uint64_t n = min((uint64_t)(uint32_t)counters[DP_C_HITS], hits_cap);
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    hits[i].reserved = 0u;
}

// One WARP per (hit, pass) (SecStructures.py:294-324 for every Duplex of the site): the window is
// decoded into two 64-bit planes + the complement strand in shared memory.  The 32 lanes are dealt
// over (variant, share of its dimers): J = 32 ... 1 lanes share one variant (V = 1 ... > 16; lane j
// of the group evaluates the dimers d = j mod J of [full, alternatives...]), so the handful of dimers
// of a non-degenerate duplex is evaluated side by side instead of one after the other, and the
// variants of a degenerate primer are spread over passes of 32 variants that different warps take.  A lane builds its variant's planes from the code bytes,
// gets the match mask with four 64-bit logic ops, walks the dimers in the reference's order,
// evaluates its share and parks the survivors in its column of shared memory; the final place
// of a survivor is its rank by (dG, construction index) among the group's survivors (the
// reference's stable sort).  One atomic per pass reserves the output range; a directory entry
// offset << 16 | count per (hit, variant) lets host or device reassemble them in order.
__global__ void __launch_bounds__(DP_DUPLEX_THREADS)
dp_duplex_kernel(const __grid_constant__ dp_duplex_args A, const __grid_constant__ dp_dg_plan plan) {
  const unsigned long long packed = A.counters[DP_C_HITS];
  const uint64_t n_hits = (uint32_t)packed;
  if (n_hits > A.hits_cap) return;                       // the host grows the buffer and rescans
  // work item = (hit, pass): a pass covers 32 / J variants of the hit, J lanes each (dp_lanes_per_variant);
  // every hit gets PP = passes of the most degenerate primer of the call, the surplus items leave at once
  uint32_t v_max = 1;
  for (int p = 0; p < DP_MAX_PRIMERS; ++p) v_max = max(v_max, plan.var_count[p]);
  const uint32_t PP = dp_passes(v_max);
  const uint64_t n_items = n_hits * PP;
  if ((uint64_t)blockIdx.x * DP_DUPLEX_WARPS >= n_items) return;
  extern __shared__ __align__(16) unsigned char dp_duplex_smem[];     // NN tables (10 KB) + one scratch per warp
  dp_nn_tables& sT = *reinterpret_cast<dp_nn_tables*>(dp_duplex_smem);
  dp_duplex_scratch* sW = reinterpret_cast<dp_duplex_scratch*>(dp_duplex_smem + DP_DUPLEX_TABLE_BYTES);
  {
    const uint64_t* src = reinterpret_cast<const uint64_t*>(A.dg.T);
    uint64_t* dst = reinterpret_cast<uint64_t*>(&sT);
    for (uint32_t i = threadIdx.x; i < sizeof(dp_nn_tables) / 8; i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  dp_dg_args G = A.dg;
  G.T = &sT;
  const uint32_t lane = threadIdx.x & 31u;
  dp_duplex_scratch& W = sW[threadIdx.x >> 5];
  unsigned long long n_tried = 0;
  const uint64_t n_warps = (uint64_t)gridDim.x * DP_DUPLEX_WARPS;
  for (uint64_t item = (uint64_t)blockIdx.x * DP_DUPLEX_WARPS + (threadIdx.x >> 5); item < n_items; item += n_warps) {
    const uint64_t hi_ = item / PP;
    const uint32_t pass = (uint32_t)(item % PP);
    const dp_raw_hit hit = A.hits[hi_];
    const int L = plan.pat_len[hit.pattern];
    const int strand = plan.pat_strand[hit.pattern];
    const uint32_t primer = plan.pat_primer[hit.pattern];
    const uint32_t V = plan.var_count[primer];
    if (pass >= dp_passes(V)) continue;                  // uniform over the warp
    const uint8_t* codes = A.codes + plan.var_off[primer];
    const uint32_t J = dp_lanes_per_variant(V);
    const uint32_t per_pass = 32u / J;                   // variants per pass
    const uint32_t j = lane & (J - 1u), first = lane & ~(J - 1u);
    // the window as the reference's strand string shows it: two planes for its ACGT letters,
    // W.r = its complement strand (3'->5') as letter indices (IUPAC letters included)
    uint64_t whi = 0, wlo = 0, wbase = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int i = (int)lane + 32 * half;
      uint32_t w = 15u;
      if (i < L) {
        const uint64_t g = hit.g + (strand ? (uint64_t)(L - 1 - i) : (uint64_t)i);
        w = dp_letter_at(A.lo, A.hi, A.exc, A.runs, A.n_runs, g);
        if (strand) w = dp_letter_complement(w);
        W.r[i] = (uint8_t)dp_letter_complement(w);
      }
      whi |= (uint64_t)__ballot_sync(0xFFFFFFFFu, w < 4u && (w & 2u)) << (32 * half);
      wlo |= (uint64_t)__ballot_sync(0xFFFFFFFFu, w < 4u && (w & 1u)) << (32 * half);
      wbase |= (uint64_t)__ballot_sync(0xFFFFFFFFu, w < 4u) << (32 * half);
    }
    __syncwarp();
    uint32_t hit_dimers = 0;
    {
      const uint32_t v0 = pass * per_pass;
      const uint32_t v = v0 + lane / J;
      uint32_t kept = 0;
      uint64_t M = 0;
      dp_seq_planes sv;
      sv.hi = sv.lo = 0;
      if (v < V) {
        const uint8_t* vc = codes + (uint64_t)v * L;
        uint32_t c = vc[0];
        for (int i = 0; i < L; ++i) {
          sv.hi |= (uint64_t)(c >> 1) << i;
          sv.lo |= (uint64_t)(c & 1u) << i;
          const uint32_t cn = i + 1 < L ? vc[i + 1] : 0u;
          if (i + 1 < L) W.idx[32 * i + lane] = (uint8_t)((c * 4u + cn) * 15u + W.r[i + 1]);
          c = cn;
        }
        M = ~(sv.hi ^ whi) & ~(sv.lo ^ wlo) & wbase;     // plain equality of letters
      }
      // Phase 1 (cheap, lanes diverge): every lane walks [full, alternatives...] of its variant and
      // notes the masks of ITS share (d = j mod J) in its column.  Phase 2 (the expensive part, lanes
      // converged): share t of every lane is evaluated in the same instruction stream -- evaluating
      // inside the walk would serialise the lanes of a group, each at its own dimer index.
      uint32_t mine_n = 0;
      bool overflow = false;                             // the lane's column does not hold the group's dimers / survivors
      if (v < V && M) {
        if (J == 1u) {
          // a lane per variant: every lane is busy anyway, evaluate inside the walk
          dp_for_each_dimer(M, [&](uint64_t sub) {
            const double dG = dp_duplex_dimer_dg(G, sv, W.r, W.idx + lane, L, sub);
            ++n_tried;
            if (dG > G.max_dimer_dG || !(dG <= G.max_dG_for_K)) return;        // SecStructures.py:302-304
            if (kept < DP_LANE_LIST) {
              W.kdg[32 * kept + lane] = dG;
              W.kmask[32 * kept + lane] = sub;
              W.kd[32 * kept + lane] = (uint16_t)kept;    // already in construction order
            }
            ++kept;
          });
          overflow = kept > DP_LANE_LIST;
        } else {
          const uint64_t S = M & ~(M << 1);
          const uint32_t n_runs = (uint32_t)__popcll(S);
          if (n_runs * (n_runs + 1u) / 2u <= DP_LANE_LIST * J) {
            uint32_t d = 0;
            dp_for_each_dimer(M, [&](uint64_t sub) {
              const uint32_t idx_d = d++;
              if ((idx_d & (J - 1u)) == j) W.kmask[32 * (mine_n++) + lane] = sub;
            });
          } else if (j == 0) {
            // too many dimers for the group's columns: its first lane counts the survivors alone
            dp_for_each_dimer(M, [&](uint64_t sub) {
              const double dG = dp_duplex_dimer_dg(G, sv, W.r, W.idx + lane, L, sub);
              ++n_tried;
              if (!(dG > G.max_dimer_dG) && dG <= G.max_dG_for_K) ++kept;
            });
            overflow = true;
          }
        }
      }
      if (J > 1u) {
        uint32_t most = mine_n;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) most = max(most, __shfl_xor_sync(0xFFFFFFFFu, most, o));
        for (uint32_t t = 0; t < most; ++t) {
          if (t < mine_n) {
            const uint64_t sub = W.kmask[32 * t + lane];
            const double dG = dp_duplex_dimer_dg(G, sv, W.r, W.idx + lane, L, sub);
            ++n_tried;
            if (!(dG > G.max_dimer_dG) && dG <= G.max_dG_for_K) {               // SecStructures.py:302-304
              W.kdg[32 * kept + lane] = dG;               // kept <= t: in place
              W.kmask[32 * kept + lane] = sub;
              W.kd[32 * kept + lane] = (uint16_t)(t * J + j);                   // construction index
              ++kept;
            }
          }
        }
      }
      __syncwarp();
      // output range of this pass: one atomic, offsets by prefix sum over the lanes
      uint32_t incl = kept;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t x = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= (uint32_t)o) incl += x;
      }
      const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
      hit_dimers += total;
      unsigned long long base = 0;
      if (total) {
        if (lane == 0) base = atomicAdd(&A.counters[DP_C_DIMERS], (unsigned long long)total);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
      }
      // the group of a variant: its range starts where its first lane's does
      const uint32_t g_excl = __shfl_sync(0xFFFFFFFFu, incl - kept, first);
      const uint32_t g_total = __shfl_sync(0xFFFFFFFFu, incl, first + J - 1u) - g_excl;
      const uint32_t big = __ballot_sync(0xFFFFFFFFu, overflow);
      const uint32_t g_mask = J == 32u ? 0xFFFFFFFFu : ((1u << J) - 1u) << first;
      const uint64_t off = base + g_excl;
      const bool fits = off + g_total <= A.out_cap;
      W.kn[lane] = (uint16_t)min(kept, (uint32_t)DP_LANE_LIST);
      __syncwarp();
      if (v < V && g_total && fits) {
        dp_raw_dimer* dst = A.out + off;
        if (!(big & g_mask)) {
          // place = rank by (dG, construction index) among the survivors of the group
          for (uint32_t e = 0; e < kept; ++e) {
            const double x = W.kdg[32 * e + lane];
            const uint32_t xd = W.kd[32 * e + lane];
            uint32_t rank = 0;
#pragma unroll 1
            for (uint32_t l = 0; l < J; ++l) {
              const uint32_t n_l = W.kn[first + l];
              for (uint32_t f = 0; f < n_l; ++f) {
                const double y = W.kdg[32 * f + first + l];
                rank += (y < x || (y == x && W.kd[32 * f + first + l] < xd)) ? 1u : 0u;
              }
            }
            dst[rank].match_mask = W.kmask[32 * e + lane];
            dst[rank].dG = x;
          }
        } else if (j == 0) {
          // more survivors than a lane's list holds (>= 5 mismatch runs): the group's first lane
          // enumerates again and inserts straight into the reserved range (stable)
          uint32_t m = 0;
          dp_for_each_dimer(M, [&](uint64_t sub) {
            const double dG = dp_duplex_dimer_dg(G, sv, W.r, W.idx + lane, L, sub);
            if (dG > G.max_dimer_dG || !(dG <= G.max_dG_for_K)) return;
            uint32_t pos = m;
            while (pos > 0 && dst[pos - 1].dG > dG) { dst[pos] = dst[pos - 1]; --pos; }
            dst[pos].match_mask = sub;
            dst[pos].dG = dG;
            ++m;
          });
        }
      }
      if (v < V && j == 0) {
        const uint64_t pair = (uint64_t)hit.pair_base + v;
        if (pair < A.dir_cap) A.dir[pair] = g_total ? (off << 16) | (uint64_t)g_total : 0ull;
      }
      __syncwarp();
    }
    if (lane == 0 && hit_dimers) atomicAdd(&A.hits[hi_].reserved, hit_dimers);    // cleared by dp_duplex_clear_kernel
    __syncwarp();                                        // W.r is rewritten for the next item
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_tried += __shfl_xor_sync(0xFFFFFFFFu, n_tried, o);
  if (lane == 0 && n_tried) atomicAdd(&A.counters[DP_C_TRIED], n_tried);
}

// ---- large results: final dp_dimer records are laid out by the device -----------------
// The duplex kernel leaves the dimers per hit slot in the hit records; after the host has
// sorted the hits and turned the counts into output offsets, every pair's dimers are copied to
// their final place with (hit rank, variant) filled in.  One warp per hit slot, lanes over variants.
struct dp_slot_info {
  uint64_t final_off;   // first output record of this hit
  uint32_t rank;        // index of the hit in the sorted hit list
  uint32_t count;       // dimers of the hit
};

__global__ void __launch_bounds__(128)
dp_dimer_gather_kernel(const dp_raw_hit* __restrict__ hits, uint64_t n_hits, const uint64_t* __restrict__ dir,
                       const dp_raw_dimer* __restrict__ raw, const __grid_constant__ dp_dg_plan plan,
                       const dp_slot_info* __restrict__ info, dp_dimer* __restrict__ out) {
  const uint64_t slot = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31u;
  if (slot >= n_hits) return;
  const dp_raw_hit h = hits[slot];
  const uint32_t V = plan.var_count[plan.pat_primer[h.pattern]];
  const dp_slot_info si = info[slot];
  uint64_t running = si.final_off;
  for (uint32_t v0 = 0; v0 < V; v0 += 32) {
    const uint32_t v = v0 + lane;
    const uint64_t e = v < V ? dir[(uint64_t)h.pair_base + v] : 0ull;
    const uint32_t c = (uint32_t)(e & 0xFFFFu);
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t x = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= (uint32_t)o) incl += x;
    }
    const uint64_t dest = running + incl - c;
    const uint64_t src = e >> 16;
    for (uint32_t k = 0; k < c; ++k) {
      dp_dimer d;
      d.hit = si.rank;
      d.variant = v;
      d.match_mask = raw[src + k].match_mask;
      d.dG = raw[src + k].dG;
      out[dest + k] = d;
    }
    running += __shfl_sync(0xFFFFFFFFu, incl, 31);
  }
}

int dp_dg_gather(dp_ctx* ctx, const dp_dg_plan& plan, uint64_t n_hits, const void* d_info, dp_dimer* d_out) {
  if (!n_hits) return DP_OK;
  const uint64_t threads = n_hits * 32;
  dp_dimer_gather_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, ctx->stream>>>(
      ctx->d_hits, n_hits, ctx->d_dir, ctx->d_dimers, plan, (const dp_slot_info*)d_info, d_out);
  DP_CUDA(ctx, cudaGetLastError());
  return DP_OK;
}

int dp_dg_enqueue(dp_ctx* ctx, const dp_template* t, const dp_dg_launch& L, uint64_t n_pairs_hint, uint32_t* launches,
                  bool hits_fresh) {
  dp_duplex_args A;
  A.dg.T = ctx->d_tables;
  A.dg.PCR_T = L.PCR_T;
  A.dg.Na_eq_log = L.Na_eq_log;
  A.dg.max_dimer_dG = L.max_dimer_dG;
  A.dg.max_dG_for_K = L.max_dG_for_K;
  A.lo = t->d_lo; A.hi = t->d_hi; A.exc = t->d_exc;
  A.runs = t->d_runs;
  A.n_runs = t->runs.size();
  A.hits = ctx->d_hits;
  A.hits_cap = ctx->hits_cap;
  A.codes = ctx->d_codes;
  A.dir = ctx->d_dir;
  A.dir_cap = ctx->dir_cap;
  A.out = ctx->d_dimers;
  A.out_cap = ctx->dimers_cap;
  A.counters = ctx->d_counters;
  // pair count unknown on the host (fused call): a persistent grid, most blocks
  // leave at once when there are few pairs
  static bool configured[16] = {false};                    // per device: opt in to > 48 KB of shared memory
  if (!configured[ctx->device & 15]) {
    DP_CUDA(ctx, cudaFuncSetAttribute(dp_duplex_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DP_DUPLEX_SMEM));
    configured[ctx->device & 15] = true;
  }
  uint64_t blocks = (uint64_t)ctx->sm_count * DP_DUPLEX_BLOCKS_PER_SM;   // 56 KB of shared memory per block
  if (n_pairs_hint) blocks = std::min<uint64_t>(blocks, (n_pairs_hint + DP_DUPLEX_WARPS - 1) / DP_DUPLEX_WARPS);
  if (!hits_fresh) dp_duplex_clear_kernel<<<ctx->sm_count, 128, 0, ctx->stream>>>(A.hits, A.hits_cap, A.counters);
  dp_duplex_kernel<<<(unsigned)blocks, DP_DUPLEX_THREADS, DP_DUPLEX_SMEM, ctx->stream>>>(A, L.plan);
  DP_CUDA(ctx, cudaGetLastError());
  *launches += hits_fresh ? 1 : 2;                         // [clear +] duplex
  return DP_OK;
}

// ---- per-variant Tm — TD_Functions.py:60-101 -------------------------------------
__global__ void dp_variant_tm_kernel(const dp_nn_tables* __restrict__ Tp, const uint8_t* __restrict__ codes,
                                     uint32_t n, uint32_t L, double conc, double cdeg, double DNA, double DMSO,
                                     double Na_eq_log, double* __restrict__ tm) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const dp_nn_tables& T = *Tp;
  const uint8_t* s = codes + (uint64_t)i * L;
  const double P = conc, D = DNA;
  const double DUP = f64_mul(cdeg, P < D ? P : D);
  const double K = __ddiv_rn(DUP, f64_mul(f64_add(P, -DUP), f64_add(D, -DUP)));
  double dH = T.ini_dH, dS = T.ini_dS;
  if (s[0] == 0 || s[0] == 2) { dH = f64_add(dH, T.ter_dH); dS = f64_add(dS, T.ter_dS); }
  if (s[L - 1] == 0 || s[L - 1] == 2) { dH = f64_add(dH, T.ter_dH); dS = f64_add(dS, T.ter_dS); }
  for (uint32_t k = 0; k + 1 < L; ++k) {
    uint32_t x = s[k], y = s[k + 1];
    uint32_t w = dp_letter_complement(y);           // perfect duplex
    dH = f64_add(dH, T.stack_dH[x][y][w]);
    dS = f64_add(dS, T.stack_dS[x][y][w]);
  }
  dS = f64_add(dS, f64_mul(f64_mul(T.dS_Na, (double)L), Na_eq_log));
  const double denom = f64_add(dS, -f64_mul(T.R, log(K)));
  const double t = f64_add(f64_add(T.K0, __ddiv_rn(f64_mul(dH, 1000.0), denom)), -f64_mul(T.T_DMSO, DMSO));
  tm[i] = t;
}

__global__ void dp_dimer_dg_kernel(dp_dg_args A, uint32_t n, const uint8_t* __restrict__ s1,
                                   const uint64_t* __restrict__ off1, const uint8_t* __restrict__ len1,
                                   const uint8_t* __restrict__ s2, const uint64_t* __restrict__ off2,
                                   const uint8_t* __restrict__ len2, const uint64_t* __restrict__ mask,
                                   const int32_t* __restrict__ offset, double* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint8_t s[DP_WIN], r[DP_WIN];
  int sl = len1[i], rl = len2[i];
  for (int k = 0; k < sl; ++k) s[k] = s1[off1[i] + k];
  for (int k = 0; k < rl; ++k) r[k] = s2[off2[i] + rl - 1 - k];   // reversed: 3'->5'
  out[i] = dp_dimer_dg_device(A, s, sl, r, rl, mask[i], offset[i]);
}

#define DP_FD_WARPS 4                    // warps per block of the primer-dimer / hairpin kernels
#define DP_FD_THREADS (32 * DP_FD_WARPS)
// ---- self / cross dimers of primer pairs -------------------------------
// SecStructures._find_dimers, DegenPrimer/SecStructures.py:609-668: slide seq2 (as its reverse
// complement) along seq1; at every relative position ("front", fwd_len+rev_len-1 of them) the
// matching letters form a Dimer; of [dimer] + dimer.alternatives the one with the lowest
// dimer_dG_corrected below max_dimer_dG is kept (first one wins a tie).  One WARP per
// (pair, front): lanes compare letters (ballot -> match mask), build the run table, then lane d
// evaluates the d-th candidate; a shuffle reduction picks the minimum (dG, d).
struct dp_fd_scratch {
  uint8_t s[DP_WIN];
  uint8_t r[DP_WIN];
  uint8_t run_start[32];
  uint8_t run_end[32];
};

__global__ void __launch_bounds__(DP_FD_THREADS)
dp_find_dimers_kernel(dp_dg_args A, const uint8_t* __restrict__ codes, const uint64_t* __restrict__ seq_off,
                      const uint8_t* __restrict__ seq_len, const uint32_t* __restrict__ pair_a,
                      const uint32_t* __restrict__ pair_b, uint32_t n_pairs, uint32_t stride,
                      dp_front_dimer* __restrict__ out) {
  __shared__ dp_nn_tables sT;
  __shared__ dp_fd_scratch sW[DP_FD_WARPS];
  {
    const uint64_t* src = reinterpret_cast<const uint64_t*>(A.T);
    uint64_t* dst = reinterpret_cast<uint64_t*>(&sT);
    for (uint32_t i = threadIdx.x; i < sizeof(dp_nn_tables) / 8; i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  dp_dg_args G = A;
  G.T = &sT;
  const uint32_t lane = threadIdx.x & 31u;
  dp_fd_scratch& W = sW[threadIdx.x >> 5];
  const uint64_t n_items = (uint64_t)n_pairs * stride;
  const uint64_t n_warps = (uint64_t)gridDim.x * DP_FD_WARPS;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  for (uint64_t item = (uint64_t)blockIdx.x * DP_FD_WARPS + (threadIdx.x >> 5); item < n_items; item += n_warps) {
    const uint32_t pair = (uint32_t)(item / stride);
    const int front = (int)(item % stride);
    const uint32_t ia = pair_a[pair], ib = pair_b[pair];
    const int la = seq_len[ia], lb = seq_len[ib];
    dp_front_dimer res;
    res.full_mask = 0; res.best_mask = 0; res.best_dG = 0.0;
    if (front >= la + lb - 1) {                          // padding of the dense [pair][front] layout
      if (lane == 0) out[item] = res;
      continue;
    }
    __syncwarp();                                        // previous item's readers are done
    const uint8_t* ca = codes + seq_off[ia];
    const uint8_t* cb = codes + seq_off[ib];
    for (int i = (int)lane; i < la; i += 32) W.s[i] = ca[i];
    for (int i = (int)lane; i < lb; i += 32) W.r[i] = cb[lb - 1 - i];      // seq2 reversed: 3'->5'
    __syncwarp();
    // margins — SecStructures.py:620-636 (fwd_right may exceed fwd_len; zip() cuts at the shorter range)
    const int fwd_left = la - front - 1 > 0 ? la - front - 1 : 0;
    int fwd_right = la;
    if (front >= lb) fwd_right -= front - lb - 1;
    const int rev_left = front >= la ? front - la + 1 : 0;
    const int rev_right = front >= lb ? lb : front + 1;
    const int cnt = min(fwd_right - fwd_left, rev_right - rev_left);
    const int off = fwd_left - rev_left;
    uint64_t M = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int i = (int)lane + 32 * half;
      bool match = false;
      if (i < cnt) match = W.s[fwd_left + i] == (uint8_t)dp_letter_complement(W.r[rev_left + i]);
      const uint64_t b = (uint64_t)__ballot_sync(0xFFFFFFFFu, match) << (32 * half);
      M |= b;
    }
    M <<= fwd_left;                                      // bit = seq1 index (fwd_left + cnt <= 64)
    if (!M) {
      if (lane == 0) out[item] = res;
      continue;
    }
    const uint64_t S = M & ~(M << 1);
    const uint64_t Eb = M & ~(M >> 1);
    const int n = __popcll(S);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int i = (int)lane + 32 * half;
      const uint64_t below = (1ull << i) - 1ull;
      if ((S >> i) & 1ull) W.run_start[__popcll(S & below)] = (uint8_t)i;
      if ((Eb >> i) & 1ull) W.run_end[__popcll(Eb & below)] = (uint8_t)i;
    }
    __syncwarp();
    const uint32_t n_dimers = (uint32_t)(n * (n + 1) / 2);
    double best = G.max_dimer_dG;                        // `if min_dG > dimer.dG` — strict
    uint32_t best_d = 0xFFFFFFFFu;
    bool bad = false;
    for (uint32_t d = lane; d < n_dimers; d += 32) {
      const uint64_t sub = dp_nth_dimer(M, n, d, W.run_start, W.run_end);
      const double dG = dp_dimer_dg_device(G, W.s, la, W.r, lb, sub, off);
      if (dG != dG) bad = true;
      if (best > dG) { best = dG; best_d = d; }
    }
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) {
      const double ob = __shfl_xor_sync(0xFFFFFFFFu, best, sh);
      const uint32_t od = __shfl_xor_sync(0xFFFFFFFFu, best_d, sh);
      if (ob < best || (ob == best && od < best_d)) { best = ob; best_d = od; }
    }
    bad = __any_sync(0xFFFFFFFFu, bad);
    if (lane == 0) {
      res.full_mask = M;
      if (bad) { res.best_mask = 0; res.best_dG = nan; }        // the reference raises (unknown NN key)
      else if (best_d != 0xFFFFFFFFu) {
        res.best_mask = dp_nth_dimer(M, n, best_d, W.run_start, W.run_end);
        res.best_dG = best;
      }
      out[item] = res;
    }
  }
}

// ---- hairpins of one primer ---------------------------------------------
// hairpin_dG_corrected, DegenPrimer/TD_Functions.py:209-276, with UnifiedNN.hairpin_loop_dG/dH/dS_37
// (DegenPrimer/UnifiedNN.py:279-325).  s: the sequence (letter codes, 5'->3'), mask: stem positions
// of the 5' arm, off: Dimer.offset of the self-dimer the hairpin folds from — position f pairs with
// L-1-(f-off).  NaN where the reference would raise.
__device__ double dp_hairpin_dg_device(const dp_dg_args& A, const dp_hairpin_tables* __restrict__ H,
                                       const uint8_t* s, int L, uint64_t mask, int off) {
  const dp_nn_tables& T = *A.T;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  if (!mask) return nan;
  const int f0 = __ffsll((long long)mask) - 1;
  const int f1 = 63 - __clzll((long long)mask);
  const int r0 = L - 1 - (f0 - off);
  const int r1 = L - 1 - (f1 - off);
  if (r0 < 0 || r0 >= L || r1 <= f1) return nan;
  double dG = 0.0, dH = T.ini_dH, dS = T.ini_dS;
  bool bad = false;
  // open end of the stem — TD_Functions.py:227-244
  if (f0 == 0 && r0 < L - 1) {
    uint32_t x = s[r0], y = s[r0 + 1], z = s[0];
    if (x > 3 || y > 3 || z > 3) bad = true;
    else { dH = f64_add(dH, T.dangling_dH[1][x][y][z]); dS = f64_add(dS, T.dangling_dS[1][x][y][z]); }
  } else if (f0 > 0 && r0 == L - 1) {
    uint32_t x = s[f0 - 1], y = s[f0], z = s[L - 1];
    if (x > 3 || y > 3 || z > 3) bad = true;
    else { dH = f64_add(dH, T.dangling_dH[0][x][y][z]); dS = f64_add(dS, T.dangling_dS[0][x][y][z]); }
  } else if (f0 > 0 && r0 < L - 1) {
    dG = f64_add(dG, T.terminal_mismatch);
  } else {
    if (s[0] == 0 || s[0] == 2) { dH = f64_add(dH, T.ter_dH); dS = f64_add(dS, T.ter_dS); }
  }
  // stacks, |x|, |xx| and internal loops — TD_Functions.py:246-267
  uint64_t m = mask & (mask - 1);
  int a = f0;
  while (m) {
    const int b = __ffsll((long long)m) - 1;
    m &= m - 1;
    const int gap = b - a;
    const int ra = L - 1 - (a - off), rb = L - 1 - (b - off);
    if (gap < 4) {
      uint32_t x = s[a], y = s[a + 1], z = s[ra], w = s[ra - 1];
      if (x > 3 || y > 3 || w > 14 || z != dp_letter_complement(x)) bad = true;
      else { dH = f64_add(dH, T.stack_dH[x][y][w]); dS = f64_add(dS, T.stack_dS[x][y][w]); }
      if (gap > 1) {
        uint32_t x2 = s[b - 1], y2 = s[b], z2 = s[rb], w2 = s[rb + 1];
        if (x2 > 3 || y2 > 3 || w2 > 14 || z2 != dp_letter_complement(y2)) bad = true;
        else { dH = f64_add(dH, T.rstack_dH[x2][y2][w2]); dS = f64_add(dS, T.rstack_dS[x2][y2][w2]); }
      }
    } else {
      dS = f64_add(dS, T.loop_dS[gap - 1]);
    }
    a = b;
  }
  // the loop: hp = s[f1 .. r1], n unpaired bases — UnifiedNN.py:279-325
  const int n = r1 - f1 - 1;
  double loop_dH = 0.0, loop_dG = H->loop_dG[n < 63 ? n : 63];
  double sp_dH = nan, sp_dG = nan;                  // special tri-/tetra-loop row, if any
  if (n == 3 || n == 4) {
    uint32_t idx = 0;
    bool plain = true;
    for (int i = f1; i <= r1; ++i) { plain &= s[i] <= 3; idx = idx * 4 + (s[i] & 3u); }
    if (plain) {
      sp_dH = n == 3 ? H->tri_dH[idx] : H->tetra_dH[idx];
      sp_dG = n == 3 ? H->tri_dG[idx] : H->tetra_dG[idx];
    }
  }
  if (n == 3) {
    if (sp_dG == sp_dG) loop_dG = f64_add(loop_dG, sp_dG);
    if (s[f1] == 0 || s[f1] == 2) loop_dG = f64_add(loop_dG, 0.5);      // AT-closing penalty
  } else if (n == 4) {
    if (sp_dG == sp_dG) loop_dG = f64_add(loop_dG, sp_dG);
    loop_dG = f64_add(loop_dG, T.terminal_mismatch);
  } else {
    loop_dG = f64_add(loop_dG, T.terminal_mismatch);
  }
  if (sp_dH == sp_dH) loop_dH = sp_dH;
  dH = f64_add(dH, loop_dH);
  dS = f64_add(dS, f64_mul(f64_add(loop_dG, -loop_dH), H->iK37));
  dS = f64_add(dS, f64_mul(f64_mul(T.dS_Na, (double)(f1 - f0)), A.Na_eq_log));
  const double tds = __ddiv_rn(f64_mul(f64_add(A.PCR_T, -T.K0), dS), 1000.0);
  const double res = f64_add(f64_add(dG, dH), -tds);
  return bad ? nan : res;
}

// SecStructures._find_hairpins, DegenPrimer/SecStructures.py:676-716, the part after the seed has been
// trimmed (a few set operations, done by the caller): of [stem] + stem.alternatives the hairpin with
// the lowest dG strictly below max_hairpin_dG.  One WARP per seed, lane d evaluates candidate d.
__global__ void __launch_bounds__(DP_FD_THREADS)
dp_find_hairpins_kernel(dp_dg_args A, const dp_hairpin_tables* __restrict__ H, double max_hairpin_dG,
                        const uint8_t* __restrict__ codes, const uint64_t* __restrict__ seq_off,
                        const uint8_t* __restrict__ seq_len, const uint32_t* __restrict__ seed_seq,
                        const uint64_t* __restrict__ seed_mask, const int32_t* __restrict__ seed_offset,
                        uint32_t n_seeds, dp_best_structure* __restrict__ out) {
  __shared__ dp_nn_tables sT;
  __shared__ dp_fd_scratch sW[DP_FD_WARPS];
  {
    const uint64_t* src = reinterpret_cast<const uint64_t*>(A.T);
    uint64_t* dst = reinterpret_cast<uint64_t*>(&sT);
    for (uint32_t i = threadIdx.x; i < sizeof(dp_nn_tables) / 8; i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  dp_dg_args G = A;
  G.T = &sT;
  const uint32_t lane = threadIdx.x & 31u;
  dp_fd_scratch& W = sW[threadIdx.x >> 5];
  const uint32_t n_warps = gridDim.x * DP_FD_WARPS;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  for (uint32_t seed = blockIdx.x * DP_FD_WARPS + (threadIdx.x >> 5); seed < n_seeds; seed += n_warps) {
    const uint32_t si = seed_seq[seed];
    const int L = seq_len[si];
    const uint64_t M = seed_mask[seed];
    const int off = seed_offset[seed];
    dp_best_structure res;
    res.best_mask = 0; res.best_dG = 0.0;
    __syncwarp();
    if (!M) {
      if (lane == 0) out[seed] = res;
      continue;
    }
    const uint8_t* cs = codes + seq_off[si];
    for (int i = (int)lane; i < L; i += 32) W.s[i] = cs[i];
    const uint64_t S = M & ~(M << 1);
    const uint64_t Eb = M & ~(M >> 1);
    const int n = __popcll(S);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int i = (int)lane + 32 * half;
      const uint64_t below = (1ull << i) - 1ull;
      if ((S >> i) & 1ull) W.run_start[__popcll(S & below)] = (uint8_t)i;
      if ((Eb >> i) & 1ull) W.run_end[__popcll(Eb & below)] = (uint8_t)i;
    }
    __syncwarp();
    const uint32_t n_cand = (uint32_t)(n * (n + 1) / 2);
    double best = max_hairpin_dG;
    uint32_t best_d = 0xFFFFFFFFu;
    bool bad = false;
    for (uint32_t d = lane; d < n_cand; d += 32) {
      const uint64_t sub = dp_nth_dimer(M, n, d, W.run_start, W.run_end);
      const double dG = dp_hairpin_dg_device(G, H, W.s, L, sub, off);
      if (dG != dG) bad = true;
      if (best > dG) { best = dG; best_d = d; }
    }
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) {
      const double ob = __shfl_xor_sync(0xFFFFFFFFu, best, sh);
      const uint32_t od = __shfl_xor_sync(0xFFFFFFFFu, best_d, sh);
      if (ob < best || (ob == best && od < best_d)) { best = ob; best_d = od; }
    }
    bad = __any_sync(0xFFFFFFFFu, bad);
    if (lane == 0) {
      if (bad) res.best_dG = nan;
      else if (best_d != 0xFFFFFFFFu) {
        res.best_mask = dp_nth_dimer(M, n, best_d, W.run_start, W.run_end);
        res.best_dG = best;
      }
      out[seed] = res;
    }
  }
}

// ---- host ----------------------------------------------------------------------------
// scratch from the context's stream-ordered pool (freed in stream order, no sync)
template <typename T>
struct dev_buf {
  T* p = nullptr;
  dp_ctx* owner = nullptr;
  ~dev_buf() { if (p) dp_pool_free(owner, p); }
  cudaError_t alloc(dp_ctx* ctx, size_t n) {
    owner = ctx;
    return dp_pool_alloc(ctx, (void**)&p, std::max<size_t>(1, n) * sizeof(T));
  }
};

static int letters_of(dp_ctx* ctx, const char* src, size_t n, std::vector<uint8_t>& out, bool bases_only) {
  out.resize(n);
  for (size_t i = 0; i < n; ++i) {
    uint32_t l = dp_letter_of((unsigned char)src[i]);
    if (l == DP_NOT_A_LETTER || (bases_only && l > 3))
      return dp_set_error(ctx, DP_ERR_ALPHABET, "byte %zu ('%c') is not %s", i, src[i],
                          bases_only ? "an unambiguous base" : "an IUPAC DNA letter");
    out[i] = (uint8_t)l;
  }
  return DP_OK;
}

static dp_dg_args make_args(dp_ctx* ctx, const dp_conditions* c) {
  dp_dg_args A;
  A.T = ctx->d_tables;
  A.PCR_T = c->PCR_T;
  A.Na_eq_log = c->Na_eq_log;
  A.max_dimer_dG = c->max_dimer_dG;
  A.max_dG_for_K = c->max_dG_for_K;
  return A;
}

extern "C" int dp_hit_windows(dp_ctx* ctx, const dp_template* t, const dp_hit* hits, uint64_t n_hits,
                              uint32_t n_primers, const uint8_t* primer_len, char* out) {
  if (!ctx || !t || (n_hits && (!hits || !out)) || !primer_len) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (n_hits == 0) return DP_OK;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  std::vector<uint64_t> hit_g(n_hits);
  std::vector<uint8_t> hit_len(n_hits), hit_strand(n_hits);
  for (uint64_t h = 0; h < n_hits; ++h) {
    const dp_hit& x = hits[h];
    if (x.primer >= n_primers || x.record >= t->records.size())
      return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: primer or record out of range", (unsigned long long)h);
    const dp_record& rec = t->records[x.record];
    uint32_t L = primer_len[x.primer];
    if (L == 0 || L > DP_MAX_PRIMER_LEN || x.pos + L > rec.len)
      return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: window outside its record", (unsigned long long)h);
    hit_g[h] = rec.start + (x.strand ? rec.len - L - x.pos : x.pos);
    hit_len[h] = (uint8_t)L;
    hit_strand[h] = x.strand ? 1 : 0;
  }
  dev_buf<uint64_t> d_g;
  dev_buf<uint8_t> d_len, d_strand, d_win;
  DP_CUDA(ctx, d_g.alloc(ctx, n_hits));
  DP_CUDA(ctx, d_len.alloc(ctx, n_hits));
  DP_CUDA(ctx, d_strand.alloc(ctx, n_hits));
  DP_CUDA(ctx, d_win.alloc(ctx, n_hits * DP_WIN));
  cudaStream_t st = ctx->stream;
  DP_CUDA(ctx, cudaMemcpyAsync(d_g.p, hit_g.data(), n_hits * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_len.p, hit_len.data(), n_hits, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_strand.p, hit_strand.data(), n_hits, cudaMemcpyHostToDevice, st));
  uint64_t n = n_hits * DP_WIN;
  dp_dg_windows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
      t->d_lo, t->d_hi, t->d_exc, t->d_runs, t->runs.size(), d_g.p, d_len.p, d_strand.p, n_hits, d_win.p);
  DP_CUDA(ctx, cudaGetLastError());
  DP_CUDA(ctx, cudaMemcpyAsync(out, d_win.p, n, cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaStreamSynchronize(st));
  for (uint64_t i = 0; i < n; ++i) {
    unsigned char l = (unsigned char)out[i];
    out[i] = l < 15 ? DP_LETTERS[l] : 0;
  }
  return DP_OK;
}

extern "C" int dp_variant_tm(dp_ctx* ctx, uint32_t n, uint32_t len, const char* seqs, double concentration,
                             double conversion_degree, const dp_conditions* cond, double* tm) {
  if (!ctx || !seqs || !cond || !tm) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_tables_load() has not been called");
  if (n == 0) return DP_OK;
  if (len == 0) return dp_set_error(ctx, DP_ERR_LENGTH, "empty sequence");
  if (!(conversion_degree > 0.0 && conversion_degree < 1.0))
    return dp_set_error(ctx, DP_ERR_ARG, "equilibrium ratio should be in the (0;1) interval");
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  std::vector<uint8_t> codes;
  int rc = letters_of(ctx, seqs, (size_t)n * len, codes, true);
  if (rc) return rc;
  dev_buf<uint8_t> d_codes;
  dev_buf<double> d_tm;
  DP_CUDA(ctx, d_codes.alloc(ctx, codes.size()));
  DP_CUDA(ctx, d_tm.alloc(ctx, n));
  DP_CUDA(ctx, cudaMemcpyAsync(d_codes.p, codes.data(), codes.size(), cudaMemcpyHostToDevice, ctx->stream));
  dp_variant_tm_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_tables, d_codes.p, n, len, concentration, conversion_degree,
                                                                  cond->DNA, cond->DMSO, cond->Na_eq_log, d_tm.p);
  DP_CUDA(ctx, cudaGetLastError());
  DP_CUDA(ctx, cudaMemcpyAsync(tm, d_tm.p, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return DP_OK;
}

extern "C" int dp_dimer_dg(dp_ctx* ctx, uint32_t n, const char* s1, const uint64_t* off1, const uint8_t* len1,
                           const char* s2, const uint64_t* off2, const uint8_t* len2, const uint64_t* match_mask,
                           const int32_t* offset, const dp_conditions* cond, double* dG) {
  if (!ctx || !s1 || !off1 || !len1 || !s2 || !off2 || !len2 || !match_mask || !offset || !cond || !dG)
    return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_tables_load() has not been called");
  if (n == 0) return DP_OK;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t b1 = 0, b2 = 0;
  for (uint32_t i = 0; i < n; ++i) {
    if (len1[i] == 0 || len1[i] > DP_WIN || len2[i] == 0 || len2[i] > DP_WIN)
      return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "dimer %u: sequence length outside [1,%d]", i, DP_WIN);
    b1 = std::max<uint64_t>(b1, off1[i] + len1[i]);
    b2 = std::max<uint64_t>(b2, off2[i] + len2[i]);
  }
  std::vector<uint8_t> c1, c2;
  int rc = letters_of(ctx, s1, b1, c1, false);
  if (rc) return rc;
  rc = letters_of(ctx, s2, b2, c2, false);
  if (rc) return rc;
  dev_buf<uint8_t> d1, d2, dl1, dl2;
  dev_buf<uint64_t> do1, do2, dm;
  dev_buf<int32_t> doff;
  dev_buf<double> dout;
  DP_CUDA(ctx, d1.alloc(ctx, b1)); DP_CUDA(ctx, d2.alloc(ctx, b2));
  DP_CUDA(ctx, dl1.alloc(ctx, n)); DP_CUDA(ctx, dl2.alloc(ctx, n));
  DP_CUDA(ctx, do1.alloc(ctx, n)); DP_CUDA(ctx, do2.alloc(ctx, n));
  DP_CUDA(ctx, dm.alloc(ctx, n)); DP_CUDA(ctx, doff.alloc(ctx, n)); DP_CUDA(ctx, dout.alloc(ctx, n));
  cudaStream_t st = ctx->stream;
  DP_CUDA(ctx, cudaMemcpyAsync(d1.p, c1.data(), b1, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d2.p, c2.data(), b2, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(dl1.p, len1, n, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(dl2.p, len2, n, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(do1.p, off1, n * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(do2.p, off2, n * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(dm.p, match_mask, n * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(doff.p, offset, n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  dp_dg_args A = make_args(ctx, cond);
  dp_dimer_dg_kernel<<<(n + 127) / 128, 128, 0, st>>>(A, n, d1.p, do1.p, dl1.p, d2.p, do2.p, dl2.p, dm.p, doff.p, dout.p);
  DP_CUDA(ctx, cudaGetLastError());
  DP_CUDA(ctx, cudaMemcpyAsync(dG, dout.p, n * sizeof(double), cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaStreamSynchronize(st));
  return DP_OK;
}

extern "C" int dp_find_dimers(dp_ctx* ctx, uint32_t n_seqs, const char* seqs, const uint64_t* off, const uint8_t* len,
                              uint32_t n_pairs, const uint32_t* pair_a, const uint32_t* pair_b,
                              const dp_conditions* cond, uint32_t stride, dp_front_dimer* out, float* ms_kernel) {
  if (!ctx || !seqs || !off || !len || !cond || (n_pairs && (!pair_a || !pair_b || !out)))
    return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_tables_load() has not been called");
  if (ms_kernel) *ms_kernel = 0.f;
  if (n_pairs == 0) return DP_OK;
  uint64_t bytes = 0;
  for (uint32_t i = 0; i < n_seqs; ++i) {
    if (len[i] == 0 || len[i] > DP_WIN)
      return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "sequence %u: length outside [1,%d]", i, DP_WIN);
    bytes = std::max<uint64_t>(bytes, off[i] + len[i]);
  }
  uint32_t need = 0;
  for (uint32_t p = 0; p < n_pairs; ++p) {
    if (pair_a[p] >= n_seqs || pair_b[p] >= n_seqs)
      return dp_set_error(ctx, DP_ERR_ARG, "pair %u: sequence index out of range", p);
    need = std::max<uint32_t>(need, (uint32_t)len[pair_a[p]] + len[pair_b[p]] - 1);
  }
  if (stride < need) return dp_set_error(ctx, DP_ERR_ARG, "stride %u is smaller than the %u fronts of the longest pair", stride, need);
  if ((uint64_t)n_pairs * stride > (1ull << 31))
    return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "more than 2^31 (pair, front) items in one call");
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  std::vector<uint8_t> codes;
  int rc = letters_of(ctx, seqs, bytes, codes, false);
  if (rc) return rc;
  const uint64_t n_items = (uint64_t)n_pairs * stride;
  dev_buf<uint8_t> d_codes, d_len;
  dev_buf<uint64_t> d_off;
  dev_buf<uint32_t> d_a, d_b;
  dev_buf<dp_front_dimer> d_out;
  DP_CUDA(ctx, d_codes.alloc(ctx, bytes)); DP_CUDA(ctx, d_len.alloc(ctx, n_seqs)); DP_CUDA(ctx, d_off.alloc(ctx, n_seqs));
  DP_CUDA(ctx, d_a.alloc(ctx, n_pairs)); DP_CUDA(ctx, d_b.alloc(ctx, n_pairs)); DP_CUDA(ctx, d_out.alloc(ctx, n_items));
  cudaStream_t st = ctx->stream;
  DP_CUDA(ctx, cudaMemcpyAsync(d_codes.p, codes.data(), bytes, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_len.p, len, n_seqs, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_off.p, off, n_seqs * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_a.p, pair_a, n_pairs * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_b.p, pair_b, n_pairs * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (ms_kernel) {
    DP_CUDA(ctx, cudaEventCreate(&e0));
    DP_CUDA(ctx, cudaEventCreate(&e1));
    DP_CUDA(ctx, cudaEventRecord(e0, st));
  }
  // one warp per (pair, front), grid-stride; 16 resident blocks of 4 warps per SM
  const uint64_t blocks_needed = (n_items + DP_FD_WARPS - 1) / DP_FD_WARPS;
  const unsigned grid = (unsigned)std::min<uint64_t>(blocks_needed, (uint64_t)ctx->sm_count * 16);
  dp_dg_args A = make_args(ctx, cond);
  dp_find_dimers_kernel<<<grid, DP_FD_THREADS, 0, st>>>(A, d_codes.p, d_off.p, d_len.p, d_a.p, d_b.p, n_pairs, stride, d_out.p);
  cudaError_t le = cudaGetLastError();
  if (ms_kernel && le == cudaSuccess) le = cudaEventRecord(e1, st);
  if (le == cudaSuccess) le = cudaMemcpyAsync(out, d_out.p, n_items * sizeof(dp_front_dimer), cudaMemcpyDeviceToHost, st);
  if (le == cudaSuccess) le = cudaStreamSynchronize(st);
  if (ms_kernel && le == cudaSuccess) le = cudaEventElapsedTime(ms_kernel, e0, e1);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  DP_CUDA(ctx, le);
  return DP_OK;
}

extern "C" int dp_hairpin_tables_load(dp_ctx* ctx, const dp_hairpin_tables* tables) {
  if (!ctx || !tables) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  if (!ctx->d_hp_tables) DP_CUDA(ctx, cudaMalloc(&ctx->d_hp_tables, sizeof(dp_hairpin_tables)));
  DP_CUDA(ctx, cudaMemcpyAsync(ctx->d_hp_tables, tables, sizeof(dp_hairpin_tables), cudaMemcpyHostToDevice, ctx->stream));
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->hp_tables_loaded = true;
  return DP_OK;
}

extern "C" int dp_find_hairpins(dp_ctx* ctx, uint32_t n_seqs, const char* seqs, const uint64_t* off, const uint8_t* len,
                                uint32_t n_seeds, const uint32_t* seed_seq, const uint64_t* seed_mask,
                                const int32_t* seed_offset, const dp_conditions* cond, double max_hairpin_dG,
                                dp_best_structure* out, float* ms_kernel) {
  if (!ctx || !seqs || !off || !len || !cond || (n_seeds && (!seed_seq || !seed_mask || !seed_offset || !out)))
    return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (!ctx->tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_tables_load() has not been called");
  if (!ctx->hp_tables_loaded) return dp_set_error(ctx, DP_ERR_STATE, "dp_hairpin_tables_load() has not been called");
  if (ms_kernel) *ms_kernel = 0.f;
  if (n_seeds == 0) return DP_OK;
  uint64_t bytes = 0;
  for (uint32_t i = 0; i < n_seqs; ++i) {
    if (len[i] == 0 || len[i] > DP_WIN)
      return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "sequence %u: length outside [1,%d]", i, DP_WIN);
    bytes = std::max<uint64_t>(bytes, off[i] + len[i]);
  }
  for (uint32_t k = 0; k < n_seeds; ++k) {
    if (seed_seq[k] >= n_seqs) return dp_set_error(ctx, DP_ERR_ARG, "seed %u: sequence index out of range", k);
    const int L = len[seed_seq[k]];
    const uint64_t m = seed_mask[k];
    if (L < 64 && (m >> L)) return dp_set_error(ctx, DP_ERR_ARG, "seed %u: stem position beyond the sequence", k);
    if (m) {
      // every stem position f must pair with L-1-(f-offset) 3' of the whole 5' arm, >= 3 bases apart
      const int f0 = __builtin_ctzll(m), f1 = 63 - __builtin_clzll(m);
      const int r0 = L - 1 - (f0 - seed_offset[k]), r1 = L - 1 - (f1 - seed_offset[k]);
      if (r0 < 0 || r0 >= L || r1 - f1 < 4)
        return dp_set_error(ctx, DP_ERR_ARG, "seed %u: stem arms overlap or leave fewer than 3 loop bases", k);
    }
  }
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  std::vector<uint8_t> codes;
  int rc = letters_of(ctx, seqs, bytes, codes, false);
  if (rc) return rc;
  dev_buf<uint8_t> d_codes, d_len;
  dev_buf<uint64_t> d_off, d_mask;
  dev_buf<uint32_t> d_seq;
  dev_buf<int32_t> d_so;
  dev_buf<dp_best_structure> d_out;
  DP_CUDA(ctx, d_codes.alloc(ctx, bytes)); DP_CUDA(ctx, d_len.alloc(ctx, n_seqs)); DP_CUDA(ctx, d_off.alloc(ctx, n_seqs));
  DP_CUDA(ctx, d_seq.alloc(ctx, n_seeds)); DP_CUDA(ctx, d_mask.alloc(ctx, n_seeds)); DP_CUDA(ctx, d_so.alloc(ctx, n_seeds));
  DP_CUDA(ctx, d_out.alloc(ctx, n_seeds));
  cudaStream_t st = ctx->stream;
  DP_CUDA(ctx, cudaMemcpyAsync(d_codes.p, codes.data(), bytes, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_len.p, len, n_seqs, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_off.p, off, n_seqs * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_seq.p, seed_seq, n_seeds * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_mask.p, seed_mask, n_seeds * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_so.p, seed_offset, n_seeds * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (ms_kernel) {
    DP_CUDA(ctx, cudaEventCreate(&e0));
    DP_CUDA(ctx, cudaEventCreate(&e1));
    DP_CUDA(ctx, cudaEventRecord(e0, st));
  }
  const uint64_t blocks_needed = ((uint64_t)n_seeds + DP_FD_WARPS - 1) / DP_FD_WARPS;
  const unsigned grid = (unsigned)std::min<uint64_t>(blocks_needed, (uint64_t)ctx->sm_count * 16);
  dp_dg_args A = make_args(ctx, cond);
  dp_find_hairpins_kernel<<<grid, DP_FD_THREADS, 0, st>>>(A, ctx->d_hp_tables, max_hairpin_dG, d_codes.p, d_off.p, d_len.p,
                                                              d_seq.p, d_mask.p, d_so.p, n_seeds, d_out.p);
  cudaError_t le = cudaGetLastError();
  if (ms_kernel && le == cudaSuccess) le = cudaEventRecord(e1, st);
  if (le == cudaSuccess) le = cudaMemcpyAsync(out, d_out.p, (size_t)n_seeds * sizeof(dp_best_structure), cudaMemcpyDeviceToHost, st);
  if (le == cudaSuccess) le = cudaStreamSynchronize(st);
  if (ms_kernel && le == cudaSuccess) le = cudaEventElapsedTime(ms_kernel, e0, e1);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  DP_CUDA(ctx, le);
  return DP_OK;
}
