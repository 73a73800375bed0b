// Scan kernel 2: bit-sliced prefilter over 32 windows at once + exact verify.
//
// Idea.  A window can only be a hit if a J-letter stretch of
// the pattern ("anchor", chosen per pattern by the host) has at most K hard
// mismatches, K = L - max(1, L-k).  That necessary condition is evaluated for
// 32 consecutive windows per thread with word-wide logic only:
//   * per plane word the thread builds the 4*J one-hot planes
//       X_j[b] = (base at 32w+b+j == X),  X in A C T G, j < J
//     (2 funnel shifts + 4 LOP3 per j) and parks them in its private column of
//     shared memory;
//   * a pattern is then just a list of NT plane slots (one per allowed base of
//     each anchor letter, padded with an all-zero plane); the number of
//     matching anchor positions is the bit-sliced sum of those planes (a
//     carry-save adder tree of LOP3 full adders), and `alive` = sum >= J-K;
//   * only alive windows (p ~ 4e-5 on random sequence) reach the exact
//     thirds-weighted popcount verify.
// Cost per (32 windows x pattern): NT LDS.32 + ~1.5*NT LOP3, no POPC, versus
// 32 x (3 LOP3 + POPC + compare) for kernel 1.
// Windows that touch a non-ACGT base are left to kernel 1 (only_exc mode).
#pragma once
#include <math.h>

#include "dp_scan.cuh"

#define DP_BS_THREADS 256
#define DP_BS_MAX_K 4
#define DP_BS_MAX_E 8
#define DP_BS_MAX_TERMS (8 + 2 * DP_BS_MAX_K + DP_BS_MAX_E)

struct dp_bs_pattern {
  uint32_t term[DP_BS_MAX_TERMS];   // byte offset of the plane column
  uint32_t anchor;                  // offset of the anchor inside the pattern
  uint32_t pattern;                 // index into c_patterns
};

// One launch group: up to NB patterns with the same K and the same number of plane
// terms.  The group carries its own copy of the patterns (local index) next to the
// prefilter entries; bs[i].pattern stays the index in the call's pattern list (that
// is what a hit reports).
#define DP_BS_GROUP_SMALL 8
#define DP_BS_GROUP_LARGE 32
template <int NB>
struct dp_bs_set {
  dp_bs_pattern bs[NB];
  dp_pattern pat[NB];
};

struct dp_bitslice_group {
  int K, E;
  std::vector<dp_bs_pattern> pats;
};

struct dp_bitslice_plan {
  uint32_t n;                       // patterns handled by the prefilter
  uint64_t pmask[2];                // the same set as a bit mask over the call's patterns
  std::vector<dp_bitslice_group> groups;
};

// one pattern against the 32 windows of the thread's current word
template <int K, int E>
__device__ __forceinline__ void dp_bs_pattern_step(const dp_scan_args& A, const dp_pattern& P, const dp_bs_pattern& bp,
                                                   const uint32_t* col, uint64_t w, unsigned long long& n_cand) {
  constexpr int J = 8 + 2 * K;
  constexpr int NT = J + E;
  constexpr int THR = J - K;
  uint32_t x[NT];
#pragma unroll
  for (int t = 0; t < NT; ++t)
    x[t] = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(col) + bp.term[t]);
  uint32_t alive = dp_count_ge<NT, THR>(x);
  while (alive) {
    const int b = __ffs(alive) - 1;
    alive &= alive - 1;
    ++n_cand;
    const uint64_t q = w * 32 + b;
    if (q < bp.anchor) continue;
    const uint64_t g = q - bp.anchor;
    if (g < A.seg_begin || g >= A.seg_end) continue;
    dp_verify_window(A, P, g, bp.pattern, false);
  }
}

// resident CTAs per SM the register budget is tuned for: shared memory allows
// 5 / 5 / 4 / 3 / 3 CTAs for K = 0..4 (4*(8+2K)+1 planes of 1 KB each)
__host__ __device__ constexpr int dp_bs_min_blocks(int K) { return K <= 1 ? 5 : K == 2 ? 4 : 3; }

template <int K, int E, int NB>
__global__ void __launch_bounds__(DP_BS_THREADS, dp_bs_min_blocks(K))
dp_scan_bitslice_kernel(const __grid_constant__ dp_scan_args A, const __grid_constant__ dp_bs_set<NB> BS, uint32_t n_bs) {
  constexpr int J = 8 + 2 * K;
  extern __shared__ uint32_t dp_bs_smem[];
  uint32_t* col = dp_bs_smem + threadIdx.x;
  col[4 * J * DP_BS_THREADS] = 0u;               // the all-zero padding plane
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  const uint64_t w_end = A.word_end + 2;          // anchors sit up to 63 bases into the window
  unsigned long long n_cand = 0;
  uint64_t w = A.word_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t h0 = 0, h1 = 0, l0 = 0, l1 = 0;
  if (w < w_end) {
    h0 = __ldg(A.hi + w); h1 = __ldg(A.hi + w + 1);
    l0 = __ldg(A.lo + w); l1 = __ldg(A.lo + w + 1);
  }
  while (w < w_end) {
    // the next word's planes are requested before this word is processed: HBM
    // latency hides behind ~300 instructions of bit-sliced work
    const uint64_t wn = w + stride;
    uint32_t nh0 = 0, nh1 = 0, nl0 = 0, nl1 = 0;
    if (wn < w_end) {
      nh0 = __ldg(A.hi + wn); nh1 = __ldg(A.hi + wn + 1);
      nl0 = __ldg(A.lo + wn); nl1 = __ldg(A.lo + wn + 1);
    }
#pragma unroll
    for (int j = 0; j < J; ++j) {
      const uint32_t hs = __funnelshift_r(h0, h1, j);
      const uint32_t ls = __funnelshift_r(l0, l1, j);
      col[(4 * j + 0) * DP_BS_THREADS] = ~(hs | ls);  // A
      col[(4 * j + 1) * DP_BS_THREADS] = ~hs & ls;    // C
      col[(4 * j + 2) * DP_BS_THREADS] = hs & ~ls;    // T
      col[(4 * j + 3) * DP_BS_THREADS] = hs & ls;     // G
    }
    // unrolled over the group, so every plane offset is a kernel-parameter constant
    // that reaches LDS through a uniform register (no per-word LDC + address add)
#pragma unroll
    for (int p = 0; p < NB; ++p)
      if (p < (int)n_bs) dp_bs_pattern_step<K, E>(A, BS.pat[p], BS.bs[p], col, w, n_cand);
    w = wn;
    h0 = nh0; h1 = nh1; l0 = nl0; l1 = nl1;
  }
  if (n_cand) atomicAdd(&A.counters[DP_C_CAND], n_cand);
}

// ---- host: anchor choice -----------------------------------------------------------
// probability that a random ACGT stretch has <= K hard mismatches against the
// J letters starting at `o` (match probability |set|/4 per position)
static double dp_bs_alive_probability(const std::vector<uint8_t>& letters, int o, int J, int K) {
  std::vector<double> d(K + 2, 0.0);   // d[m] = P(m mismatches so far), m = K+1 absorbs "more"
  d[0] = 1.0;
  for (int j = 0; j < J; ++j) {
    double pm = __builtin_popcount(dp_letter_set(letters[o + j])) / 4.0;
    for (int m = K + 1; m >= 0; --m) {
      double stay = d[m] * pm;
      double from = m > 0 ? d[m - 1] * (1.0 - pm) : 0.0;
      if (m == K + 1) stay = d[m];   // absorbing
      d[m] = stay + from;
    }
  }
  double s = 0.0;
  for (int m = 0; m <= K; ++m) s += d[m];
  return s;
}

// Decides which patterns take the prefilter and how they are grouped into launches.
// Returns false if none does.
static bool dp_bitslice_make_plan(const std::vector<dp_pattern>& pats,
                                  const std::vector<std::vector<uint8_t>>& letters,
                                  dp_bitslice_plan& plan) {
  plan.n = 0;
  plan.pmask[0] = plan.pmask[1] = 0;
  plan.groups.clear();
  // (K, E class) -> prefilter entries
  std::vector<dp_bs_pattern> bucket[DP_BS_MAX_K + 1][4];
  static const int E_CLASS[4] = {0, 2, 4, 8};
  for (size_t p = 0; p < pats.size(); ++p) {
    const int need = (pats[p].thr3 + 1) / 3;
    const int K = pats[p].len - need;
    if (K < 0 || K > DP_BS_MAX_K) continue;
    const int J = 8 + 2 * K;
    if ((int)pats[p].len < J) continue;
    const std::vector<uint8_t>& lt = letters[p];
    double best_p = 2.0;
    int best_o = -1, best_terms = 0;
    for (int o = 0; o + J <= (int)pats[p].len; ++o) {
      int terms = 0;
      for (int j = 0; j < J; ++j) terms += __builtin_popcount(dp_letter_set(lt[o + j]));
      if (terms > J + DP_BS_MAX_E) continue;
      double pa = dp_bs_alive_probability(lt, o, J, K);
      if (pa < best_p - 1e-15 || (fabs(pa - best_p) <= 1e-15 && terms < best_terms)) {
        best_p = pa; best_o = o; best_terms = terms;
      }
    }
    if (best_o < 0 || best_p > 2e-3) continue;   // too unselective: kernel 1 is cheaper
    const int extra = best_terms - J;
    const int ec = extra == 0 ? 0 : extra <= 2 ? 1 : extra <= 4 ? 2 : 3;
    const int NT = J + E_CLASS[ec];
    dp_bs_pattern bp;
    memset(&bp, 0, sizeof bp);
    bp.anchor = (uint32_t)best_o;
    bp.pattern = (uint32_t)p;
    int n = 0;
    for (int j = 0; j < J; ++j) {
      uint32_t set = dp_letter_set(lt[best_o + j]);
      for (int c = 0; c < 4; ++c)
        if (set & (1u << c)) bp.term[n++] = (uint32_t)((4 * j + c) * DP_BS_THREADS * sizeof(uint32_t));
    }
    for (; n < NT; ++n) bp.term[n] = (uint32_t)(4 * J * DP_BS_THREADS * sizeof(uint32_t));   // the all-zero plane
    bucket[K][ec].push_back(bp);
    plan.pmask[p >> 6] |= 1ull << (p & 63);
    ++plan.n;
  }
  // launches of at most DP_BS_GROUP_LARGE patterns, evenly filled (every launch rebuilds
  // the planes of a word once, so fewer and fuller groups are cheaper)
  for (int K = 0; K <= DP_BS_MAX_K; ++K)
    for (int ec = 0; ec < 4; ++ec) {
      const std::vector<dp_bs_pattern>& b = bucket[K][ec];
      if (b.empty()) continue;
      const size_t n_groups = (b.size() + DP_BS_GROUP_LARGE - 1) / DP_BS_GROUP_LARGE;
      const size_t per = (b.size() + n_groups - 1) / n_groups;
      for (size_t at = 0; at < b.size(); at += per) {
        dp_bitslice_group g;
        g.K = K;
        g.E = E_CLASS[ec];
        g.pats.assign(b.begin() + at, b.begin() + std::min(b.size(), at + per));
        plan.groups.push_back(g);
      }
    }
  return plan.n > 0;
}

template <int K, int E, int NB>
static int dp_bitslice_launch_one(dp_ctx* ctx, const dp_scan_args& A, const std::vector<dp_pattern>& pats,
                                  const dp_bitslice_group& g) {
  constexpr int J = 8 + 2 * K;
  const size_t smem = (size_t)(4 * J + 1) * DP_BS_THREADS * sizeof(uint32_t);
  static bool configured[16] = {false};      // per device
  static int per_sm_cached[16] = {0};
  const int dev = ctx->device & 15;
  if (!configured[dev]) {
    DP_CUDA(ctx, cudaFuncSetAttribute(dp_scan_bitslice_kernel<K, E, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    DP_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dp_scan_bitslice_kernel<K, E, NB>, DP_BS_THREADS, smem));
    per_sm_cached[dev] = per_sm < 1 ? 1 : per_sm;
    configured[dev] = true;
  }
  static thread_local dp_bs_set<NB> BS;
  memset(&BS, 0, sizeof BS);
  for (size_t i = 0; i < g.pats.size(); ++i) {
    BS.bs[i] = g.pats[i];
    BS.pat[i] = pats[g.pats[i].pattern];
  }
  const uint64_t words = A.word_end + 2 - A.word_begin;
  unsigned blocks = (unsigned)std::min<uint64_t>((words + DP_BS_THREADS - 1) / DP_BS_THREADS,
                                                 (uint64_t)ctx->sm_count * per_sm_cached[dev]);
  dp_scan_bitslice_kernel<K, E, NB><<<blocks, DP_BS_THREADS, smem, ctx->stream>>>(A, BS, (uint32_t)g.pats.size());
  DP_CUDA(ctx, cudaGetLastError());
  return DP_OK;
}

static int dp_bitslice_launch(dp_ctx* ctx, const dp_scan_args& A, const std::vector<dp_pattern>& pats,
                              const dp_bitslice_plan& plan, uint32_t* launches) {
  for (const dp_bitslice_group& g : plan.groups) {
    int rc = DP_ERR_ARG;
    const bool small = g.pats.size() <= DP_BS_GROUP_SMALL;
#define DP_BS_CASE(KK, EE)                                                                        \
    if (g.K == KK && g.E == EE)                                                                   \
      rc = small ? dp_bitslice_launch_one<KK, EE, DP_BS_GROUP_SMALL>(ctx, A, pats, g)             \
                 : dp_bitslice_launch_one<KK, EE, DP_BS_GROUP_LARGE>(ctx, A, pats, g);
#define DP_BS_ROW(KK) DP_BS_CASE(KK, 0) DP_BS_CASE(KK, 2) DP_BS_CASE(KK, 4) DP_BS_CASE(KK, 8)
    DP_BS_ROW(0) DP_BS_ROW(1) DP_BS_ROW(2) DP_BS_ROW(3) DP_BS_ROW(4)
#undef DP_BS_ROW
#undef DP_BS_CASE
    if (rc == DP_ERR_ARG) return dp_set_error(ctx, DP_ERR_STATE, "internal: no bit-slice kernel for K=%d E=%d", g.K, g.E);
    if (rc) return rc;
    ++*launches;
  }
  return DP_OK;
}
