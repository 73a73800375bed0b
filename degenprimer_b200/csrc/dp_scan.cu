// Annealing-site scan (dp_scan): host driver + the per-window popcount kernel.
//
// Replaces SearchEngine._find / _find_mp / _find_in_slice / _find_in_chunk
// (DegenPrimer/SearchEngine.py:190-211,257-359).  The reference correlates
// root-of-unity maps with FFTs; the value it thresholds is exactly
//   score3(q) = 3*#full + 2*#partial + #nonACGT   (thirds)
// and a window is a hit iff score3 >= 3*max(1, L-k) - 1.  Both strands are
// scanned from the forward planes: the reverse strand uses reverse-complemented
// masks and reports pos_rev = T - L - pos_fwd.
#include <algorithm>
#include <stdlib.h>
#include <string.h>

#include "dp_scan.cuh"
#include "dp_scan_bitslice.cuh"

// ---- kernel 1: every window, XOR/AND-style mask select + popcount ---------------
// One thread owns one plane word = 32 consecutive window starts.
//   only_exc = 0 : all windows
//   only_exc = 1 : only windows that touch a non-ACGT base (the bit-sliced
//                  kernel skips exactly those)
template <int NW>
__global__ void __launch_bounds__(256)
dp_scan_popc_kernel(dp_scan_args A, uint64_t pmask0, uint64_t pmask1, int only_exc) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t w = A.word_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < A.word_end; w += stride) {
    uint32_t H[NW + 1], Lw[NW + 1], E[NW + 1];
#pragma unroll
    for (int k = 0; k <= NW; ++k) {
      H[k] = __ldg(A.hi + w + k);
      Lw[k] = __ldg(A.lo + w + k);
      E[k] = __ldg(A.exc + w + k);
    }
    if (only_exc) {
      uint32_t any = 0;
#pragma unroll
      for (int k = 0; k <= NW; ++k) any |= E[k];
      if (!any) continue;
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
        const dp_pattern& P = c_patterns[p];
        if (only_exc) {
          uint32_t touch = 0;
#pragma unroll
          for (int k = 0; k < NW; ++k) touch |= wex[k] & P.lmask[k];
          if (!touch) continue;
        }
        uint32_t s = dp_score3<NW>(P, whi, wlo, wex);
        if (s >= P.thr3) dp_emit_hit(A, g, p, P.len, s);
      }
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
                          const int32_t* kmax, int strand_mask, std::vector<dp_pattern>& pats,
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

static int ensure_hits(dp_ctx* ctx, uint64_t cap) {
  if (ctx->hits_cap >= cap) return DP_OK;
  if (ctx->d_hits) { cudaFree(ctx->d_hits); ctx->d_hits = nullptr; ctx->hits_cap = 0; }
  DP_CUDA(ctx, cudaMalloc(&ctx->d_hits, cap * sizeof(dp_raw_hit)));
  ctx->hits_cap = cap;
  return DP_OK;
}

extern "C" int dp_scan(dp_ctx* ctx, const dp_template* t, uint32_t n_primers,
                       const char* const* masters, const int32_t* max_mismatches,
                       uint64_t seg_begin, uint64_t seg_end, int strand_mask, int algo,
                       dp_hit** hits_out, uint64_t* n_hits_out, dp_scan_stats* stats) {
  if (!ctx || !t || !masters || !hits_out || !n_hits_out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *hits_out = nullptr;
  *n_hits_out = 0;
  if (stats) memset(stats, 0, sizeof *stats);
  if (n_primers == 0 || n_primers > DP_MAX_PRIMERS)
    return dp_set_error(ctx, DP_ERR_ARG, "n_primers must be in [1,%d]", DP_MAX_PRIMERS);
  if (!(strand_mask & DP_STRAND_BOTH)) return dp_set_error(ctx, DP_ERR_ARG, "empty strand mask");
  if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  std::vector<dp_pattern> pats;
  std::vector<std::vector<uint8_t>> letters;
  int rc = build_patterns(ctx, n_primers, masters, max_mismatches, strand_mask, pats, letters);
  if (rc) return rc;
  uint32_t max_len = 0;
  for (auto& p : pats) max_len = std::max<uint32_t>(max_len, p.len);
  if (t->records.size() == 1 && t->records[0].len < max_len)
    return dp_set_error(ctx, DP_ERR_LENGTH, "template (%llu) is shorter than the primer (%u)",
                        (unsigned long long)t->records[0].len, max_len);
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaEvent_t ev_begin = ctx->ev0, ev_k0 = ctx->ev1, ev_k1 = ctx->ev2;
  DP_CUDA(ctx, cudaEventRecord(ev_begin, ctx->stream));
  DP_CUDA(ctx, cudaMemcpyToSymbolAsync(c_patterns, pats.data(), pats.size() * sizeof(dp_pattern), 0,
                                       cudaMemcpyHostToDevice, ctx->stream));
  const uint32_t n_pat = (uint32_t)pats.size();
  uint64_t pmask_all[2] = {0, 0};
  for (uint32_t p = 0; p < n_pat; ++p) pmask_all[p >> 6] |= 1ull << (p & 63);

  dp_scan_args A;
  A.lo = t->d_lo; A.hi = t->d_hi; A.exc = t->d_exc;
  A.rec_start = t->d_rec_start; A.rec_len = t->d_rec_len;
  A.n_records = (uint32_t)t->records.size();
  A.n_patterns = n_pat;
  A.n_words = t->n_words;
  if (seg_end > t->n_words * 32) seg_end = t->n_words * 32;
  if (seg_begin > seg_end) seg_begin = seg_end;
  A.seg_begin = seg_begin; A.seg_end = seg_end;
  A.word_begin = seg_begin / 32;
  A.word_end = (seg_end + 31) / 32;
  A.counters = ctx->d_counters;

  // plan: which patterns go through the bit-sliced prefilter
  dp_bitslice_plan plan;
  int used_algo = DP_ALGO_POPC;
  if (algo == DP_ALGO_AUTO || algo == DP_ALGO_BITSLICE) {
    if (dp_bitslice_make_plan(pats, letters, plan)) used_algo = DP_ALGO_BITSLICE;
  } else if (algo != DP_ALGO_POPC) {
    return dp_set_error(ctx, DP_ERR_ARG, "unknown algo %d", algo);
  }
  const int NW = max_len > 32 ? 2 : 1;
  const uint64_t n_scan_words = A.word_end - A.word_begin;
  uint32_t launches = 0;
  float ms_scan = 0.f, ms_main = 0.f;
  unsigned long long h_counters[4] = {0, 0, 0, 0};
  uint64_t cap = std::max<uint64_t>(ctx->hits_cap, 1u << 20);
  for (int attempt = 0; attempt < 3; ++attempt) {
    rc = ensure_hits(ctx, cap);
    if (rc) return rc;
    A.hits = (dp_raw_hit*)ctx->d_hits;
    A.hits_cap = ctx->hits_cap;
    DP_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, 4 * sizeof(unsigned long long), ctx->stream));
    DP_CUDA(ctx, cudaEventRecord(ev_k0, ctx->stream));
    if (n_scan_words) {
      unsigned blocks = (unsigned)std::min<uint64_t>((n_scan_words + 255) / 256, (uint64_t)ctx->sm_count * 8);
      if (used_algo == DP_ALGO_BITSLICE) {
        rc = dp_bitslice_launch(ctx, A, plan, &launches);
        if (rc) return rc;
        DP_CUDA(ctx, cudaEventRecord(ctx->ev3, ctx->stream));
        // windows touching non-ACGT bases, for the prefiltered patterns
        if (plan.pmask[0] | plan.pmask[1]) {
          if (NW == 1) dp_scan_popc_kernel<1><<<blocks, 256, 0, ctx->stream>>>(A, plan.pmask[0], plan.pmask[1], 1);
          else dp_scan_popc_kernel<2><<<blocks, 256, 0, ctx->stream>>>(A, plan.pmask[0], plan.pmask[1], 1);
          ++launches;
        }
        uint64_t rest[2] = {pmask_all[0] & ~plan.pmask[0], pmask_all[1] & ~plan.pmask[1]};
        if (rest[0] | rest[1]) {
          if (NW == 1) dp_scan_popc_kernel<1><<<blocks, 256, 0, ctx->stream>>>(A, rest[0], rest[1], 0);
          else dp_scan_popc_kernel<2><<<blocks, 256, 0, ctx->stream>>>(A, rest[0], rest[1], 0);
          ++launches;
        }
      } else {
        if (NW == 1) dp_scan_popc_kernel<1><<<blocks, 256, 0, ctx->stream>>>(A, pmask_all[0], pmask_all[1], 0);
        else dp_scan_popc_kernel<2><<<blocks, 256, 0, ctx->stream>>>(A, pmask_all[0], pmask_all[1], 0);
        ++launches;
        DP_CUDA(ctx, cudaEventRecord(ctx->ev3, ctx->stream));
      }
      DP_CUDA(ctx, cudaGetLastError());
    }
    DP_CUDA(ctx, cudaEventRecord(ev_k1, ctx->stream));
    DP_CUDA(ctx, cudaMemcpyAsync(h_counters, ctx->d_counters, sizeof h_counters, cudaMemcpyDeviceToHost, ctx->stream));
    DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev_k0, ev_k1);
    ms_scan += ms;
    if (n_scan_words) cudaEventElapsedTime(&ms_main, ev_k0, ctx->ev3);
    if (h_counters[0] <= ctx->hits_cap) break;
    cap = h_counters[0] + (h_counters[0] >> 3) + 1024;     // overflow: grow and rescan
    if (attempt == 2) return dp_set_error(ctx, DP_ERR_STATE, "internal: hit buffer overflow persists");
    if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  }
  const uint64_t n = h_counters[0];
  std::vector<dp_raw_hit> raw(n);
  if (n) DP_CUDA(ctx, cudaMemcpy(raw.data(), ctx->d_hits, n * sizeof(dp_raw_hit), cudaMemcpyDeviceToHost));
  dp_hit* out = (dp_hit*)malloc(std::max<uint64_t>(1, n) * sizeof(dp_hit));
  if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu hits", (unsigned long long)n);
  for (uint64_t i = 0; i < n; ++i) {
    const dp_raw_hit& h = raw[i];
    const dp_pattern& P = pats[h.pattern];
    const dp_record& rec = t->records[h.record];
    uint64_t off = h.g - rec.start;
    out[i].pos = P.strand ? rec.len - P.len - off : off;
    out[i].record = h.record;
    out[i].primer = P.primer;
    out[i].strand = P.strand;
    out[i].score3 = (uint8_t)h.score3;
  }
  std::sort(out, out + n, [](const dp_hit& a, const dp_hit& b) {
    if (a.primer != b.primer) return a.primer < b.primer;
    if (a.strand != b.strand) return a.strand < b.strand;
    if (a.record != b.record) return a.record < b.record;
    return a.pos < b.pos;
  });
  *hits_out = out;
  *n_hits_out = n;
  if (stats) {
    uint64_t windows = 0;
    for (auto& p : pats)
      for (auto& rec : t->records)
        if (rec.len >= p.len) windows += rec.len - p.len + 1;
    stats->windows = windows;
    stats->candidates = h_counters[1];
    stats->hits = n;
    stats->ms_scan = ms_scan;
    stats->ms_main = ms_main;
    cudaEventRecord(ev_k1, ctx->stream);
    cudaEventSynchronize(ev_k1);
    float ms_total = 0.f;
    cudaEventElapsedTime(&ms_total, ev_begin, ev_k1);
    stats->ms_total = ms_total;
    stats->launches = launches;
    stats->algo = (uint32_t)used_algo;
  }
  return DP_OK;
}
