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
__device__ double dp_dimer_dg_device(const dp_dg_args& A, const uint8_t* s, int s_len,
                                     const uint8_t* r, int r_len, uint64_t mask, int off) {
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

// ---- window letters of every hit ------------------------------------------------
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
    uint64_t wi = g >> 5;
    uint32_t b = (uint32_t)(g & 31);
    uint32_t letter;
    if ((exc[wi] >> b) & 1u) {
      uint64_t lo_i = 0, hi_i = n_runs;
      while (lo_i < hi_i) {
        uint64_t mid = (lo_i + hi_i) >> 1;
        if (runs[mid].start <= g) lo_i = mid + 1; else hi_i = mid;
      }
      letter = 14u;
      if (lo_i) { const dp_exc_run& rr = runs[lo_i - 1]; if (g < rr.end) letter = rr.letter; }
    } else {
      letter = ((hi[wi] >> b) & 1u) * 2u + ((lo[wi] >> b) & 1u);
    }
    out = (uint8_t)(strand ? dp_letter_complement(letter) : letter);
  }
  win[idx] = out;
}

// ---- per (hit, variant) duplex ----------------------------------------------------
// pass 0: counts[pair] = surviving dimers;  pass 1: write them at out_off[pair]
// in the reference's order [full, alternatives...] and stable-sort by dG.
template <int PASS>
__global__ void __launch_bounds__(128)
dp_duplex_kernel(dp_dg_args A, const uint8_t* __restrict__ win, uint64_t hit0, uint64_t n_hits_p,
                 uint32_t V, int L, const uint8_t* __restrict__ var_codes /* V x L */,
                 uint64_t pair0, uint32_t* __restrict__ counts, const uint64_t* __restrict__ out_off,
                 dp_dimer* __restrict__ out, unsigned long long* __restrict__ tried) {
  uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_hits_p * V) return;
  const uint64_t h = hit0 + idx / V;
  const uint32_t v = (uint32_t)(idx % V);
  uint8_t s[DP_WIN], r[DP_WIN];
  const uint8_t* wl = win + h * DP_WIN;
  const uint8_t* vc = var_codes + (uint64_t)v * L;
  uint64_t M = 0;
  for (int i = 0; i < L; ++i) {
    uint8_t w = wl[i];
    s[i] = vc[i];
    r[i] = (uint8_t)dp_letter_complement(w);        // rev_str = complement(window)
    if (w == s[i]) M |= 1ull << i;                   // plain character equality
  }
  uint32_t kept = 0;
  unsigned long long n_tried = 0;
  dp_dimer* dst = PASS ? out + out_off[pair0 + idx] : nullptr;
  auto consider = [&](uint64_t sub) {
    ++n_tried;
    double dG = dp_dimer_dg_device(A, s, L, r, L, sub, 0);
    if (dG > A.max_dimer_dG) return;        // SecStructures.py:302
    if (!(dG <= A.max_dG_for_K)) return;    // K < min_K, SecStructures.py:303-304
    if (PASS) {
      dp_dimer d;
      d.hit = (uint32_t)h; d.variant = v; d.match_mask = sub; d.dG = dG;
      // stable insertion by dG (ties keep construction order)
      uint32_t pos = kept;
      while (pos > 0 && dst[pos - 1].dG > dG) { dst[pos] = dst[pos - 1]; --pos; }
      dst[pos] = d;
    }
    ++kept;
  };
  if (M) {
    consider(M);
    // Dimer.alternatives (SecStructures.py:126-152): split M into maximal runs
    // of consecutive indices and emit every contiguous range of runs except
    // the full one, in the reference's order: for each left cut l, first the
    // right cuts from the far right inwards, then the next left cut.
    const uint64_t S = M & ~(M << 1);      // first index of every run
    const uint64_t Eb = M & ~(M >> 1);     // last index of every run
    const uint64_t top_end = 1ull << (63 - __clzll((long long)Eb));
    uint64_t sl = S;
    while (sl & (sl - 1)) {                // at least one more run to the right
      const int st = __ffsll((long long)sl) - 1;
      const uint64_t lo_m = ~0ull << st;
      uint64_t eh = Eb & lo_m & ~top_end;  // ends of runs l .. last-1
      while (eh) {
        const int e = 63 - __clzll((long long)eh);
        eh &= ~(1ull << e);
        const uint64_t hi_m = (2ull << e) - 1;   // e <= 62 here
        consider(M & lo_m & hi_m);
      }
      sl &= sl - 1;
      const int st2 = __ffsll((long long)sl) - 1;
      consider(M & (~0ull << st2));
    }
  }
  if (!PASS) counts[pair0 + idx] = kept;
  if (!PASS && n_tried) atomicAdd(tried, n_tried);
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

// ---- host ----------------------------------------------------------------------------
template <typename T>
struct dev_buf {
  T* p = nullptr;
  ~dev_buf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t n) { return cudaMalloc(&p, std::max<size_t>(1, n) * sizeof(T)); }
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
  if (n_hits == 0) { *dimers_out = (dp_dimer*)malloc(sizeof(dp_dimer)); return DP_OK; }
  if (n_hits > 0xFFFFFFFFull) return dp_set_error(ctx, DP_ERR_ARG, "more than 2^32 hits in one call");
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  // hits must be grouped by primer (dp_scan returns them sorted that way)
  std::vector<uint64_t> hit_g(n_hits);
  std::vector<uint8_t> hit_len(n_hits), hit_strand(n_hits);
  std::vector<uint64_t> pair0(n_hits + 1, 0);
  struct Group { uint64_t h0, n; uint32_t primer; };
  std::vector<Group> groups;
  for (uint64_t h = 0; h < n_hits; ++h) {
    const dp_hit& x = hits[h];
    if (x.primer >= n_primers) return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: primer %u out of range", (unsigned long long)h, x.primer);
    if (x.record >= t->records.size()) return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: record out of range", (unsigned long long)h);
    const dp_record& rec = t->records[x.record];
    uint32_t L = primer_len[x.primer];
    if (L == 0 || L > DP_MAX_PRIMER_LEN || x.pos + L > rec.len)
      return dp_set_error(ctx, DP_ERR_ARG, "hit %llu: window outside its record", (unsigned long long)h);
    uint64_t fwd = x.strand ? rec.len - L - x.pos : x.pos;
    hit_g[h] = rec.start + fwd;
    hit_len[h] = (uint8_t)L;
    hit_strand[h] = x.strand ? 1 : 0;
    pair0[h + 1] = pair0[h] + var_count[x.primer];
    if (groups.empty() || groups.back().primer != x.primer) groups.push_back(Group{h, 1, x.primer});
    else ++groups.back().n;
  }
  const uint64_t n_pairs = pair0[n_hits];
  // variants -> plane codes
  uint64_t var_bytes = 0;
  for (uint32_t p = 0; p < n_primers; ++p)
    var_bytes = std::max<uint64_t>(var_bytes, var_offset[p] + (uint64_t)var_count[p] * primer_len[p]);
  std::vector<uint8_t> codes;
  int rc = letters_of(ctx, variants, var_bytes, codes, true);
  if (rc) return rc;

  cudaStream_t st = ctx->stream;
  dev_buf<uint64_t> d_g, d_off;
  dev_buf<uint8_t> d_len, d_strand, d_win, d_codes;
  dev_buf<uint32_t> d_counts;
  dev_buf<unsigned long long> d_tried;
  dev_buf<dp_dimer> d_out;
  DP_CUDA(ctx, d_g.alloc(n_hits));
  DP_CUDA(ctx, d_len.alloc(n_hits));
  DP_CUDA(ctx, d_strand.alloc(n_hits));
  DP_CUDA(ctx, d_win.alloc(n_hits * DP_WIN));
  DP_CUDA(ctx, d_codes.alloc(var_bytes));
  DP_CUDA(ctx, d_counts.alloc(n_pairs));
  DP_CUDA(ctx, d_off.alloc(n_pairs + 1));
  DP_CUDA(ctx, d_tried.alloc(1));
  DP_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_g.p, hit_g.data(), n_hits * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_len.p, hit_len.data(), n_hits, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_strand.p, hit_strand.data(), n_hits, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemcpyAsync(d_codes.p, codes.data(), var_bytes, cudaMemcpyHostToDevice, st));
  DP_CUDA(ctx, cudaMemsetAsync(d_tried.p, 0, sizeof(unsigned long long), st));
  uint32_t launches = 0;
  {
    uint64_t n = n_hits * DP_WIN;
    dp_dg_windows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
        t->d_lo, t->d_hi, t->d_exc, t->d_runs, t->runs.size(), d_g.p, d_len.p, d_strand.p, n_hits, d_win.p);
    DP_CUDA(ctx, cudaGetLastError());
    ++launches;
  }
  dp_dg_args A = make_args(ctx, cond);
  DP_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
  for (const Group& g : groups) {
    uint32_t V = var_count[g.primer];
    uint64_t n = g.n * V;
    if (!n) continue;
    dp_duplex_kernel<0><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(
        A, d_win.p, g.h0, g.n, V, primer_len[g.primer], d_codes.p + var_offset[g.primer], pair0[g.h0],
        d_counts.p, nullptr, nullptr, d_tried.p);
    DP_CUDA(ctx, cudaGetLastError());
    ++launches;
  }
  std::vector<uint32_t> counts(n_pairs);
  unsigned long long tried = 0;
  if (n_pairs) DP_CUDA(ctx, cudaMemcpyAsync(counts.data(), d_counts.p, n_pairs * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaMemcpyAsync(&tried, d_tried.p, sizeof tried, cudaMemcpyDeviceToHost, st));
  DP_CUDA(ctx, cudaStreamSynchronize(st));
  if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
  std::vector<uint64_t> offs(n_pairs + 1, 0);
  for (uint64_t i = 0; i < n_pairs; ++i) offs[i + 1] = offs[i] + counts[i];
  const uint64_t n_out = offs[n_pairs];
  dp_dimer* out = (dp_dimer*)malloc(std::max<uint64_t>(1, n_out) * sizeof(dp_dimer));
  if (!out) return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for %llu dimers", (unsigned long long)n_out);
  if (n_out) {
    cudaError_t e = d_out.alloc(n_out);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_off.p, offs.data(), (n_pairs + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, st);
    for (const Group& g : groups) {
      if (e != cudaSuccess) break;
      uint32_t V = var_count[g.primer];
      uint64_t n = g.n * V;
      if (!n) continue;
      dp_duplex_kernel<1><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(
          A, d_win.p, g.h0, g.n, V, primer_len[g.primer], d_codes.p + var_offset[g.primer], pair0[g.h0],
          nullptr, d_off.p, d_out.p, d_tried.p);
      e = cudaGetLastError();
      ++launches;
    }
    if (e == cudaSuccess) e = cudaEventRecord(ctx->ev2, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out.p, n_out * sizeof(dp_dimer), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { free(out); return dp_set_error(ctx, DP_ERR_CUDA, "duplex pass 2: %s", cudaGetErrorString(e)); }
  } else {
    DP_CUDA(ctx, cudaEventRecord(ctx->ev2, st));
    DP_CUDA(ctx, cudaStreamSynchronize(st));
  }
  *dimers_out = out;
  *n_dimers_out = n_out;
  if (stats) {
    stats->duplexes = n_pairs;
    stats->dimers_tried = tried;
    stats->dimers_kept = n_out;
    cudaEventElapsedTime(&stats->ms_kernel, ctx->ev1, ctx->ev2);
    cudaEventRecord(ctx->ev1, st);
    cudaEventSynchronize(ctx->ev1);
    cudaEventElapsedTime(&stats->ms_total, ctx->ev0, ctx->ev1);
    stats->launches = launches;
  }
  return DP_OK;
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
  DP_CUDA(ctx, d_g.alloc(n_hits));
  DP_CUDA(ctx, d_len.alloc(n_hits));
  DP_CUDA(ctx, d_strand.alloc(n_hits));
  DP_CUDA(ctx, d_win.alloc(n_hits * DP_WIN));
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
  DP_CUDA(ctx, d_codes.alloc(codes.size()));
  DP_CUDA(ctx, d_tm.alloc(n));
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
  DP_CUDA(ctx, d1.alloc(b1)); DP_CUDA(ctx, d2.alloc(b2));
  DP_CUDA(ctx, dl1.alloc(n)); DP_CUDA(ctx, dl2.alloc(n));
  DP_CUDA(ctx, do1.alloc(n)); DP_CUDA(ctx, do2.alloc(n));
  DP_CUDA(ctx, dm.alloc(n)); DP_CUDA(ctx, doff.alloc(n)); DP_CUDA(ctx, dout.alloc(n));
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
