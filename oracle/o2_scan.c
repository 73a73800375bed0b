/* ORACLE O2 in plain C (TEST INFRASTRUCTURE, see oracle/README.md).
 *
 * Brute-force restatement of what DegenPrimer/SearchEngine.py:190-211,321-323
 * computes through FFTs: for every window q of a strand,
 *     score3(q) = sum_j w3(P[j], t[q+j])      (thirds)
 * and q is a hit iff score3 >= 3*max(1, L-k) - 1.  Both strands are evaluated
 * on explicit strings (the reverse strand is materialised exactly like
 * reverse_complement(), DegenPrimer/SecStructures.py:50-55), so this shares no
 * trick with the CUDA kernels.  Used for templates too long for numpy O2.
 *
 * build: gcc -O3 -fopenmp -shared -fPIC oracle/o2_scan.c -o oracle/_build/libo2scan.so
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static int base_index(unsigned char c) {
  switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}

/* w3[letter][base A,C,G,T] */
static int letter_weights(unsigned char l, int w[4]) {
  static const char* letters = "ACGTRYSWKMBDHVN";
  static const int tab[15][4] = {
    {3,0,0,0},{0,3,0,0},{0,0,3,0},{0,0,0,3},
    {3,0,3,0} /*R=AG*/, {0,3,0,3} /*Y=CT*/, {0,2,2,0} /*S*/, {2,0,0,2} /*W*/,
    {0,0,3,3} /*K=GT*/, {3,3,0,0} /*M=AC*/,
    {0,2,2,3} /*B: T full*/, {2,0,3,2} /*D: G full*/, {2,3,0,2} /*H: C full*/, {3,2,2,0} /*V: A full*/,
    {2,2,2,2} /*N*/ };
  const char* p = strchr(letters, l);
  if (!p || !l) return -1;
  memcpy(w, tab[p - letters], sizeof(int) * 4);
  return 0;
}

static unsigned char complement(unsigned char c) {
  static const char* a = "ATGCRYSWKMBVDHN";
  static const char* b = "TACGYRSWMKVBHDN";
  const char* p = strchr(a, c);
  return (p && c) ? (unsigned char)b[p - a] : 0;
}

/* Scans one strand string.  Writes up to cap hit positions (ascending) to out,
 * returns the number of hits found (may exceed cap), or -1 on a bad letter. */
static int64_t scan_strand(const unsigned char* t, int64_t T, const unsigned char* primer, int L,
                           int k, int64_t* out, int64_t cap) {
  int lut[64][256];
  if (L > 64) return -1;
  for (int j = 0; j < L; ++j) {
    int w[4];
    if (letter_weights(primer[j], w)) return -1;
    for (int c = 0; c < 256; ++c) { int b = base_index((unsigned char)c); lut[j][c] = b < 0 ? 1 : w[b]; }
  }
  int need = L - k; if (need < 1) need = 1;
  const int thr = 3 * need - 1;
  const int64_t n = T - L + 1;
  if (n <= 0) return 0;
  /* every thread scans one contiguous stretch of window starts and keeps its own ascending hit
   * list; the lists are concatenated in thread order */
  int64_t found = 0;
  int failed = 0;
#pragma omp parallel
  {
#ifdef _OPENMP
    const int nt = omp_get_num_threads(), me = omp_get_thread_num();
#else
    const int nt = 1, me = 0;
#endif
    const int64_t lo = n * me / nt, hi = n * (me + 1) / nt;
    int64_t cap_l = 1024, n_l = 0;
    int64_t* mine = (int64_t*)malloc(sizeof(int64_t) * (size_t)cap_l);
    for (int64_t q = lo; mine && q < hi; ++q) {
      /* plain sum over the window; gives up once even full scores for the rest cannot reach thr */
      int s = 0, left = 3 * L;
      for (int j = 0; j < L; ++j) {
        s += lut[j][t[q + j]];
        left -= 3;
        if (s + left < thr) break;
      }
      if (s >= thr) {
        if (n_l == cap_l) {
          cap_l *= 2;
          int64_t* grown = (int64_t*)realloc(mine, sizeof(int64_t) * (size_t)cap_l);
          if (!grown) { free(mine); mine = NULL; break; }
          mine = grown;
        }
        mine[n_l++] = q;
      }
    }
    if (!mine) {
#pragma omp atomic write
      failed = 1;
    }
    /* ordered hand-over: thread i appends after threads 0..i-1 */
    for (int turn = 0; turn < nt; ++turn) {
      if (turn == me && mine) {
        for (int64_t i = 0; i < n_l; ++i) { if (found < cap) out[found] = mine[i]; ++found; }
      }
#pragma omp barrier
    }
    free(mine);
  }
  if (failed) return -2;
  return found;
}

/* fwd: the template (upper case).  Returns 0, or -1 bad letter, -2 no memory. */
int o2_scan(const unsigned char* fwd, int64_t T, const unsigned char* primer, int L, int k,
            int64_t* fwd_out, int64_t fwd_cap, int64_t* n_fwd,
            int64_t* rev_out, int64_t rev_cap, int64_t* n_rev) {
  unsigned char* rev = (unsigned char*)malloc((size_t)(T > 0 ? T : 1));
  if (!rev) return -2;
  unsigned char comp[256];
  for (int c = 0; c < 256; ++c) comp[c] = complement((unsigned char)c);
  int bad = 0;
#pragma omp parallel for schedule(static) reduction(|:bad)
  for (int64_t i = 0; i < T; ++i) {
    unsigned char c = comp[fwd[T - 1 - i]];
    if (!c) bad = 1;
    rev[i] = c;
  }
  if (bad) { free(rev); return -1; }
  int64_t a = scan_strand(fwd, T, primer, L, k, fwd_out, fwd_cap);
  int64_t b = a < 0 ? a : scan_strand(rev, T, primer, L, k, rev_out, rev_cap);
  free(rev);
  if (a < 0 || b < 0) return (int)(a < 0 ? a : b);
  *n_fwd = a; *n_rev = b;
  return 0;
}
