// Template loader: ASCII / FASTA -> three bit-planes resident in HBM.
//
// Replaces BioUtils/BioUtils/SeqUtils.py:103-271 (SeqView over SeqIO.index_db)
// and the str()/upper()/reverse_complement() host copies of
// DegenPrimer/SearchEngine.py:287-290,333-336.  The ASCII bytes are streamed to
// the device through two pinned staging buffers (or straight from the caller's
// buffer when that is already pinned) and packed there; the host never touches
// a base twice.
#include <algorithm>
#include <stdexcept>
#include <stdlib.h>
#include <string.h>

#include "dp_internal.cuh"

#define DP_PAD_BYTE 0xFFu
#define DP_STAGE_BYTES (16u << 20)   // per staging buffer, multiple of 32
#define DP_GUARD_WORDS 8             // readable plane words past the data

// ---- pack kernel --------------------------------------------------------------
// One thread packs 32 consecutive bases into one word of each plane.
//   src        chunk of ASCII, src[0] is base 32*word0 of the padded layout
//   n_valid    bytes of src that exist; everything after is padding
//   prev_byte  the byte in front of src[0] (DP_PAD_BYTE at the very start)
//   bounds     class-change list: entry = pos*16 + class, class 0 = ACGT,
//              4..14 = IUPAC letter index, 15 = padding
//   counters   [0] entries wanted (may exceed cap), [1] min position of a
//              byte outside the alphabet (init UINT64_MAX)
__device__ __forceinline__ uint32_t base_class(uint32_t c) {
  if (c == DP_PAD_BYTE) return 15u;
  uint32_t u = c & 0xDFu;
  if (u == 'A' || u == 'C' || u == 'G' || u == 'T') return 0u;
  uint32_t l = dp_letter_of((unsigned char)c);
  return l == DP_NOT_A_LETTER ? 16u : l;   // 16 = foreign byte
}

__global__ void __launch_bounds__(256)
dp_pack_kernel(const uint8_t* __restrict__ src, uint64_t n_valid, uint64_t n_words,
               uint64_t word0, uint32_t prev_byte,
               uint32_t* __restrict__ lo, uint32_t* __restrict__ hi, uint32_t* __restrict__ exc,
               unsigned long long* __restrict__ bounds, uint64_t bounds_cap,
               unsigned long long* __restrict__ counters) {
  uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= n_words) return;
  uint64_t i0 = w * 32;
  uint32_t bytes[8];
  if (i0 + 32 <= n_valid) {
    const uint4* p = reinterpret_cast<const uint4*>(src + i0);
    uint4 a = __ldg(p), b = __ldg(p + 1);
    bytes[0] = a.x; bytes[1] = a.y; bytes[2] = a.z; bytes[3] = a.w;
    bytes[4] = b.x; bytes[5] = b.y; bytes[6] = b.z; bytes[7] = b.w;
  } else {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      uint32_t v = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint64_t idx = i0 + 4 * k + j;
        uint32_t c = idx < n_valid ? src[idx] : DP_PAD_BYTE;
        v |= c << (8 * j);
      }
      bytes[k] = v;
    }
  }
  uint32_t pc = (i0 == 0) ? prev_byte : (i0 - 1 < n_valid ? (uint32_t)src[i0 - 1] : DP_PAD_BYTE);
  uint32_t prev_cls = base_class(pc);
  if (prev_cls == 16u) prev_cls = 15u;
  uint32_t wl = 0, wh = 0, we = 0;
  uint64_t g0 = (word0 + w) * 32;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int b = 4 * k + j;
      uint32_t c = (bytes[k] >> (8 * j)) & 0xFFu;
      uint32_t cls = base_class(c);
      uint32_t code;
      if (cls == 0u) {
        code = (c >> 1) & 3u;
      } else {
        if (cls == 16u) {
          atomicMin(&counters[1], (unsigned long long)(g0 + b));
          cls = 15u;
        }
        code = dp_filler(g0 + b);
        we |= 1u << b;
      }
      wl |= (code & 1u) << b;
      wh |= (code >> 1) << b;
      if (cls != prev_cls) {
        unsigned long long slot = atomicAdd(&counters[0], 1ull);
        if (slot < bounds_cap) bounds[slot] = (unsigned long long)(g0 + b) * 16ull + cls;
        prev_cls = cls;
      }
    }
  }
  lo[word0 + w] = wl;
  hi[word0 + w] = wh;
  exc[word0 + w] = we;
}

__global__ void dp_fill_guard_kernel(uint32_t* lo, uint32_t* hi, uint32_t* exc, uint64_t from, uint64_t to) {
  uint64_t w = from + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < to) { lo[w] = 0; hi[w] = 0; exc[w] = 0xFFFFFFFFu; }
}

// ---- exception plane from the tables ----------------------------------------------
// The exception plane is redundant with the record table and the exception-run table: a bit is set
// where the position lies outside every record (padding, spacers, tail guard) or inside a run of
// non-ACGT letters.  dp_template_load_mem rebuilds it here for an image in page-locked memory instead
// of moving it over PCIe (a third of the image).
__global__ void __launch_bounds__(256)
dp_exc_records_kernel(uint32_t* __restrict__ exc, uint64_t n_words_alloc, const uint64_t* __restrict__ rec_start,
                      const uint64_t* __restrict__ rec_len, uint32_t n_records) {
  const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= n_words_alloc) return;
  const uint64_t g = w * 32;
  uint32_t a = 0, b = n_records;            // records with start <= g
  while (a < b) {
    const uint32_t m = (a + b) >> 1;
    if (rec_start[m] <= g) a = m + 1; else b = m;
  }
  uint32_t mask = 0xFFFFFFFFu;
  if (a) {
    const uint64_t end = rec_start[a - 1] + rec_len[a - 1];      // records start on word boundaries
    if (g + 32 <= end) mask = 0u;
    else if (g < end) mask = 0xFFFFFFFFu << (uint32_t)(end - g);
  }
  exc[w] = mask;
}

// one warp per run; whole words are stored, the two edge words are OR-ed in (a neighbouring run or
// the end of the record may share them)
__global__ void __launch_bounds__(256)
dp_exc_runs_kernel(uint32_t* __restrict__ exc, const dp_exc_run* __restrict__ runs, uint64_t n_runs) {
  const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_runs) return;
  const uint32_t lane = threadIdx.x & 31u;
  const uint64_t s = runs[r].start, e = runs[r].end;           // [s, e)
  const uint64_t w0 = s >> 5, w1 = (e - 1) >> 5;
  for (uint64_t w = w0 + lane; w <= w1; w += 32) {
    uint32_t mask = 0xFFFFFFFFu;
    if (w == w0) mask &= 0xFFFFFFFFu << (uint32_t)(s & 31);
    if (w == w1) mask &= 0xFFFFFFFFu >> (31u - (uint32_t)((e - 1) & 31));
    if (mask == 0xFFFFFFFFu) exc[w] = mask; else atomicOr(exc + w, mask);
  }
}

// ---- unpack kernel ------------------------------------------------------------
__device__ __forceinline__ uint32_t dp_run_letter(const dp_exc_run* runs, uint64_t n_runs, uint64_t g) {
  uint64_t a = 0, b = n_runs;            // last run with start <= g
  while (a < b) {
    uint64_t m = (a + b) >> 1;
    if (runs[m].start <= g) a = m + 1; else b = m;
  }
  if (a == 0) return 14u;
  const dp_exc_run& r = runs[a - 1];
  return g < r.end ? r.letter : 14u;
}

__global__ void dp_unpack_kernel(const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi,
                                 const uint32_t* __restrict__ exc, const dp_exc_run* __restrict__ runs,
                                 uint64_t n_runs, uint64_t rec_start, uint64_t rec_len,
                                 uint64_t begin, uint64_t n, int reverse, char* __restrict__ out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t p = reverse ? rec_len - 1 - (begin + i) : begin + i;
  uint64_t g = rec_start + p;
  uint32_t w = (uint32_t)(g & 31);
  uint64_t wi = g >> 5;
  uint32_t letter;
  if ((exc[wi] >> w) & 1u) letter = dp_run_letter(runs, n_runs, g);
  else letter = ((hi[wi] >> w) & 1u) * 2u + ((lo[wi] >> w) & 1u);
  if (reverse) letter = dp_letter_complement(letter);
  out[i] = DP_LETTERS[letter];
}

// ---- host side ------------------------------------------------------------------
static int ensure_stage(dp_ctx* ctx) {
  if (ctx->stage_bytes) return DP_OK;
  for (int i = 0; i < 2; ++i) {
    DP_CUDA(ctx, cudaMallocHost(&ctx->h_stage[i], DP_STAGE_BYTES));
    DP_CUDA(ctx, cudaMalloc(&ctx->d_stage[i], DP_STAGE_BYTES));
    DP_CUDA(ctx, cudaEventCreateWithFlags(&ctx->stage_ev[i], cudaEventDisableTiming));
    DP_CUDA(ctx, cudaEventCreateWithFlags(&ctx->stage_copied[i], cudaEventDisableTiming));
  }
  ctx->stage_bytes = DP_STAGE_BYTES;
  return DP_OK;
}

static bool is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost;
}

static void free_template_device(dp_template* t) {
  // stream-ordered: safe while earlier kernels on ctx->stream still read the planes
  dp_ctx* ctx = t->ctx;
  dp_pool_free(ctx, t->d_lo);
  dp_pool_free(ctx, t->d_hi);
  dp_pool_free(ctx, t->d_exc);
  dp_pool_free(ctx, t->d_rec_start);
  dp_pool_free(ctx, t->d_rec_len);
  dp_pool_free(ctx, t->d_runs);
  t->d_lo = t->d_hi = t->d_exc = nullptr;
  t->d_rec_start = t->d_rec_len = nullptr;
  t->d_runs = nullptr;
}

// Streams the records of `t` (layout already decided: t->records[r].start/len)
// to the device and packs them.  One attempt with a given boundary capacity.
// Chunks travel on ctx->copy_stream while the previous chunk is packed on
// ctx->stream (two staging slots, events both ways); pinned sources are never
// touched by the host.
static int pack_once(dp_ctx* ctx, dp_template* t, const char* const* seqs, uint64_t bounds_cap,
                     unsigned long long* d_bounds, unsigned long long* h_counters) {
  const uint32_t n_records = (uint32_t)t->records.size();
  unsigned long long init[2] = {0ull, ~0ull};
  unsigned long long* d_cnt = ctx->d_counters + DP_C_PACK0;
  DP_CUDA(ctx, cudaMemcpyAsync(d_cnt, init, sizeof init, cudaMemcpyHostToDevice, ctx->stream));
  // the copy stream must not overwrite a staging slot (or start) before what is
  // already queued on the main stream has finished with it
  DP_CUDA(ctx, cudaEventRecord(ctx->stage_ev[0], ctx->stream));
  DP_CUDA(ctx, cudaEventRecord(ctx->stage_ev[1], ctx->stream));
  int slot = 0;
  bool h_busy[2] = {false, false};   // h_stage[slot] has an H2D in flight
  uint64_t chunk_word0 = 0;       // padded word offset of the current gather chunk
  size_t chunk_fill = 0;          // bytes placed so far (padded layout)
  uint32_t chunk_prev = DP_PAD_BYTE;
  char* hbuf = nullptr;
  // src (pinned host memory) -> d_stage[slot] -> planes
  auto ship = [&](const char* src, uint64_t n_bytes, uint64_t word0, uint32_t prev, bool from_stage) -> int {
    DP_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->stage_ev[slot], 0));
    DP_CUDA(ctx, cudaMemcpyAsync(ctx->d_stage[slot], src, n_bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    DP_CUDA(ctx, cudaEventRecord(ctx->stage_copied[slot], ctx->copy_stream));
    if (from_stage) h_busy[slot] = true;
    DP_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->stage_copied[slot], 0));
    uint64_t nw = (n_bytes + 31) / 32;
    dp_pack_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, ctx->stream>>>(
        (const uint8_t*)ctx->d_stage[slot], n_bytes, nw, word0, prev,
        t->d_lo, t->d_hi, t->d_exc, d_bounds, bounds_cap, d_cnt);
    DP_CUDA(ctx, cudaGetLastError());
    DP_CUDA(ctx, cudaEventRecord(ctx->stage_ev[slot], ctx->stream));
    slot ^= 1;
    return DP_OK;
  };
  auto flush_gather = [&](void) -> int {
    if (!chunk_fill) return DP_OK;
    size_t n_bytes = (chunk_fill + 31) & ~(size_t)31;
    memset(hbuf + chunk_fill, DP_PAD_BYTE, n_bytes - chunk_fill);
    uint32_t last = (unsigned char)hbuf[n_bytes - 1];
    int rc = ship(hbuf, n_bytes, chunk_word0, chunk_prev, true);
    if (rc) return rc;
    chunk_word0 += n_bytes / 32;
    chunk_fill = 0;
    chunk_prev = last;
    hbuf = nullptr;
    return DP_OK;
  };
  auto acquire = [&](void) -> int {
    if (h_busy[slot]) { DP_CUDA(ctx, cudaEventSynchronize(ctx->stage_copied[slot])); h_busy[slot] = false; }
    hbuf = ctx->h_stage[slot];
    return DP_OK;
  };
  for (uint32_t r = 0; r < n_records; ++r) {
    if (ctx->abort_flag) return dp_set_error(ctx, DP_ERR_ABORTED, "aborted");
    const dp_record& rec = t->records[r];
    const char* src = seqs[r];
    uint64_t rec_words = (rec.len + 31) / 32;
    if (rec.len >= (1u << 20) && is_pinned(src)) {
      int rc = flush_gather();
      if (rc) return rc;
      uint64_t done = 0;
      uint32_t prev = DP_PAD_BYTE;
      while (done < rec.len) {
        uint64_t n = std::min<uint64_t>(DP_STAGE_BYTES, rec.len - done);
        rc = ship(src + done, n, rec.start / 32 + done / 32, prev, false);
        if (rc) return rc;
        prev = (unsigned char)src[done + n - 1];
        done += n;
      }
      chunk_word0 = rec.start / 32 + rec_words;
      chunk_prev = DP_PAD_BYTE;
      continue;
    }
    // gather path
    if (chunk_fill == 0) { chunk_word0 = rec.start / 32; }
    uint64_t done = 0;
    while (done < rec.len || (rec.len == 0 && done == 0)) {
      if (!hbuf) { int rc = acquire(); if (rc) return rc; }
      // position of the next byte inside the chunk (padded layout)
      size_t at = (size_t)((rec.start + done) - chunk_word0 * 32);
      if (at > chunk_fill) { memset(hbuf + chunk_fill, DP_PAD_BYTE, at - chunk_fill); chunk_fill = at; }
      size_t room = DP_STAGE_BYTES - chunk_fill;
      size_t n = (size_t)std::min<uint64_t>(room, rec.len - done);
      memcpy(hbuf + chunk_fill, src + done, n);
      chunk_fill += n;
      done += n;
      if (chunk_fill == DP_STAGE_BYTES) { int rc = flush_gather(); if (rc) return rc; }
      if (rec.len == 0) break;
    }
  }
  {
    if (chunk_fill && !hbuf) return dp_set_error(ctx, DP_ERR_STATE, "internal: gather buffer lost");
    int rc = flush_gather();
    if (rc) return rc;
  }
  uint64_t guard_to = t->n_words_alloc;
  uint64_t nguard = guard_to - t->n_words;
  dp_fill_guard_kernel<<<(unsigned)((nguard + 255) / 256), 256, 0, ctx->stream>>>(t->d_lo, t->d_hi, t->d_exc, t->n_words, guard_to);
  DP_CUDA(ctx, cudaGetLastError());
  DP_CUDA(ctx, cudaMemcpyAsync(ctx->h_res, d_cnt, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(h_counters, ctx->h_res, 2 * sizeof(unsigned long long));
  return DP_OK;
}

int dp_upload_and_pack(dp_ctx* ctx, dp_template* t, uint32_t n_records,
                       const char* const* seqs, const uint64_t* lens) {
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = ensure_stage(ctx);
  if (rc) return rc;
  // layout: every record starts on a word boundary
  uint64_t w = 0;
  t->total_bases = 0;
  for (uint32_t r = 0; r < n_records; ++r) {
    t->records[r].start = w * 32;
    t->records[r].len = lens[r];
    w += (lens[r] + 31) / 32;
    t->total_bases += lens[r];
  }
  t->n_words = w;
  t->n_words_alloc = w + DP_GUARD_WORDS;
  size_t plane_bytes = (size_t)t->n_words_alloc * sizeof(uint32_t);
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_lo, plane_bytes));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_hi, plane_bytes));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_exc, plane_bytes));
  t->device_bytes = 3 * plane_bytes;

  uint64_t cap = std::max<uint64_t>(1u << 16, t->total_bases / 64);
  std::vector<unsigned long long> bounds;
  for (int attempt = 0;; ++attempt) {
    unsigned long long* d_bounds = nullptr;
    DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_bounds, cap * sizeof(unsigned long long)));
    unsigned long long h_counters[2] = {0, 0};
    rc = pack_once(ctx, t, seqs, cap, d_bounds, h_counters);
    if (rc) { dp_pool_free(ctx, d_bounds); return rc; }
    if (h_counters[1] != ~0ull) {
      dp_pool_free(ctx, d_bounds);
      uint64_t g = h_counters[1];
      uint32_t r = 0;
      while (r + 1 < n_records && t->records[r + 1].start <= g) ++r;
      uint64_t off = g - t->records[r].start;
      unsigned char c = off < lens[r] ? (unsigned char)seqs[r][off] : 0;
      return dp_set_error(ctx, DP_ERR_ALPHABET, "record %u offset %llu: byte 0x%02x ('%c') is not an IUPAC DNA letter",
                          r, (unsigned long long)off, c, (c >= 32 && c < 127) ? c : '?');
    }
    if (h_counters[0] > cap) {
      if (attempt) { dp_pool_free(ctx, d_bounds); return dp_set_error(ctx, DP_ERR_STATE, "internal: boundary list overflow twice"); }
      dp_pool_free(ctx, d_bounds);
      cap = h_counters[0];
      continue;
    }
    bounds.resize(h_counters[0]);
    if (!bounds.empty())
      DP_CUDA(ctx, cudaMemcpy(bounds.data(), d_bounds, bounds.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    dp_pool_free(ctx, d_bounds);
    break;
  }
  std::sort(bounds.begin(), bounds.end());
  bounds.push_back((unsigned long long)(t->n_words * 32) * 16ull);   // sentinel closes an open run
  t->runs.clear();
  for (size_t i = 0; i + 1 < bounds.size(); ++i) {
    uint32_t cls = (uint32_t)(bounds[i] & 15ull);
    if (cls == 0u || cls == 15u) continue;
    dp_exc_run run;
    run.start = bounds[i] >> 4;
    run.end = bounds[i + 1] >> 4;
    run.letter = cls;
    run.pad = 0;
    if (run.end > run.start) t->runs.push_back(run);
  }
  size_t nr = t->runs.size();
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_runs, std::max<size_t>(1, nr) * sizeof(dp_exc_run)));
  std::vector<uint64_t> starts(n_records), rl(n_records);
  for (uint32_t r = 0; r < n_records; ++r) { starts[r] = t->records[r].start; rl[r] = t->records[r].len; }
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_rec_start, std::max<size_t>(1, n_records) * sizeof(uint64_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_rec_len, std::max<size_t>(1, n_records) * sizeof(uint64_t)));
  if (nr) DP_CUDA(ctx, cudaMemcpyAsync(t->d_runs, t->runs.data(), nr * sizeof(dp_exc_run), cudaMemcpyHostToDevice, ctx->stream));
  if (n_records) {
    DP_CUDA(ctx, cudaMemcpyAsync(t->d_rec_start, starts.data(), n_records * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    DP_CUDA(ctx, cudaMemcpyAsync(t->d_rec_len, rl.data(), n_records * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
  }
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));    // starts/rl/runs are pageable host memory
  t->device_bytes += nr * sizeof(dp_exc_run) + 2 * n_records * sizeof(uint64_t);
  return DP_OK;
}

// live templates by serial number: a context that keeps the hits of its last dp_find resident
// (dp_find_again) must not touch a template that has been freed since
#include <mutex>
#include <unordered_set>
static std::mutex g_live_mutex;
static std::unordered_set<uint64_t> g_live;
static uint64_t g_next_uid = 1;

bool dp_template_alive(uint64_t uid) {
  std::lock_guard<std::mutex> lock(g_live_mutex);
  return g_live.count(uid) != 0;
}

static void delete_template(dp_template* t) {
  {
    std::lock_guard<std::mutex> lock(g_live_mutex);
    g_live.erase(t->uid);
  }
  delete t;
}

static dp_template* new_template(dp_ctx* ctx, uint32_t n_records) {
  dp_template* t = new dp_template();
  {
    std::lock_guard<std::mutex> lock(g_live_mutex);
    t->uid = g_next_uid++;
    g_live.insert(t->uid);
  }
  t->ctx = ctx;
  t->n_words = t->n_words_alloc = 0;
  t->d_lo = t->d_hi = t->d_exc = nullptr;
  t->d_rec_start = t->d_rec_len = nullptr;
  t->d_runs = nullptr;
  t->total_bases = 0;
  t->device_bytes = 0;
  t->records.resize(n_records);
  return t;
}

extern "C" {

int dp_template_from_records(dp_ctx* ctx, uint32_t n_records, const char* const* seqs,
                             const uint64_t* lens, dp_template** out) {
  if (!ctx || !out || (n_records && (!seqs || !lens))) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *out = nullptr;
  dp_template* t = new_template(ctx, n_records);
  for (uint32_t r = 0; r < n_records; ++r) {
    char name[32];
    snprintf(name, sizeof name, "record_%u", r);
    t->records[r].id = name;
  }
  int rc = dp_upload_and_pack(ctx, t, n_records, seqs, lens);
  if (rc) { free_template_device(t); delete_template(t); return rc; }
  *out = t;
  return DP_OK;
}

int dp_template_from_ascii(dp_ctx* ctx, const char* seq, uint64_t len, dp_template** out) {
  const char* seqs[1] = {seq};
  uint64_t lens[1] = {len};
  return dp_template_from_records(ctx, 1, seqs, lens, out);
}

static int template_from_fasta(dp_ctx* ctx, const char* path, dp_template** out);

int dp_template_from_fasta(dp_ctx* ctx, const char* path, dp_template** out) {
  if (!ctx || !path || !out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *out = nullptr;
  try {                                   // no C++ exception may cross the C ABI (a file larger than memory ...)
    return template_from_fasta(ctx, path, out);
  } catch (const std::exception& e) {
    return dp_set_error(ctx, DP_ERR_NOMEM, "%s: %s", path, e.what());
  }
}

}  // extern "C"

static int template_from_fasta(dp_ctx* ctx, const char* path, dp_template** out) {
  FILE* fh = fopen(path, "rb");
  if (!fh) return dp_set_error(ctx, DP_ERR_IO, "cannot open %s", path);
  fseek(fh, 0, SEEK_END);
  long fsize = ftell(fh);
  fseek(fh, 0, SEEK_SET);
  std::string data;
  data.resize(fsize > 0 ? (size_t)fsize : 0);
  if (fsize > 0 && fread(&data[0], 1, (size_t)fsize, fh) != (size_t)fsize) {
    fclose(fh);
    return dp_set_error(ctx, DP_ERR_IO, "short read on %s", path);
  }
  fclose(fh);
  struct Rec { std::string id, descr, seq; };
  std::vector<Rec> recs;
  size_t p = 0, n = data.size();
  while (p < n) {
    size_t eol = data.find('\n', p);
    if (eol == std::string::npos) eol = n;
    size_t end = eol;
    while (end > p && (data[end - 1] == '\r' || data[end - 1] == ' ' || data[end - 1] == '\t')) --end;
    if (end > p && data[p] == '>') {
      Rec r;
      r.descr = data.substr(p + 1, end - p - 1);
      size_t sp = r.descr.find_first_of(" \t");
      r.id = sp == std::string::npos ? r.descr : r.descr.substr(0, sp);
      recs.push_back(r);
    } else if (end > p && !recs.empty()) {
      std::string& s = recs.back().seq;
      for (size_t i = p; i < end; ++i) {
        char c = data[i];
        if (c != ' ' && c != '\t') s.push_back(c);
      }
    }
    p = eol + 1;
  }
  data.clear();
  data.shrink_to_fit();
  if (recs.empty()) return dp_set_error(ctx, DP_ERR_IO, "no FASTA records in %s", path);
  std::sort(recs.begin(), recs.end(), [](const Rec& a, const Rec& b) { return a.id < b.id; });
  for (size_t i = 1; i < recs.size(); ++i)
    if (recs[i].id == recs[i - 1].id) return dp_set_error(ctx, DP_ERR_ARG, "duplicate record id '%s' in %s", recs[i].id.c_str(), path);
  uint32_t nr = (uint32_t)recs.size();
  std::vector<const char*> seqs(nr);
  std::vector<uint64_t> lens(nr);
  dp_template* t = new_template(ctx, nr);
  for (uint32_t r = 0; r < nr; ++r) {
    seqs[r] = recs[r].seq.data();
    lens[r] = recs[r].seq.size();
    t->records[r].id = recs[r].id;
    t->records[r].descr = recs[r].descr;
  }
  int rc = dp_upload_and_pack(ctx, t, nr, seqs.data(), lens.data());
  if (rc) { free_template_device(t); delete_template(t); return rc; }
  *out = t;
  return DP_OK;
}

extern "C" {

void dp_template_free(dp_template* t) {
  if (!t) return;
  if (t->ctx) cudaSetDevice(t->ctx->device);
  free_template_device(t);
  delete_template(t);
}

int dp_template_get_info(const dp_template* t, dp_template_info* info) {
  if (!t || !info) return DP_ERR_ARG;
  info->n_records = (uint32_t)t->records.size();
  info->reserved = 0;
  info->total_bases = t->total_bases;
  info->padded_bases = t->n_words * 32;
  info->n_exception_runs = t->runs.size();
  info->device_bytes = t->device_bytes;
  return DP_OK;
}

int dp_template_record(const dp_template* t, uint32_t r, uint64_t* start, uint64_t* len,
                       const char** id, const char** descr) {
  if (!t || r >= t->records.size()) return DP_ERR_ARG;
  const dp_record& rec = t->records[r];
  if (start) *start = rec.start;
  if (len) *len = rec.len;
  if (id) *id = rec.id.c_str();
  if (descr) *descr = rec.descr.c_str();
  return DP_OK;
}

int dp_template_unpack(dp_ctx* ctx, const dp_template* t, uint32_t r, uint64_t begin, uint64_t n,
                       int reverse, char* out) {
  if (!ctx || !t || !out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (r >= t->records.size()) return dp_set_error(ctx, DP_ERR_ARG, "record %u out of range", r);
  const dp_record& rec = t->records[r];
  if (begin > rec.len || n > rec.len - begin) return dp_set_error(ctx, DP_ERR_ARG, "range outside record");
  if (n == 0) return DP_OK;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  char* d_out = nullptr;
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&d_out, n));
  dp_unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(
      t->d_lo, t->d_hi, t->d_exc, t->d_runs, t->runs.size(), rec.start, rec.len, begin, n, reverse, d_out);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, n, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  dp_pool_free(ctx, d_out);
  if (e != cudaSuccess) return dp_set_error(ctx, DP_ERR_CUDA, "unpack: %s", cudaGetErrorString(e));
  return DP_OK;
}

}  // extern "C"

// ---- on-disk packed template ---------------------------------------------
// Replaces the SQLite index the reference builds for every run (SeqIO.index_db into a temp
// file, BioUtils/SeqUtils.py:140-172) and the parse + pack of the FASTA text: the three
// bit-planes, the exception-run table and the record table exactly as they sit in HBM.
//   header  64 B : magic "DPB200PK", u32 version, u32 n_records, u64 n_words, u64 n_words_alloc,
//                  u64 total_bases, u64 n_runs, u64 names_bytes, u64 checksum (of all that follows)
//   records      : n_records x {u64 start, u64 len, u32 id_len, u32 descr_len}
//   names        : id and description bytes of every record, back to back
//   runs         : n_runs x dp_exc_run {u64 start, u64 end, u32 letter, u32 pad}
//   planes       : lo, hi, exc — n_words_alloc little-endian u32 each (tail guard included)
// 0.375 B/bp instead of >= 1 B/bp of text, nothing to parse, and the load is one H2D stream.
#define DP_PACK_MAGIC "DPB200PK"
#define DP_PACK_VERSION 3u

struct dp_pack_header {
  char magic[8];
  uint32_t version;
  uint32_t n_records;
  uint64_t n_words;
  uint64_t n_words_alloc;
  uint64_t total_bases;
  uint64_t n_runs;
  uint64_t names_bytes;
  uint64_t checksum;
};
static_assert(sizeof(dp_pack_header) == 64, "packed-template header is 64 bytes");

struct dp_pack_record {
  uint64_t start, len;
  uint32_t id_len, descr_len;
};

// Checksum: the six sections after the header, each zero-padded to a multiple of 8 bytes, read as
// one sequence of little-endian 64-bit words w[0..N); lane l = i & 3 takes the words i == l (mod 4)
// (four independent multiply chains, so the sum runs at memory speed instead of at the latency of
// one multiply per word); h_l starts at SEED + l, h = (h ^ w) * P, h ^= h >> 29; the lanes are
// folded at the end.  Only the last call of a section may pass n that is not a multiple of 8.
#define DP_PACK_SEED 0xCBF29CE484222325ull
#define DP_PACK_PRIME 0x100000001B3ull
struct dp_pack_sum {
  uint64_t h[4];
  uint64_t i;      // words taken so far
  dp_pack_sum() : i(0) { for (int l = 0; l < 4; ++l) h[l] = DP_PACK_SEED + (uint64_t)l; }
  static inline uint64_t step(uint64_t h, uint64_t w) { h = (h ^ w) * DP_PACK_PRIME; return h ^ (h >> 29); }
  void mix(const void* data, size_t n) {
    const unsigned char* p = (const unsigned char*)data;
    size_t k = 0;
    while ((i & 3) && k + 8 <= n) { uint64_t w; memcpy(&w, p + k, 8); h[i & 3] = step(h[i & 3], w); ++i; k += 8; }
    uint64_t a = h[0], b = h[1], c = h[2], d = h[3];
    for (; k + 32 <= n; k += 32) {
      uint64_t w[4];
      memcpy(w, p + k, 32);
      a = step(a, w[0]); b = step(b, w[1]); c = step(c, w[2]); d = step(d, w[3]);
      i += 4;
    }
    h[0] = a; h[1] = b; h[2] = c; h[3] = d;
    for (; k + 8 <= n; k += 8) { uint64_t w; memcpy(&w, p + k, 8); h[i & 3] = step(h[i & 3], w); ++i; }
    if (k < n) { uint64_t w = 0; memcpy(&w, p + k, n - k); h[i & 3] = step(h[i & 3], w); ++i; }
  }
  uint64_t value() const { uint64_t r = h[0]; for (int l = 1; l < 4; ++l) r = (r * DP_PACK_PRIME) ^ h[l]; return r; }
};

// Builds the file image of a template: header, tables and (when `planes` is given) the three planes
// copied from the device straight to their place.  `image` must hold dp_template_image_size() bytes.
static uint64_t image_bytes(const dp_template* t) {
  uint64_t names = 0;
  for (const dp_record& x : t->records) names += x.id.size() + x.descr.size();
  names += t->origin.size();
  return sizeof(dp_pack_header) + t->records.size() * sizeof(dp_pack_record) + names +
         t->runs.size() * sizeof(dp_exc_run) + 3ull * t->n_words_alloc * 4ull;
}

static int template_save_to(dp_ctx* ctx, const dp_template* t, char* image) {
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint32_t nrec = (uint32_t)t->records.size();
  dp_pack_header h;
  memset(&h, 0, sizeof h);
  memcpy(h.magic, DP_PACK_MAGIC, 8);
  h.version = DP_PACK_VERSION;
  h.n_records = nrec;
  h.n_words = t->n_words;
  h.n_words_alloc = t->n_words_alloc;
  h.total_bases = t->total_bases;
  h.n_runs = t->runs.size();
  char* p = image + sizeof h;
  dp_pack_record* recs = (dp_pack_record*)p;
  p += nrec * sizeof(dp_pack_record);
  char* names = p;
  for (uint32_t r = 0; r < nrec; ++r) {
    const dp_record& x = t->records[r];
    dp_pack_record pr;
    memset(&pr, 0, sizeof pr);
    pr.start = x.start;
    pr.len = x.len;
    pr.id_len = (uint32_t)x.id.size();
    pr.descr_len = (uint32_t)x.descr.size();
    memcpy(&recs[r], &pr, sizeof pr);
    memcpy(p, x.id.data(), x.id.size());
    p += x.id.size();
    memcpy(p, x.descr.data(), x.descr.size());
    p += x.descr.size();
  }
  // the origin string (what the image was packed from) rides behind the record names
  memcpy(p, t->origin.data(), t->origin.size());
  p += t->origin.size();
  h.names_bytes = (uint64_t)(p - names);
  char* runs = p;
  if (!t->runs.empty()) memcpy(runs, t->runs.data(), t->runs.size() * sizeof(dp_exc_run));
  p += t->runs.size() * sizeof(dp_exc_run);
  const size_t plane_bytes = (size_t)t->n_words_alloc * 4;
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  DP_CUDA(ctx, cudaMemcpy(p, t->d_lo, plane_bytes, cudaMemcpyDeviceToHost));
  DP_CUDA(ctx, cudaMemcpy(p + plane_bytes, t->d_hi, plane_bytes, cudaMemcpyDeviceToHost));
  DP_CUDA(ctx, cudaMemcpy(p + 2 * plane_bytes, t->d_exc, plane_bytes, cudaMemcpyDeviceToHost));
  dp_pack_sum c;
  c.mix(&h, sizeof h);                                      // checksum field still zero
  c.mix(recs, nrec * sizeof(dp_pack_record));
  c.mix(names, (size_t)h.names_bytes);
  c.mix(runs, t->runs.size() * sizeof(dp_exc_run));
  for (int pl = 0; pl < 3; ++pl) c.mix(p + (size_t)pl * plane_bytes, plane_bytes);   // one section per plane
  h.checksum = c.value();
  memcpy(image, &h, sizeof h);
  return DP_OK;
}

extern "C" int dp_template_image_size(const dp_template* t, uint64_t* bytes) {
  if (!t || !bytes) return DP_ERR_ARG;
  *bytes = image_bytes(t);
  return DP_OK;
}

extern "C" int dp_template_save_mem(dp_ctx* ctx, const dp_template* t, void* image, uint64_t capacity) {
  if (!ctx || !t || !image) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  if (capacity < image_bytes(t))
    return dp_set_error(ctx, DP_ERR_ARG, "image buffer holds %llu bytes, the template needs %llu",
                        (unsigned long long)capacity, (unsigned long long)image_bytes(t));
  return template_save_to(ctx, t, (char*)image);
}

extern "C" int dp_template_save(dp_ctx* ctx, const dp_template* t, const char* path) {
  if (!ctx || !t || !path) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  std::vector<char> image;
  try {
    image.resize((size_t)image_bytes(t));
  } catch (const std::exception&) {
    return dp_set_error(ctx, DP_ERR_NOMEM, "out of host memory for the %llu-byte image of the template", (unsigned long long)image_bytes(t));
  }
  int rc = template_save_to(ctx, t, image.data());
  if (rc) return rc;
  std::string tmp = std::string(path) + ".tmp";
  FILE* fh = fopen(tmp.c_str(), "wb");
  if (!fh) return dp_set_error(ctx, DP_ERR_IO, "cannot create %s", tmp.c_str());
  bool ok = fwrite(image.data(), 1, image.size(), fh) == image.size();
  ok = (fclose(fh) == 0) && ok;
  if (!ok || rename(tmp.c_str(), path) != 0) {
    remove(tmp.c_str());
    return dp_set_error(ctx, DP_ERR_IO, "cannot write %s", path);
  }
  return DP_OK;
}

// A packed template is read from a file or from an image of that file in host memory
// (dp_template_load_mem: the cache kept resident by a long-running caller, or mmap()ed).
struct dp_pack_source {
  FILE* fh;                  // file source, or
  const char* mem;           // memory source
  uint64_t size, at;
  const char* name;
  bool read(void* dst, size_t n) {
    if (n > size - at) return false;
    if (fh) { if (n && fread(dst, 1, n, fh) != n) return false; }
    else if (n) memcpy(dst, mem + at, n);
    at += n;
    return true;
  }
};

// header arithmetic on untrusted input: every term is bounded before it is used as a size
static bool pack_expected_size(const dp_pack_header& h, uint64_t* want) {
  uint64_t a, b, c, total = sizeof(dp_pack_header);
  if (__builtin_mul_overflow((uint64_t)h.n_records, (uint64_t)sizeof(dp_pack_record), &a)) return false;
  if (__builtin_mul_overflow(h.n_runs, (uint64_t)sizeof(dp_exc_run), &b)) return false;
  if (__builtin_mul_overflow(h.n_words_alloc, (uint64_t)12, &c)) return false;
  if (__builtin_add_overflow(total, a, &total) || __builtin_add_overflow(total, h.names_bytes, &total) ||
      __builtin_add_overflow(total, b, &total) || __builtin_add_overflow(total, c, &total))
    return false;
  *want = total;
  return true;
}

static int template_load_from(dp_ctx* ctx, dp_pack_source& S, int verify, dp_template** out) {
  const char* path = S.name;
  dp_pack_header h;
  if (!S.read(&h, sizeof h)) return dp_set_error(ctx, DP_ERR_IO, "%s: shorter than a packed-template header", path);
  if (memcmp(h.magic, DP_PACK_MAGIC, 8) != 0) return dp_set_error(ctx, DP_ERR_IO, "%s: not a packed template (bad magic)", path);
  if (h.version != DP_PACK_VERSION) return dp_set_error(ctx, DP_ERR_UNSUPPORTED, "%s: packed-template version %u, expected %u", path, h.version, DP_PACK_VERSION);
  if (h.n_words > (1ull << 58) || h.n_words_alloc != h.n_words + DP_GUARD_WORDS || h.total_bases > h.n_words * 32)
    return dp_set_error(ctx, DP_ERR_IO, "%s: inconsistent header", path);
  uint64_t want = 0;
  if (!pack_expected_size(h, &want) || S.size != want)
    return dp_set_error(ctx, DP_ERR_IO, "%s: %llu bytes, header says %llu (truncated?)", path,
                        (unsigned long long)S.size, (unsigned long long)want);
  // from here on every table size is bounded by the size of the source
  std::vector<dp_pack_record> recs(h.n_records);
  std::string names((size_t)h.names_bytes, '\0');
  dp_template* t = new_template(ctx, h.n_records);
  struct dropper {     // error paths: no copy may still be writing into planes that go back to the pool
    dp_template* t;
    ~dropper() { if (t) { cudaStreamSynchronize(t->ctx->copy_stream); free_template_device(t); delete_template(t); } }
  } drop{t};
  t->runs.resize((size_t)h.n_runs);
  if (!S.read(recs.data(), recs.size() * sizeof(dp_pack_record)) || !S.read(&names[0], names.size()) ||
      !S.read(t->runs.data(), t->runs.size() * sizeof(dp_exc_run)))
    return dp_set_error(ctx, DP_ERR_IO, "short read on %s", path);
  dp_pack_sum c;
  // the header (without its checksum field) is covered too
  dp_pack_header hc = h;
  hc.checksum = 0;
  c.mix(&hc, sizeof hc);
  c.mix(recs.data(), recs.size() * sizeof(dp_pack_record));
  c.mix(names.data(), names.size());
  c.mix(t->runs.data(), t->runs.size() * sizeof(dp_exc_run));
  size_t np = 0;
  uint64_t prev_end = 0;
  for (uint32_t r = 0; r < h.n_records; ++r) {
    const dp_pack_record& x = recs[r];
    if ((x.start & 31) || x.start < prev_end || x.len > h.n_words * 32 || x.start > h.n_words * 32 - x.len ||
        (uint64_t)x.id_len + x.descr_len > names.size() - np)
      return dp_set_error(ctx, DP_ERR_IO, "%s: record table is inconsistent at record %u", path, r);
    prev_end = x.start + x.len;
    t->records[r].start = x.start;
    t->records[r].len = x.len;
    t->records[r].id = names.substr(np, x.id_len);
    np += x.id_len;
    t->records[r].descr = names.substr(np, x.descr_len);
    np += x.descr_len;
  }
  t->origin = names.substr(np);
  for (size_t i = 0; i < t->runs.size(); ++i)
    if (t->runs[i].end <= t->runs[i].start || t->runs[i].end > h.n_words * 32 || (i && t->runs[i].start < t->runs[i - 1].end) ||
        t->runs[i].letter < 4 || t->runs[i].letter > 14)
      return dp_set_error(ctx, DP_ERR_IO, "%s: exception-run table is inconsistent at run %zu", path, i);
  t->n_words = h.n_words;
  t->n_words_alloc = h.n_words_alloc;
  t->total_bases = h.total_bases;
  DP_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = ensure_stage(ctx);
  if (rc) return rc;
  const size_t plane_bytes = (size_t)h.n_words_alloc * 4;
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_lo, plane_bytes));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_hi, plane_bytes));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_exc, plane_bytes));
  // the planes were allocated in ctx->stream order; the copies run on copy_stream behind an event
  DP_CUDA(ctx, cudaEventRecord(ctx->stage_ev[0], ctx->stream));
  DP_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->stage_ev[0], 0));
  uint32_t* dst[3] = {t->d_lo, t->d_hi, t->d_exc};
  bool rebuild_exc = false;
  if (S.mem && is_pinned(S.mem)) {
    // an image in pinned host memory: the planes go out as three DMA transfers, the host touches
    // them only when asked to verify the checksum (which then overlaps the copies)
    // Unverified, the exception plane does not travel: it is rebuilt on the device from the tables
    // below (dp_exc_records_kernel / dp_exc_runs_kernel) while the other two planes are on their way.
    rebuild_exc = !verify;
    for (int pl = 0; pl < (rebuild_exc ? 2 : 3); ++pl)
      DP_CUDA(ctx, cudaMemcpyAsync(dst[pl], S.mem + S.at + (size_t)pl * plane_bytes, plane_bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    if (verify)
      for (int pl = 0; pl < 3; ++pl) c.mix(S.mem + S.at + (size_t)pl * plane_bytes, plane_bytes);
    S.at += 3 * plane_bytes;
  } else {
    bool slot_busy[2] = {false, false};
    int slot = 0;
    for (int pl = 0; pl < 3; ++pl) {
      size_t done = 0;
      while (done < plane_bytes) {
        const size_t n = std::min<size_t>(DP_STAGE_BYTES, plane_bytes - done);
        if (slot_busy[slot]) DP_CUDA(ctx, cudaEventSynchronize(ctx->stage_copied[slot]));   // pinned slot free again
        if (!S.read(ctx->h_stage[slot], n)) return dp_set_error(ctx, DP_ERR_IO, "short read on %s", path);
        if (verify) c.mix(ctx->h_stage[slot], n);  // DP_STAGE_BYTES is a multiple of 32: only a plane's last chunk has a tail
        DP_CUDA(ctx, cudaMemcpyAsync((char*)dst[pl] + done, ctx->h_stage[slot], n, cudaMemcpyHostToDevice, ctx->copy_stream));
        DP_CUDA(ctx, cudaEventRecord(ctx->stage_copied[slot], ctx->copy_stream));
        slot_busy[slot] = true;
        slot ^= 1;
        done += n;
      }
    }
  }
  if (verify && c.value() != h.checksum) {
    DP_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    return dp_set_error(ctx, DP_ERR_IO, "%s: checksum mismatch (file is corrupt)", path);
  }
  const size_t nr = t->runs.size();
  const uint32_t nrec = h.n_records;
  // small tables go through the pinned result area (async copies from pageable memory would block)
  const size_t tab_bytes = nr * sizeof(dp_exc_run) + 2 * (size_t)nrec * sizeof(uint64_t);
  rc = dp_ws_pinned(ctx, tab_bytes + 64);
  if (rc) return rc;
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));       // nothing queued earlier still reads the pinned area
  uint64_t* starts = (uint64_t*)ctx->h_res;
  uint64_t* rl = starts + nrec;
  dp_exc_run* hruns = (dp_exc_run*)(rl + nrec);
  for (uint32_t r = 0; r < nrec; ++r) { starts[r] = t->records[r].start; rl[r] = t->records[r].len; }
  if (nr) memcpy(hruns, t->runs.data(), nr * sizeof(dp_exc_run));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_runs, std::max<size_t>(1, nr) * sizeof(dp_exc_run)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_rec_start, std::max<size_t>(1, nrec) * sizeof(uint64_t)));
  DP_CUDA(ctx, dp_pool_alloc(ctx, (void**)&t->d_rec_len, std::max<size_t>(1, nrec) * sizeof(uint64_t)));
  if (nr) DP_CUDA(ctx, cudaMemcpyAsync(t->d_runs, hruns, nr * sizeof(dp_exc_run), cudaMemcpyHostToDevice, ctx->stream));
  if (nrec) {
    DP_CUDA(ctx, cudaMemcpyAsync(t->d_rec_start, starts, nrec * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    DP_CUDA(ctx, cudaMemcpyAsync(t->d_rec_len, rl, nrec * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
  }
  if (rebuild_exc) {
    dp_exc_records_kernel<<<(unsigned)((h.n_words_alloc + 255) / 256), 256, 0, ctx->stream>>>(
        t->d_exc, h.n_words_alloc, t->d_rec_start, t->d_rec_len, nrec);
    DP_CUDA(ctx, cudaGetLastError());
    if (nr) {
      dp_exc_runs_kernel<<<(unsigned)((nr * 32 + 255) / 256), 256, 0, ctx->stream>>>(t->d_exc, t->d_runs, nr);
      DP_CUDA(ctx, cudaGetLastError());
    }
  }
  // later kernels on ctx->stream wait for the planes on the device, the host waits once
  DP_CUDA(ctx, cudaEventRecord(ctx->stage_copied[0], ctx->copy_stream));
  DP_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->stage_copied[0], 0));
  DP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  t->device_bytes = 3 * plane_bytes + nr * sizeof(dp_exc_run) + 2 * nrec * sizeof(uint64_t);
  drop.t = nullptr;
  *out = t;
  return DP_OK;
}

extern "C" int dp_template_load(dp_ctx* ctx, const char* path, dp_template** out) {
  if (!ctx || !path || !out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *out = nullptr;
  FILE* fh = fopen(path, "rb");
  if (!fh) return dp_set_error(ctx, DP_ERR_IO, "cannot open %s", path);
  struct closer { FILE* f; ~closer() { if (f) fclose(f); } } guard{fh};
  fseek(fh, 0, SEEK_END);
  const long fsize = ftell(fh);
  fseek(fh, 0, SEEK_SET);
  dp_pack_source S{fh, nullptr, fsize > 0 ? (uint64_t)fsize : 0, 0, path};
  try {
    return template_load_from(ctx, S, 1, out);
  } catch (const std::exception& e) {
    return dp_set_error(ctx, DP_ERR_NOMEM, "%s: %s", path, e.what());
  }
}

extern "C" int dp_template_load_mem(dp_ctx* ctx, const void* image, uint64_t bytes, int verify, dp_template** out) {
  if (!ctx || !image || !out) return dp_set_error(ctx, DP_ERR_ARG, "null argument");
  *out = nullptr;
  dp_pack_source S{nullptr, (const char*)image, bytes, 0, "packed image"};
  try {
    return template_load_from(ctx, S, verify, out);
  } catch (const std::exception& e) {
    return dp_set_error(ctx, DP_ERR_NOMEM, "packed image: %s", e.what());
  }
}

// id / description of a record (records made by dp_template_from_records are named record_<r>):
// what a GenBank / EMBL loader calls after it has handed the sequences over.  Host-side only.
extern "C" int dp_template_set_origin(dp_template* t, const char* origin) {
  if (!t) return DP_ERR_ARG;
  t->origin = origin ? origin : "";
  return DP_OK;
}

extern "C" const char* dp_template_origin(const dp_template* t) { return t ? t->origin.c_str() : ""; }

extern "C" int dp_template_set_record_name(dp_template* t, uint32_t r, const char* id, const char* descr) {
  if (!t || r >= t->records.size() || !id) return DP_ERR_ARG;
  t->records[r].id = id;
  t->records[r].descr = descr ? descr : "";
  return DP_OK;
}
