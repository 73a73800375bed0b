// Internal declarations shared by the translation units of libdegenprimer_b200.
// Nothing in here is part of the ABI (include/degenprimer_b200.h is).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/degenprimer_b200.h"
#include "dp_device.cuh"

// ---- letters -----------------------------------------------------------------
// 2-bit plane code: A=0 C=1 T=2 G=3  (hi = ASCII bit 2, lo = ASCII bit 1).
// Letter index: A C T G R Y S W K M B D H V N  (0..14); 15 = not a DNA letter.
#define DP_LETTERS "ACTGRYSWKMBDHVN"
#define DP_NOT_A_LETTER 15

struct dp_record {
  uint64_t start;    // first base in the padded concatenation (multiple of 32)
  uint64_t len;
  std::string id;
  std::string descr;
};

struct dp_template {
  dp_ctx* ctx;
  uint64_t n_words;            // plane words that hold record data
  uint64_t n_words_alloc;      // + tail guard so kernels may read i+1, i+2, i+3
  uint32_t* d_lo;
  uint32_t* d_hi;
  uint32_t* d_exc;
  std::vector<dp_record> records;
  uint64_t total_bases;
  // record table on the device: start[r], len[r]
  uint64_t* d_rec_start;
  uint64_t* d_rec_len;
  // exception runs, sorted by start (host + device)
  std::vector<dp_exc_run> runs;
  dp_exc_run* d_runs;
  uint64_t device_bytes;
  std::string origin;          // free text saved with the packed image: what it was packed from
  uint64_t uid;                // process-wide serial number; dp_template_alive(uid) is false once freed
};
bool dp_template_alive(uint64_t uid);

// one surviving dimer as written by the duplex kernel; (hit, variant) are
// implied by the pair directory
struct dp_raw_dimer {
  uint64_t match_mask;
  double dG;
};

// What the duplex kernel needs to know about the patterns/primers of a call
// (kernel parameter, lives in the constant bank).
#define DP_MAX_PATTERNS (2 * DP_MAX_PRIMERS)
#define DP_MAX_RANKS 16
struct dp_dg_plan {
  uint8_t pat_len[DP_MAX_PATTERNS];
  uint8_t pat_strand[DP_MAX_PATTERNS];
  uint8_t pat_primer[DP_MAX_PATTERNS];
  uint32_t var_count[DP_MAX_PRIMERS];
  uint32_t var_off[DP_MAX_PRIMERS];      // byte offset of the primer's variants in d_codes
};


// dp_comm_set_layout: local record of a sender -> where it was cut from
struct dp_comm_piece {
  uint64_t local_start;   // padded start of the local record in the sender's template
  uint64_t offset;        // first base of the piece inside the whole record
  uint32_t record;        // index of the whole record
  uint64_t record_len;    // its length
};

struct dp_ctx {
  int device;
  cudaStream_t stream;
  cudaStream_t copy_stream;    // H2D of template chunks (overlaps the pack kernel)
  cudaEvent_t ev0, ev1, ev2, ev3, ev4, ev5;
  int sm_count;
  int cc_major, cc_minor;
  volatile int abort_flag;
  bool tables_loaded;
  dp_nn_tables* d_tables;      // device copy
  bool hp_tables_loaded;
  dp_hairpin_tables* d_hp_tables;   // hairpin-loop tables (global memory, 81 KB), allocated on first load
  cudaMemPool_t pool;          // stream-ordered pool: template planes, scratch
  // grow-only workspace of the scan / duplex stages
  dp_raw_hit* d_hits;
  uint64_t hits_cap;
  uint64_t* d_dir;             // per (hit, variant) pair: offset << 16 | count
  uint64_t dir_cap;
  dp_raw_dimer* d_dimers;
  uint64_t dimers_cap;
  uint8_t* d_codes;            // variant letters as plane codes
  uint64_t codes_cap;
  unsigned long long* d_counters;
  bool counters_clean;         // slots 0..7 are zero (the publish kernel leaves them so)
  char* h_res;                 // pinned (device-visible) staging for results
  size_t h_res_bytes;
  // pinned staging for uploads
  char* h_stage[2];
  char* d_stage[2];
  cudaEvent_t stage_ev[2];
  cudaEvent_t stage_copied[2];
  size_t stage_bytes;
  // fused call in flight (dp_find_begin .. dp_find_end); owned by dp_scan.cu
  void* pending;
  void (*pending_free)(void*);
  // optional device mirror of {counters, raw hits} written by the publish kernel
  void* wire;
  uint64_t wire_bytes;
  // multi-GPU hit exchange (dp_comm.cu): the publish kernel stores this rank's wire slot
  // straight into every peer's receive buffer over NVLink (CUDA IPC peer memory)
  std::vector<uint8_t> codes_shadow;   // host copy of what d_codes holds (variants of the last dp_find)
  bool timing;                 // per-kernel CUDA-event timing of every call (dp_set_timing)
  int comm_world, comm_rank;   // world == 0: no exchange attached
  uint64_t comm_seq;           // calls exchanged so far; call c uses receive buffer c & 1
  char* d_recv[2];             // this rank's receive buffers: world x wire_bytes each (cudaMalloc, IPC exported)
  char** d_peer[2];            // device table: d_peer[b][r] = rank r's receive buffer b as seen from this GPU
  void* peer_mapped[2][DP_MAX_RANKS];   // host copy of the table (what cudaIpcOpenMemHandle returned)
  char* h_gather[2];           // pinned host copies of the receive buffers, written by the publish kernel
  uint64_t* h_sorted[2];       // the same calls converted to dp_hit and sorted by the kernel (32-byte header first)
  unsigned long long* comm_err;         // pinned host flag (device-visible): a wait for a peer timed out
  uint64_t comm_slot_bytes;             // bytes per rank in a receive buffer: 64-byte header + raw hits
  std::vector<dp_comm_piece> comm_layout;      // multi-record shards (dp_comm_set_layout); empty: one segment per rank
  std::vector<uint32_t> comm_layout_base;      // entries of rank r: [base[r], base[r+1])
  dp_dg_plan comm_plan[4];     // (length, strand, primer) of the patterns of the last calls (call & 3)
  bool wire_owned;
  dp_dg_plan last_plan;        // (length, strand, primer) of the patterns of the last scan
  uint64_t last_rec0_len;      // its template: length of record 0 and record count
  uint64_t last_n_records;
  std::string err;
};

int dp_set_error(dp_ctx* ctx, int code, const char* fmt, ...);

#define DP_CUDA(ctx, call)                                                     \
  do {                                                                         \
    cudaError_t e__ = (call);                                                  \
    if (e__ != cudaSuccess)                                                    \
      return dp_set_error((ctx), DP_ERR_CUDA, "%s failed at %s:%d: %s", #call, \
                          __FILE__, __LINE__, cudaGetErrorString(e__));        \
  } while (0)

// letter index of an ASCII byte (case-folded), DP_NOT_A_LETTER if foreign
__host__ __device__ inline uint32_t dp_letter_of(unsigned char c) {
  switch (c & 0xDF) {
    case 'A': return 0;  case 'C': return 1;  case 'T': return 2;  case 'G': return 3;
    case 'R': return 4;  case 'Y': return 5;  case 'S': return 6;  case 'W': return 7;
    case 'K': return 8;  case 'M': return 9;  case 'B': return 10; case 'D': return 11;
    case 'H': return 12; case 'V': return 13; case 'N': return 14;
    default: return DP_NOT_A_LETTER;
  }
}

// 4-bit base set of a letter: bit c set <=> plane code c is in the expansion
// (DegenPrimer/SeqUtils.py:30-40).  A=1 C=2 T=4 G=8.
__host__ __device__ inline uint32_t dp_letter_set(uint32_t letter) {
  const uint32_t A = 1, C = 2, T = 4, G = 8;
  switch (letter) {
    case 0: return A;  case 1: return C;  case 2: return T;  case 3: return G;
    case 4: return A | G;      // R
    case 5: return C | T;      // Y
    case 6: return G | C;      // S
    case 7: return A | T;      // W
    case 8: return G | T;      // K
    case 9: return A | C;      // M
    case 10: return C | G | T; // B
    case 11: return A | G | T; // D
    case 12: return A | C | T; // H
    case 13: return A | C | G; // V
    case 14: return A | C | G | T;
    default: return 0;
  }
}

// complement of a letter index (DegenPrimer/SecStructures.py:34-48)
__host__ __device__ inline uint32_t dp_letter_complement(uint32_t l) {
  // A C T G R Y S W K M B D H V N -> T G A C Y R S W M K V H D B N
  const unsigned char m[16] = {2, 3, 0, 1, 5, 4, 6, 7, 9, 8, 13, 12, 11, 10, 14, 15};
  return m[l & 15];
}

// deterministic 2-bit filler for non-ACGT positions, so that exception runs
// look like random sequence to the prefilter instead of a homopolymer
__host__ __device__ inline uint32_t dp_filler(uint64_t pos) {
  uint64_t x = pos * 0x9E3779B97F4A7C15ull;
  x ^= x >> 29;
  x *= 0xBF58476D1CE4E5B9ull;
  return (uint32_t)(x >> 62);
}

// host helpers implemented in dp_template.cu / dp_scan.cu
int dp_upload_and_pack(dp_ctx* ctx, dp_template* t, uint32_t n_records,
                       const char* const* seqs, const uint64_t* lens);

// host memory of result arrays handed to the caller (released with dp_free)
void* dp_result_alloc(size_t bytes);
void dp_result_free(void* p);
bool dp_result_is_pinned(const void* p);     // big blocks are page-locked: D2H copies go straight in

// stream-ordered allocation from the context's pool (no cudaMalloc on hot calls)
cudaError_t dp_pool_alloc(dp_ctx* ctx, void** p, size_t bytes);
void dp_pool_free(dp_ctx* ctx, void* p);
// grow-only workspace; contents are lost when a buffer grows
int dp_ws_hits(dp_ctx* ctx, uint64_t cap);
int dp_ws_dir(dp_ctx* ctx, uint64_t cap);
int dp_ws_dimers(dp_ctx* ctx, uint64_t cap);
int dp_ws_codes(dp_ctx* ctx, uint64_t cap);
int dp_ws_pinned(dp_ctx* ctx, size_t bytes);

// dp_comm.cu: what the publish kernel needs for the exchange (world == 0: none)
struct dp_exchange_args {
  int world, rank;
  uint64_t seq;                // this call's sequence number c
  uint64_t slot_bytes;         // bytes per rank in a receive buffer
  uint64_t slot_hits;          // raw hit records a slot holds
  char* const* peers;          // peers[r] = rank r's receive buffer (c & 1)
  const char* recv_prev;       // this rank's receive buffer ((c-1) & 1): call c-1 of every rank
  char* h_prev;                // pinned host copy of it
  unsigned long long* err;
  uint64_t rec0_len, n_records;
  // converted + sorted copy of the landed call (dp_exchange_to_host): header {n, valid, seq, 0}
  // then dp_hit records; (length, strand, primer) of that call's patterns
  uint64_t* h_sorted;
  uint8_t prev_len[DP_MAX_PATTERNS], prev_strand[DP_MAX_PATTERNS], prev_primer[DP_MAX_PATTERNS];
};
#define DP_EXCHANGE_SORT_MAX 2048     // hits of all ranks the publish kernel converts and sorts itself
void dp_comm_fill(dp_ctx* ctx, dp_exchange_args* X);     // for the call being queued; advances the sequence
void dp_comm_release(dp_ctx* ctx);
void dp_exchange_drain_launch(const dp_exchange_args& X, cudaStream_t st);   // dp_scan.cu

// duplex stage on the hits that sit in ctx->d_hits (dp_dg.cu).  n_pairs_hint = 0:
// the pair count is only known on the device (fused dp_find).
struct dp_dg_launch {
  dp_dg_plan plan;
  double PCR_T, Na_eq_log, max_dimer_dG, max_dG_for_K;
};
int dp_dg_enqueue(dp_ctx* ctx, const dp_template* t, const dp_dg_launch& L, uint64_t n_pairs_hint, uint32_t* launches,
                  bool hits_fresh);     // hits_fresh: straight from a scan / an upload, dimer counts are zero
// large results: the duplex kernel leaves the number of dimers of every hit in the hit record
// (dp_raw_hit.reserved); the gather copies every pair's dimers to d_info[slot].final_off with
// (rank, variant) filled in (dp_dg.cu)
struct dp_slot_info_host { uint64_t final_off; uint32_t rank; uint32_t count; };
int dp_dg_gather(dp_ctx* ctx, const dp_dg_plan& plan, uint64_t n_hits, const void* d_info, dp_dimer* d_out);
