/* degenprimer_b200 — C ABI of the B200-native DegenPrimer hot path.
 *
 * The reference (allista/DegenPrimer, pure Python 2.7) has no FFI layer; the
 * boundary it offers is the Python API of DegenPrimer/SearchEngine.py as called
 * from DegenPrimer/PCR_ProductsFinder.py:36,74,88,100,118-120.  This header is
 * the thin C-ABI the Python facade (degenprimer_b200/SearchEngine.py) binds with
 * ctypes; every entry point names the reference code it replaces.
 *
 * Conventions
 *   - every function returns DP_OK (0) or a negative DP_ERR_* code;
 *     dp_last_error(ctx) holds the message of the last failure on that ctx.
 *   - one dp_ctx per host thread; a ctx owns one CUDA stream and is NOT
 *     re-entrant.  The CUDA context is created inside dp_create(), i.e. in the
 *     calling process (never before a fork — SURVEY §7 "CUDA + Manager").
 *   - result arrays returned through `T** out` are malloc()ed by the library
 *     and released with dp_free().
 *   - no torch / C++ types cross this boundary: plain pointers and sizes only.
 *   - there is no CPU implementation behind these calls: without a usable
 *     sm_100 device dp_create() fails.
 *
 * Base encoding in HBM: three bit-planes, one uint32 word per
 * 32 bases each, bit b of word i = base 32*i+b.  hi = ASCII bit 2, lo = ASCII
 * bit 1 (A=00 C=01 T=10 G=11, complement = flip hi); exc = 1 for every byte
 * outside ACGT (IUPAC codes, record padding).  Exception letters live in a
 * sorted run table.
 */
#ifndef DEGENPRIMER_B200_H
#define DEGENPRIMER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DP_ABI_VERSION 1

enum {
  DP_OK = 0,
  DP_ERR_CUDA = -1,        /* a CUDA runtime call failed / no usable device     */
  DP_ERR_ARG = -2,         /* bad argument                                        */
  DP_ERR_ALPHABET = -3,    /* byte outside the IUPAC DNA alphabet: the reference */
                           /* raises KeyError in reverse_complement()            */
                           /* (DegenPrimer/SecStructures.py:50-55)               */
  DP_ERR_LENGTH = -4,      /* template shorter than primer / empty primer:       */
                           /* ValueError of SearchEngine.py:232-238              */
  DP_ERR_UNSUPPORTED = -5, /* primer longer than DP_MAX_PRIMER_LEN               */
  DP_ERR_IO = -6,          /* cannot read a FASTA file                            */
  DP_ERR_NOMEM = -7,
  DP_ERR_ABORTED = -8,     /* abort flag was raised (reference returns None)     */
  DP_ERR_STATE = -9        /* tables not loaded, template freed, ...             */
};

#define DP_MAX_PRIMER_LEN 64   /* configs use 18-35 nt                          */
#define DP_MAX_PRIMERS    64   /* masters per dp_scan call (x2 strands)         */

typedef struct dp_ctx dp_ctx;
typedef struct dp_template dp_template;

/* ---- context --------------------------------------------------------------- */
int dp_abi_version(void);
int dp_device_count(int* n);
/* Creates the CUDA context on `device_ordinal`.  NULL on failure; the reason is
 * returned by dp_create_error(). */
dp_ctx* dp_create(int device_ordinal);
const char* dp_create_error(void);
void dp_destroy(dp_ctx* ctx);
const char* dp_last_error(const dp_ctx* ctx);
void dp_free(void* p);
/* Cooperative abort (BioUtils/Tools/AbortableBase.py:16-35): once raised, the
 * next kernel-launch boundary returns DP_ERR_ABORTED.  Returns previous flag. */
int dp_abort_requested(dp_ctx* ctx, int flag);
/* Device properties for the bench / roofline (name is a 256-byte buffer). */
int dp_device_info(dp_ctx* ctx, char* name, int* sm_count, int* cc_major,
                   int* cc_minor, int* clock_khz, uint64_t* hbm_bytes);
/* The CUDA stream (cudaStream_t) every kernel and copy of this ctx is queued on:
 * lets a caller bracket calls with its own CUDA events (bench.py wraps it in
 * torch.cuda.ExternalStream).  dp_synchronize() waits for it. */
void* dp_stream(dp_ctx* ctx);
int dp_synchronize(dp_ctx* ctx);

/* ---- template loader: FASTA / ASCII -> bit-planes in HBM ---------------------
 * Replaces BioUtils/BioUtils/SeqUtils.py:103-271 (SeqView over
 * Bio.SeqIO.index_db) + the str(template.seq) / reverse_complement() copies of
 * DegenPrimer/SearchEngine.py:287-290,333-336.  Input is case-folded like
 * SeqView(upper=True) (DegenPrimer/iPCR_Base.py:47-52). */
int dp_template_from_records(dp_ctx* ctx, uint32_t n_records,
                             const char* const* seqs, const uint64_t* lens,
                             dp_template** out);
int dp_template_from_ascii(dp_ctx* ctx, const char* seq, uint64_t len,
                           dp_template** out);
/* Parses every record of a FASTA file; records are stored sorted by id, like
 * SeqView._ids (BioUtils/BioUtils/SeqUtils.py:170). */
int dp_template_from_fasta(dp_ctx* ctx, const char* path, dp_template** out);
void dp_template_free(dp_template* t);

/* SURVEY §8(f) N4 — on-disk packed template.  Replaces the SQLite index the reference
 * builds on every run (SeqIO.index_db into a temp file, BioUtils/SeqUtils.py:140-172)
 * together with parsing and packing the text: the file holds the three bit-planes, the
 * exception-run table and the record table (ids, descriptions) exactly as they sit in
 * HBM — 0.375 B/bp, nothing to parse, one H2D stream through pinned staging.
 * Layout (little endian): 64-byte header {magic "DPB200PK", u32 version = 3,
 * u32 n_records, u64 n_words, u64 n_words_alloc, u64 total_bases, u64 n_runs,
 * u64 names_bytes, u64 checksum of the header (checksum field zero) and the six sections
 * after it (four interleaved multiply-xor lanes over their 64-bit words, see dp_pack_sum)};
 * n_records x {u64 start, u64 len, u32 id_len, u32 descr_len}; the id and description
 * bytes, followed by the origin string (names_bytes covers both);
 * n_runs x {u64 start, u64 end, u32 letter, u32 pad}; planes lo, hi, exc of
 * n_words_alloc u32 each.  dp_template_load checks magic, version, sizes (overflow-safe), the
 * tables' consistency (run letters included) and the checksum (DP_ERR_IO /
 * DP_ERR_UNSUPPORTED otherwise; DP_ERR_NOMEM instead of an exception across the ABI). */
/* id / description of record r (dp_template_from_records names them record_<r>); host-side,
 * used by loaders of formats other than FASTA (GenBank, EMBL) before dp_template_save. */
int dp_template_set_record_name(dp_template* t, uint32_t r, const char* id, const char* descr);
int dp_template_save(dp_ctx* ctx, const dp_template* t, const char* path);
int dp_template_load(dp_ctx* ctx, const char* path, dp_template** out);
/* The same image in host memory instead of a file: a long-running caller keeps its genomes
 * packed in (pinned) host memory, 0.375 B/bp, and brings one to the device with three DMA
 * transfers -- the counterpart of the reference holding SeqRecords in RAM
 * (BioUtils/BioUtils/SeqUtils.py:216-225) at 1 B/bp and re-encoding them on every search
 * (DegenPrimer/SearchEngine.py:287-290,333-336).  dp_template_save_mem writes
 * dp_template_image_size() bytes.  dp_template_load_mem: `verify` != 0 checks the checksum
 * (one pass of the host over the image); an image that is not pinned is staged like a file. */
int dp_template_image_size(const dp_template* t, uint64_t* bytes);
int dp_template_save_mem(dp_ctx* ctx, const dp_template* t, void* image, uint64_t capacity);
int dp_template_load_mem(dp_ctx* ctx, const void* image, uint64_t bytes, int verify, dp_template** out);
/* Free text stored with the image (covered by its checksum): what the template was packed
 * from.  degenprimer_b200/SeqView.py records path, size and mtime of the source file there and
 * rebuilds a cache whose origin differs -- the reference's SeqIO.index_db raises when the
 * file list of an index differs (BioUtils/BioUtils/SeqUtils.py:169). */
int dp_template_set_origin(dp_template* t, const char* origin);
const char* dp_template_origin(const dp_template* t);

typedef struct dp_template_info {
  uint32_t n_records;
  uint32_t reserved;
  uint64_t total_bases;      /* sum of record lengths                           */
  uint64_t padded_bases;     /* 32 * number of plane words                      */
  uint64_t n_exception_runs; /* maximal runs of one non-ACGT letter             */
  uint64_t device_bytes;     /* HBM held by this template                       */
} dp_template_info;
int dp_template_get_info(const dp_template* t, dp_template_info* info);
/* Record r: offset of its first base in the padded concatenation (multiple of
 * 32), its length, its id / description (NUL-terminated, owned by `t`). */
int dp_template_record(const dp_template* t, uint32_t r, uint64_t* start,
                       uint64_t* len, const char** id, const char** descr);
/* Decode bases [begin, begin+n) of record r back to upper-case ASCII
 * (n bytes, no NUL).  reverse!=0 returns the same stretch of the
 * reverse-complement strand, i.e. reverse_complement(record)[begin:begin+n]. */
int dp_template_unpack(dp_ctx* ctx, const dp_template* t, uint32_t r,
                       uint64_t begin, uint64_t n, int reverse, char* out);

/* ---- scan: mismatch-tolerant annealing-site search ----------------------------
 * Replaces SearchEngine._find / _find_mp / _find_in_slice / _find_in_chunk
 * (DegenPrimer/SearchEngine.py:190-211,257-359): for every master primer and
 * both strands, every window whose thirds-weighted score reaches
 * 3*max(1, L-k)-1 (the integer form of `score >= max(1,L-k)-0.5`).
 *
 * A hit is reported in the coordinates the reference uses:
 *   strand 0: pos = 0-based window start on the record (fwd_matches)
 *   strand 1: pos = 0-based window start on reverse_complement(record)
 *             (rev_matches) = record_len - L - forward_start
 * Hits come back sorted by (primer, strand, record, pos).
 */
typedef struct dp_hit {
  uint64_t pos;
  uint32_t record;
  uint16_t primer;   /* index into `masters`                                    */
  uint8_t strand;    /* 0 forward, 1 reverse                                     */
  uint8_t score3;    /* 3 * reference score (exact integer)                      */
} dp_hit;

enum { DP_STRAND_FWD = 1, DP_STRAND_REV = 2, DP_STRAND_BOTH = 3 };
enum {
  DP_ALGO_AUTO = 0,
  DP_ALGO_POPC = 1,     /* per-window XOR/AND + popcount                        */
  DP_ALGO_BITSLICE = 2, /* bit-sliced prefilter over 32 windows + exact verify  */
  DP_ALGO_JIT = 3       /* the same prefilter, compiled at run time (NVRTC) for
                           this pattern set: compiles now if it has to.  AUTO
                           uses such a kernel when dp_scan_prepare_kernels made
                           one before, else the generic BITSLICE kernel        */
};

typedef struct dp_scan_stats {
  uint64_t windows;        /* window x pattern pairs examined                    */
  uint64_t candidates;     /* pairs that reached the exact verify stage          */
  uint64_t hits;
  float ms_scan;           /* CUDA-event time of all scan kernels of the call    */
  float ms_total;          /* incl. hit download and sort                        */
  float ms_main;           /* the dominant kernel alone (bit-slice, or popcount  */
                           /* when that is the only one)                         */
  float reserved;
  uint32_t launches;       /* kernels launched by this call                      */
  uint32_t algo;           /* DP_ALGO_* actually used                            */
} dp_scan_stats;

/* Per-kernel timing of every call (the ms_* fields of the stats: six CUDA event
 * records and four elapsed-time queries, ~15 us of host time per call).  On by
 * default; a throughput loop that does not read the ms_* fields switches it off. */
int dp_set_timing(dp_ctx* ctx, int on);

/* Compiles (NVRTC, ~0.2-0.5 s once per process and pattern set) the scan kernel
 * specialised for these primers; later dp_scan / dp_find calls with the same
 * masters, mismatch limits and strand mask use it under DP_ALGO_AUTO.  Call it
 * for a query that will be scanned more than once or over a large template.
 * Fails (DP_ERR_STATE) if libnvrtc cannot be loaded or the compile fails -- the
 * generic kernel then keeps serving DP_ALGO_AUTO; results are identical. */
int dp_scan_prepare_kernels(dp_ctx* ctx, uint32_t n_primers, const char* const* masters,
                            const int32_t* max_mismatches, int strand_mask);

/* Build check that needs no device: generates the specialised kernel source for
 * a primer set, compiles it with NVRTC and optionally writes the source / cubin
 * of the first launch unit to files.  Returns the number of launch units (0: no
 * pattern takes the prefilter) or a negative DP_ERR code (dp_last_error(NULL)). */
int dp_jit_selftest(uint32_t n_primers, const char* const* masters, const int32_t* max_mismatches,
                    int strand_mask, const char* source_path, const char* cubin_path);

/* The prefilter plan of the specialised kernel of a primer set, as text: one line
 *   //PLAN pattern <i> K <k> need <t> p_alive <p> groups <j,j,..> <j,j,..> ...
 * per (primer, strand) pattern (window positions of every group; a hit has at least `need`
 * groups that match completely), then "//PLAN mode ...".  Host-only (tests, documentation).
 * Returns the number of launch units or a negative DP_ERR code. */
int dp_jit_describe(uint32_t n_primers, const char* const* masters, const int32_t* max_mismatches,
                    int strand_mask, char* out, uint64_t capacity);

/* seg_begin/seg_end restrict window STARTS to [seg_begin, seg_end) of the
 * padded concatenation (multi-GPU sharding); pass 0, UINT64_MAX for
 * everything.  `stats` may be NULL. */
int dp_scan(dp_ctx* ctx, const dp_template* t, uint32_t n_primers,
            const char* const* masters, const int32_t* max_mismatches,
            uint64_t seg_begin, uint64_t seg_end, int strand_mask, int algo,
            dp_hit** hits, uint64_t* n_hits, dp_scan_stats* stats);

/* ---- duplex thermodynamics ---------------------------------------------------
 * Tables: DegenPrimer/UnifiedNN.py:208-246 + data/internal-NN.csv, loops.csv,
 * dangling-ends.csv, flattened by the host (degenprimer_b200/UnifiedNN.py).
 * 2-bit base index everywhere = the plane code: A=0 C=1 T=2 G=3.
 * Letter index (15 letters): A C T G R Y S W K M B D H V N.
 */
typedef struct dp_nn_tables {
  /* stack[x][y][w] = int_dPar_37(seq = X Y, rev = comp(X) W): dH, dS
   * (DegenPrimer/TD_Functions.py:187-191).  x = primer[a], y = primer[a+1]
   * (bases), w = letter of the 3'->5' complement strand under primer[a+1]
   * (15 letters; ambiguous -> expansion average, UnifiedNN.py:231-245). */
  double stack_dH[4][4][15];
  double stack_dS[4][4][15];
  /* rstack[x][y][w] = int_dPar_37(seq = comp(Y) W, rev = Y X): the right-hand
   * stack of |x| and |xx| (DegenPrimer/TD_Functions.py:193-197).
   * x = primer[b-1], y = primer[b], w = complement-strand letter under
   * primer[b-1]. */
  double rstack_dH[4][4][15];
  double rstack_dS[4][4][15];
  /* dangling ends (DegenPrimer/UnifiedNN.py:248-252): [kind][x][y][z],
   * kind 0 = "XY/-Z", kind 1 = "XY/Z-"; NaN where the table has no row. */
  double dangling_dH[2][4][4][4];
  double dangling_dS[2][4][4][4];
  /* internal_loop_dS_37 for a loop of n unpaired bases per strand, n=0..63
   * (DegenPrimer/UnifiedNN.py:264-276,317-319). */
  double loop_dS[64];
  double ini_dH, ini_dS, ter_dH, ter_dS;
  double terminal_mismatch;   /* -0.72, DegenPrimer/UnifiedNN.py:139            */
  double dS_Na;               /* 0.368                                           */
  double R;                   /* 1.9872                                          */
  double K0;                  /* -273.15                                         */
  double T_DMSO;              /* 0.75                                            */
} dp_nn_tables;
int dp_tables_load(dp_ctx* ctx, const dp_nn_tables* tables);

/* PCR conditions read by the hot path (DegenPrimer/PCR_Parameters.py:66-97). */
typedef struct dp_conditions {
  double PCR_T;        /* Celsius                                                */
  double Na_eq_log;    /* ln(Na + sqrt(14.4*(Mg-dNTP)))                          */
  double max_dimer_dG; /* keep dimers with dG <= this (-3)                       */
  double max_dG_for_K; /* largest dG whose exp(-1000 dG / (R (T-K0))) >= min_K,  */
                       /* found by the host with the same libm exp()             */
  double DNA;          /* mol/L, only for dp_variant_tm                          */
  double DMSO;         /* %,    only for dp_variant_tm                           */
} dp_conditions;

/* (dp_find, the fused scan + duplex call, is declared after dp_duplex_dg.) */

/* One surviving dimer of one (site, variant) duplex — Dimer of
 * DegenPrimer/SecStructures.py:58-192 after Duplex._recalculate :294-324. */
typedef struct dp_dimer {
  uint32_t hit;         /* index into the `hits` array passed in                */
  uint32_t variant;     /* index into the primer's variant list                 */
  uint64_t match_mask;  /* bit i = primer index i is paired (fwd_matches)       */
  double dG;            /* kcal/mol at PCR_T                                    */
} dp_dimer;

typedef struct dp_dg_stats {
  uint64_t duplexes;     /* (site, variant) pairs evaluated                      */
  uint64_t dimers_tried; /* dimers whose dG was computed                         */
  uint64_t dimers_kept;
  float ms_kernel;
  float ms_total;
  uint32_t launches;
  uint32_t reserved;
} dp_dg_stats;

/* Replaces SearchEngine._compile_duplexes_for_position + Duplex.__init__ /
 * _recalculate + Dimer.from_sequences / alternatives + dimer_dG_corrected
 * (DegenPrimer/SearchEngine.py:106-118, SecStructures.py:126-152,231-324,
 * TD_Functions.py:124-206) for every hit x every unambiguous variant of the
 * hit's primer.
 *   variants: all variants of all primers, concatenated, variant v of primer p
 *   is the primer_len[p] bytes at variants + var_offset[p] + v*primer_len[p];
 *   var_count[p] variants each.
 * Dimers come back grouped by (hit, variant) in input order, each group sorted
 * by dG ascending, ties in the reference's construction order (stable sort,
 * SecStructures.py:314). */
int dp_duplex_dg(dp_ctx* ctx, const dp_template* t, const dp_hit* hits,
                 uint64_t n_hits, uint32_t n_primers, const uint8_t* primer_len,
                 const uint32_t* var_count, const uint64_t* var_offset,
                 const char* variants, const dp_conditions* cond,
                 dp_dimer** dimers, uint64_t* n_dimers, dp_dg_stats* stats);

/* SearchEngine.find (DegenPrimer/SearchEngine.py:375-386) = find_matches +
 * compile_duplexes_mp for every primer of the call, as ONE stream submission:
 * scan kernels -> duplex kernel over every (hit, variant) pair (sizes stay on
 * the device) -> one synchronise.  Same results as dp_scan followed by
 * dp_duplex_dg: hits sorted by (primer, strand, record, pos); dimers grouped by
 * (hit index in that order, variant), each group sorted by dG (stable).
 * variants/var_count/var_offset as in dp_duplex_dg; primer lengths are those of
 * `masters`. */
int dp_find(dp_ctx* ctx, const dp_template* t, uint32_t n_primers,
            const char* const* masters, const int32_t* max_mismatches,
            uint64_t seg_begin, uint64_t seg_end, int strand_mask, int algo,
            const uint32_t* var_count, const uint64_t* var_offset,
            const char* variants, const dp_conditions* cond,
            dp_hit** hits, uint64_t* n_hits, dp_dimer** dimers, uint64_t* n_dimers,
            dp_scan_stats* scan_stats, dp_dg_stats* dg_stats);

/* dp_find in two halves, so that a caller can queue its own work on dp_stream()
 * behind the kernels, or do host work (collecting the multi-GPU hit exchange of
 * an earlier call, dp_comm_hits with age 2) before the host waits.  dp_find == begin + end. */
int dp_find_begin(dp_ctx* ctx, const dp_template* t, uint32_t n_primers,
                  const char* const* masters, const int32_t* max_mismatches,
                  uint64_t seg_begin, uint64_t seg_end, int strand_mask, int algo,
                  const uint32_t* var_count, const uint64_t* var_offset,
                  const char* variants, const dp_conditions* cond);
int dp_find_end(dp_ctx* ctx, dp_hit** hits, uint64_t* n_hits, dp_dimer** dimers,
                uint64_t* n_dimers, dp_scan_stats* scan_stats, dp_dg_stats* dg_stats);

/* Device-side mirror of a call's hits for the multi-GPU exchange: when set, the
 * last kernel of every dp_scan / dp_find writes
 *     uint64 header[8]   header[0] = n_hits | n_pairs << 32 (n_hits may exceed
 *                        what fits: then the mirror is truncated)
 *     dp_wire_hit hit[]  unsorted, sender coordinates
 * into `device_ptr` (caller-owned device memory, 8-byte aligned, `bytes` long).
 * pattern = index of the (primer, strand) pair in call order (2*primer + strand
 * with DP_STRAND_BOTH); g = forward window start in the sender's padded
 * concatenation (record r starts at dp_template_record().start).
 * Pass NULL to switch the mirror off. */
typedef struct dp_wire_hit {
  uint64_t g;
  uint32_t record;
  uint16_t pattern;
  uint16_t score3;
  uint32_t pair_base;
  uint32_t reserved;
} dp_wire_hit;
int dp_set_wire(dp_ctx* ctx, void* device_ptr, uint64_t bytes);

/* Hit exchange between the GPUs of one node over NVLink peer memory (SURVEY §8e;
 * the reference's counterpart is the pickling of slice results through
 * multiprocessing Queues, BioUtils/Tools/Multiprocessing.py:201-211).  One
 * process per GPU.  Every rank calls dp_comm_export (allocates its two receive
 * buffers, world slots of 64 + max_hits*24 bytes each, and returns their two
 * 64-byte CUDA IPC handles); the host side moves the handles between the ranks
 * (torch.distributed all_gather in degenprimer_b200/distributed.py) and every
 * rank calls dp_comm_attach with world*128 bytes, rank-major.  From then on the
 * publish kernel of every dp_find / dp_find_begin of this ctx -- the last kernel
 * of the call -- stores this rank's slot (layout of dp_set_wire) into the
 * receive buffer of every rank and raises its sequence word: the gather is the
 * epilogue of the compute, there is no separate collective launch.  All ranks
 * must make the same sequence of calls; ranks run at most one call apart (a call
 * first waits, with a 2 s timeout, until every peer's previous call has landed).
 * After a timeout (a rank made fewer calls -- e.g. it failed before it queued one)
 * the ranks' sequence numbers no longer agree: every later dp_comm_gathered /
 * dp_comm_hits of this ctx returns DP_ERR_STATE until the exchange is set up again
 * (dp_comm_destroy, dp_comm_export, dp_comm_attach on every rank).
 * The gathered hits are fetched with `age`:
 *   age 0  the call that just returned (one small kernel waits for the slowest
 *          rank's slot),
 *   age 1  the call before it: in pinned host memory once the last call's
 *          kernels have finished (waits for the stream),
 *   age 2  two calls back: in pinned host memory already, nothing is waited for
 *          -- valid between dp_find_begin and dp_find_end of the newest call, so
 *          the conversion overlaps the kernels of the call in flight.
 * dp_comm_gathered() gives `world` slots of `slot_bytes`, slot r = rank r's hits
 * of that call.  A slot holds max_hits records; header[0] tells when a rank had
 * more, and bit 1 of header[4] says that the slot carries fewer records than header[0]
 * (then fall back to a host-side gather).  max_hits <= 2^20.  dp_comm_destroy only after every
 * rank has collected its last call (barrier on the host side). */
int dp_comm_export(dp_ctx* ctx, int world, int rank, uint32_t max_hits, char handles[128]);
int dp_comm_attach(dp_ctx* ctx, const char* all_handles);
int dp_comm_destroy(dp_ctx* ctx);
int dp_comm_gathered(dp_ctx* ctx, int age, const void** host_ptr, uint64_t* slot_bytes, int* world);
/* The gathered hits of a call (age as above), converted: one segment (one
 * record) per rank, hit.record = sender rank, hit.pos in that rank's segment
 * coordinates (strand 1: on its reverse complement), sorted by (primer, strand,
 * record, pos).  header[5] / header[6] of a slot carry the sender's record length
 * and record count.  DP_ERR_UNSUPPORTED if a slot overflowed or a rank scanned a
 * multi-record template (use dp_comm_gathered then). */
int dp_comm_hits(dp_ctx* ctx, int age, dp_hit** hits, uint64_t* n_hits);
/* Shards of a multi-record genome (one contiguous range of the padded concatenation per
 * rank, cut inside records where it has to be): tells dp_comm_hits where every rank's LOCAL
 * records come from, so that it returns hits in whole-genome coordinates -- record = index
 * of the whole record, pos on that record (strand 1: on its reverse complement).  Rank r's
 * local record j is entry e = base[r] + j: it starts at local_start[e] of r's padded
 * concatenation and holds the bases from offset[e] on of record record[e], which is
 * record_len[e] long.  base has world+1 entries.  NULL base: back to one segment per rank. */
int dp_comm_set_layout(dp_ctx* ctx, const uint32_t* base, const uint64_t* local_start,
                       const uint64_t* offset, const uint32_t* record, const uint64_t* record_len);

/* The duplex stage of the LAST dp_find / dp_find_end of this context again under
 * other conditions (the optimiser's inner loop, DegenPrimer/PCR_Optimizer.py:
 * 162-214 and PCR_ProductsFinder.matches_to_mixture :67-81, re-evaluates every
 * duplex for each trial point): the hits are still on the device, nothing is
 * scanned or uploaded.  dimers[].hit indexes the hit array that dp_find returned.
 * The template of that call must still exist.  DP_ERR_STATE if another call has
 * used the context since (dp_scan, dp_duplex_dg, another dp_find_begin). */
int dp_find_again(dp_ctx* ctx, const dp_conditions* cond, dp_dimer** dimers, uint64_t* n_dimers,
                  dp_dg_stats* dg_stats);

/* The template window under every hit, as the reference's strand strings show
 * it (fwd_seq[pos:pos+L] / rev_seq[pos:pos+L], DegenPrimer/SearchEngine.py:113):
 * out receives n_hits rows of 64 bytes, upper-case ASCII, NUL padded. */
int dp_hit_windows(dp_ctx* ctx, const dp_template* t, const dp_hit* hits,
                   uint64_t n_hits, uint32_t n_primers, const uint8_t* primer_len,
                   char* out);

/* Per-variant perfect-match Tr/Tm — TD_Functions.primer_template_Tr
 * (DegenPrimer/TD_Functions.py:60-101) for `n` sequences of length `len`
 * stored back to back; `concentration` is per variant (Primer.py:180-186),
 * conversion_degree in (0,1): 0.5 gives Tm. */
int dp_variant_tm(dp_ctx* ctx, uint32_t n, uint32_t len, const char* seqs,
                  double concentration, double conversion_degree,
                  const dp_conditions* cond, double* tm);

/* dp_variant_tm for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments, results bit-identical
 * to dp_variant_tm; the staging memory dp_variant_tm takes from the context's pool is ONE
 * file-scope `static __device__` array of 296 KiB (csrc/dp_static.cu) that every GPU which loads
 * the library holds until the process exits.  Limits per call: n <= 4096 sequences and
 * n * len <= 262144 bytes (dp_variant_tm_static_limits reports them); larger calls are refused
 * with DP_ERR_UNSUPPORTED and a message naming the limits.  The call makes no cudaMalloc /
 * cudaFree / stream-ordered or pool allocation and no host allocation on its success path.
 * All calls on a device share the array, which keeps the last call's contents (the call clears
 * what must start clean): calls that OVERLAP on one GPU -- two contexts, two threads -- race on
 * it; issue them one after the other. */
int dp_variant_tm_static(dp_ctx* ctx, uint32_t n, uint32_t len, const char* seqs,
                         double concentration, double conversion_degree,
                         const dp_conditions* cond, double* tm);
int dp_variant_tm_static_limits(uint32_t* max_seqs, uint64_t* max_seq_bytes, uint64_t* scratch_bytes);

/* dimer_dG_corrected for arbitrary dimers (offset != 0, unequal lengths:
 * dangling ends reachable) — DegenPrimer/TD_Functions.py:124-206.
 * Dimer i: seq1 = s1 + off1[i] (len1[i] bytes, 5'->3'), seq2 likewise,
 * match_mask over seq1 indices, offset as in SecStructures.Dimer. */
int dp_dimer_dg(dp_ctx* ctx, uint32_t n, const char* s1, const uint64_t* off1,
                const uint8_t* len1, const char* s2, const uint64_t* off2,
                const uint8_t* len2, const uint64_t* match_mask,
                const int32_t* offset, const dp_conditions* cond, double* dG);

/* SURVEY §8(f) N3 — primer self- and cross-dimers: SecStructures._find_dimers
 * (DegenPrimer/SecStructures.py:609-668) for a batch of sequence pairs.
 * Sequence i = seqs + off[i], len[i] <= 64 IUPAC letters, 5'->3'.  For pair p =
 * (pair_a[p], pair_b[p]) and every front f in [0, len_a + len_b - 1) — seq2's reverse
 * complement slid along seq1, Dimer.offset = max(0, len_a-f-1) - max(0, f-len_a+1) —
 * out[p*stride + f] holds the matching seq1 indices (full_mask; 0 = no match at this
 * front) and, of [dimer] + dimer.alternatives, the structure with the lowest
 * dimer_dG_corrected strictly below cond->max_dimer_dG (best_mask, best_dG; best_mask = 0
 * when none; best_dG = NaN where the reference raises for an unknown NN key).  Entries
 * f >= len_a + len_b - 1 are zero.  stride >= the largest front count of the batch.
 * The de-duplication by structure hash and the 3'-dimer counters of the reference are a
 * few dictionary operations per pair and stay with the caller (SecStructures.find_dimers).
 * ms_kernel (optional): device time of the kernel by CUDA events. */
typedef struct dp_front_dimer {
  uint64_t full_mask;
  uint64_t best_mask;
  double best_dG;
} dp_front_dimer;
int dp_find_dimers(dp_ctx* ctx, uint32_t n_seqs, const char* seqs, const uint64_t* off,
                   const uint8_t* len, uint32_t n_pairs, const uint32_t* pair_a,
                   const uint32_t* pair_b, const dp_conditions* cond, uint32_t stride,
                   dp_front_dimer* out, float* ms_kernel);

/* SURVEY §8(f) N3 — hairpins of a primer: SecStructures._find_hairpins
 * (DegenPrimer/SecStructures.py:676-716) with hairpin_dG_corrected
 * (DegenPrimer/TD_Functions.py:209-276) and UnifiedNN.hairpin_loop_dG/dH/dS_37
 * (DegenPrimer/UnifiedNN.py:279-325).  Tables: data/loops.csv (hairpin column) and
 * data/3-4-hairpin-loops.csv, flattened by the host (degenprimer_b200/UnifiedNN.py). */
typedef struct dp_hairpin_tables {
  /* loops[n].hairpin, or its extrapolation (UnifiedNN.py:254-261), n = unpaired bases
   * (index 63 also serves longer loops); NaN for n < 3 */
  double loop_dG[64];
  /* special tri-/tetra-loops, key = closing pair + loop letters (5 / 6 bases) read as a
   * base-4 number, first letter most significant, A=0 C=1 T=2 G=3; NaN = no row */
  double tri_dH[1024], tri_dG[1024];
  double tetra_dH[4096], tetra_dG[4096];
  double iK37;                /* -1000 / 310.15, UnifiedNN.py:135 */
} dp_hairpin_tables;
int dp_hairpin_tables_load(dp_ctx* ctx, const dp_hairpin_tables* tables);

typedef struct dp_best_structure {
  uint64_t best_mask;         /* 0 = nothing below the threshold */
  double best_dG;             /* NaN where the reference raises */
} dp_best_structure;
/* Seed k: sequence seed_seq[k] (index into seqs/off/len as for dp_find_dimers), stem
 * positions of the 5' arm seed_mask[k] and Dimer.offset seed_offset[k] of the symmetric
 * self-dimer it folds from — position f pairs with len-1-(f-offset); the caller has
 * already dropped the innermost pairs until >= 3 bases are left in the loop
 * (SecStructures.py:684-690).  out[k]: of [stem] + stem.alternatives the hairpin with the
 * lowest dG strictly below max_hairpin_dG (3 kcal/mol in the reference). */
int dp_find_hairpins(dp_ctx* ctx, uint32_t n_seqs, const char* seqs, const uint64_t* off,
                     const uint8_t* len, uint32_t n_seeds, const uint32_t* seed_seq,
                     const uint64_t* seed_mask, const int32_t* seed_offset,
                     const dp_conditions* cond, double max_hairpin_dG,
                     dp_best_structure* out, float* ms_kernel);

/* SURVEY §8(f) N1 — product pairing: the forward/reverse windowed join of
 * MixtureFactory.create_PCR_mixture (DegenPrimer/PCR_Mixture.py:216-236).  fwd_pos / rev_pos:
 * the positions (1-based 3' ends, as SearchEngine reports them) of the sites that can be
 * extended, each sorted ascending.  Returns every (forward site, reverse site) pair the
 * reference's loop turns into a product -- start = fwd_pos+1, end = rev_pos-1,
 * min_amplicon <= end-start+1 <= max_amplicon -- in the loop's order (forward index, then
 * reverse index), INCLUDING its quirks: the scan of reverse sites starts at `rev_start`, which the
 * reference sets to an index RELATIVE to the previous slice, and a first candidate shorter than
 * min_amplicon ends the scan for that forward site.  rev_start (optional, n_fwd entries) receives
 * the slice start used for every forward site: the reference remembers reverse annealings by
 * rev - rev_start.  The recurrence is walked by one device thread, everything else is parallel. */
typedef struct dp_pair {
  uint32_t fwd;   /* index into fwd_pos */
  uint32_t rev;   /* index into rev_pos */
} dp_pair;
int dp_pair_products(dp_ctx* ctx, uint32_t n_fwd, const int64_t* fwd_pos, uint32_t n_rev,
                     const int64_t* rev_pos, int64_t min_amplicon, int64_t max_amplicon,
                     dp_pair** pairs, uint64_t* n_pairs, uint32_t* rev_start);

/* ---- roofline calibration -------------------------------------------------------
 * Measured issue rate of the pipes the scan kernels live on (the
 * integer roofline's denominators must be measured on the box).
 * which: 0 = LOP3 (ALU pipe), 1 = POPC, 2 = conflict-free LDS.32.
 * lane_ops_per_s = warp instructions x 32 per second over the whole chip. */
int dp_calibrate(dp_ctx* ctx, int which, int reps, double* lane_ops_per_s,
                 double* ms_best);

#ifdef __cplusplus
}
#endif
#endif /* DEGENPRIMER_B200_H */
