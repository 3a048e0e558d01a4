/*
 * danbing_b200.h -- C ABI of libdanbing_b200.so, the B200-native (sm_100a) implementation of the
 * `danbing-tk align` hot path (read pair -> VNTR locus -> per-locus k-mer counts).
 *
 * The reference (ChaissonLab/danbing-tk @ 324930d) has no plugin / FFI interface for this path: it
 * sits behind a process boundary (argv + files + stdin/stdout, src/aQueryFasta_thread.cpp:2286-2660).
 * The drop-in is therefore (1) the `danbing-tk` CLI built from host/ on top of this ABI and (2) this
 * ABI itself, which is what a maintainer's binding (cgo / JNI / ctypes / a patched CountWords) would
 * call instead of the per-pair loop of CountWords (src/aQueryFasta_thread.cpp:2002-2249).
 * INTEGRATION.md shows that binding.  Every entry point below cites the reference code it replaces;
 * "cpp" = src/aQueryFasta_thread.cpp, "h" = src/aQueryFasta_thread.h, "bio" = src/binaryKmerIO.hpp,
 * "kio" = src/kmerIO.hpp.
 *
 * Conventions: plain C, no exceptions, no C++/torch types.  Return 0 = OK, negative = error
 * (dbt_strerror / dbt_last_error).  The caller owns every host pointer; a dbt_ctx owns all device
 * memory, streams and events.  A context is driven by one host thread at a time; different
 * contexts (GPUs) may be driven concurrently.  There is NO CPU fallback: without a CUDA device
 * dbt_create fails with DBT_E_CUDA.
 */
#ifndef DANBING_B200_H_
#define DANBING_B200_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DBT_ABI_VERSION 6

/* ---- error codes -------------------------------------------------------------------------- */
#define DBT_OK            0
#define DBT_E_CUDA       (-1)  /* CUDA runtime/driver error (message in dbt_last_error)         */
#define DBT_E_NOMEM      (-2)  /* host or device allocation failed                             */
#define DBT_E_BADARG     (-3)  /* NULL / out-of-range argument                                 */
#define DBT_E_CAPACITY   (-4)  /* batch, read length or record buffer exceeds the context's     */
#define DBT_E_STATE      (-5)  /* call order violated (e.g. submit before load_index)          */
#define DBT_E_INDEX      (-6)  /* inconsistent RPGG index (tr/fl k-mer missing from kmers.dbi) */
#define DBT_E_IO         (-7)  /* file missing/short (host loaders); the reference asserts or, */
                               /* for .fl.kdb/.tre.kdb, reads garbage (h:681) -- we fail loudly */
#define DBT_E_NCCL       (-8)  /* NCCL unavailable or a collective failed                      */

/* ---- run flags (dbt_params.flags) ---------------------------------------------------------- */
#define DBT_F_KAM        0x001u /* kam records wanted (reference default; `-ka` clears, cpp:2368) */
#define DBT_F_QC         0x002u /* `-qc FILE` locus mask active (cpp:2381, 2059-2062)           */
#define DBT_F_TRACE      0x004u /* test/diagnostic: emit a record for EVERY pair, not only kam   */
#define DBT_F_THREADING  0x008u /* `-g/-gc/-gcc` (cpp:2360).  With a graph in dbt_index: restored graph threading
                                   (isThreadFeasible, cpp:1114-1260, pair glue = SURVEY.md §8(c) decision D1).
                                   Without one: the reference's HEAD behaviour (caller commented out, cpp:2072) */
#define DBT_F_CORRECTION 0x010u /* `-gc` error-correcting threading                            */
#define DBT_F_ALN        0x020u /* `-a`                                                        */
#define DBT_F_ALN_MIN    0x040u /* `-ae`                                                       */
#define DBT_F_SIM        0x100u /* `-s 1|2` (cpp:2169): kam records also for pairs whose both mates assignTRkmc removed (dst = nloci); the
                                   caller prints them when the read name carries a source locus */
#define DBT_F_BUBBLES    0x200u /* `-bu` (cpp:1559-1606): novel (k+1)-mer edges inside the observed TR segment of every kept mate of a counted
                                   pair are handed back as events (dbt_bubble_events); needs tre_n / tre_kmers in dbt_index; not with threading */
#define DBT_F_BAIT       0x400u /* `-b` (cpp:1377-1419, 2110-2125; with qualities in the batch the quality-masked form): a pair is removed when a bait k-mer of its locus occurs in a mate fewer
                                   than min or more than max times; needs bt_n / bt_kmers / bt_vals in dbt_index */
#define DBT_F_EXTRACT    0x080u /* `-e 1|2` (cpp:2094-2099, 1618-1644): a pair that passes the locus vote and the QC gate is
                                   handed out as a record (status COUNTED, dst = its locus, no annotation) and NOTHING is counted */

/* ---- per-pair status (dbt_pair_record.status); order = the order of the exits in cpp:2035-2145 */
#define DBT_ST_SHORT       0   /* a mate has no valid k-mer window            (cpp:2037)       */
#define DBT_ST_SUBFILTERED 1   /* rejected by subfilter                       (cpp:2046)       */
#define DBT_ST_KFILTERED   2   /* both mates removed by kfilter               (cpp:2054)       */
#define DBT_ST_UNASSIGNED  3   /* countHit returned nloci                     (cpp:2058)       */
#define DBT_ST_QCFILTERED  4   /* locus fails QC                              (cpp:2059)       */
#define DBT_ST_ASGN_REMOVED 5  /* both mates removed by assignTRkmc           (cpp:2145)       */
#define DBT_ST_COUNTED     6   /* counts accumulated, kam line written        (cpp:2146-2175)  */
#define DBT_ST_THREAD_FAILED 8 /* restored threading: neither mate threads through graphDB[destLocus]
                                  (commented cpp:2087 "removed by threading"); COUNTED = at least one does    */
#define DBT_ST_NOT_THREADED 7  /* DBT_F_THREADING at the reference's HEAD: the pair enters the threading step (cpp:2070)
                                  and nothing passes it -- the only caller of isThreadFeasible is commented
                                  out (cpp:2072-2088), so `alned` stays false and cpp:2090 skips counting and output */

/* ---- counters (dbt_download / dbt_batch_result.counters); cpp:1820, 2617-2626 --------------- */
enum {
  DBT_C_NREADS = 0,            /* reads (2 per pair) submitted                                 */
  DBT_C_NSHORT,                /* PAIRS skipped as short (cpp:2038 counts pairs)               */
  DBT_C_NSUBFILTERED,          /* reads                                                        */
  DBT_C_NKMERFILTERED,
  DBT_C_NBAITFILTERED,
  DBT_C_NQUALFILTERED,
  DBT_C_NLOCUSASSIGNFILTERED,
  DBT_C_NQCFILTERED,
  DBT_C_NTHREADINGREADS,
  DBT_C_NFEASIBLEREADS,
  DBT_C_NASGNREADS,
  DBT_C_NHASH0,                /* subfilter probe count, counted as cpp:176,182 do              */
  DBT_C_NHASH1,                /* kfilter probe count, cpp:206,217                              */
  DBT_C_NINC,                  /* TR k-mer occurrences incremented (= sum of counts delta)      */
  DBT_C_NCANDIDATES,           /* pairs that reached kfilter                                    */
  DBT_C_NRECORDS,              /* records emitted                                               */
  /* measurement only (no reference counterpart): share of the work done by the generic candidate kernel */
  DBT_C_GEN_PAIRS,             /* candidate pairs processed by the generic (exact replay) kernel */
  DBT_C_GEN_NHASH1,            /* their part of DBT_C_NHASH1                                    */
  DBT_C_GEN_NTHREADINGREADS,   /* their part of DBT_C_NTHREADINGREADS                           */
  DBT_C_GEN_NINC,              /* their part of DBT_C_NINC                                      */
  /* measurement only: graph probes of restored threading as the REFERENCE algorithm issues them      */
  DBT_C_THR_WALK,              /* k-mers of the mates handed to isThreadFeasible (one node look-up per step, cpp:1170) */
  DBT_C_THR_EC,                /* look-ups inside errorCorrection_forward/backward: expansion (cpp:907-920) + extension walks */
  DBT_C_THR_CLS,               /* valid k-mers of the edited vectors looked up in trResults (cpp:2203-2208) */
  DBT_C_THR_REFRESH,           /* implementation detail: per-position probes the two threading kernels actually issued */
  /* measurement only: the part of the above that belongs to thread_kernel (mates with errors); the rest is thread_fast_kernel's */
  DBT_C_THR2_PAIRS,            /* pairs with at least one mate handed to thread_kernel            */
  DBT_C_THR2_MATES,            /* mates threaded by thread_kernel                                 */
  DBT_C_THR2_WALK,             /* their part of DBT_C_THR_WALK                                    */
  DBT_C_THR2_CLS,              /* thread_kernel's part of DBT_C_THR_CLS                           */
  DBT_C_THR2_NINC,             /* thread_kernel's part of DBT_C_NINC                              */
  DBT_C_THR2_REFRESH,          /* thread_kernel's part of DBT_C_THR_REFRESH                       */
  DBT_C_THR1_FIXED,            /* mates with one substitution that thread_fast_kernel corrected itself (alned = 2) */
  DBT_N_COUNTERS = 32
};

typedef struct dbt_ctx dbt_ctx;

/* Run parameters == the reference's globals/flags (cpp:25-34, 2336-2339, 2344-2429). */
typedef struct {
  int32_t  device;             /* CUDA device ordinal                                          */
  uint32_t k;                  /* -k   ksize, 1..31 (cpp:2376)                                  */
  uint32_t n_filter;           /* -kf INT1  N_FILTER  (cpp:2370); 0 disables subfilter          */
  uint32_t nm_filter;          /* -kf INT2  NM_FILTER (cpp:2371); 0 disables subfilter          */
  uint32_t cth;                /* -cth Cthreshold (cpp:2422)                                    */
  uint32_t nm_tr;              /* -c   NM_TR (cpp:2374)                                         */
  uint32_t max_nt;             /* MAX_NT, fixed 2 in the reference (cpp:31)                     */
  uint32_t thread_cth;         /* -g/-gc INT1 (cpp:2363)                                        */
  uint32_t max_ncorrection;    /* -gc INT2 (cpp:2364)                                           */
  uint32_t flags;              /* DBT_F_*                                                       */
  uint32_t max_read_len;       /* longest read a batch may contain (bases)                      */
  uint32_t max_batch_pairs;    /* largest npairs of one dbt_submit                              */
  uint32_t max_records;        /* record buffer capacity per slot; 0 = max_batch_pairs          */
  float    table_load;         /* hash table load factor; 0 = default                           */
  uint32_t bucket_slots;       /* 16-byte slots per probe bucket: 2 (one 32 B sector) or 4; 0 = default (2) */
  uint32_t qth;                /* `-qth` [20]: with DBT_F_BAIT and qualities in the batch, only k-mers whose bases all have phred >= qth enter
                                  the bait histogram (qString2qMask, h:1038-1071) */
  uint32_t reserved[2];
} dbt_params;

/*
 * Host view of the RPGG index == the containers the reference builds at start-up (cpp:2465-2504):
 *   dbi_keys/dbi_vals/vv : PREF.kmers.dbi  (readBinaryIndex, h:654-673; value encoding kio:59-74)
 *   tr_n/tr_kmers        : PREF.tr.kmers   (readKmersWithZeroCount, h:469-480), per-locus sizes +
 *                          keys ALREADY in the reference's output order (the libstdc++
 *                          unordered_map iteration order of each locus map; dbt_host_load_index
 *                          reproduces it) -- counts[] comes back in exactly this order
 *   fl_n/fl_kmers        : PREF.fl.kdb     (readBinaryKmerSetDB, h:675-698)
 *   qc                   : -qc FILE after `- '0'` (readQCFile, kio:111-120); NULL unless DBT_F_QC
 *   g_n/g_kmers/g_masks  : PREF.graph.kmers (readGraphKmers, h:549-575; GraphType h:32): per-locus node
 *                          counts, strand-specific k-mers, 4-bit out-edge masks (bit b: successor
 *                          ((node & rmask) << 2) + b exists, cpp:533-539).  NULL = no graph.
 * The caller may free everything once dbt_load_index returns.
 */
typedef struct {
  uint64_t nloci;
  uint64_t nk_dbi;  const uint64_t* dbi_keys;  const uint32_t* dbi_vals;
  uint64_t nvv;     const uint32_t* vv;
  const uint64_t* tr_n;   const uint64_t* tr_kmers;
  const uint64_t* fl_n;   const uint64_t* fl_kmers;
  const uint8_t*  qc;
  const uint64_t* g_n;    const uint64_t* g_kmers;  const uint8_t* g_masks;
  const uint64_t* bt_n;   const uint64_t* bt_kmers;  const uint16_t* bt_vals;   /* PREF[.qc].bt.kmdb: per-locus bait k-mers and (min << 8 | max), readBinaryBaitDB h:542-547; NULL unless DBT_F_BAIT */
  const uint64_t* tre_n;  const uint64_t* tre_kmers;  /* PREF.tre.kdb: per-locus TR (k+1)-mer edges (readBinaryKmerSetDB, h:675-698); NULL unless DBT_F_BUBBLES */
} dbt_index;

#define DBT_LAYOUT_ASCII   0   /* concatenated ASCII read sequences + byte offsets               */
#define DBT_LAYOUT_PACKED  1   /* output of dbt_pack_reads (2-bit bases + invalid mask, SoA)    */

/*
 * One batch of read pairs == seqs[] of one CountWords batch (cpp:1857, 1972-1974).
 * Read 2i is the reference's "read 1" (seq1, the mate that ARRIVED SECOND), read 2i+1 its "read 2".
 *   ASCII : bases = concatenated sequences, offsets[2*npairs+1] = byte offset of each read.
 *   PACKED: bases = buffer filled by dbt_pack_reads for exactly this npairs; offsets = NULL.
 * Host buffers should be pinned (dbt_host_alloc) for asynchronous copies; pageable works, slower.
 */
typedef struct {
  uint64_t first_pair_index;   /* global index of pair 0 (only echoed into records)             */
  uint64_t npairs;
  int32_t  layout;
  const uint8_t*  bases;
  const uint64_t* offsets;
  const uint8_t*  quals;       /* ASCII layout only, may be NULL: phred+33 quality strings laid out like `bases` (same offsets); used by DBT_F_BAIT */
} dbt_batch;

/* per-mate part of a record == km_asgn_read_t (cpp:93-108) */
typedef struct {
  int16_t  si, ei, si_, ei_, ti;
  int8_t   nt, bs;
  uint8_t  flags;              /* bit0 kf, bit1 hf, bit2 bf, bit3 qf, bit4 af, bit5 rm          */
  uint8_t  pad;
  uint16_t nk;                 /* length of the annotation `as` (0 when the reference leaves it empty) */
} dbt_mate_record;
#define DBT_MF_KF 0x01
#define DBT_MF_HF 0x02
#define DBT_MF_BF 0x04
#define DBT_MF_QF 0x08
#define DBT_MF_AF 0x10
#define DBT_MF_RM 0x20

/*
 * One record == one km_asgn_t / kam line (cpp:110-144, 1646-1681) plus trace fields.
 * A record occupies dbt_record_stride(ctx) bytes: this 64-byte header, then the 2-bit k-mer
 * annotation (`as`: 0 '*', 1 '.', 2 '=', cpp:122) of mate[0] and of mate[1], each
 * dbt_record_as_stride(ctx) bytes, k-mer i in bits 2*(i%4) of byte i/4.
 * mate[0] is the reference's r1 (read 2i), mate[1] its r2.
 */
typedef struct {
  uint64_t pair_index;
  uint32_t dst, dst0;          /* destLocus (nloci = unassigned), destLocus0 (cpp:2056)         */
  uint32_t fc, rc;             /* top.fc / top.rc when voting stopped (nm1, nm2, cpp:437-438)   */
  uint16_t nhash0, nhash1;     /* probes this pair added to nhash0 / nhash1                     */
  uint8_t  status;             /* DBT_ST_*                                                      */
  uint8_t  pad[3];
  dbt_mate_record mate[2];
} dbt_pair_record;

/*
 * Restored threading (DBT_F_THREADING with a graph): a record is the same 64-byte header (mate[].flags
 * carry kf/hf/rm; the asgn fields stay at their defaults) followed, per mate, by one alignment blob of
 * dbt_record_aln_stride(ctx) bytes == sam_t.r1 / sam_t.r2 (cpp:53-75):
 *   dbt_aln_mate header, then es[nes] (CIGAR entries, one byte each: bits 0-2 type 0 '*' 1 '=' 2 'X'
 *   3 'D' 4 'I' (edit_t.t, cpp:46-51), bits 3-5 graph base 0 none 1 'A' 2 'C' 3 'G' 4 'T' 5 '*' (edit_t.g)),
 *   starting at byte sizeof(dbt_aln_mate); then tr[ntr] (k-mer annotation, 0 '*' 1 '.' 2 '=', cpp:60)
 *   starting at byte sizeof(dbt_aln_mate) + dbt_record_es_cap(ctx).
 * Records are emitted for the pairs the reference prints an .aln line for (cpp:2232-2239): with
 * DBT_F_ALN every pair that entered threading, with DBT_F_ALN_MIN only threaded ones; DBT_F_TRACE: all.
 */
#define DBT_ES_TYPE(b)  ((b) & 7u)          /* index into "*=XDI"  */
#define DBT_ES_BASE(b)  (((b) >> 3) & 7u)   /* index into "\0ACGT*" */
/* capacities of es[] / tr[] inside a blob: the read plus what deletions can insert */
#define DBT_ES_CAP_FOR(max_read_len, max_ncorrection) ((((uint32_t)(max_read_len) + 2u * (uint32_t)(max_ncorrection) + 8u + 7u) / 8u) * 8u)
#define DBT_TR_CAP_FOR(max_read_len, k, max_ncorrection) \
  (((((uint32_t)(max_read_len) >= (uint32_t)(k) ? (uint32_t)(max_read_len) - (uint32_t)(k) + 1u : 1u) + 2u * (uint32_t)(max_ncorrection) + 8u + 7u) / 8u) * 8u)
typedef struct {
  uint8_t  alned;              /* isThreadFeasible's return: 0 infeasible, 1 threaded, 2 threaded with edits */
  uint8_t  err;                /* 1 = the graph is not closed / capacity exceeded (reference: assert or UB) */
  uint16_t nes, ntr, nkm;      /* lengths of es, tr and of the edited k-mer vector                        */
  uint64_t kmers_hash;         /* FNV-1a over the edited k-mer vector (u64 little-endian), for parity tests */
} dbt_aln_mate;

typedef struct {
  uint64_t n_records;          /* out */
  uint64_t capacity;           /* in: records the buffer can hold                               */
  uint8_t* records;            /* in: caller buffer, capacity * dbt_record_stride bytes (pinned) */
  uint64_t counters[DBT_N_COUNTERS]; /* out: this batch's counter deltas                        */
} dbt_batch_result;

/* ---- lifecycle ----------------------------------------------------------------------------- */
int  dbt_abi_version(void);
int  dbt_device_count(void);                         /* <0 on error; 0 = no GPU                  */
void dbt_default_params(dbt_params* p);              /* the reference's defaults (cpp:26-33, 2338) */
int  dbt_create(dbt_ctx** out, const dbt_params* p); /* replaces Counts/Threads set-up, cpp:2509-2565 */
void dbt_destroy(dbt_ctx*);
const char* dbt_strerror(int code);
const char* dbt_last_error(void);                    /* thread-local detail of the last failure  */

/* Builds the device-resident index (replaces cpp:2484-2504 + the per-pair container lookups). */
int  dbt_load_index(dbt_ctx*, const dbt_index*);

/* ---- per batch (replaces the pair loop of CountWords, cpp:2002-2249) ------------------------ */
int  dbt_upload (dbt_ctx*, const dbt_batch*, int slot);       /* async H2D (+ device 2-bit pack) */
int  dbt_process(dbt_ctx*, int slot);                         /* async kernels on the uploaded slot */
int  dbt_fetch  (dbt_ctx*, dbt_batch_result*, int slot);      /* async D2H of records + counters */
int  dbt_submit (dbt_ctx*, const dbt_batch*, dbt_batch_result*, int slot); /* upload+process+fetch */
int  dbt_wait   (dbt_ctx*, int slot);                         /* blocks; finalises dbt_batch_result */
#define DBT_N_SLOTS 2

/* ---- raw FASTA/FASTQ ingestion ----------------------------------------
 * Replaces, for name-collated input, the serial read loop of CountWords (cpp:1918-1976: four / two getline calls per
 * record, prunePEinfo cpp:455-462, pairing through a hash map, the late mate = read 1, pairs with a mate shorter than
 * min_read_len forgotten).  `text` = whole records of the file as they are (FASTA: title line + sequence line; FASTQ: four
 * lines), last byte '\n', < 4 GiB.  The device finds the lines, pairs every record with the record before it when their
 * pruned titles are equal, packs the sequences and leaves the slot in the state dbt_upload leaves it in: dbt_process /
 * dbt_fetch / dbt_wait follow.  Pair i of the batch = the i-th such pair in stream order (the reference's order).
 * The call returns after the split (it synchronises the slot's stream; the other slot keeps running):
 *   info            counts, flags, the longest read
 *   pair_offset[i]  byte offset in `text` of the first record of pair i (its mate is the next record) -- kam / aln
 *                   lines are formatted from the text itself
 *   pair_hash[i]    dbt_title_hash of the pair's pruned title (may be NULL)
 *   single_offset[] byte offsets of the records left unpaired, ascending: the caller keeps the reference's map for them
 * flags != 0 (three or more equal titles in a row, or more lines / pairs / singletons than the capacities): nothing was
 * uploaded, the caller pairs this text on the host and submits it with dbt_upload.  DBT_E_CAPACITY with info filled: a read
 * is longer than max_read_len, or more pairs than max_batch_pairs. */
#define DBT_RAW_SLOWPATH 1u
#define DBT_RAW_OVERFLOW 2u
typedef struct {
  uint32_t nlines, nrecords, npairs, nsingle, nshort_pairs, flags, max_read_len, max_record_bytes;
} dbt_raw_info;
typedef struct {
  uint64_t first_pair_index;
  const uint8_t* text; uint64_t nbytes;
  int32_t  is_fastq; uint32_t min_read_len;     /* Cthreshold + k - 1 (cpp:1892) */
  uint32_t* pair_offset; uint64_t* pair_hash;   /* [max_batch_pairs] */
  uint32_t* single_offset; uint32_t single_cap;
} dbt_raw_batch;
int  dbt_upload_raw(dbt_ctx*, const dbt_raw_batch*, dbt_raw_info* info, int slot);
/* optional: allocate both slots (and, raw_bytes > 0, the splitter's buffers for texts of that size) now instead of on first use */
int  dbt_prepare(dbt_ctx*, uint64_t raw_bytes);
uint64_t dbt_title_hash(const uint8_t* title, size_t len);     /* FNV-1a 64 of the pruned title, as the device computes it */
/* `-bu`: after dbt_wait(slot), the novel edges of the batch (unordered; the caller sorts by key and replays them into the reference's
   containers, cpp:1565, 1599-1606).  key = pair index within the batch << 10 | mate << 9 | k-mer position; locus = the pair's locus.
   Returns DBT_E_CAPACITY when the batch produced more events than the slot holds (4 per pair + 2^20): submit smaller batches. */
int  dbt_bubble_events(dbt_ctx*, int slot, uint64_t* key, uint64_t* edge, uint32_t* locus, uint64_t cap, uint64_t* n);
/* [0] H2D + split kernels [1] pack_kernel of the slot's last dbt_upload_raw, milliseconds */
int  dbt_raw_times(dbt_ctx*, int slot, float ms[2]);

/* ---- results (replaces dumpTRKmers/writeTRKmerSummary inputs, cpp:2631-2641) ---------------- */
int  dbt_download(dbt_ctx*, uint64_t* trkmc /*[nk_tr]*/, uint64_t* kmc /*[nloci]*/,
                  uint32_t* nmapread /*[nloci]*/, uint64_t counters[DBT_N_COUNTERS]);
int  dbt_reset_counts(dbt_ctx*);
uint64_t dbt_nk_tr(const dbt_ctx*);
uint64_t dbt_nloci(const dbt_ctx*);
uint32_t dbt_record_stride(const dbt_ctx*);
uint32_t dbt_record_as_stride(const dbt_ctx*);
uint32_t dbt_record_aln_stride(const dbt_ctx*);   /* 0 unless restored threading is active            */
uint32_t dbt_record_es_cap(const dbt_ctx*);

/* ---- multi-GPU: one context per GPU, reads sharded, one sum-reduce at the end --------------- */
#define DBT_NCCL_ID_BYTES 128
int  dbt_nccl_unique_id(uint8_t id[DBT_NCCL_ID_BYTES]);
int  dbt_comm_init(dbt_ctx*, int nranks, int rank, const uint8_t id[DBT_NCCL_ID_BYTES]);
/* One grouped ncclAllReduce(sum) of counts, kmc, nmapread and the counters, OUT OF PLACE: the rank's own arrays are left
   untouched; dbt_download hands out the all-rank sums until the next dbt_process / dbt_reset_counts.  Calling it again (or
   after more batches) reduces the current local state afresh -- nothing is ever multiplied by the number of ranks.
   Requires every dbt_fetch to have been finished with dbt_wait.  Replaces "one address space, one set of atomics" of
   cpp:2598-2656. */
int  dbt_allreduce(dbt_ctx*);

/* ---- measurement helpers (CUDA events on the context's own streams) ------------------------- */
int  dbt_timer_start(dbt_ctx*, int slot);
int  dbt_timer_stop (dbt_ctx*, int slot, float* ms);          /* synchronises the slot           */
/* [0] pack (the slot's last ASCII upload) [1] subfilter [2] compaction [3] candidate_fast [4] candidate (generic)
   [5] thread_fast [6] thread (mates with errors); the rest 0 */
int  dbt_kernel_times(dbt_ctx*, int slot, float ms[8], uint32_t launches[8]); /* last dbt_process */
uint64_t dbt_launch_count(const dbt_ctx*);                    /* kernels launched since create   */
/* after dbt_wait(slot): [0] candidate pairs, [1] pairs the closed-form kernel left to the generic one,
   [2] pairs handed to the threading kernel */
int  dbt_batch_stats(dbt_ctx*, int slot, uint64_t stats[4]);
/* [0] main buckets [1] slots per bucket [2] keys [3] used slots [4] main buckets that overflowed
   [5] keys in the stash [6] table bytes [7] stash buckets that overflowed */
int  dbt_table_stats(const dbt_ctx*, uint64_t stats[8]);
int  dbt_graph_merged(const dbt_ctx*);                         /* 1: the graphs are merged into the index entries (one probe round feeds the vote and the threading); 0: no graph / switched off; < 0: refused, -(1 a node is no index key | 2 key of another locus | 4 not in its shared list | 8, 16 a TR k-mer lost its count position to the flank class) */
uint64_t dbt_graph_table_bytes(const dbt_ctx*);                /* bytes of the graph table in HBM (0 without restored threading) */
#define DBT_GATHER_COOPERATIVE 0x10000u /* OR into sector_bytes (32 / 64 / 128): sector_bytes/16 adjacent lanes fetch one chunk as ONE coalesced request */
#define DBT_GATHER_WIDE_LOAD   0x20000u /* OR into sector_bytes = 32: each thread fetches its chunk with one 256-bit load (what the probe functions issue) */
int  dbt_random_gather_bench(int device, uint64_t table_bytes, uint32_t sector_bytes,
                             uint64_t n_gathers, int iters, float* best_ms); /* R_rand roofline    */

/* ---- bias correction downstream of OUT.trkmc.ar ---------------------------
 * dbt_pred replaces load_eachBinGT + norm_rd + bias_correction of `danbing-tk-pred` (src/pred.cpp:58-80, src/pred.h:141-195).
 * All matrices are column-major float32 as the reference's Eigen arrays are: element (sample s, k-mer k) of gt at s + k*ns,
 * (sample s, locus t) of bias at s + t*ns.  Results are bit-identical to the reference's (same IEEE operations in the same order).
 * Columns of `bias` (and blocks of gt_corrected = gt_raw) of loci without k-mers or without invariant k-mers are left as the caller
 * passed them / uncorrected (pred.h:181).  Every invariant k-mer index must lie inside its own locus (DBT_E_INDEX otherwise). */
typedef struct {
  uint64_t ns, nk, n_tr, n_ik;     /* samples, k-mers, loci, invariant k-mers                                          */
  const uint64_t* const* kmc;      /* ns arrays of nk counts: the payload of each sample's OUT.trkmc.ar (pred.h:141-164) */
  const float*    rd;              /* ns read depths (column 2 of the metadata file, pred.h:41-50)                     */
  const uint32_t* nk_cum;          /* n_tr: k-mers up to and including locus t (ikmer.meta, pred.h:69-123)             */
  const uint32_t* nik_cum;         /* n_tr: invariant k-mers up to and including locus t                               */
  const uint32_t* iki;             /* n_ik: k-mer index of each invariant k-mer                                        */
  const uint8_t*  ikmc;            /* n_ik: its expected count                                                         */
} dbt_pred_input;
/* kernel_ms (may be NULL): [0] normalisation + transpose, [1] bias correction, CUDA events */
int  dbt_pred(int device, const dbt_pred_input* in, float* gt_raw, float* gt_corrected, float* bias, float kernel_ms[2]);

/* ---- allocation-free packing ------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * dbt_pack_reads_device_noalloc = dbt_pack_reads (same arguments minus the thread count, same output layout and bytes), computed
 * by pack_kernel on GPU `device` inside a fixed, file-scope device array of DBT_NOALLOC_SCRATCH_BYTES that every GPU loading the
 * library holds until exit.  It calls no cudaMalloc / cudaFree / allocating routine and works on the legacy default stream.
 * Limits: npairs <= DBT_NOALLOC_MAX_PAIRS and dbt_noalloc_scratch_needed(max_read_len, npairs, bases in the batch) <=
 * DBT_NOALLOC_SCRATCH_BYTES; anything larger returns DBT_E_CAPACITY (message in dbt_last_error), nothing is allocated instead.
 * All calls share the array (it keeps the previous call's contents; the call clears what must start at zero): calls that overlap
 * on one GPU race -- callers inside one process are serialised by the library, other users of the GPU are the caller's business. */
#define DBT_NOALLOC_SCRATCH_BYTES (8u << 20)
#define DBT_NOALLOC_MAX_PAIRS     16384u
size_t dbt_noalloc_scratch_needed(uint32_t max_read_len, uint64_t npairs, uint64_t nbases);
int  dbt_pack_reads_device_noalloc(int device, uint32_t k, uint32_t max_read_len, uint64_t npairs, const uint8_t* ascii,
                                   const uint64_t* offsets, uint8_t* packed_out);

/* pinned host -> device copy rate of this box (what bounds `e2e` once the kernels are faster than the bus): best of `iters`
   copies of `bytes`, and the same for device -> host */
int  dbt_bus_bench(int device, uint64_t bytes, int iters, float* h2d_ms, float* d2h_ms);

/* Creates the CUDA context of `device` (nothing else), so that a later dbt_create / dbt_pred does not pay for it: meant to be
   called from a helper thread while the host parses its inputs (bin/danbing-tk-pred does). */
int  dbt_init_device(int device);

/* ---- host helpers --------------------------------------------------------------------------- */
void* dbt_host_alloc(size_t bytes);                           /* pinned */
void  dbt_host_free(void*);
/* packed layout sizes for npairs reads of <= max_read_len bases */
size_t dbt_packed_bytes(uint32_t max_read_len, uint64_t npairs);
/* Multi-threaded host packer (north_star item 1). nthreads<=0: hardware concurrency. */
int  dbt_pack_reads(uint32_t k, uint32_t max_read_len, uint64_t npairs, const uint8_t* ascii,
                    const uint64_t* offsets, uint8_t* packed_out, int nthreads);

#ifdef __cplusplus
}
#endif
#endif /* DANBING_B200_H_ */
