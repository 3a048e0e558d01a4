// Device-side views shared by the kernels and the C ABI implementation.
#pragma once
#include <stdint.h>
#include "../../include/danbing_b200.h"
#include "kmer.cuh"
#include "table.cuh"

namespace dbt {

constexpr uint32_t NAN32 = 0xFFFFFFFFu;
constexpr int VOTE_SMEM_SLOTS = 64;          // per-pair shared-memory (locus -> hits1,hits2) entries
constexpr uint32_t VOTE_SPILL_AT = VOTE_SMEM_SLOTS - 8;  // loci held in shared memory before a pair spills to the global pool
                                                        // (RunParams::vote_spill_at; the DBT_VOTE_SPILL_AT environment variable lowers it in tests)
constexpr uint32_t VOTE_POOL = 64;           // global spill arrays [nloci+1], shared under locks by all in-flight kernels of a context
constexpr int CAND_WARPS_SLOW = 4;             // generic candidate_kernel: warps per block

// resident RPGG index
struct DevIndex {
	Table    table;
	const uint32_t* vv;         // kmerDBi_vv (cpp:2470): [n, locus_0 .. locus_{n-1}]*
	const uint32_t* vv_aux;     // parallel to vv: (class<<30)|idx of the k-mer in that locus
	uint64_t nvv;
	uint64_t* counts;           // trResults flattened in output order (cpp:2465, h:968-973)
	uint64_t* kmc;              // cpp:2476
	uint32_t* nmapread;         // cpp:2475
	const uint8_t* qc;          // qcFilter (cpp:2471) or nullptr
	const uint16_t* permtab;    // all-ties std::sort permutations, perm_U at offset U*(U-1)/2
	uint32_t permtab_maxU;
	uint64_t* vote_pool;        // VOTE_POOL epoch-tagged hits arrays [(nloci+1)] (spill target of the smem vote table)
	uint32_t* vote_lock;        // [VOTE_POOL] 0 free / 1 held
	uint32_t* vote_epoch;       // [VOTE_POOL] last epoch used in each array
	uint64_t nloci, nk_tr;
	Table    gtable;            // graphDB of all loci (graph_table.cuh); slots == nullptr when no graph is loaded
	Table    btable;            // `-b`: bait k-mers of all loci, keyed (k-mer, locus), payload = min << 8 | max (bait.cuh); slots == nullptr: no bait filter
	const uint32_t* vv_g;       // parallel to vv: graph entry (ten bits, see table.cuh slot_val) of the k-mer in that locus
	const uint32_t* pre;        // presence bitmap over the index keys (one hash, 2^pre_bits bits, >= 30 bits per key): a clear bit proves a
	uint32_t pre_bits;          // k-mer absent without touching the table -- subfilter_kernel's probes of background reads.  0: none
	uint32_t gmerged;           // 1: every graph node is an index key of its locus and every TR k-mer keeps its count position in the
	                            // index entry -> candidate_fast_kernel threads the pairs itself from the entries it already holds
};

struct RunParams {
	uint32_t k, n_filter, nm_filter, cth, nm_tr, max_nt, flags;
	ReadLayout lay;
	uint32_t max_read_len;
	uint32_t nk_max;            // max_read_len - k + 1
	uint32_t nkp;               // nk_max rounded up to a multiple of 32
	uint32_t umax;              // 2*nk_max rounded up to 4
	uint32_t sortn;             // power of two >= 2*nk_max
	uint32_t as_stride, rec_stride;
	uint32_t warp_smem;         // bytes of shared memory per warp in candidate_kernel
	uint32_t vote_spill_at;     // VOTE_SPILL_AT, or the test override
	uint32_t max_records;
	// restored graph threading (P1); all 0 unless DBT_F_THREADING and a graph is loaded
	uint32_t thr_on;            // 1 = hand candidates to thread_kernel instead of the HEAD no-op
	uint32_t thread_cth, max_ncorrection, correction;
	uint32_t es_cap, tr_cap, aln_stride;   // record blob geometry (include/danbing_b200.h, dbt_aln_mate)
	uint32_t qth;               // `-qth`
	uint32_t coop_stage;        // the tables are far beyond L2: stage buckets with one coalesced request each (table.cuh, stage_buckets_coop)
};

// one in-flight batch
struct SlotView {
	const uint64_t* packed[2];  // [mate][pair][W64]
	const uint16_t* meta[2];    // [mate][pair]
	uint32_t* passmask;         // bit per pair: passed subfilter
	uint32_t* blocksum;         // compaction scratch
	uint32_t* cand;             // candidate pair ids, ascending
	uint32_t* ncand;            // [0] = number of candidates, [1] = error flags, [2] = length of slow[], [3] = length of thr[], [4] = length of thr2[]
	uint32_t* slow;             // candidate ordinals left to the generic kernel
	uint4*    thr;              // work items of thread_kernel (ncand[3] entries): x = candidate ordinal, y = pair index within the
	                            // batch, z = dst | kf bits << 28 | rm bits << 30 -- one 16-byte load per pair instead of a chain of three
	uint4*    thr2;             // work items of thread_kernel (ncand[4] entries), appended by thread_fast_kernel: x, y, z as thr,
	                            // w = mates to thread (THR2_MATE0 | THR2_MATE1) | THR2_ALNED: the other mate threaded cleanly, the pair already counts
	uint8_t*  records;
	unsigned long long* counters;   // DBT_N_COUNTERS
	uint64_t npairs, first_pair_index;
	const uint8_t*  qual;       // phred+33 quality strings (FASTQ + `-b`), or nullptr: read r (2 * pair + mate) starts at qual[qoff[r]]
	const uint64_t* qoff;
};

// device buffers of the raw FASTA/FASTQ splitter (fastx.cu), one set per slot
struct RawView {
	const uint8_t* raw;         // the text (padded to a multiple of 32 bytes + 64)
	uint32_t* nlmask;           // bit per byte: '\n'
	uint32_t* blocksum;         // compaction scratch
	uint32_t* nl;               // newline offsets, ascending (max_lines)
	uint32_t* hdr;              // dbt_raw_info in the making: nlines, nrecords, npairs, nsingle, nshort_pairs, flags, max_read_len, max_record_bytes
	uint64_t* hash; uint32_t* tlen; uint32_t* slen; uint8_t* eq;       // per record (max_recs, eq: +2)
	uint32_t* pairmask; uint32_t* singlemask; uint32_t* pairlist; uint32_t* singlelist;
	uint64_t* rstart; uint64_t* rend;                                   // per read of the pairs formed: [start, end) of the sequence in raw
	uint64_t* rqual;                                                    // ... and, FASTQ, the start of its quality line
	uint32_t check_qual;                                                // the quality strings will be read per base (`-b`): a record whose quality line is not
	                                                                    // as long as its sequence sends the chunk to the host pairer, which refuses it by name
	uint32_t* pairoff; uint64_t* pairhash; uint32_t* singleoff;
	uint32_t max_lines, max_recs, max_pairs, max_single;
};

constexpr uint32_t THR2_MATE0 = 1u, THR2_MATE1 = 2u, THR2_ALNED = 4u;
constexpr int DBT_ST_THREAD_PENDING = 9;       // internal: handed to thread_kernel, which writes the final status
constexpr uint32_t ERR_RECORD_OVERFLOW = 1u;
constexpr uint32_t ERR_READ_TOO_LONG   = 2u;
constexpr uint32_t ERR_INDEX_MISSING   = 4u;   // tr/fl k-mer not in kmers.dbi for its locus
constexpr uint32_t ERR_INDEX_DUPKEY    = 8u;
constexpr uint32_t ERR_INDEX_DUPTR     = 16u;
constexpr uint32_t ERR_INDEX_FULL      = 32u;  // stash exhausted (table_load too high)

}  // namespace dbt
