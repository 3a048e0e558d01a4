// C ABI of libdanbing_b200.so (include/danbing_b200.h): context, device-resident index, batch
// pipeline (upload -> subfilter -> compaction -> candidate kernel -> fetch), results, NCCL reduce.
// No CPU fallback: every entry point that computes needs a CUDA device.
#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>
#include <dlfcn.h>
#include <sched.h>

#include "launchers.h"
#include "host_pack.h"

#ifndef DBT_DEFAULT_L2_FETCH
#define DBT_DEFAULT_L2_FETCH 0
#endif
#include "graph_table.cuh"

using namespace dbt;

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
	va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
	return code;
}
namespace dbt {   // for the other translation units of the library (noalloc.cu): sets dbt_last_error(), returns code
int abi_fail(int code, const char* fmt, ...) {
	va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
	return code;
}
}
#define CU(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return fail(DBT_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); } while (0)

extern "C" const char* dbt_strerror(int code) {
	switch (code) {
	case DBT_OK: return "ok";
	case DBT_E_CUDA: return "CUDA error";
	case DBT_E_NOMEM: return "out of memory";
	case DBT_E_BADARG: return "bad argument";
	case DBT_E_CAPACITY: return "capacity exceeded";
	case DBT_E_STATE: return "invalid call order";
	case DBT_E_INDEX: return "inconsistent RPGG index";
	case DBT_E_IO: return "I/O error";
	case DBT_E_NCCL: return "NCCL error";
	default: return "unknown error";
	}
}
extern "C" const char* dbt_last_error(void) { return g_err; }
extern "C" int dbt_abi_version(void) { return DBT_ABI_VERSION; }
extern "C" int dbt_device_count(void) {
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) { cudaGetLastError(); return 0; }
	if (e != cudaSuccess) { fail(DBT_E_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e)); cudaGetLastError(); return -1; }
	return n;
}

extern "C" void dbt_default_params(dbt_params* p) {
	memset(p, 0, sizeof(*p));
	p->device = 0; p->k = 21; p->n_filter = 4; p->nm_filter = 1; p->cth = 10; p->nm_tr = 40; p->max_nt = 2;   // cpp:26-33, 2338
	p->thread_cth = 100; p->max_ncorrection = 4; p->flags = DBT_F_KAM; p->qth = 20;
	p->max_read_len = 150; p->max_batch_pairs = 1u << 20; p->max_records = 0; p->table_load = 0.f; p->bucket_slots = 0;
}

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
enum { KT_PACK = 0, KT_SUBFILTER, KT_COMPACT, KT_CANDIDATE, KT_GENERIC, KT_THREAD, KT_THREAD2, KT_N };   // KT_CANDIDATE = closed-form kernel, KT_GENERIC = generic kernel,
                                                                                                   // KT_THREAD = thread_fast_kernel, KT_THREAD2 = thread_kernel (mates with errors)

struct Slot {
	bool ready = false;
	cudaStream_t st = nullptr;
	uint8_t* d_ascii = nullptr; uint64_t* d_offsets = nullptr;
	uint64_t* d_packed[2] = {nullptr, nullptr}; uint16_t* d_meta[2] = {nullptr, nullptr};
	uint32_t* d_passmask = nullptr; uint32_t* d_blocksum = nullptr; uint32_t* d_cand = nullptr; uint32_t* d_ncand = nullptr; uint32_t* d_slow = nullptr; uint4* d_thr = nullptr; uint4* d_thr2 = nullptr;
	uint8_t* d_records = nullptr; unsigned long long* d_counters = nullptr;
	uint8_t* h_records = nullptr;            // pinned staging (DBT_F_TRACE, or a pageable caller buffer)
	// record emission: emit bitmap -> ordered list -> dense copy of the records that become kam / aln records
	uint8_t* d_records_out = nullptr; uint32_t* d_emitmask = nullptr; uint32_t* d_emitsum = nullptr; uint32_t* d_emitlist = nullptr; uint32_t* d_nemit = nullptr;
	bool emitted = false;
	uint64_t* h_counters = nullptr;          // pinned: DBT_N_COUNTERS u64 + 4 u32 (ncand, err, nslow, nthr) + u32 nemit
	uint64_t last_ncand = 0, last_nslow = 0, last_nthr = 0;
	uint64_t npairs = 0, first_pair_index = 0;
	bool uploaded = false, processed = false, fetch_pending = false;
	dbt_batch_result* result = nullptr;
	cudaEvent_t ev[KT_N + 1] = {};
	cudaEvent_t t0 = nullptr, t1 = nullptr, evp0 = nullptr, evp1 = nullptr;   // evp0/evp1 bracket pack_kernel (dbt_upload, ASCII layout)
	bool ev_pack = false, ev_proc = false;
	// raw FASTA/FASTQ ingestion (dbt_upload_raw): device text + splitter buffers, grown on demand
	uint8_t* d_raw = nullptr; uint64_t raw_cap = 0;
	uint32_t *d_nlmask = nullptr, *d_rblocksum = nullptr, *d_nl = nullptr, *d_rhdr = nullptr, *d_tlen = nullptr, *d_slen = nullptr;
	uint64_t* d_rhash = nullptr; uint8_t* d_eq = nullptr;
	uint32_t *d_pairmask = nullptr, *d_singlemask = nullptr, *d_pairlist = nullptr, *d_singlelist = nullptr, *d_pairoff = nullptr, *d_singleoff = nullptr;
	uint64_t *d_rstart = nullptr, *d_rend = nullptr, *d_pairhash = nullptr, *d_rqual = nullptr;
	uint8_t* d_qual = nullptr;               // quality strings of an ASCII batch (`-b` with FASTQ)
	const uint8_t* q_base = nullptr; const uint64_t* q_off = nullptr;      // where the current batch's qualities are (nullptr: none)
	uint32_t raw_max_lines = 0, raw_max_recs = 0, raw_max_single = 0;
	uint32_t* h_rhdr = nullptr;              // pinned: dbt_raw_info
	cudaEvent_t evr0 = nullptr, evr1 = nullptr;
	bool ev_raw = false;
	// `-bu`: novel-edge events of the batch
	uint64_t *d_evkey = nullptr, *d_evedge = nullptr; uint32_t *d_evlocus = nullptr, *d_nev = nullptr; uint32_t ev_cap = 0;
	uint32_t* h_nev = nullptr; bool ev_pending = false;
};

struct Id128 { char b[128]; };       // ncclUniqueId is passed BY VALUE to ncclCommInitRank
struct NcclApi {
	void* lib = nullptr;
	int (*GetUniqueId)(void*) = nullptr;
	int (*CommInitRank)(void**, int, Id128, int) = nullptr;
	int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
	int (*CommDestroy)(void*) = nullptr;
	int (*GroupStart)() = nullptr;
	int (*GroupEnd)() = nullptr;
	const char* (*GetErrorString)(int) = nullptr;
};

struct dbt_ctx {
	dbt_params p;
	RunParams rp;
	int B = 2, nsm = 0;
	size_t smem_optin = 0;
	unsigned cand_grid = 0, cand_grid_fast = 0;
	bool index_loaded = false;
	DevIndex ix{};
	// owned device memory of the index
	uint4* d_gslots = nullptr; uint64_t gtable_bytes = 0, gnodes = 0;   // graph table (restored threading)
	uint32_t* d_pre = nullptr;                                           // presence bitmap of the index keys (subfilter_kernel)
	uint4* d_tslots = nullptr; Table ttable{}; bool bubbles = false;    // TR edge table (`-bu`)
	uint32_t gmerge_err = 0;             // why the merge was refused: 1 node not an index key, 2 key of another locus, 4 not in the shared list, 8/16 TR k-mer without its count position
	uint4* d_bslots = nullptr;                                           // bait table (`-b`)
	uint32_t* d_vv_g = nullptr;                                          // graph entries of multi-locus k-mers (graphs merged into the index)
	uint4* d_slots = nullptr; uint32_t* d_vv = nullptr; uint32_t* d_vv_aux = nullptr;
	// results: ONE allocation [counts nk_tr | kmc nloci | totals DBT_N_COUNTERS] u64 + [nmapread nloci] u32, so that the
	// end-of-run exchange is one grouped collective (a u64 part and a u32 part); d_red = the same layout, filled by dbt_allreduce
	uint64_t* d_results = nullptr; uint64_t* d_red = nullptr; uint64_t n_res64 = 0;
	uint64_t* d_counts = nullptr; uint64_t* d_kmc = nullptr; uint32_t* d_nmapread = nullptr; uint8_t* d_qc = nullptr;
	uint16_t* d_permtab = nullptr; uint64_t* d_vote_pool = nullptr; uint32_t* d_vote_lock = nullptr;
	uint64_t* d_totals = nullptr;
	uint64_t* h_totals = nullptr;            // pinned: local totals in, reduced totals out
	bool reduced_valid = false;              // d_red holds the all-rank sums of the current local state
	uint64_t totals_red[DBT_N_COUNTERS] = {};
	cudaStream_t red_st = nullptr; cudaEvent_t red_ev = nullptr;
	uint64_t table_bytes = 0, nkeys = 0;
	Slot slot[DBT_N_SLOTS];
	uint64_t totals[DBT_N_COUNTERS] = {};
	std::atomic<uint64_t> launches{0};
	void* nccl_comm = nullptr; int nranks = 1, rank = 0;
};

static NcclApi g_nccl;

static int set_device(const dbt_ctx* c) { CU(cudaSetDevice(c->p.device)); return DBT_OK; }

// L2 fetch granularity for this device.  The path is random 32-byte bucket probes; with the default granularity an L2 miss
// pulls 128 bytes from HBM (ncu at 20 000 loci, profiles/r1/s41: 120 bytes of DRAM read per probe in all three kernels).
// DBT_L2_FETCH_GRANULARITY = 32 | 64 | 128 overrides; 0 leaves the driver default.
static void set_l2_fetch_granularity() {
	size_t g = DBT_DEFAULT_L2_FETCH;
	if (const char* e = getenv("DBT_L2_FETCH_GRANULARITY")) g = (size_t)atoi(e);
	if (g == 32 || g == 64 || g == 128) { if (cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, g) != cudaSuccess) cudaGetLastError(); }
}

static uint32_t round_up(uint32_t v, uint32_t m) { return (v + m - 1) / m * m; }

extern "C" int dbt_create(dbt_ctx** out, const dbt_params* p) {
	if (!out || !p) return fail(DBT_E_BADARG, "null argument");
	*out = nullptr;
	if (p->k < 1 || p->k > 27) return fail(DBT_E_BADARG, "k=%u outside 1..27 (sort keys pack k-mer<<10|mate|pos into 64 bits)", p->k);
	if (p->max_read_len < p->k || p->max_read_len - p->k + 1 > 512) return fail(DBT_E_BADARG, "max_read_len=%u must satisfy k <= len <= k+511", p->max_read_len);
	if (p->n_filter == 1 && p->nm_filter) return fail(DBT_E_BADARG, "-kf 1 x divides by zero in the reference (cpp:174)");
	if (!p->max_batch_pairs) return fail(DBT_E_BADARG, "max_batch_pairs == 0");
	if (p->bucket_slots && p->bucket_slots != 2 && p->bucket_slots != 4) return fail(DBT_E_BADARG, "bucket_slots must be 2 (32 B sector) or 4 (64 B)");
	int ndev = dbt_device_count();
	if (ndev <= 0) return fail(DBT_E_CUDA, "no CUDA device available (this library has no CPU fallback)");
	if (p->device < 0 || p->device >= ndev) return fail(DBT_E_BADARG, "device %d out of range (%d devices)", p->device, ndev);
	dbt_ctx* c = new (std::nothrow) dbt_ctx();
	if (!c) return fail(DBT_E_NOMEM, "host allocation");
	c->p = *p;
	if (c->p.max_nt == 0) c->p.max_nt = 2;
	if (c->p.max_nt != 2) { delete c; return fail(DBT_E_BADARG, "max_nt must be 2: MAX_NT is a constant of the reference (cpp:31)"); }
	if (c->p.max_records == 0) c->p.max_records = c->p.max_batch_pairs;
	if (c->p.flags & DBT_F_TRACE) c->p.max_records = c->p.max_batch_pairs;
	if (c->p.table_load <= 0.f || c->p.table_load > 0.9f) c->p.table_load = 0.4f;
	c->B = c->p.bucket_slots ? (int)c->p.bucket_slots : 2;
	cudaError_t e = cudaSetDevice(p->device);
	if (e == cudaSuccess) set_l2_fetch_granularity();
	cudaDeviceProp prop;
	if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, p->device);
	if (e != cudaSuccess) { delete c; return fail(DBT_E_CUDA, "cudaSetDevice/GetDeviceProperties: %s", cudaGetErrorString(e)); }
	c->nsm = prop.multiProcessorCount;
	c->smem_optin = prop.sharedMemPerBlockOptin;

	RunParams& rp = c->rp;
	rp.k = p->k; rp.n_filter = p->n_filter; rp.nm_filter = p->nm_filter; rp.cth = p->cth; rp.nm_tr = p->nm_tr; rp.max_nt = c->p.max_nt;
	rp.flags = p->flags; rp.qth = p->qth;
	rp.lay = make_layout(p->max_read_len);
	rp.max_read_len = p->max_read_len;
	rp.nk_max = p->max_read_len - p->k + 1;
	rp.nkp = round_up(rp.nk_max, 32);
	rp.nkp = rp.nkp <= 5 * 32 ? 5 * 32 : (rp.nkp <= 8 * 32 ? 8 * 32 : 16 * 32);     // = 32 * MAXJ of the candidate_kernel instantiation
	rp.umax = round_up(2 * rp.nk_max, 4);
	rp.sortn = (rp.nkp / 32 <= 8) ? 512 : 1024;       // = 32*E of candidate_kernel<B, MAXJ>: E = 16 for MAXJ <= 8, else 32
	rp.as_stride = round_up((rp.nk_max + 3) / 4, 8);
	rp.rec_stride = (uint32_t)sizeof(dbt_pair_record) + 2 * rp.as_stride;
	rp.max_records = c->p.max_records;
	rp.vote_spill_at = VOTE_SPILL_AT;
	if (const char* e = getenv("DBT_VOTE_SPILL_AT")) { const int v = atoi(e); if (v >= 0 && v < (int)VOTE_SPILL_AT) rp.vote_spill_at = (uint32_t)v; }   // test hook
	// restored threading: geometry of the alignment blobs; switched on by dbt_load_index when the index carries a graph
	rp.thr_on = 0; rp.thread_cth = p->thread_cth; rp.max_ncorrection = p->max_ncorrection; rp.correction = (p->flags & DBT_F_CORRECTION) ? 1u : 0u;
	rp.es_cap = DBT_ES_CAP_FOR(p->max_read_len, p->max_ncorrection);
	rp.tr_cap = DBT_TR_CAP_FOR(p->max_read_len, p->k, p->max_ncorrection);
	rp.aln_stride = (uint32_t)sizeof(dbt_aln_mate) + rp.es_cap + rp.tr_cap;
	if ((p->flags & DBT_F_THREADING) && rp.tr_cap > 16 * 32) { delete c; return fail(DBT_E_CAPACITY, "threading: read length + 2*max_ncorrection exceeds 512 k-mers"); }
	// must mirror the carve-up in candidate_kernel
	size_t ws = 0;
	ws += (size_t)2 * rp.lay.W64 * 8;                // pk
	ws += (size_t)2 * rp.nkp * 8;                    // pkey
	ws += (size_t)4 * rp.nkp * 4;                    // pval, paux
	ws += (size_t)3 * rp.umax * 4;                   // uval, uaux, unml
	ws += (size_t)3 * VOTE_SMEM_SLOTS * 4;           // vote table
	ws += (size_t)2 * rp.umax * 2;                   // utf, utr
	ws += (size_t)(rp.umax + 2) * 2;                 // ustart
	ws += (size_t)2 * rp.umax * 2;                   // umid, perm
	rp.warp_smem = round_up((uint32_t)ws, 16);
	size_t block_smem = (size_t)rp.warp_smem * CAND_WARPS_SLOW;
	if (block_smem > c->smem_optin) {
		const size_t lim = c->smem_optin;
		delete c;
		return fail(DBT_E_CAPACITY, "max_read_len=%u needs %zu B shared memory per block (> %zu)", p->max_read_len, block_smem, lim);
	}
	const size_t slow_smem = (size_t)rp.warp_smem * CAND_WARPS_SLOW;
	unsigned blocks_per_sm = (unsigned)std::max<size_t>(1, std::min<size_t>(4, (size_t)(prop.sharedMemPerMultiprocessor - 1024) / (slow_smem + 1024)));
	c->cand_grid = (unsigned)c->nsm * blocks_per_sm;
	c->cand_grid_fast = (unsigned)c->nsm;                // candidate_fast_kernel: the launcher multiplies by the resident blocks per SM
	*out = c;
	return DBT_OK;
}

static void free_slot(Slot& s) {
	if (s.st) cudaStreamSynchronize(s.st);
	cudaFree(s.d_ascii); cudaFree(s.d_offsets);
	for (int m = 0; m < 2; ++m) { cudaFree(s.d_packed[m]); cudaFree(s.d_meta[m]); }
	cudaFree(s.d_passmask); cudaFree(s.d_blocksum); cudaFree(s.d_cand); cudaFree(s.d_ncand); cudaFree(s.d_slow); cudaFree(s.d_thr); cudaFree(s.d_thr2); cudaFree(s.d_records); cudaFree(s.d_counters);
	cudaFree(s.d_records_out); cudaFree(s.d_emitmask); cudaFree(s.d_emitsum); cudaFree(s.d_emitlist); cudaFree(s.d_nemit);
	cudaFree(s.d_raw); cudaFree(s.d_nlmask); cudaFree(s.d_rblocksum); cudaFree(s.d_nl); cudaFree(s.d_rhdr); cudaFree(s.d_tlen); cudaFree(s.d_slen);
	cudaFree(s.d_rhash); cudaFree(s.d_eq); cudaFree(s.d_pairmask); cudaFree(s.d_singlemask); cudaFree(s.d_pairlist); cudaFree(s.d_singlelist);
	cudaFree(s.d_pairoff); cudaFree(s.d_singleoff); cudaFree(s.d_rstart); cudaFree(s.d_rend); cudaFree(s.d_pairhash); cudaFree(s.d_rqual); cudaFree(s.d_qual);
	cudaFree(s.d_evkey); cudaFree(s.d_evedge); cudaFree(s.d_evlocus); cudaFree(s.d_nev);
	if (s.h_nev) cudaFreeHost(s.h_nev);
	if (s.h_rhdr) cudaFreeHost(s.h_rhdr);
	if (s.evr0) cudaEventDestroy(s.evr0);
	if (s.evr1) cudaEventDestroy(s.evr1);
	if (s.h_records) cudaFreeHost(s.h_records);
	if (s.h_counters) cudaFreeHost(s.h_counters);
	for (auto& e : s.ev) if (e) cudaEventDestroy(e);
	if (s.t0) cudaEventDestroy(s.t0);
	if (s.t1) cudaEventDestroy(s.t1);
	if (s.evp0) cudaEventDestroy(s.evp0);
	if (s.evp1) cudaEventDestroy(s.evp1);
	if (s.st) cudaStreamDestroy(s.st);
	s = Slot();
}

extern "C" void dbt_destroy(dbt_ctx* c) {
	if (!c) return;
	cudaSetDevice(c->p.device);
	for (auto& s : c->slot) free_slot(s);
	if (c->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl_comm);
	cudaFree(c->d_gslots); cudaFree(c->d_pre); cudaFree(c->d_tslots); cudaFree(c->d_vv_g); cudaFree(c->d_bslots);
	cudaFree(c->d_slots); cudaFree(c->d_vv); cudaFree(c->d_vv_aux); cudaFree(c->d_results); cudaFree(c->d_red);
	cudaFree(c->d_qc); cudaFree(c->d_permtab); cudaFree(c->d_vote_pool); cudaFree(c->d_vote_lock);
	if (c->h_totals) cudaFreeHost(c->h_totals);
	if (c->red_ev) cudaEventDestroy(c->red_ev);
	if (c->red_st) cudaStreamDestroy(c->red_st);
	delete c;
}

template <typename T>
static int dev_alloc(T** p, size_t n, const char* what) {
	*p = nullptr;
	cudaError_t e = cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(T));
	if (e != cudaSuccess) { cudaGetLastError(); return fail(DBT_E_NOMEM, "cudaMalloc(%s, %zu B): %s", what, n * sizeof(T), cudaGetErrorString(e)); }
	return DBT_OK;
}
#define TRY(x) do { int r__ = (x); if (r__ != DBT_OK) return r__; } while (0)

static int ensure_slot(dbt_ctx* c, int si) {
	Slot& s = c->slot[si];
	if (s.ready) return DBT_OK;
	const RunParams& rp = c->rp;
	const uint64_t P = c->p.max_batch_pairs;
	CU(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
	for (int m = 0; m < 2; ++m) { TRY(dev_alloc(&s.d_packed[m], P * rp.lay.W64, "packed")); TRY(dev_alloc(&s.d_meta[m], P, "meta")); }
	const uint64_t nwords = (P + 31) / 32;
	TRY(dev_alloc(&s.d_passmask, nwords, "passmask"));
	TRY(dev_alloc(&s.d_blocksum, (nwords + COMPACT_WORDS_H - 1) / COMPACT_WORDS_H + 1, "blocksum"));
	TRY(dev_alloc(&s.d_cand, P, "cand"));
	TRY(dev_alloc(&s.d_ncand, 8, "ncand"));
	TRY(dev_alloc(&s.d_slow, P, "slow list"));
	if (rp.thr_on) { TRY(dev_alloc(&s.d_thr, P, "thread list")); TRY(dev_alloc(&s.d_thr2, P, "thread list 2")); }
	TRY(dev_alloc(&s.d_counters, DBT_N_COUNTERS, "counters"));
	const bool want_records = rp.thr_on ? (c->p.flags & (DBT_F_ALN | DBT_F_ALN_MIN | DBT_F_TRACE)) != 0 : (c->p.flags & (DBT_F_KAM | DBT_F_EXTRACT | DBT_F_BUBBLES | DBT_F_TRACE)) != 0;
	if (want_records) {
		TRY(dev_alloc(&s.d_records, (size_t)c->p.max_records * rp.rec_stride, "records"));
		if (!(c->p.flags & DBT_F_TRACE)) {
			const uint64_t ew = ((uint64_t)c->p.max_records + 31) / 32;
			TRY(dev_alloc(&s.d_records_out, (size_t)c->p.max_records * rp.rec_stride, "records (dense)"));
			TRY(dev_alloc(&s.d_emitmask, ew + 8, "emit mask"));
			TRY(dev_alloc(&s.d_emitsum, (ew + COMPACT_WORDS_H - 1) / COMPACT_WORDS_H + 1, "emit blocksum"));
			TRY(dev_alloc(&s.d_emitlist, (size_t)c->p.max_records + 32, "emit list"));
			TRY(dev_alloc(&s.d_nemit, 4, "nemit"));
		}
	}
	if (c->bubbles) {
		s.ev_cap = (uint32_t)std::min<uint64_t>(4 * P + (1u << 20), 0xFFFFFFF0ull);
		TRY(dev_alloc(&s.d_evkey, s.ev_cap, "bubble event keys")); TRY(dev_alloc(&s.d_evedge, s.ev_cap, "bubble event edges"));
		TRY(dev_alloc(&s.d_evlocus, s.ev_cap, "bubble event loci")); TRY(dev_alloc(&s.d_nev, 1, "bubble event count"));
		CU(cudaHostAlloc((void**)&s.h_nev, 16, cudaHostAllocDefault));
	}
	CU(cudaHostAlloc((void**)&s.h_counters, (DBT_N_COUNTERS + 4) * sizeof(uint64_t), cudaHostAllocDefault));
	for (auto& e : s.ev) CU(cudaEventCreate(&e));
	CU(cudaEventCreate(&s.t0)); CU(cudaEventCreate(&s.t1)); CU(cudaEventCreate(&s.evp0)); CU(cudaEventCreate(&s.evp1));
	s.ready = true;
	return DBT_OK;
}

// ------------------------------------------------------------------------------------------
// index
// ------------------------------------------------------------------------------------------
// graphDB of all loci -> one device table (graph_table.cuh); replaces readGraphKmers (h:549-575) + cpp:2468
static int load_graph(dbt_ctx* c, const dbt_index* in, uint64_t ntr) {
	const uint64_t nloci = in->nloci;
	if (nloci >= G_LOCUS_MASK) return fail(DBT_E_CAPACITY, "threading supports < %u loci", G_LOCUS_MASK);
	uint64_t ng = 0;
	for (uint64_t l = 0; l < nloci; ++l) ng += in->g_n[l];
	if (ntr >= 0xFFFFFFFFull) return fail(DBT_E_CAPACITY, "too many TR k-mers");
	const int B = c->B;
	// entries <= canonical nodes (ng/2 when both strands are stored) + TR k-mers missing from the graph (normally none)
	const uint64_t nent = ng / 2 + ng % 2 + 1024;
	uint64_t nbuckets = (uint64_t)((double)nent / c->p.table_load) / B + 64;
	uint64_t nstash = std::max<uint64_t>(64, (nent / 2 + ntr / 8 + B - 1) / B);
	if (nbuckets >= (1ull << 32) || nstash >= (1ull << 32)) return fail(DBT_E_CAPACITY, "graph table needs %llu buckets (limit 2^32)", (unsigned long long)nbuckets);
	c->gtable_bytes = (nbuckets + nstash) * B * sizeof(uint4); c->gnodes = ng;
	cudaStream_t st; CU(cudaStreamCreate(&st));
	TRY(dev_alloc(&c->d_gslots, (nbuckets + nstash) * B, "graph table"));
	CU(cudaMemsetAsync(c->d_gslots, 0xFF, c->gtable_bytes, st));
	Table t{c->d_gslots, nbuckets, nstash};
	uint32_t* d_err; TRY(dev_alloc(&d_err, 1, "err")); CU(cudaMemsetAsync(d_err, 0, 4, st));
	{
		std::vector<uint32_t> loc(std::max<uint64_t>(ng, ntr));
		uint64_t j = 0;
		for (uint64_t l = 0; l < nloci; ++l) for (uint64_t i = 0; i < in->g_n[l]; ++i) loc[j++] = (uint32_t)l;
		uint64_t* dk; uint8_t* dm; uint32_t* dl;
		TRY(dev_alloc(&dk, std::max<uint64_t>(ng, ntr), "graph kmers")); TRY(dev_alloc(&dm, ng, "graph masks")); TRY(dev_alloc(&dl, std::max<uint64_t>(ng, ntr), "graph loci"));
		if (ng) {
			CU(cudaMemcpyAsync(dk, in->g_kmers, ng * 8, cudaMemcpyHostToDevice, st));
			CU(cudaMemcpyAsync(dm, in->g_masks, ng, cudaMemcpyHostToDevice, st));
			CU(cudaMemcpyAsync(dl, loc.data(), ng * 4, cudaMemcpyHostToDevice, st));
			for (int pass = 0; pass < 2; ++pass) { CU(launch_graph_build(st, B, t, dk, dm, dl, ng, c->p.k, pass, d_err)); ++c->launches; }
			CU(launch_graph_build(st, B, t, dk, dm, dl, ng, c->p.k, 3, d_err)); ++c->launches;
			CU(cudaStreamSynchronize(st));
		}
		if (ntr) {                                     // TR k-mers attach their count-array position (trResults[locus], cpp:2465)
			j = 0;
			for (uint64_t l = 0; l < nloci; ++l) for (uint64_t i = 0; i < in->tr_n[l]; ++i) loc[j++] = (uint32_t)l;
			CU(cudaMemcpyAsync(dk, in->tr_kmers, ntr * 8, cudaMemcpyHostToDevice, st));
			CU(cudaMemcpyAsync(dl, loc.data(), ntr * 4, cudaMemcpyHostToDevice, st));
			CU(launch_graph_build(st, B, t, dk, nullptr, dl, ntr, c->p.k, 2, d_err)); ++c->launches;
			CU(cudaStreamSynchronize(st));
		}
		cudaFree(dk); cudaFree(dm); cudaFree(dl);
	}
	uint32_t herr = 0;
	CU(cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost));
	cudaFree(d_err);
	CU(cudaStreamDestroy(st));
	if (herr & ERR_INDEX_FULL) return fail(DBT_E_CAPACITY, "graph table overflow: lower table_load");
	if (herr & ERR_INDEX_MISSING) return fail(DBT_E_INDEX, "graph.kmers holds a node that does not fit k=%u", c->p.k);
	if (herr & ERR_INDEX_DUPKEY) return fail(DBT_E_INDEX, "duplicate node within a locus of the graph (merge masks before dbt_load_index, h:569)");
	if (herr & ERR_INDEX_DUPTR) return fail(DBT_E_INDEX, "duplicate TR k-mer within a locus");
	c->ix.gtable = t;
	c->rp.thr_on = 1;
	c->rp.rec_stride = (uint32_t)sizeof(dbt_pair_record) + 2 * c->rp.aln_stride;
	return DBT_OK;
}

// baitDB of all loci (PREF[.qc].bt.kmdb, cpp:2495-2499) -> one device table keyed (canonical k-mer, locus), payload = min << 8 | max
static int load_baits(dbt_ctx* c, const dbt_index* in) {
	const uint64_t nloci = in->nloci;
	if (nloci >= G_LOCUS_MASK) return fail(DBT_E_CAPACITY, "-b supports < %u loci", G_LOCUS_MASK);
	uint64_t nb = 0;
	for (uint64_t l = 0; l < nloci; ++l) nb += in->bt_n[l];
	const int B = c->B;
	const uint64_t nbuckets = (uint64_t)((double)(nb + 1024) / c->p.table_load) / B + 64, nstash = std::max<uint64_t>(64, ((nb + 1024) / 2 + B - 1) / B);
	if (nbuckets >= (1ull << 32) || nstash >= (1ull << 32)) return fail(DBT_E_CAPACITY, "bait table needs %llu buckets (limit 2^32)", (unsigned long long)nbuckets);
	cudaStream_t st; CU(cudaStreamCreate(&st));
	TRY(dev_alloc(&c->d_bslots, (nbuckets + nstash) * B, "bait table"));
	CU(cudaMemsetAsync(c->d_bslots, 0xFF, (nbuckets + nstash) * B * sizeof(uint4), st));
	Table t{c->d_bslots, nbuckets, nstash};
	uint32_t herr = 0;
	if (nb) {
		uint32_t* d_err; TRY(dev_alloc(&d_err, 1, "err")); CU(cudaMemsetAsync(d_err, 0, 4, st));
		std::vector<uint32_t> loc(nb);
		uint64_t j = 0;
		for (uint64_t l = 0; l < nloci; ++l) for (uint64_t i = 0; i < in->bt_n[l]; ++i) loc[j++] = (uint32_t)l;
		uint64_t* dk; uint16_t* dv; uint32_t* dl;
		TRY(dev_alloc(&dk, nb, "bait kmers")); TRY(dev_alloc(&dv, nb, "bait bounds")); TRY(dev_alloc(&dl, nb, "bait loci"));
		CU(cudaMemcpyAsync(dk, in->bt_kmers, nb * 8, cudaMemcpyHostToDevice, st));
		CU(cudaMemcpyAsync(dv, in->bt_vals, nb * 2, cudaMemcpyHostToDevice, st));
		CU(cudaMemcpyAsync(dl, loc.data(), nb * 4, cudaMemcpyHostToDevice, st));
		CU(launch_bait_insert(st, B, t, dk, dv, dl, nb, d_err)); ++c->launches;
		CU(cudaMemcpyAsync(&herr, d_err, 4, cudaMemcpyDeviceToHost, st));
		CU(cudaStreamSynchronize(st));
		cudaFree(dk); cudaFree(dv); cudaFree(dl); cudaFree(d_err);
	}
	CU(cudaStreamDestroy(st));
	if (herr & ERR_INDEX_FULL) return fail(DBT_E_CAPACITY, "bait table overflow: lower table_load");
	c->ix.btable = t;
	return DBT_OK;
}

// trEdgeDB of all loci (PREF.tre.kdb, cpp:2492) -> one device table keyed (canonical (k+1)-mer, locus)
static int load_edges(dbt_ctx* c, const dbt_index* in) {
	const uint64_t nloci = in->nloci;
	if (nloci >= G_LOCUS_MASK) return fail(DBT_E_CAPACITY, "-bu supports < %u loci", G_LOCUS_MASK);
	if (c->p.k + 1 > 31) return fail(DBT_E_CAPACITY, "-bu: (k+1)-mers must fit 62 bits");
	uint64_t ne = 0;
	for (uint64_t l = 0; l < nloci; ++l) ne += in->tre_n[l];
	const int B = c->B;
	const uint64_t nbuckets = (uint64_t)((double)(ne + 1024) / c->p.table_load) / B + 64, nstash = std::max<uint64_t>(64, ((ne + 1024) / 2 + B - 1) / B);
	if (nbuckets >= (1ull << 32) || nstash >= (1ull << 32)) return fail(DBT_E_CAPACITY, "edge table needs %llu buckets (limit 2^32)", (unsigned long long)nbuckets);
	cudaStream_t st; CU(cudaStreamCreate(&st));
	TRY(dev_alloc(&c->d_tslots, (nbuckets + nstash) * B, "edge table"));
	CU(cudaMemsetAsync(c->d_tslots, 0xFF, (nbuckets + nstash) * B * sizeof(uint4), st));
	Table t{c->d_tslots, nbuckets, nstash};
	uint32_t* d_err; TRY(dev_alloc(&d_err, 1, "err")); CU(cudaMemsetAsync(d_err, 0, 4, st));
	uint32_t herr = 0;
	if (ne) {
		std::vector<uint32_t> loc(ne);
		uint64_t j = 0;
		for (uint64_t l = 0; l < nloci; ++l) for (uint64_t i = 0; i < in->tre_n[l]; ++i) loc[j++] = (uint32_t)l;
		uint64_t* dk; uint32_t* dl;
		TRY(dev_alloc(&dk, ne, "edges")); TRY(dev_alloc(&dl, ne, "edge loci"));
		CU(cudaMemcpyAsync(dk, in->tre_kmers, ne * 8, cudaMemcpyHostToDevice, st));
		CU(cudaMemcpyAsync(dl, loc.data(), ne * 4, cudaMemcpyHostToDevice, st));
		CU(launch_edge_insert(st, B, t, dk, dl, ne, c->p.k + 1, d_err)); ++c->launches;
		CU(cudaMemcpyAsync(&herr, d_err, 4, cudaMemcpyDeviceToHost, st));
		CU(cudaStreamSynchronize(st));
		cudaFree(dk); cudaFree(dl);
	}
	cudaFree(d_err);
	CU(cudaStreamDestroy(st));
	if (herr & ERR_INDEX_FULL) return fail(DBT_E_CAPACITY, "edge table overflow: lower table_load");
	if (herr & ERR_INDEX_MISSING) return fail(DBT_E_INDEX, "tre.kdb holds an edge that does not fit k+1=%u bases", c->p.k + 1);
	c->ttable = t; c->bubbles = true;
	return DBT_OK;
}

// Graph entries into the index entries (kernels.cu, graph_merge_kernel): when every graph node is an index key of its locus and
// every TR k-mer keeps its count position there -- true for RPGGs whose graph, TR and flank k-mers come from the same haplotype
// sequences, as fa2kmers writes them -- candidate_fast_kernel reads the graph entry of every k-mer of a pair out of the index
// buckets it has probed anyway and threads the pair itself; thread_fast_kernel then only sees what the generic kernel hands on.
// Otherwise (or with DBT_NO_GRAPH_MERGE set: the test hook for the two-table path) nothing changes.
static int merge_graph(dbt_ctx* c, const dbt_index* in, uint64_t ntr) {
	const uint64_t nloci = in->nloci;
	if (getenv("DBT_NO_GRAPH_MERGE") || nloci >= (1ull << (VAL_LOCUS_BITS - 1)) || c->rp.nk_max > 255) return DBT_OK;
	uint64_t ng = 0;
	for (uint64_t l = 0; l < nloci; ++l) ng += in->g_n[l];
	cudaStream_t st; CU(cudaStreamCreate(&st));
	TRY(dev_alloc(&c->d_vv_g, in->nvv + 1, "vv_g"));
	CU(cudaMemsetAsync(c->d_vv_g, 0, (in->nvv + 1) * 4, st));
	uint32_t* d_err; TRY(dev_alloc(&d_err, 1, "err")); CU(cudaMemsetAsync(d_err, 0, 4, st));
	std::vector<uint32_t> loc(std::max<uint64_t>(ng, ntr));
	uint64_t* dk; uint8_t* dm; uint32_t* dl;
	TRY(dev_alloc(&dk, std::max<uint64_t>(ng, ntr), "graph kmers")); TRY(dev_alloc(&dm, ng, "graph masks")); TRY(dev_alloc(&dl, std::max<uint64_t>(ng, ntr), "graph loci"));
	uint64_t j = 0;
	for (uint64_t l = 0; l < nloci; ++l) for (uint64_t i = 0; i < in->g_n[l]; ++i) loc[j++] = (uint32_t)l;
	if (ng) {
		CU(cudaMemcpyAsync(dk, in->g_kmers, ng * 8, cudaMemcpyHostToDevice, st));
		CU(cudaMemcpyAsync(dm, in->g_masks, ng, cudaMemcpyHostToDevice, st));
		CU(cudaMemcpyAsync(dl, loc.data(), ng * 4, cudaMemcpyHostToDevice, st));
		CU(launch_graph_merge(st, c->B, c->ix.table, c->d_vv, c->d_vv_g, dk, dm, dl, ng, c->p.k, d_err)); ++c->launches;
		CU(cudaStreamSynchronize(st));
	}
	if (ntr) {
		j = 0;
		for (uint64_t l = 0; l < nloci; ++l) for (uint64_t i = 0; i < in->tr_n[l]; ++i) loc[j++] = (uint32_t)l;
		CU(cudaMemcpyAsync(dk, in->tr_kmers, ntr * 8, cudaMemcpyHostToDevice, st));
		CU(cudaMemcpyAsync(dl, loc.data(), ntr * 4, cudaMemcpyHostToDevice, st));
		CU(launch_tr_class_check(st, c->B, c->ix.table, c->d_vv, c->d_vv_aux, dk, dl, ntr, d_err)); ++c->launches;
		CU(cudaStreamSynchronize(st));
	}
	uint32_t herr = 0;
	CU(cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost));
	cudaFree(dk); cudaFree(dm); cudaFree(dl); cudaFree(d_err);
	CU(cudaStreamDestroy(st));
	c->ix.vv_g = c->d_vv_g;
	c->ix.gmerged = herr ? 0u : 1u;       // (bits already OR-ed into value words are masked off by every reader: slot_val)
	c->gmerge_err = herr;
	return DBT_OK;
}

extern "C" int dbt_load_index(dbt_ctx* c, const dbt_index* in) {
	if (!c || !in) return fail(DBT_E_BADARG, "null argument");
	if (c->index_loaded) return fail(DBT_E_STATE, "index already loaded");
	if (!in->nloci || !in->tr_n || (in->nk_dbi && (!in->dbi_keys || !in->dbi_vals))) return fail(DBT_E_BADARG, "incomplete dbt_index");
	if ((c->p.flags & DBT_F_QC) && !in->qc) return fail(DBT_E_BADARG, "DBT_F_QC set but qc == NULL");
	TRY(set_device(c));
	const uint64_t nloci = in->nloci;
	uint64_t ntr = 0, nfl = 0;
	for (uint64_t l = 0; l < nloci; ++l) { ntr += in->tr_n[l]; nfl += in->fl_n ? in->fl_n[l] : 0; }
	if (ntr >= (1ull << AUX_CLASS_SHIFT)) return fail(DBT_E_CAPACITY, "%llu TR k-mers exceed the 30-bit count index", (unsigned long long)ntr);
	if (nloci >= 0x7FFFFFFFull) return fail(DBT_E_CAPACITY, "too many loci");
	const uint64_t kmax = 1ull << (2 * c->p.k);
	for (uint64_t i = 0; i < in->nk_dbi; ++i) if (in->dbi_keys[i] >= kmax) return fail(DBT_E_INDEX, "kmers.dbi key %llu does not fit k=%u", (unsigned long long)in->dbi_keys[i], c->p.k);

	cudaStream_t st;
	CU(cudaStreamCreate(&st));
	// table
	const int B = c->B;
	uint64_t nslots_min = (uint64_t)((double)std::max<uint64_t>(in->nk_dbi, 1) / c->p.table_load) + B;
	uint64_t nbuckets = (nslots_min + B - 1) / B;
	if (nbuckets < 64) nbuckets = 64;
	// stash: slots for half the keys, i.e. load ~0.15 at the default table_load (7 % of the keys overflow at B = 2)
	uint64_t nstash = std::max<uint64_t>(64, (in->nk_dbi / 2 + B - 1) / B);
	if (nbuckets >= (1ull << 32) || nstash >= (1ull << 32)) return fail(DBT_E_CAPACITY, "index table needs %llu buckets (limit 2^32)", (unsigned long long)nbuckets);
	c->table_bytes = (nbuckets + nstash) * B * sizeof(uint4);
	c->nkeys = in->nk_dbi;
	TRY(dev_alloc(&c->d_slots, (nbuckets + nstash) * B, "index table"));
	CU(cudaMemsetAsync(c->d_slots, 0xFF, c->table_bytes, st));
	TRY(dev_alloc(&c->d_vv, in->nvv + 1, "vv"));
	TRY(dev_alloc(&c->d_vv_aux, in->nvv + 1, "vv_aux"));
	CU(cudaMemsetAsync(c->d_vv_aux, 0, (in->nvv + 1) * 4, st));
	if (in->nvv) CU(cudaMemcpyAsync(c->d_vv, in->vv, in->nvv * 4, cudaMemcpyHostToDevice, st));
	uint32_t* d_err; TRY(dev_alloc(&d_err, 1, "err"));
	CU(cudaMemsetAsync(d_err, 0, 4, st));
	Table t{c->d_slots, nbuckets, nstash};
	{	// bulk insert of kmers.dbi (readBinaryIndex, h:671)
		uint64_t* dk; uint32_t* dv;
		TRY(dev_alloc(&dk, in->nk_dbi, "dbi keys")); TRY(dev_alloc(&dv, in->nk_dbi, "dbi vals"));
		if (in->nk_dbi) {
			CU(cudaMemcpyAsync(dk, in->dbi_keys, in->nk_dbi * 8, cudaMemcpyHostToDevice, st));
			CU(cudaMemcpyAsync(dv, in->dbi_vals, in->nk_dbi * 4, cudaMemcpyHostToDevice, st));
			CU(launch_insert(st, B, t, dk, dv, in->nk_dbi, d_err)); ++c->launches;
			// presence bitmap for subfilter_kernel: >= 30 bits per key, at most 32 MB so that it stays in L2 beside the streamed reads
			// (a genome-scale index gets none: its bitmap would live in HBM like the table and save nothing).  DBT_PREFILTER=0: off
			uint32_t pb = 16;
			while ((1ull << pb) < 30 * in->nk_dbi) ++pb;
			const char* pe = getenv("DBT_PREFILTER");
			if ((1ull << pb) / 8 <= (32ull << 20) && !(pe && atoi(pe) == 0)) {
				TRY(dev_alloc(&c->d_pre, (1ull << pb) / 32, "presence bitmap"));
				CU(cudaMemsetAsync(c->d_pre, 0, (1ull << pb) / 8, st));
				CU(launch_prefilter_build(st, dk, in->nk_dbi, c->d_pre, pb)); ++c->launches;
				c->ix.pre = c->d_pre; c->ix.pre_bits = pb;
			}
		}
		CU(cudaStreamSynchronize(st));
		cudaFree(dk); cudaFree(dv);
	}
	{	// per-locus classes: TR first, flank second (flank wins, cpp:1467-1468)
		for (int pass = 0; pass < 2; ++pass) {
			const uint64_t n = pass == 0 ? ntr : nfl;
			const uint64_t* hn = pass == 0 ? in->tr_n : in->fl_n;
			const uint64_t* hk = pass == 0 ? in->tr_kmers : in->fl_kmers;
			if (!n) continue;
			if (!hk || !hn) return fail(DBT_E_BADARG, "missing k-mer array");
			std::vector<uint32_t> loc(n);
			uint64_t j = 0;
			for (uint64_t l = 0; l < nloci; ++l) for (uint64_t i = 0; i < hn[l]; ++i) loc[j++] = (uint32_t)l;
			uint64_t* dk; uint32_t* dl;
			TRY(dev_alloc(&dk, n, "locus kmers")); TRY(dev_alloc(&dl, n, "locus ids"));
			CU(cudaMemcpyAsync(dk, hk, n * 8, cudaMemcpyHostToDevice, st));
			CU(cudaMemcpyAsync(dl, loc.data(), n * 4, cudaMemcpyHostToDevice, st));
			CU(launch_fill_aux(st, B, t, c->d_vv, c->d_vv_aux, dk, dl, n, pass == 0 ? 2u : 1u, d_err)); ++c->launches;
			CU(cudaStreamSynchronize(st));
			cudaFree(dk); cudaFree(dl);
		}
	}
	uint32_t herr = 0;
	CU(cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost));
	cudaFree(d_err);
	if (herr & ERR_INDEX_DUPKEY) return fail(DBT_E_INDEX, "duplicate key in kmers.dbi");
	if (herr & ERR_INDEX_FULL) return fail(DBT_E_CAPACITY, "index table overflow: lower table_load");
	if (herr & ERR_INDEX_MISSING) return fail(DBT_E_INDEX, "a tr/fl k-mer is missing from kmers.dbi for its locus (index files out of sync)");
	if (herr & ERR_INDEX_DUPTR) return fail(DBT_E_INDEX, "duplicate TR k-mer within a locus");

	// results
	c->n_res64 = ntr + nloci + DBT_N_COUNTERS;
	TRY(dev_alloc(&c->d_results, c->n_res64 + (nloci + 1) / 2, "results"));
	c->d_counts = c->d_results; c->d_kmc = c->d_counts + ntr; c->d_totals = c->d_kmc + nloci;
	c->d_nmapread = reinterpret_cast<uint32_t*>(c->d_totals + DBT_N_COUNTERS);
	CU(cudaMemsetAsync(c->d_results, 0, (c->n_res64 + (nloci + 1) / 2) * 8, st));
	if (in->qc) { TRY(dev_alloc(&c->d_qc, nloci, "qc")); CU(cudaMemcpyAsync(c->d_qc, in->qc, nloci, cudaMemcpyHostToDevice, st)); }
	// all-ties visiting orders: what std::sort does with a comparator that never says "less"
	// (getSortedIndex on nmappedloci == 1 everywhere, cpp:247-250, 320-321).  Computed with the host's
	// own std::sort, i.e. the libstdc++ the reference would be built with on this machine.
	{
		const uint32_t maxU = c->rp.umax;
		std::vector<uint16_t> tab((size_t)maxU * (maxU + 1) / 2 + 1);
		std::vector<uint64_t> idx;
		for (uint32_t U = 1; U <= maxU; ++U) {
			idx.resize(U);
			for (uint32_t i = 0; i < U; ++i) idx[i] = i;
			std::sort(idx.begin(), idx.end(), [](uint64_t, uint64_t) { return false; });
			uint16_t* o = tab.data() + (size_t)U * (U - 1) / 2;
			for (uint32_t i = 0; i < U; ++i) o[i] = (uint16_t)idx[i];
		}
		TRY(dev_alloc(&c->d_permtab, tab.size(), "permtab"));
		CU(cudaMemcpyAsync(c->d_permtab, tab.data(), tab.size() * 2, cudaMemcpyHostToDevice, st));
		CU(cudaStreamSynchronize(st));
		c->ix.permtab_maxU = maxU;
	}
	// vote spill pool: VOTE_POOL epoch-tagged hits arrays + their locks and epochs (candidate.cuh, VoteHits)
	TRY(dev_alloc(&c->d_vote_pool, (uint64_t)VOTE_POOL * (nloci + 1), "vote pool"));
	TRY(dev_alloc(&c->d_vote_lock, 2 * VOTE_POOL, "vote locks"));
	CU(cudaMemsetAsync(c->d_vote_pool, 0, (uint64_t)VOTE_POOL * (nloci + 1) * 8, st));
	CU(cudaMemsetAsync(c->d_vote_lock, 0, 2 * VOTE_POOL * 4, st));
	CU(cudaStreamSynchronize(st));
	CU(cudaStreamDestroy(st));

	c->ix.table = t; c->ix.vv = c->d_vv; c->ix.vv_aux = c->d_vv_aux; c->ix.nvv = in->nvv;
	c->ix.counts = c->d_counts; c->ix.kmc = c->d_kmc; c->ix.nmapread = c->d_nmapread; c->ix.qc = c->d_qc;
	c->ix.permtab = c->d_permtab; c->ix.vote_pool = c->d_vote_pool; c->ix.vote_lock = c->d_vote_lock; c->ix.vote_epoch = c->d_vote_lock + VOTE_POOL;
	c->ix.nloci = nloci; c->ix.nk_tr = ntr;
	c->ix.gtable = Table{nullptr, 0, 0};
	c->ix.vv_g = nullptr; c->ix.gmerged = 0;
	if ((c->p.flags & DBT_F_THREADING) && in->g_n && in->g_kmers && in->g_masks) { TRY(load_graph(c, in, ntr)); TRY(merge_graph(c, in, ntr)); }
	c->ix.btable = Table{nullptr, 0, 0};
	if (c->p.flags & DBT_F_BAIT) {
		if (c->rp.thr_on) return fail(DBT_E_BADARG, "DBT_F_BAIT with graph threading: the reference filters baits only without threading (cpp:2110)");
		if (!in->bt_n || ((!in->bt_kmers || !in->bt_vals) && nloci)) return fail(DBT_E_BADARG, "DBT_F_BAIT needs bt_n / bt_kmers / bt_vals (PREF.bt.kmdb)");
		TRY(load_baits(c, in));
	}
	if (c->p.flags & DBT_F_BUBBLES) {
		if (c->rp.thr_on) return fail(DBT_E_BADARG, "DBT_F_BUBBLES with graph threading: the reference counts bubbles only without threading (cpp:2110, 2160)");
		if (!in->tre_n || (!in->tre_kmers && nloci)) return fail(DBT_E_BADARG, "DBT_F_BUBBLES needs tre_n / tre_kmers (PREF.tre.kdb)");
		TRY(load_edges(c, in));
	}
	{	// tables far beyond L2: the probe rounds are bound by the random-sector rate of HBM -> one coalesced request per bucket
		int l2 = 0;
		if (cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, c->p.device) != cudaSuccess) { cudaGetLastError(); l2 = 0; }
		c->rp.coop_stage = (l2 > 0 && c->table_bytes + c->gtable_bytes > 2 * (uint64_t)l2) ? 1u : 0u;
		if (const char* e = getenv("DBT_COOP_STAGE")) c->rp.coop_stage = atoi(e) ? 1u : 0u;
	}
	c->index_loaded = true;
	return DBT_OK;
}

extern "C" uint64_t dbt_nk_tr(const dbt_ctx* c) { return c ? c->ix.nk_tr : 0; }
extern "C" uint64_t dbt_nloci(const dbt_ctx* c) { return c ? c->ix.nloci : 0; }
extern "C" uint32_t dbt_record_stride(const dbt_ctx* c) { return c ? c->rp.rec_stride : 0; }
extern "C" uint32_t dbt_record_as_stride(const dbt_ctx* c) { return c ? c->rp.as_stride : 0; }
extern "C" uint32_t dbt_record_aln_stride(const dbt_ctx* c) { return c ? c->rp.aln_stride : 0; }
extern "C" uint32_t dbt_record_es_cap(const dbt_ctx* c) { return c ? c->rp.es_cap : 0; }
extern "C" uint64_t dbt_graph_table_bytes(const dbt_ctx* c) { return c ? c->gtable_bytes : 0; }
extern "C" int dbt_graph_merged(const dbt_ctx* c) { return !c ? 0 : (c->ix.gmerged ? 1 : -(int)c->gmerge_err); }
extern "C" uint64_t dbt_launch_count(const dbt_ctx* c) { return c ? c->launches.load() : 0; }

// ------------------------------------------------------------------------------------------
// packed host layout: [packed mate0 | packed mate1 | meta mate0 | meta mate1], sized for npairs
// ------------------------------------------------------------------------------------------
extern "C" size_t dbt_packed_bytes(uint32_t max_read_len, uint64_t npairs) {
	ReadLayout l = make_layout(max_read_len);
	return (size_t)npairs * (2 * (size_t)l.W64 * 8 + 2 * 2);
}

// CPUs this thread may run on (affinity mask: what a NUMA-bound or cgroup-pinned process really has)
static int usable_cpus() {
	cpu_set_t set;
	if (sched_getaffinity(0, sizeof(set), &set) == 0) { const int n = CPU_COUNT(&set); if (n > 0) return n; }
	return (int)std::max(1u, std::thread::hardware_concurrency());
}

extern "C" int dbt_pack_reads(uint32_t k, uint32_t max_read_len, uint64_t npairs, const uint8_t* ascii,
                              const uint64_t* offsets, uint8_t* packed_out, int nthreads) {
	if (!ascii || !offsets || !packed_out) return fail(DBT_E_BADARG, "null argument");
	const ReadLayout lay = make_layout(max_read_len);
	uint64_t* p0 = reinterpret_cast<uint64_t*>(packed_out);
	uint64_t* p1 = p0 + npairs * lay.W64;
	uint16_t* m0 = reinterpret_cast<uint16_t*>(p1 + npairs * lay.W64);
	uint16_t* m1 = m0 + npairs;
	const int impl = getenv("DBT_HOST_PACK_SWAR") ? HOST_PACK_SWAR : HOST_PACK_AUTO;   // test hook: force the non-AVX2 path
	if (nthreads <= 0) nthreads = usable_cpus();
	nthreads = (int)std::min<uint64_t>((uint64_t)nthreads, std::max<uint64_t>(1, npairs / 4096));
	std::atomic<int> bad{0};
	const uint64_t end = offsets[2 * npairs];
	auto work = [&](uint64_t a, uint64_t b) {
		for (uint64_t p = a; p < b; ++p) {
			for (int m = 0; m < 2; ++m) {
				uint64_t off = offsets[2 * p + m], len = offsets[2 * p + m + 1] - off;
				if (len > max_read_len) { bad = 1; len = max_read_len; }
				host_pack_read(ascii + off, (uint32_t)len, end - off, k, lay, (m ? p1 : p0) + p * lay.W64, (m ? m1 : m0) + p, impl);
			}
		}
	};
	std::vector<std::thread> th;
	uint64_t chunk = (npairs + nthreads - 1) / nthreads;
	for (int t = 1; t < nthreads; ++t) th.emplace_back(work, std::min<uint64_t>(npairs, t * chunk), std::min<uint64_t>(npairs, (t + 1) * chunk));
	work(0, std::min<uint64_t>(npairs, chunk));
	for (auto& x : th) x.join();
	if (bad) return fail(DBT_E_CAPACITY, "a read is longer than max_read_len=%u", max_read_len);
	return DBT_OK;
}

extern "C" void* dbt_host_alloc(size_t bytes) {
	void* p = nullptr;
	if (cudaHostAlloc(&p, std::max<size_t>(bytes, 1), cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
	return p;
}
extern "C" void dbt_host_free(void* p) { if (p) cudaFreeHost(p); }

// ------------------------------------------------------------------------------------------
// batches
// ------------------------------------------------------------------------------------------
static SlotView slot_view(const dbt_ctx* c, const Slot& s) {
	SlotView v;
	v.packed[0] = s.d_packed[0]; v.packed[1] = s.d_packed[1]; v.meta[0] = s.d_meta[0]; v.meta[1] = s.d_meta[1];
	v.passmask = s.d_passmask; v.blocksum = s.d_blocksum; v.cand = s.d_cand; v.ncand = s.d_ncand; v.slow = s.d_slow; v.thr = s.d_thr; v.thr2 = s.d_thr2; v.records = s.d_records;
	v.counters = s.d_counters; v.npairs = s.npairs; v.first_pair_index = s.first_pair_index;
	v.qual = s.q_base; v.qoff = s.q_off;
	(void)c;
	return v;
}

// pinned staging for records: only a pageable caller buffer or DBT_F_TRACE needs it
static int ensure_h_records(dbt_ctx* c, Slot& s) {
	if (s.h_records) return DBT_OK;
	cudaError_t e = cudaHostAlloc((void**)&s.h_records, std::max<size_t>((size_t)c->p.max_records * c->rp.rec_stride, 16), cudaHostAllocDefault);
	if (e != cudaSuccess) { cudaGetLastError(); return fail(DBT_E_NOMEM, "cudaHostAlloc(records): %s", cudaGetErrorString(e)); }
	return DBT_OK;
}

static int check_slot(dbt_ctx* c, int slot) {
	if (!c) return fail(DBT_E_BADARG, "null context");
	if (slot < 0 || slot >= DBT_N_SLOTS) return fail(DBT_E_BADARG, "slot %d out of range", slot);
	if (!c->index_loaded) return fail(DBT_E_STATE, "dbt_load_index has not been called");
	TRY(set_device(c));
	return ensure_slot(c, slot);
}

extern "C" int dbt_upload(dbt_ctx* c, const dbt_batch* b, int slot) {
	TRY(check_slot(c, slot));
	if (!b) return fail(DBT_E_BADARG, "null batch");
	Slot& s = c->slot[slot];
	if (s.fetch_pending) return fail(DBT_E_STATE, "slot %d has an unfinished fetch; call dbt_wait", slot);
	if (b->npairs > c->p.max_batch_pairs) return fail(DBT_E_CAPACITY, "npairs=%llu > max_batch_pairs=%u", (unsigned long long)b->npairs, c->p.max_batch_pairs);
	s.npairs = b->npairs; s.first_pair_index = b->first_pair_index; s.processed = false; s.ev_pack = false;
	s.q_base = nullptr; s.q_off = nullptr;
	const RunParams& rp = c->rp;
	CU(cudaMemsetAsync(s.d_ncand, 0, 32, s.st));
	if (b->npairs) {
		if (!b->bases) return fail(DBT_E_BADARG, "null bases");
		if (b->layout == DBT_LAYOUT_ASCII) {
			if (!b->offsets) return fail(DBT_E_BADARG, "null offsets");
			const uint64_t nbytes = b->offsets[2 * b->npairs] - b->offsets[0];
			if (nbytes > (uint64_t)2 * c->p.max_batch_pairs * rp.max_read_len) return fail(DBT_E_CAPACITY, "batch holds %llu bases, more than max_batch_pairs*2*max_read_len", (unsigned long long)nbytes);
			if (!s.d_ascii) {                                   // first ASCII batch of the slot (raw-text submissions never need these)
				TRY(dev_alloc(&s.d_ascii, (size_t)2 * c->p.max_batch_pairs * rp.max_read_len + 16, "ascii"));
				TRY(dev_alloc(&s.d_offsets, 2 * (size_t)c->p.max_batch_pairs + 1, "offsets"));
			}
			CU(cudaMemcpyAsync(s.d_ascii, b->bases + b->offsets[0], nbytes, cudaMemcpyHostToDevice, s.st));
			CU(cudaMemcpyAsync(s.d_offsets, b->offsets, (2 * b->npairs + 1) * 8, cudaMemcpyHostToDevice, s.st));
			if (b->quals && (c->p.flags & DBT_F_BAIT)) {            // quality strings, laid out like the bases
				if (!s.d_qual) TRY(dev_alloc(&s.d_qual, (size_t)2 * c->p.max_batch_pairs * rp.max_read_len + 16, "qualities"));
				CU(cudaMemcpyAsync(s.d_qual, b->quals + b->offsets[0], nbytes, cudaMemcpyHostToDevice, s.st));
				s.q_base = s.d_qual - b->offsets[0]; s.q_off = s.d_offsets;
			}
			// offsets are used relative to bases: shift by offsets[0] inside the kernel via pointer arithmetic
			CU(cudaEventRecord(s.evp0, s.st));
			CU(launch_pack(s.st, s.d_ascii - b->offsets[0], s.d_offsets, s.d_offsets + 1, b->npairs, rp, 0, s.d_packed[0], s.d_packed[1], s.d_meta[0], s.d_meta[1], s.d_ncand + 1)); ++c->launches;
			CU(cudaEventRecord(s.evp1, s.st));
			s.ev_pack = true;
		} else if (b->layout == DBT_LAYOUT_PACKED) {
			const uint64_t* p0 = reinterpret_cast<const uint64_t*>(b->bases);
			const uint64_t* p1 = p0 + b->npairs * rp.lay.W64;
			const uint16_t* m0 = reinterpret_cast<const uint16_t*>(p1 + b->npairs * rp.lay.W64);
			const uint16_t* m1 = m0 + b->npairs;
			CU(cudaMemcpyAsync(s.d_packed[0], p0, b->npairs * rp.lay.W64 * 8, cudaMemcpyHostToDevice, s.st));
			CU(cudaMemcpyAsync(s.d_packed[1], p1, b->npairs * rp.lay.W64 * 8, cudaMemcpyHostToDevice, s.st));
			CU(cudaMemcpyAsync(s.d_meta[0], m0, b->npairs * 2, cudaMemcpyHostToDevice, s.st));
			CU(cudaMemcpyAsync(s.d_meta[1], m1, b->npairs * 2, cudaMemcpyHostToDevice, s.st));
		} else return fail(DBT_E_BADARG, "unknown layout %d", b->layout);
	}
	s.uploaded = true;
	return DBT_OK;
}

// ---- raw FASTA/FASTQ text: split into records and pair neighbours on the device (fastx.cu), then pack ----
static int ensure_raw(dbt_ctx* c, Slot& s, uint64_t nbytes) {
	if (!s.h_rhdr) {
		CU(cudaHostAlloc((void**)&s.h_rhdr, 64, cudaHostAllocDefault));
		CU(cudaEventCreate(&s.evr0)); CU(cudaEventCreate(&s.evr1));
		TRY(dev_alloc(&s.d_rhdr, 16, "raw header"));
		const uint64_t P = c->p.max_batch_pairs;
		TRY(dev_alloc(&s.d_rstart, 2 * P, "raw read starts")); TRY(dev_alloc(&s.d_rend, 2 * P, "raw read ends")); TRY(dev_alloc(&s.d_rqual, 2 * P, "raw quality starts"));
		TRY(dev_alloc(&s.d_pairoff, P, "raw pair offsets")); TRY(dev_alloc(&s.d_pairhash, P, "raw pair hashes"));
		s.raw_max_single = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(P / 8, 1024), 1u << 22);
		TRY(dev_alloc(&s.d_singleoff, s.raw_max_single, "raw singleton offsets"));
	}
	if (nbytes <= s.raw_cap) return DBT_OK;
	CU(cudaStreamSynchronize(s.st));
	cudaFree(s.d_raw); cudaFree(s.d_nlmask); cudaFree(s.d_rblocksum); cudaFree(s.d_nl); cudaFree(s.d_tlen); cudaFree(s.d_slen); cudaFree(s.d_rhash); cudaFree(s.d_eq);
	cudaFree(s.d_pairmask); cudaFree(s.d_singlemask); cudaFree(s.d_pairlist); cudaFree(s.d_singlelist);
	s.d_raw = nullptr; s.d_nlmask = s.d_rblocksum = s.d_nl = s.d_tlen = s.d_slen = s.d_pairmask = s.d_singlemask = s.d_pairlist = s.d_singlelist = nullptr;
	s.d_rhash = nullptr; s.d_eq = nullptr; s.raw_cap = 0;
	const uint64_t cap = (std::max<uint64_t>(nbytes + nbytes / 8, 1 << 20) + 31) & ~31ull;
	// a record is at least two lines; lines of real read files are tens of bytes: room for one line per 8 bytes, one record per 16
	const uint64_t max_lines = cap / 8, max_recs = cap / 16;
	TRY(dev_alloc(&s.d_raw, cap + 64, "raw text"));
	TRY(dev_alloc(&s.d_nlmask, cap / 32 + 8, "newline bitmap"));
	TRY(dev_alloc(&s.d_rblocksum, cap / 32 / COMPACT_WORDS_H + 8, "raw compaction scratch"));
	TRY(dev_alloc(&s.d_nl, max_lines + 8, "newline offsets"));
	TRY(dev_alloc(&s.d_rhash, max_recs + 8, "record hashes")); TRY(dev_alloc(&s.d_tlen, max_recs + 8, "title lengths")); TRY(dev_alloc(&s.d_slen, max_recs + 8, "sequence lengths"));
	TRY(dev_alloc(&s.d_eq, max_recs + 8, "record equality"));
	TRY(dev_alloc(&s.d_pairmask, max_recs / 32 + 8, "pair bitmap")); TRY(dev_alloc(&s.d_singlemask, max_recs / 32 + 8, "singleton bitmap"));
	TRY(dev_alloc(&s.d_pairlist, max_recs + 8, "pair list")); TRY(dev_alloc(&s.d_singlelist, max_recs + 8, "singleton list"));
	s.raw_cap = cap; s.raw_max_lines = (uint32_t)max_lines; s.raw_max_recs = (uint32_t)max_recs;
	return DBT_OK;
}

extern "C" int dbt_prepare(dbt_ctx* c, uint64_t raw_bytes) {
	if (!c) return fail(DBT_E_BADARG, "null context");
	for (int i = 0; i < DBT_N_SLOTS; ++i) {
		TRY(check_slot(c, i));
		if (raw_bytes) TRY(ensure_raw(c, c->slot[i], raw_bytes));
	}
	return DBT_OK;
}

extern "C" uint64_t dbt_title_hash(const uint8_t* t, size_t len) {
	if (len >= 2 && t[len - 2] == '/' && (t[len - 1] == '1' || t[len - 1] == '2')) len -= 2;
	uint64_t h = 0xcbf29ce484222325ull;
	for (size_t i = 0; i < len; ++i) { h ^= t[i]; h *= 0x100000001b3ull; }
	return h;
}

extern "C" int dbt_upload_raw(dbt_ctx* c, const dbt_raw_batch* b, dbt_raw_info* info, int slot) {
	TRY(check_slot(c, slot));
	if (!b || !info) return fail(DBT_E_BADARG, "null argument");
	Slot& s = c->slot[slot];
	if (s.fetch_pending) return fail(DBT_E_STATE, "slot %d has an unfinished fetch; call dbt_wait", slot);
	memset(info, 0, sizeof *info);
	s.uploaded = false; s.processed = false; s.ev_pack = false; s.ev_raw = false; s.npairs = 0; s.first_pair_index = b->first_pair_index;
	s.q_base = nullptr; s.q_off = nullptr;
	if (b->nbytes >= (1ull << 32) - 64) return fail(DBT_E_CAPACITY, "raw text of %llu bytes: a batch must stay below 4 GiB", (unsigned long long)b->nbytes);
	if (b->nbytes && (!b->text || b->text[b->nbytes - 1] != '\n')) return fail(DBT_E_BADARG, "raw text must hold whole lines (last byte '\\n')");
	CU(cudaMemsetAsync(s.d_ncand, 0, 32, s.st));
	if (!b->nbytes) { s.uploaded = true; return DBT_OK; }
	if (!b->pair_offset || (!b->single_offset && b->single_cap)) return fail(DBT_E_BADARG, "null output array");
	TRY(ensure_raw(c, s, b->nbytes));
	const RunParams& rp = c->rp;
	RawView v{};
	v.raw = s.d_raw; v.nlmask = s.d_nlmask; v.blocksum = s.d_rblocksum; v.nl = s.d_nl; v.hdr = s.d_rhdr; v.hash = s.d_rhash; v.tlen = s.d_tlen; v.slen = s.d_slen; v.eq = s.d_eq;
	v.pairmask = s.d_pairmask; v.singlemask = s.d_singlemask; v.pairlist = s.d_pairlist; v.singlelist = s.d_singlelist;
	v.check_qual = (b->is_fastq && (c->p.flags & DBT_F_BAIT)) ? 1u : 0u;
	v.rstart = s.d_rstart; v.rend = s.d_rend; v.rqual = s.d_rqual; v.pairoff = s.d_pairoff; v.pairhash = s.d_pairhash; v.singleoff = s.d_singleoff;
	v.max_lines = s.raw_max_lines; v.max_recs = s.raw_max_recs; v.max_pairs = c->p.max_batch_pairs; v.max_single = std::min(s.raw_max_single, b->single_cap);
	if (b->is_fastq && (c->p.flags & DBT_F_BAIT)) { s.q_base = s.d_raw; s.q_off = s.d_rqual; }
	CU(cudaEventRecord(s.evr0, s.st));
	CU(cudaMemcpyAsync(s.d_raw, b->text, b->nbytes, cudaMemcpyHostToDevice, s.st));
	CU(cudaMemsetAsync(s.d_raw + b->nbytes, 0, 64 + (32 - b->nbytes % 32) % 32, s.st));      // nl_mask_kernel reads 32-byte words, pack_kernel stages 16-byte words
	CU(launch_fastx_split(s.st, c->nsm, v, b->nbytes, b->is_fastq ? 4u : 2u, b->min_read_len)); c->launches += 14;
	CU(cudaMemcpyAsync(s.h_rhdr, s.d_rhdr, 32, cudaMemcpyDeviceToHost, s.st));
	CU(cudaEventRecord(s.evr1, s.st));
	CU(cudaStreamSynchronize(s.st));
	s.ev_raw = true;
	memcpy(info, s.h_rhdr, sizeof *info);
	if (info->flags) return DBT_OK;                          // the caller takes the host path for this text
	if (info->npairs > c->p.max_batch_pairs) return fail(DBT_E_CAPACITY, "raw text holds %u pairs > max_batch_pairs=%u", info->npairs, c->p.max_batch_pairs);
	const uint64_t P = info->npairs;
	if (P) {
		CU(cudaMemcpyAsync(b->pair_offset, s.d_pairoff, P * 4, cudaMemcpyDeviceToHost, s.st));
		if (b->pair_hash) CU(cudaMemcpyAsync(b->pair_hash, s.d_pairhash, P * 8, cudaMemcpyDeviceToHost, s.st));
	}
	if (info->nsingle) CU(cudaMemcpyAsync(b->single_offset, s.d_singleoff, (size_t)info->nsingle * 4, cudaMemcpyDeviceToHost, s.st));
	// a pair shorter than min_read_len never reaches here; a pair LONGER than the context was built for is the caller's to handle
	if (P && info->max_read_len > rp.max_read_len) {
		CU(cudaStreamSynchronize(s.st));
		return fail(DBT_E_CAPACITY, "a read of %u bases is longer than max_read_len=%u", info->max_read_len, rp.max_read_len);
	}
	if (P) {
		CU(cudaEventRecord(s.evp0, s.st));
		CU(launch_pack(s.st, s.d_raw, s.d_rstart, s.d_rend, P, rp, 32u * info->max_record_bytes, s.d_packed[0], s.d_packed[1], s.d_meta[0], s.d_meta[1], s.d_ncand + 1)); ++c->launches;
		CU(cudaEventRecord(s.evp1, s.st));
		s.ev_pack = true;
	}
	CU(cudaStreamSynchronize(s.st));                         // pair_offset / single_offset are the caller's to read now
	s.npairs = P; s.uploaded = true;
	return DBT_OK;
}

extern "C" int dbt_bubble_events(dbt_ctx* c, int slot, uint64_t* key, uint64_t* edge, uint32_t* locus, uint64_t cap, uint64_t* n) {
	TRY(check_slot(c, slot));
	if (!n) return fail(DBT_E_BADARG, "null argument");
	*n = 0;
	if (!c->bubbles) return fail(DBT_E_STATE, "the context was not created with DBT_F_BUBBLES");
	Slot& s = c->slot[slot];
	if (s.fetch_pending) return fail(DBT_E_STATE, "slot %d has an unfinished fetch; call dbt_wait", slot);
	CU(cudaStreamSynchronize(s.st));
	const uint64_t nev = s.ev_pending ? *s.h_nev : 0;
	if (nev > s.ev_cap) return fail(DBT_E_CAPACITY, "%llu novel-edge events in one batch, the slot holds %u: submit smaller batches", (unsigned long long)nev, s.ev_cap);
	if (nev > cap) return fail(DBT_E_CAPACITY, "%llu novel-edge events, the caller's arrays hold %llu", (unsigned long long)nev, (unsigned long long)cap);
	if (nev) {
		if (!key || !edge || !locus) return fail(DBT_E_BADARG, "null output array");
		CU(cudaMemcpy(key, s.d_evkey, nev * 8, cudaMemcpyDeviceToHost));
		CU(cudaMemcpy(edge, s.d_evedge, nev * 8, cudaMemcpyDeviceToHost));
		CU(cudaMemcpy(locus, s.d_evlocus, nev * 4, cudaMemcpyDeviceToHost));
	}
	*n = nev;
	return DBT_OK;
}

extern "C" int dbt_raw_times(dbt_ctx* c, int slot, float ms[2]) {
	TRY(check_slot(c, slot));
	Slot& s = c->slot[slot];
	ms[0] = ms[1] = 0.f;
	CU(cudaStreamSynchronize(s.st));
	if (s.ev_raw) CU(cudaEventElapsedTime(&ms[0], s.evr0, s.evr1));
	if (s.ev_pack) CU(cudaEventElapsedTime(&ms[1], s.evp0, s.evp1));
	return DBT_OK;
}

extern "C" int dbt_process(dbt_ctx* c, int slot) {
	TRY(check_slot(c, slot));
	Slot& s = c->slot[slot];
	if (!s.uploaded) return fail(DBT_E_STATE, "slot %d has no uploaded batch", slot);
	c->reduced_valid = false;                  // the local arrays move on; dbt_download hands them out again until the next reduce
	SlotView sv = slot_view(c, s);
	CU(cudaMemsetAsync(s.d_counters, 0, DBT_N_COUNTERS * 8, s.st));
	CU(cudaMemsetAsync(s.d_ncand, 0, 4, s.st));
	CU(cudaMemsetAsync(s.d_ncand + 2, 0, 12, s.st));
	if (c->rp.thr_on && (c->p.flags & DBT_F_TRACE) && s.d_records)      // blobs of pairs that never reach threading read as zeros
		CU(cudaMemsetAsync(s.d_records, 0, (size_t)c->p.max_records * c->rp.rec_stride, s.st));
	CU(cudaEventRecord(s.ev[1], s.st));
	CU(launch_subfilter(s.st, c->B, c->nsm, c->ix, c->rp, sv)); if (s.npairs) ++c->launches;
	CU(cudaEventRecord(s.ev[2], s.st));
	if (s.npairs) { CU(launch_compact(s.st, sv)); c->launches += 3; }
	CU(cudaEventRecord(s.ev[3], s.st));
	CU(launch_candidate(s.st, c->B, c->cand_grid_fast, c->cand_grid, c->ix, c->rp, sv, s.ev[4]));
	if (s.npairs) c->launches += ((c->p.flags & DBT_F_TRACE) || c->rp.nk_max > 255) ? 1 : 2;   // generic kernel alone, or closed-form + generic
	if (c->bubbles && s.npairs) {
		CU(cudaMemsetAsync(s.d_nev, 0, 4, s.st));
		CU(launch_bubbles(s.st, c->B, c->nsm, c->ttable, c->rp, sv, (uint32_t)c->ix.nloci, s.d_evkey, s.d_evedge, s.d_evlocus, s.d_nev, s.ev_cap)); ++c->launches;
	}
	CU(cudaEventRecord(s.ev[5], s.st));
	if (c->rp.thr_on) { CU(launch_thread(s.st, c->B, c->nsm, c->ix, c->rp, sv, s.ev[6])); if (s.npairs) c->launches += 2; }
	else CU(cudaEventRecord(s.ev[6], s.st));
	CU(cudaEventRecord(s.ev[7], s.st));
	s.processed = true; s.ev_proc = true;
	return DBT_OK;
}

extern "C" int dbt_fetch(dbt_ctx* c, dbt_batch_result* r, int slot) {
	TRY(check_slot(c, slot));
	Slot& s = c->slot[slot];
	if (!s.processed) return fail(DBT_E_STATE, "slot %d has not been processed", slot);
	if (!r) return fail(DBT_E_BADARG, "null result");
	s.emitted = false;
	if (s.d_records_out && r->records && s.npairs) {          // kam / aln records: compact on the device, in slot (= pair) order
		const bool thr = c->rp.thr_on != 0, aln_min = (c->p.flags & DBT_F_ALN_MIN) != 0;
		const bool okam = thr ? (c->p.flags & (DBT_F_ALN | DBT_F_ALN_MIN)) != 0 : (c->p.flags & (DBT_F_KAM | DBT_F_EXTRACT)) != 0;
		if (okam) {
			CU(launch_record_emit(s.st, c->nsm, s.d_records, c->rp.rec_stride, s.d_ncand, c->p.max_records, ((thr && !aln_min) ? 1 : 0) | ((!thr && (c->p.flags & DBT_F_SIM)) ? 2 : 0),
			                      s.d_emitmask, s.d_emitsum, s.d_emitlist, s.d_nemit, s.d_records_out));
			c->launches += 5;
			CU(cudaMemcpyAsync(reinterpret_cast<uint32_t*>(s.h_counters + DBT_N_COUNTERS + 2), s.d_nemit, 4, cudaMemcpyDeviceToHost, s.st));
			s.emitted = true;
		}
	}
	s.ev_pending = false;
	if (c->bubbles) { *s.h_nev = 0; if (s.npairs) { CU(cudaMemcpyAsync(s.h_nev, s.d_nev, 4, cudaMemcpyDeviceToHost, s.st)); s.ev_pending = true; } }
	CU(cudaMemcpyAsync(s.h_counters, s.d_counters, DBT_N_COUNTERS * 8, cudaMemcpyDeviceToHost, s.st));
	CU(cudaMemcpyAsync(s.h_counters + DBT_N_COUNTERS, s.d_ncand, 16, cudaMemcpyDeviceToHost, s.st));
	s.result = r; s.fetch_pending = true;
	return DBT_OK;
}

extern "C" int dbt_submit(dbt_ctx* c, const dbt_batch* b, dbt_batch_result* r, int slot) {
	TRY(dbt_upload(c, b, slot));
	TRY(dbt_process(c, slot));
	return dbt_fetch(c, r, slot);
}

extern "C" int dbt_wait(dbt_ctx* c, int slot) {
	TRY(check_slot(c, slot));
	Slot& s = c->slot[slot];
	CU(cudaStreamSynchronize(s.st));
	if (!s.fetch_pending) return DBT_OK;
	s.fetch_pending = false;
	dbt_batch_result* r = s.result; s.result = nullptr;
	const uint32_t* tail = reinterpret_cast<const uint32_t*>(s.h_counters + DBT_N_COUNTERS);
	const uint32_t ncand = tail[0], err = tail[1];
	s.last_ncand = ncand; s.last_nslow = tail[2]; s.last_nthr = tail[3];
	if (err & ERR_READ_TOO_LONG) return fail(DBT_E_CAPACITY, "a read is longer than max_read_len=%u", c->p.max_read_len);
	for (int i = 0; i < DBT_N_COUNTERS; ++i) r->counters[i] = s.h_counters[i];
	r->counters[DBT_C_NREADS] = 2 * s.npairs;
	r->n_records = 0;
	const bool thr = c->rp.thr_on != 0;
	const bool aln_min = (c->p.flags & DBT_F_ALN_MIN) != 0, aln = (c->p.flags & (DBT_F_ALN | DBT_F_ALN_MIN)) != 0;
	const bool trace = (c->p.flags & DBT_F_TRACE) != 0, okam = thr ? aln : (c->p.flags & (DBT_F_KAM | DBT_F_EXTRACT)) != 0;
	int rc = DBT_OK;
	if (s.emitted) {
		// dense records are already in slot order on the device: one copy straight into the caller's buffer
		// (DMA when it is pinned -- dbt_host_alloc -- else staged through this slot's pinned buffer)
		if (err & ERR_RECORD_OVERFLOW) rc = fail(DBT_E_CAPACITY, "%u candidate pairs exceed max_records=%u", ncand, c->p.max_records);
		uint64_t n = *reinterpret_cast<const uint32_t*>(s.h_counters + DBT_N_COUNTERS + 2);
		if (n > r->capacity) { rc = fail(DBT_E_CAPACITY, "result buffer holds %llu records, more are pending", (unsigned long long)r->capacity); n = r->capacity; }
		if (n) {
			const size_t bytes = (size_t)n * c->rp.rec_stride;
			cudaPointerAttributes pa{};
			const bool pinned = cudaPointerGetAttributes(&pa, r->records) == cudaSuccess && pa.type == cudaMemoryTypeHost;
			if (!pinned) cudaGetLastError();
			if (!pinned) TRY(ensure_h_records(c, s));
			CU(cudaMemcpyAsync(pinned ? r->records : s.h_records, s.d_records_out, bytes, cudaMemcpyDeviceToHost, s.st));
			CU(cudaStreamSynchronize(s.st));
			if (!pinned) memcpy(r->records, s.h_records, bytes);
		}
		r->n_records = n;
	} else if ((trace || okam) && r->records) {
		if (err & ERR_RECORD_OVERFLOW) rc = fail(DBT_E_CAPACITY, "%u candidate pairs exceed max_records=%u", ncand, c->p.max_records);
		const uint64_t nslots = trace ? s.npairs : std::min<uint64_t>(ncand, c->p.max_records);
		const uint32_t stride = c->rp.rec_stride;
		if (nslots) {
			TRY(ensure_h_records(c, s));
			CU(cudaMemcpyAsync(s.h_records, s.d_records, nslots * stride, cudaMemcpyDeviceToHost, s.st));
			CU(cudaStreamSynchronize(s.st));
		}
		uint64_t n = 0;
		for (uint64_t i = 0; i < nslots; ++i) {
			const uint8_t* src = s.h_records + i * stride;
			const uint8_t stt = reinterpret_cast<const dbt_pair_record*>(src)->status;
			// kam line / `-ae`: counted pairs only; `-a`: every pair that entered threading (cpp:2232-2239)
			if (!trace && stt != DBT_ST_COUNTED && !(thr && !aln_min && stt == DBT_ST_THREAD_FAILED) && !(!thr && (c->p.flags & DBT_F_SIM) && stt == DBT_ST_ASGN_REMOVED)) continue;
			if (n >= r->capacity) { rc = fail(DBT_E_CAPACITY, "result buffer holds %llu records, more are pending", (unsigned long long)r->capacity); break; }
			memcpy(r->records + n * stride, src, stride);
			++n;
		}
		r->n_records = n;
	}
	r->counters[DBT_C_NRECORDS] = r->n_records;
	for (int i = 0; i < DBT_N_COUNTERS; ++i) c->totals[i] += r->counters[i];
	return rc;
}

// ------------------------------------------------------------------------------------------
// results
// ------------------------------------------------------------------------------------------
static int sync_all(dbt_ctx* c) {
	for (auto& s : c->slot) if (s.ready) CU(cudaStreamSynchronize(s.st));
	return DBT_OK;
}

extern "C" int dbt_download(dbt_ctx* c, uint64_t* trkmc, uint64_t* kmc, uint32_t* nmapread, uint64_t counters[DBT_N_COUNTERS]) {
	if (!c) return fail(DBT_E_BADARG, "null context");
	if (!c->index_loaded) return fail(DBT_E_STATE, "no index");
	TRY(set_device(c)); TRY(sync_all(c));
	// after dbt_allreduce (and until the next dbt_process / dbt_reset_counts): the all-rank sums
	const bool red = c->reduced_valid;
	const uint64_t* base = red ? c->d_red : c->d_results;
	if (trkmc && c->ix.nk_tr) CU(cudaMemcpy(trkmc, base, c->ix.nk_tr * 8, cudaMemcpyDeviceToHost));
	if (kmc) CU(cudaMemcpy(kmc, base + c->ix.nk_tr, c->ix.nloci * 8, cudaMemcpyDeviceToHost));
	if (nmapread) CU(cudaMemcpy(nmapread, base + c->n_res64, c->ix.nloci * 4, cudaMemcpyDeviceToHost));
	if (counters) for (int i = 0; i < DBT_N_COUNTERS; ++i) counters[i] = red ? c->totals_red[i] : c->totals[i];
	return DBT_OK;
}

extern "C" int dbt_reset_counts(dbt_ctx* c) {
	if (!c) return fail(DBT_E_BADARG, "null context");
	if (!c->index_loaded) return fail(DBT_E_STATE, "no index");
	TRY(set_device(c)); TRY(sync_all(c));
	CU(cudaMemset(c->d_results, 0, (c->n_res64 + (c->ix.nloci + 1) / 2) * 8));
	memset(c->totals, 0, sizeof(c->totals));
	c->reduced_valid = false;
	return DBT_OK;
}

// ------------------------------------------------------------------------------------------
// measurement helpers
// ------------------------------------------------------------------------------------------
extern "C" int dbt_timer_start(dbt_ctx* c, int slot) {
	TRY(check_slot(c, slot));
	CU(cudaEventRecord(c->slot[slot].t0, c->slot[slot].st));
	return DBT_OK;
}
extern "C" int dbt_timer_stop(dbt_ctx* c, int slot, float* ms) {
	TRY(check_slot(c, slot));
	Slot& s = c->slot[slot];
	CU(cudaEventRecord(s.t1, s.st));
	CU(cudaEventSynchronize(s.t1));
	if (ms) CU(cudaEventElapsedTime(ms, s.t0, s.t1));
	return DBT_OK;
}
extern "C" int dbt_kernel_times(dbt_ctx* c, int slot, float ms[8], uint32_t launches[8]) {
	TRY(check_slot(c, slot));
	Slot& s = c->slot[slot];
	CU(cudaStreamSynchronize(s.st));
	for (int i = 0; i < 8; ++i) { ms[i] = 0; launches[i] = 0; }
	if (s.ev_pack) { CU(cudaEventElapsedTime(&ms[KT_PACK], s.evp0, s.evp1)); launches[KT_PACK] = 1; }   // the slot's last ASCII upload
	if (s.ev_proc) {
		CU(cudaEventElapsedTime(&ms[KT_SUBFILTER], s.ev[1], s.ev[2])); launches[KT_SUBFILTER] = 1;
		CU(cudaEventElapsedTime(&ms[KT_COMPACT], s.ev[2], s.ev[3])); launches[KT_COMPACT] = 3;
		const bool all = (c->p.flags & DBT_F_TRACE) || c->rp.nk_max > 255;
		CU(cudaEventElapsedTime(&ms[KT_CANDIDATE], s.ev[3], s.ev[4])); launches[KT_CANDIDATE] = all ? 0 : 1;
		CU(cudaEventElapsedTime(&ms[KT_GENERIC], s.ev[4], s.ev[5])); launches[KT_GENERIC] = 1;
		CU(cudaEventElapsedTime(&ms[KT_THREAD], s.ev[5], s.ev[6])); launches[KT_THREAD] = c->rp.thr_on ? 1 : 0;
		CU(cudaEventElapsedTime(&ms[KT_THREAD2], s.ev[6], s.ev[7])); launches[KT_THREAD2] = c->rp.thr_on ? 1 : 0;
	}
	return DBT_OK;
}

extern "C" int dbt_batch_stats(dbt_ctx* c, int slot, uint64_t stats[4]) {
	if (!c || !stats) return fail(DBT_E_BADARG, "null argument");
	if (slot < 0 || slot >= DBT_N_SLOTS) return fail(DBT_E_BADARG, "slot %d out of range", slot);
	stats[0] = c->slot[slot].last_ncand; stats[1] = c->slot[slot].last_nslow; stats[2] = c->slot[slot].last_nthr; stats[3] = 0;
	return DBT_OK;
}

extern "C" int dbt_table_stats(const dbt_ctx* cc, uint64_t stats[8]) {
	dbt_ctx* c = const_cast<dbt_ctx*>(cc);
	if (!c || !stats) return fail(DBT_E_BADARG, "null argument");
	if (!c->index_loaded) return fail(DBT_E_STATE, "no index");
	TRY(set_device(c));
	unsigned long long* d; TRY(dev_alloc(&d, 4, "stats"));
	CU(cudaMemset(d, 0, 32));
	CU(launch_table_stats(0, c->B, c->ix.table, d));
	unsigned long long h[4];
	CU(cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost));
	cudaFree(d);
	stats[0] = c->ix.table.nbuckets; stats[1] = (uint64_t)c->B; stats[2] = c->nkeys; stats[3] = h[0]; stats[4] = h[1]; stats[5] = h[2];
	stats[6] = c->table_bytes; stats[7] = h[3];
	return DBT_OK;
}

extern "C" int dbt_random_gather_bench(int device, uint64_t table_bytes, uint32_t sector_bytes, uint64_t n_gathers, int iters, float* best_ms) {
	if (!best_ms) return fail(DBT_E_BADARG, "null argument");
	CU(cudaSetDevice(device));
	set_l2_fetch_granularity();
	cudaDeviceProp prop; CU(cudaGetDeviceProperties(&prop, device));
	uint4* t; TRY(dev_alloc(&t, table_bytes / 16, "gather table"));
	uint32_t* sink; TRY(dev_alloc(&sink, 1, "sink"));
	CU(cudaMemset(t, 1, table_bytes));
	cudaEvent_t a, b; CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
	float best = 1e30f;
	for (int it = 0; it < iters + 1; ++it) {
		CU(cudaEventRecord(a, 0));
		CU(launch_random_gather(0, prop.multiProcessorCount, t, table_bytes, sector_bytes, n_gathers, 0x9E3779B97F4A7C15ull * (it + 1), sink));
		CU(cudaEventRecord(b, 0));
		CU(cudaEventSynchronize(b));
		float ms; CU(cudaEventElapsedTime(&ms, a, b));
		if (it > 0 && ms < best) best = ms;
	}
	cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(t); cudaFree(sink);
	*best_ms = best;
	return DBT_OK;
}

extern "C" int dbt_init_device(int device) {
	CU(cudaSetDevice(device));
	CU(cudaFree(nullptr));
	return DBT_OK;
}

extern "C" int dbt_bus_bench(int device, uint64_t bytes, int iters, float* h2d_ms, float* d2h_ms) {
	if (!h2d_ms || !d2h_ms || !bytes) return fail(DBT_E_BADARG, "null argument");
	CU(cudaSetDevice(device));
	void* h = nullptr; uint8_t* d = nullptr;
	CU(cudaMallocHost(&h, bytes));
	memset(h, 1, bytes);
	if (int rc = dev_alloc(&d, bytes, "bus bench")) { cudaFreeHost(h); return rc; }
	cudaEvent_t a, b; CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
	float best[2] = {1e30f, 1e30f};
	for (int dir = 0; dir < 2; ++dir)
		for (int it = 0; it < iters + 1; ++it) {
			CU(cudaEventRecord(a, 0));
			CU(dir ? cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, 0) : cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, 0));
			CU(cudaEventRecord(b, 0));
			CU(cudaEventSynchronize(b));
			float ms; CU(cudaEventElapsedTime(&ms, a, b));
			if (it > 0 && ms < best[dir]) best[dir] = ms;
		}
	cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(d); cudaFreeHost(h);
	*h2d_ms = best[0]; *d2h_ms = best[1];
	return DBT_OK;
}

// ------------------------------------------------------------------------------------------
// NCCL (dlopen'ed: the library loads and runs single-GPU without libnccl)
// ------------------------------------------------------------------------------------------
static int nccl_load() {
	if (g_nccl.lib) return DBT_OK;
	const char* names[] = {"libnccl.so.2", "libnccl.so"};
	for (const char* n : names) { g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (g_nccl.lib) break; }
	if (!g_nccl.lib) return fail(DBT_E_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
	g_nccl.GetUniqueId = (int (*)(void*))dlsym(g_nccl.lib, "ncclGetUniqueId");
	g_nccl.CommInitRank = (int (*)(void**, int, Id128, int))dlsym(g_nccl.lib, "ncclCommInitRank");
	g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(g_nccl.lib, "ncclAllReduce");
	g_nccl.CommDestroy = (int (*)(void*))dlsym(g_nccl.lib, "ncclCommDestroy");
	g_nccl.GetErrorString = (const char* (*)(int))dlsym(g_nccl.lib, "ncclGetErrorString");
	g_nccl.GroupStart = (int (*)())dlsym(g_nccl.lib, "ncclGroupStart");
	g_nccl.GroupEnd = (int (*)())dlsym(g_nccl.lib, "ncclGroupEnd");
	if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) return fail(DBT_E_NCCL, "libnccl lacks a required symbol");
	return DBT_OK;
}
#define NC(call) do { int r__ = (call); if (r__ != 0) return fail(DBT_E_NCCL, "%s: %s", #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(r__) : "?"); } while (0)

extern "C" int dbt_nccl_unique_id(uint8_t id[DBT_NCCL_ID_BYTES]) {
	TRY(nccl_load());
	NC(g_nccl.GetUniqueId(id));
	return DBT_OK;
}

extern "C" int dbt_comm_init(dbt_ctx* c, int nranks, int rank, const uint8_t id[DBT_NCCL_ID_BYTES]) {
	if (!c || !id || nranks < 1 || rank < 0 || rank >= nranks) return fail(DBT_E_BADARG, "bad communicator arguments");
	TRY(set_device(c));
	c->nranks = nranks; c->rank = rank;
	if (nranks == 1) return DBT_OK;
	TRY(nccl_load());
	Id128 u; memcpy(u.b, id, 128);
	NC(g_nccl.CommInitRank(&c->nccl_comm, nranks, u, rank));
	return DBT_OK;
}

// One exchange step at the end of a run: integer sums are order-independent, so
// N-GPU outputs are bit-identical to the single-GPU ones.  OUT OF PLACE: the rank's own arrays stay as they are and
// the sums land in a second block of the same layout, which dbt_download hands out until the next dbt_process /
// dbt_reset_counts -- so a second call (or more batches followed by another call) reduces the local state afresh
// instead of multiplying already-reduced counts by the number of ranks.  One grouped collective (u64 part + u32 part)
// on the context's own stream; no blocking copies.
extern "C" int dbt_allreduce(dbt_ctx* c) {
	if (!c) return fail(DBT_E_BADARG, "null context");
	if (!c->index_loaded) return fail(DBT_E_STATE, "no index");
	TRY(set_device(c));
	if (c->nranks == 1) { TRY(sync_all(c)); return DBT_OK; }
	if (!c->nccl_comm) return fail(DBT_E_STATE, "dbt_comm_init has not been called");
	enum { ncclUint32 = 3, ncclUint64 = 5, ncclSum = 0 };       // nccl.h: ncclDataType_t / ncclRedOp_t
	const uint64_t nloci = c->ix.nloci;
	if (!c->red_st) {
		CU(cudaStreamCreateWithFlags(&c->red_st, cudaStreamNonBlocking));
		CU(cudaEventCreateWithFlags(&c->red_ev, cudaEventDisableTiming));
		TRY(dev_alloc(&c->d_red, c->n_res64 + (nloci + 1) / 2, "reduced results"));
		CU(cudaHostAlloc((void**)&c->h_totals, 2 * sizeof(c->totals), cudaHostAllocDefault));
	}
	for (auto& s : c->slot) if (s.ready) { CU(cudaEventRecord(c->red_ev, s.st)); CU(cudaStreamWaitEvent(c->red_st, c->red_ev, 0)); }
	// the host-side totals ride in the same u64 block (they are complete: dbt_wait adds a batch's counters before it returns)
	for (auto& s : c->slot) if (s.fetch_pending) return fail(DBT_E_STATE, "dbt_allreduce with an unfinished fetch; call dbt_wait first");
	memcpy(c->h_totals, c->totals, sizeof(c->totals));
	CU(cudaMemcpyAsync(c->d_totals, c->h_totals, sizeof(c->totals), cudaMemcpyHostToDevice, c->red_st));
	uint32_t* red32 = reinterpret_cast<uint32_t*>(c->d_red + c->n_res64);
	if (g_nccl.GroupStart) NC(g_nccl.GroupStart());
	NC(g_nccl.AllReduce(c->d_results, c->d_red, c->n_res64, ncclUint64, ncclSum, c->nccl_comm, c->red_st));
	NC(g_nccl.AllReduce(c->d_nmapread, red32, nloci, ncclUint32, ncclSum, c->nccl_comm, c->red_st));
	if (g_nccl.GroupEnd) NC(g_nccl.GroupEnd());
	CU(cudaMemcpyAsync(c->h_totals + DBT_N_COUNTERS, c->d_red + (c->n_res64 - DBT_N_COUNTERS), sizeof(c->totals), cudaMemcpyDeviceToHost, c->red_st));
	CU(cudaStreamSynchronize(c->red_st));
	memcpy(c->totals_red, c->h_totals + DBT_N_COUNTERS, sizeof(c->totals));
	c->reduced_valid = true;
	return DBT_OK;
}
