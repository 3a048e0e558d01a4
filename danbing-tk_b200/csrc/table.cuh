// GPU-resident RPGG index (SURVEY.md §8(a) types Y1-Y3): ONE open-addressing table in HBM.
//
// What it replaces in the reference:
//   kmerDBi      unordered_map<size_t,size_t>   canonical k-mer -> val         (h:35, kio:59-74)
//   trResults[l] unordered_map<size_t,atomic>   TR k-mers of locus l -> count  (h:31, cpp:2465)
//   flankDB[l]   unordered_set<uint64_t>        flank k-mers of locus l        (h:47, cpp:2466)
// Every TR / flank k-mer of locus l is also a key of kmerDBi that lists l (ktools serialize builds
// kmerDBi as the union of .tr.kmers and .fl.kmers, kmertools.cpp:236-237), so the per-locus class
// ("flank" / "TR" + position in the count array) is folded into the index entry itself: one probe
// answers kfilter (cpp:207), the locus vote (cpp:367) AND assignTRkmc (cpp:1466-1468).
//
// Slot (16 B): { u64 key | bit63 = bucket-overflow flag (slot 0 only) ; u32 val ; u32 aux }
//   val : the reference's value -- even: locus<<1, odd: (offset into vv)<<1|1
//   aux : (class<<30) | idx for single-locus keys; class 0 none, 1 flank, 2 TR (= the `as` states of
//         cpp:78-92); idx = position in the output-ordered count array.  Multi-locus keys keep their
//         per-locus aux in vv_aux[], parallel to vv[].
// Bucket = B consecutive slots (B*16 bytes: 32 B = one L2 sector for B=2, 64 B for B=4).  A key lives in its
// home bucket = mulhi(hash(key), nbuckets) or, when that bucket was full at insertion time, in the STASH: a
// second, sparsely loaded bucket array behind the main one, entered at stash_bucket(key) and probed linearly.
// A full bucket whose slot 0 carries the overflow flag tells a lookup to go on; so a miss -- the dominant
// case for background reads in subfilter -- costs ONE sector unless more than B keys hashed to that bucket,
// and ~99.7 % of all look-ups end after at most two independent buckets (home, stash entry), which is what
// lets the candidate kernels batch probes in rounds instead of walking chains lane by lane.
// hash = high half of key * odd constant (Fibonacci hashing, 4 integer instructions; bucket occupancy on
// the synthetic RPGGs matches the Poisson expectation to three digits, same as a murmur3 finaliser).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace dbt {

constexpr uint64_t SLOT_EMPTY = 0xFFFFFFFFFFFFFFFFull;
constexpr uint64_t SLOT_OVF   = 0x8000000000000000ull;
constexpr uint64_t SLOT_KEYMASK = ~SLOT_OVF;
constexpr uint32_t AUX_CLASS_SHIFT = 30;
constexpr uint32_t AUX_IDX_MASK = (1u << AUX_CLASS_SHIFT) - 1;
constexpr uint32_t AUX_FLANK_TR = 1u << 29;               // class 1 (flank wins, cpp:1467-1468) entry of a k-mer that is ALSO a TR k-mer of the locus:
                                                          // bits 0-28 keep its count-array position (graph threading counts by trResults membership alone)
constexpr uint32_t AUX_MISS = 3u << AUX_CLASS_SHIFT;      // per-position marker: k-mer not in the index

struct Table {
	uint4*   slots;      // (nbuckets + nstash) * B slots
	uint64_t nbuckets;   // main buckets, < 2^32
	uint64_t nstash;     // stash buckets, < 2^32, located behind the main ones
};

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {            // used by the R_rand microbenchmark only
	x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
	return x;
}

__device__ __forceinline__ uint64_t home_bucket(const Table& t, uint64_t key) {
	return __umulhi((uint32_t)((key * 0x9E3779B97F4A7C15ull) >> 32), (uint32_t)t.nbuckets);
}
__device__ __forceinline__ uint64_t stash_bucket(const Table& t, uint64_t key) {
	return t.nbuckets + __umulhi((uint32_t)((key * 0xD6E8FEB86659FD93ull) >> 32), (uint32_t)t.nstash);
}
__device__ __forceinline__ uint64_t stash_next(const Table& t, uint64_t b) {
	return (b + 1 == t.nbuckets + t.nstash) ? t.nbuckets : b + 1;
}

__device__ __forceinline__ uint64_t slot_key(const uint4& s) { return ((uint64_t)s.y << 32) | s.x; }

// read-only 128-bit load through the non-coherent path (the table is immutable after load_index)
#ifdef DBT_PROBE_LDCG
__device__ __forceinline__ uint4 ldg128(const uint4* p) { return __ldcg(p); }   // L2 only: random probe sectors do not displace the L1 lines of the stack
#else
__device__ __forceinline__ uint4 ldg128(const uint4* p) { return __ldg(p); }
#endif

// One 32-byte bucket half (two slots) in ONE 256-bit load (LDG.E.256, new on sm_100): one request to L2 per bucket instead of two.
// Random gathers issued as one request per 32-byte sector run at 45 G/s on B200, as two 16-byte requests of one thread at 28 G/s
// (profiles/r2/k4_rrand_sweep.json).  p must be 32-byte aligned (buckets are: B * 16 bytes from a 256-byte aligned base).
__device__ __forceinline__ void ldg256(const uint4* p, uint4& a, uint4& b) {
	uint64_t r0, r1, r2, r3;
#ifdef DBT_PROBE_LDCG
	asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r0), "=l"(r1), "=l"(r2), "=l"(r3) : "l"(p));
#else
	asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r0), "=l"(r1), "=l"(r2), "=l"(r3) : "l"(p));
#endif
	a = make_uint4((uint32_t)r0, (uint32_t)(r0 >> 32), (uint32_t)r1, (uint32_t)(r1 >> 32));
	b = make_uint4((uint32_t)r2, (uint32_t)(r2 >> 32), (uint32_t)r3, (uint32_t)(r3 >> 32));
}

// bucket staging: 16-byte asynchronous copies HBM/L2 -> shared memory, so that probe loops can be ROLLED (small code,
// the kernels are instruction-fetch sensitive) and still keep every probe of a read in flight at once
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
	const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
#ifdef DBT_CPASYNC_CA
	asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gsrc) : "memory");   // allocates in L1: the correction code re-visits the read's own nodes
#else
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gsrc) : "memory");
#endif
}
// Stages the home bucket of every lane's key (valid lanes; hb = bucket number < 2^32) into dst[slot][lane], with B ADJACENT LANES
// fetching the B slots of one bucket, so that a bucket is one coalesced request to L2 instead of B 16-byte requests of one lane.
// Random 32-byte sectors: 43-45 G/s as one request, 27.5 G/s as two (profiles/r2/k4_rrand_sweep.json) -- this is what bounds the
// probe rounds when the tables live in HBM.  It costs a ballot and B shuffles per 32 buckets, so kernels that are bound by
// instruction issue with L2-resident tables keep the per-lane form (RunParams::coop_stage).  All 32 lanes must call it.
template <int B>
__device__ __forceinline__ void stage_buckets_coop(const uint4* __restrict__ slots, bool valid, uint32_t hb, uint4 (*dst)[32], uint32_t lane) {
	const uint32_t vm = __ballot_sync(0xffffffffu, valid);
	const uint32_t q = lane % B;
#pragma unroll
	for (int h = 0; h < B; ++h) {
		const uint32_t src = h * (32 / B) + lane / B;
		const uint32_t b = __shfl_sync(0xffffffffu, hb, src);
		if ((vm >> src) & 1u) cp_async16(&dst[q][src], slots + (uint64_t)b * B + q);
	}
}
__device__ __forceinline__ void cp_async_wait_all() {
	asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// presence bitmap (DevIndex::pre): bit of a key
__host__ __device__ __forceinline__ uint32_t pre_bit(uint64_t key, uint32_t bits) { return (uint32_t)((key * 0xA24BAED4963EE407ull) >> (64 - bits)); }

// Looks key up. Returns true on hit and sets val/aux. key must be < 2^63 and != invalid.
// Index value word (slot .z): even = locus << 1 for a k-mer unique to one locus, odd = (offset into vv) << 1 | 1.  When the de Bruijn
// graphs are merged into the index (dbt_load_index, graph_merge_kernel) the top ten bits of an EVEN value carry the graph entry of
// (k-mer, that locus): present + out-edge mask of the canonical orientation (bits 22-26) and of its reverse complement (27-31).
constexpr uint32_t VAL_LOCUS_BITS = 22;
constexpr uint32_t VAL_LOCUS_MASK = (1u << VAL_LOCUS_BITS) - 1;
__host__ __device__ __forceinline__ uint32_t slot_val(uint32_t z) { return (z & 1u) ? z : (z & VAL_LOCUS_MASK); }

template <int B>
__device__ __forceinline__ bool table_lookup(const Table& t, uint64_t key, uint32_t& val, uint32_t& aux) {
	uint64_t b = home_bucket(t, key);
	bool home = true;
	for (;;) {
		const uint4* bp = t.slots + b * B;
		uint4 s[B];
#pragma unroll
		for (int i = 0; i < B; i += 2) ldg256(bp + i, s[i], s[i + 1]);
		bool full = true;
#pragma unroll
		for (int i = 0; i < B; ++i) {
			uint64_t k = slot_key(s[i]);
			if (k == SLOT_EMPTY) full = false;
			else if ((k & SLOT_KEYMASK) == key) { val = slot_val(s[i].z); aux = s[i].w; return true; }
		}
		if (!full || !(slot_key(s[0]) & SLOT_OVF)) return false;
		b = home ? stash_bucket(t, key) : stash_next(t, b);
		home = false;
	}
}

// Same walk, but returns the slot address (used by the aux-fill kernels at load time).
template <int B>
__device__ __forceinline__ uint4* table_find_slot(const Table& t, uint64_t key) {
	uint64_t b = home_bucket(t, key);
	bool home = true;
	for (;;) {
		uint4* bp = t.slots + b * B;
		bool full = true;
		uint64_t k0 = 0;
#pragma unroll
		for (int i = 0; i < B; ++i) {
			uint64_t k = *reinterpret_cast<volatile uint64_t*>(bp + i);
			if (i == 0) k0 = k;
			if (k == SLOT_EMPTY) full = false;
			else if ((k & SLOT_KEYMASK) == key) return bp + i;
		}
		if (!full || !(k0 & SLOT_OVF)) return nullptr;
		b = home ? stash_bucket(t, key) : stash_next(t, b);
		home = false;
	}
}

// Inserts (key,val); aux starts at 0. Returns 0 ok, 1 duplicate key, 2 table full.
template <int B>
__device__ __forceinline__ int table_insert(const Table& t, uint64_t key, uint32_t val) {
	uint64_t b = home_bucket(t, key);
	bool home = true;
	for (uint64_t tries = 0; tries <= t.nstash; ++tries) {
		uint4* bp = t.slots + b * B;
#pragma unroll
		for (int i = 0; i < B; ++i) {
			unsigned long long* kp = reinterpret_cast<unsigned long long*>(bp + i);
			unsigned long long old = atomicCAS(kp, (unsigned long long)SLOT_EMPTY, (unsigned long long)key);
			if (old == SLOT_EMPTY) {
				uint32_t* w = reinterpret_cast<uint32_t*>(bp + i);
				w[2] = val; w[3] = 0;
				return 0;
			}
			if ((old & SLOT_KEYMASK) == key) return 1;
		}
		// bucket full: leave the overflow mark so lookups continue past it
		atomicOr(reinterpret_cast<unsigned long long*>(bp), (unsigned long long)SLOT_OVF);
		b = home ? stash_bucket(t, key) : stash_next(t, b);
		home = false;
	}
	return 2;
}

}  // namespace dbt
