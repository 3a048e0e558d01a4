// GPU-resident de Bruijn graphs of all loci (SURVEY.md §8(a) type Y4): ONE open-addressing table in HBM.
//
// What it replaces in the reference:
//   graphDB[l]   unordered_map<size_t,uint8_t>   strand-specific k-mer -> 4-bit out-edge mask   (h:32, cpp:2468)
//   trResults[l].count(toCaKmer(kmer))            TR membership used for the `tr` annotation       (cpp:885, 1175)
//   trResults[l].find(canonical k-mer)            count increments of the threaded k-mers         (cpp:2195-2209, D1)
// fa2kmers stores both strands of every haplotype (buildKmerGraph, h:239-243), so a node and its reverse
// complement always come as a pair: the table is keyed by (CANONICAL k-mer, locus) and one 16-byte entry carries
// the masks of both orientations, which halves the footprint (L2 residency) and gives errorCorrection_backward
// (cpp:1091-1106, getOutNodes_rc) its reverse-strand mask from the probe the forward walk already made.
//
// Slot (16 B): { u64 key | bit63 = bucket-overflow flag (slot 0 only) ; u32 lw ; u32 tr }
//   lw : locus (bits 0-21) | canonical orientation: present (22), out-edge mask (23-26)
//        | rc orientation: present (27), out-edge mask (28-31)
//        -- the two 5-bit fields are what a probe hands out (GNode.bits), swapped for a k-mer met in its rc orientation
//   tr : position in the output-ordered count array + 1 when the canonical k-mer is a TR k-mer of the locus, else 0.
//        TR k-mers that are not graph nodes get an entry with both present bits clear, so that the annotation
//        and the counting stay exact for arbitrary input files.
// Buckets, home/stash addressing and the overflow flag are those of table.cuh (32 B bucket = one L2 sector); the
// hash mixes the locus in so that a k-mer shared by several loci spreads over independent buckets.
#pragma once
#include "table.cuh"
#include "kmer.cuh"

namespace dbt {

constexpr uint32_t G_LOCUS_BITS = 22;
constexpr uint32_t G_LOCUS_MASK = (1u << G_LOCUS_BITS) - 1;
constexpr uint32_t G_PRESENT_CA = 1u << 22, G_PRESENT_RC = 1u << 27;
constexpr uint32_t G_MASK_CA_SHIFT = 23, G_MASK_RC_SHIFT = 28;

// what one probe tells about a STRAND-SPECIFIC k-mer
struct GNode {
	uint32_t bits;    // bit0 present; bits 1-4 out-edge mask of this orientation; bit 5 opposite orientation present;
	                  // bits 6-9 mask of the opposite orientation; bit 10 canonical k-mer is a TR k-mer of the locus
	uint32_t tr;      // count-array position + 1, or 0
};
constexpr uint32_t GN_PRESENT = 1u, GN_RCPRESENT = 1u << 5, GN_TR = 1u << 10;
__device__ __forceinline__ uint32_t gn_mask(uint32_t bits) { return (bits >> 1) & 0xFu; }
__device__ __forceinline__ uint32_t gn_rcmask(uint32_t bits) { return (bits >> 6) & 0xFu; }

__device__ __forceinline__ uint64_t g_hash_key(uint64_t canon, uint32_t locus) {
	return canon ^ ((uint64_t)locus * 0xC2B2AE3D27D4EB4Full);
}

// reverse complement of a k-mer in the reference's encoding (first base most significant): getNuRC, h:165-178
__device__ __forceinline__ uint64_t kmer_rc(uint64_t x, uint32_t k) {
	return rev2(~x) >> (64 - 2 * k);
}

__device__ __forceinline__ GNode g_decode(uint32_t lw, uint32_t tr, bool is_rc) {
	// x = [canonical orientation: present, mask | rc orientation: present, mask]; met in rc orientation: swap the halves
	uint32_t x = lw >> 22;
	if (is_rc) x = ((x & 31u) << 5) | (x >> 5);
	GNode n;
	n.bits = x | (tr ? GN_TR : 0u);
	n.tr = tr;
	return n;
}

// Looks the strand-specific k-mer `fw` (not KMER_INVALID) of `locus` up.
template <int B>
__device__ __forceinline__ GNode g_lookup(const Table& t, uint64_t fw, uint32_t locus, uint32_t k) {
	const uint64_t rc = kmer_rc(fw, k);
	const bool is_rc = rc < fw;
	const uint64_t canon = is_rc ? rc : fw;
	const uint64_t hk = g_hash_key(canon, locus);
	uint64_t b = home_bucket(t, hk);
	bool home = true;
	for (;;) {
		const uint4* bp = t.slots + b * B;
		// named registers, not an array: inside divergent callers nvcc kept `uint4 s[B]` in local memory (ncu: 50 M LDL/STL)
		uint4 s0, s1;
		ldg256(bp, s0, s1);
		uint4 s2 = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0, 0), s3 = s2;
		if (B == 4) ldg256(bp + 2, s2, s3);
		const uint64_t k0 = slot_key(s0), k1 = slot_key(s1), k2 = slot_key(s2), k3 = slot_key(s3);
		if ((k0 & SLOT_KEYMASK) == canon && k0 != SLOT_EMPTY && (s0.z & G_LOCUS_MASK) == locus) return g_decode(s0.z, s0.w, is_rc);
		if ((k1 & SLOT_KEYMASK) == canon && k1 != SLOT_EMPTY && (s1.z & G_LOCUS_MASK) == locus) return g_decode(s1.z, s1.w, is_rc);
		if (B == 4) {
			if ((k2 & SLOT_KEYMASK) == canon && k2 != SLOT_EMPTY && (s2.z & G_LOCUS_MASK) == locus) return g_decode(s2.z, s2.w, is_rc);
			if ((k3 & SLOT_KEYMASK) == canon && k3 != SLOT_EMPTY && (s3.z & G_LOCUS_MASK) == locus) return g_decode(s3.z, s3.w, is_rc);
		}
		const bool full = k0 != SLOT_EMPTY && k1 != SLOT_EMPTY && (B == 2 || (k2 != SLOT_EMPTY && k3 != SLOT_EMPTY));
		if (!full || !(k0 & SLOT_OVF)) { GNode n; n.bits = 0; n.tr = 0; return n; }
		b = home ? stash_bucket(t, hk) : stash_next(t, b);
		home = false;
	}
}

// ---- build (load time) ------------------------------------------------------------------------------------
// Finds the slot of (canon, locus) or nullptr.  Only called when no other thread can be inserting THIS pair.
template <int B>
__device__ __forceinline__ uint4* g_find_slot(const Table& t, uint64_t canon, uint32_t locus) {
	const uint64_t hk = g_hash_key(canon, locus);
	uint64_t b = home_bucket(t, hk);
	bool home = true;
	for (;;) {
		uint4* bp = t.slots + b * B;
		bool full = true;
		uint64_t k0 = 0;
#pragma unroll
		for (int i = 0; i < B; ++i) {
			const uint64_t kk = *reinterpret_cast<volatile uint64_t*>(bp + i);
			if (i == 0) k0 = kk;
			if (kk == SLOT_EMPTY) full = false;
			else if ((kk & SLOT_KEYMASK) == canon) {
				const uint32_t lw = *(reinterpret_cast<volatile uint32_t*>(bp + i) + 2);
				if ((lw & G_LOCUS_MASK) == locus) return bp + i;
			}
		}
		if (!full || !(k0 & SLOT_OVF)) return nullptr;
		b = home ? stash_bucket(t, hk) : stash_next(t, b);
		home = false;
	}
}

// Claims a fresh slot for (canon, locus) -- the pair must not be in the table and no other thread may insert it
// concurrently -- and publishes lw/tr.  Returns false when the stash is exhausted.
template <int B>
__device__ __forceinline__ bool g_claim(const Table& t, uint64_t canon, uint32_t locus, uint32_t lw, uint32_t tr) {
	const uint64_t hk = g_hash_key(canon, locus);
	uint64_t b = home_bucket(t, hk);
	bool home = true;
	for (uint64_t tries = 0; tries <= t.nstash; ++tries) {
		uint4* bp = t.slots + b * B;
#pragma unroll
		for (int i = 0; i < B; ++i) {
			unsigned long long* kp = reinterpret_cast<unsigned long long*>(bp + i);
			// the locus word is written BEFORE the key becomes visible is impossible with one CAS; instead a reader that
			// sees the key with a stale locus word (0xFFFFFFFF & mask = all ones) cannot mistake it for its own locus
			if (*reinterpret_cast<volatile unsigned long long*>(kp) == SLOT_EMPTY) {
				const unsigned long long old = atomicCAS(kp, (unsigned long long)SLOT_EMPTY, (unsigned long long)canon);
				if (old == SLOT_EMPTY) {
					uint32_t* w = reinterpret_cast<uint32_t*>(bp + i);
					w[3] = tr;
					__threadfence();
					*reinterpret_cast<volatile uint32_t*>(w + 2) = lw;
					return true;
				}
			}
		}
		atomicOr(reinterpret_cast<unsigned long long*>(bp), (unsigned long long)SLOT_OVF);
		b = home ? stash_bucket(t, hk) : stash_next(t, b);
		home = false;
	}
	return false;
}

}  // namespace dbt
