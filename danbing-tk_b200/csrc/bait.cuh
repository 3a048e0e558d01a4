// `-b`: the bait filter of the pair loop, bfilter_FPSv1 (src/aQueryFasta_thread.cpp:1377-1396), FASTA form.
//
// baitDB[locus] (PREF.bt.kmdb / PREF.qc.bt.kmdb, readBinaryBaitDB h:542-547) maps a canonical k-mer to (min << 8 | max): the range of
// occurrences seen in true-positive reads.  For a mate, the reference builds the histogram of its canonical k-mers (kc8_t: 8-bit
// counters that wrap) and raises bf when a k-mer that is a bait k-mer of the voted locus occurs fewer than min or more than max
// times; either mate raising it removes the pair (cpp:2119-2124).  Which k-mer it meets first matters only for `-tb`, which the
// host replays (host/bait.hpp).  On the device all baits of all loci sit in one table keyed (k-mer, locus), the bucket / stash scheme
// of graph_table.cuh with the bounds in the payload word; lanes take the positions of the mate, probe, and only a hit counts its
// occurrences (straight from the packed words: hits are rare).
#pragma once
#include "graph_table.cuh"

namespace dbt {

// qString2qMask (src/aQueryFasta_thread.h:1038-1071), on ONE lane: which k-mers enter the histogram of a FASTQ read.  A k-mer
// counts when all its k bases have phred >= qth -- as far as the reference's scan gets: it walks the bases with one index that it
// compares with the NUMBER OF K-MERS nk = nq - k + 1, so the scan stops k-1 bases before the end of the read and k-mers that
// start after the last scanned stretch stay masked out.  mask: bit i of word i/32 = k-mer i counts (nw words, zeroed here).
__device__ __forceinline__ void quality_mask(const uint8_t* __restrict__ q, int nq, int K, int qth, uint32_t* mask, int nw) {
	for (int w = 0; w < nw; ++w) mask[w] = 0;
	const int nk = nq - K + 1;
	if (nk < 1) return;
	auto lowq = [&](int i) { return (int)q[i] - 33 < qth; };
	auto set = [&](int i) { if ((i >> 5) < nw) mask[i >> 5] |= 1u << (i & 31); };
	int qi = 0, ki = 0;
	while (lowq(qi)) { ++qi; ++ki; if (qi >= nk) return; }
	while (qi < nk) {
		bool pass = true;
		for (const int qj = qi; qi < qj + K; ++qi) {
			if (lowq(qi)) {
				pass = false; ki = qi;
				while (lowq(qi)) { ++qi; ++ki; if (qi >= nk) return; }
				break;
			}
		}
		if (pass) {                                           // qi sits just behind the k-mer that starts at ki
			set(ki); ++ki;
			if (qi >= nk) return;
			while (!lowq(qi)) { set(ki); ++qi; ++ki; if (qi >= nk) return; }
			ki = qi;
			while (lowq(qi)) { ++qi; ++ki; if (qi >= nk) return; }
		}
	}
}

// rec: packed record of the mate (WB base words, then the invalid-mask words) in shared or global memory
template <int B>
// qmask: nullptr (FASTA: every k-mer counts, cpp:1377-1396) or the quality mask of the mate (cpp:1398-1419)
__device__ __noinline__ bool bait_filtered(const Table bt, const uint64_t* rec, uint32_t WB, uint32_t nkm, uint32_t K, uint32_t dst, int lane, const uint32_t* qmask) {
	bool viol = false;
	for (uint32_t i = (uint32_t)lane; i < nkm; i += 32) {
		if (qmask && !((qmask[i >> 5] >> (i & 31)) & 1u)) continue;
		const uint64_t kk = canonical_kmer_at(rec, WB, i, K);
		if (kk == KMER_INVALID) continue;                     // (counted under the key -1 by the reference; -1 is no bait)
		const uint4* s = g_find_slot<B>(bt, kk, dst);
		if (!s) continue;
		const uint32_t th = reinterpret_cast<const uint32_t*>(s)[3], mi = (th >> 8) & 0xFFu, ma = th & 0xFFu;
		uint32_t cnt = 0;
		for (uint32_t q = 0; q < nkm; ++q) cnt += (!qmask || ((qmask[q >> 5] >> (q & 31)) & 1u)) && canonical_kmer_at(rec, WB, q, K) == kk;
		cnt &= 0xFFu;                                          // kc8_t: uint8_t counters
		if (cnt < mi || cnt > ma) viol = true;
	}
	return __any_sync(0xffffffffu, viol);
}

// load time: one thread per bait k-mer
template <int B>
__global__ void bait_insert_kernel(Table t, const uint64_t* __restrict__ kmers, const uint16_t* __restrict__ vals, const uint32_t* __restrict__ locus_of, uint64_t n, uint32_t* err) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (!g_claim<B>(t, kmers[i], locus_of[i], locus_of[i] | G_PRESENT_CA, (uint32_t)vals[i])) atomicOr(err, ERR_INDEX_FULL);
}

}  // namespace dbt
