// GPU-side FASTA/FASTQ ingestion (the device half of the reader/pairer A1, src/aQueryFasta_thread.cpp:1918-1976).
//
// The reference reads a record as exactly L lines (FASTA: title, sequence; FASTQ: title, sequence, '+', quality -- four
// getline calls, cpp:1920-1925 / 1948-1950; it never looks at the '>' / '@'), strips a trailing "/1" or "/2" from the title
// (prunePEinfo, cpp:455-462) and pairs records by title through a hash map; the mate that arrives second becomes read 1
// (cpp:1972-1974) and a pair with a mate shorter than Cthreshold + k - 1 is forgotten (cpp:1957).  In a name-collated file
// the two mates are neighbours, so on the raw text of a batch (whole records, one H2D copy):
//
//   nl_mask_kernel     one bit per byte: is it '\n'          (128-bit coalesced loads, one 32-bit word per thread)
//   compact_*          bitmap -> ascending newline offsets   (the ordered stream compaction of kernels.cu)
//   rec_hash_kernel    per record: pruned title length + FNV-1a hash, sequence length
//   rec_eq_kernel      per record: same pruned title as the record before it (hash, then bytes)
//   rec_class_kernel   pair = a record that equals its predecessor and neither neighbour relation continues the run;
//                      singleton = a record equal to neither neighbour; three or more equal titles in a row raise
//                      DBT_RAW_SLOWPATH (the map semantics then depend on the order of arrival: host pairer)
//   compact_* (x2)     pair bitmap -> late-mate record numbers in stream order (= the reference's pair order); singletons
//   pair_fill_kernel   per pair: [start, end) of both sequences inside the raw text for pack_kernel (read 2i = late mate),
//                      the byte offset of the pair's first record (the host formats kam / aln lines from the raw text) and
//                      the title hash (the host checks it against the stragglers it holds)
//
// Singletons (a mate whose partner is not its neighbour) go back to the host, which keeps the reference's map for them.
#include <algorithm>
#include "device_types.cuh"
#include "launchers.h"

namespace dbt {

__global__ void __launch_bounds__(256) nl_mask_kernel(const uint4* __restrict__ raw, uint64_t nbytes, uint32_t* __restrict__ mask) {
	const uint64_t nwords = (nbytes + 31) / 32;
	for (uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += (uint64_t)gridDim.x * blockDim.x) {
		const uint4 a = raw[2 * w], b = raw[2 * w + 1];                  // the buffer is padded to a multiple of 32 bytes
		const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
		uint32_t bits = 0;
#pragma unroll
		for (int q = 0; q < 8; ++q) {
			const uint32_t x = v[q] ^ 0x0A0A0A0Au;                      // a zero byte where the text has '\n'
			const uint32_t z = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu);   // 0x80 in exactly those bytes
			bits |= ((((z >> 7) * 0x01020408u) >> 24) & 0xFu) << (4 * q);  // gather the four flags into a nibble (bit i = byte i)
		}
		const uint64_t left = nbytes - w * 32;
		if (left < 32) bits &= (1u << left) - 1u;
		mask[w] = bits;
	}
}

constexpr uint64_t FNV_OFF = 0xcbf29ce484222325ull, FNV_PRIME = 0x100000001b3ull;

// hdr: [0] nlines [1] nrecords [2] npairs [3] nsingle [4] short pairs [5] flags [6] max sequence length [7] max record bytes
__global__ void __launch_bounds__(256) rec_hash_kernel(const uint8_t* __restrict__ raw, const uint32_t* __restrict__ nl, uint32_t* __restrict__ hdr,
                                                      uint32_t L, uint32_t max_lines, uint32_t max_recs, uint64_t* __restrict__ hash, uint32_t* __restrict__ tlen, uint32_t* __restrict__ slen) {
	const uint32_t nrec = min(min(hdr[0], max_lines) / L, max_recs);
	if (blockIdx.x == 0 && threadIdx.x == 0) { hdr[1] = nrec; if (hdr[0] > max_lines || hdr[0] / L > max_recs) atomicOr(hdr + 5, DBT_RAW_OVERFLOW); }
	uint32_t mx_s = 0, mx_r = 0;
	for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < nrec; r += gridDim.x * blockDim.x) {
		const uint32_t t0 = r ? nl[L * r - 1] + 1 : 0, t1 = nl[L * r], s1 = nl[L * r + 1], r1 = nl[L * r + L - 1];
		uint32_t n = t1 - t0;
		if (n >= 2 && raw[t1 - 2] == '/' && (raw[t1 - 1] == '1' || raw[t1 - 1] == '2')) n -= 2;   // prunePEinfo
		uint64_t h = FNV_OFF;
		for (uint32_t i = 0; i < n; ++i) { h ^= raw[t0 + i]; h *= FNV_PRIME; }
		hash[r] = h; tlen[r] = n; slen[r] = s1 - t1 - 1;
		mx_s = max(mx_s, s1 - t1 - 1); mx_r = max(mx_r, r1 + 1 - t0);
	}
	mx_s = __reduce_max_sync(0xffffffffu, mx_s); mx_r = __reduce_max_sync(0xffffffffu, mx_r);
	if ((threadIdx.x & 31) == 0) { if (mx_s) atomicMax(hdr + 6, mx_s); if (mx_r) atomicMax(hdr + 7, mx_r); }
}

__global__ void __launch_bounds__(256) rec_eq_kernel(const uint8_t* __restrict__ raw, const uint32_t* __restrict__ nl, const uint32_t* __restrict__ hdr, uint32_t L,
                                                    const uint64_t* __restrict__ hash, const uint32_t* __restrict__ tlen, uint8_t* __restrict__ eq) {
	const uint32_t nrec = hdr[1];
	for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r <= nrec; r += gridDim.x * blockDim.x) {   // eq[nrec] = 0: sentinel
		bool e = false;
		if (r >= 1 && r < nrec && hash[r] == hash[r - 1] && tlen[r] == tlen[r - 1]) {
			const uint32_t a = nl[L * r - 1] + 1, b = r >= 2 ? nl[L * (r - 1) - 1] + 1 : 0, n = tlen[r];
			e = true;
			for (uint32_t i = 0; i < n; ++i) if (raw[a + i] != raw[b + i]) { e = false; break; }
		}
		eq[r] = e ? 1 : 0;
	}
}

// one thread per record, 32 records per bitmap word
__global__ void __launch_bounds__(256) rec_class_kernel(uint32_t* hdr, const uint8_t* __restrict__ eq,
                                                       const uint32_t* __restrict__ slen, uint32_t min_len, uint32_t max_recs, uint32_t* __restrict__ pairmask, uint32_t* __restrict__ singlemask) {
	const uint32_t nrec = hdr[1];
	const uint32_t nw = (max_recs + 31) / 32;                         // every word of the bitmaps is rewritten: the compaction scans all of them
	uint32_t nshort = 0; bool slow = false;
	for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < nw; w += (gridDim.x * blockDim.x) >> 5) {
		const uint32_t r = w * 32 + (threadIdx.x & 31);
		bool pair = false, single = false;
		if (r < nrec) {
			const bool e0 = r ? eq[r - 1] != 0 : false, e1 = eq[r] != 0, e2 = eq[r + 1] != 0;
			if (e1 && (e0 || e2)) slow = true;                       // a run of three or more equal titles
			else if (e1) { if (slen[r] >= min_len && slen[r - 1] >= min_len) pair = true; else ++nshort; }
			else if (!e2) single = true;
		}
		const uint32_t pm = __ballot_sync(0xffffffffu, pair), sm = __ballot_sync(0xffffffffu, single);
		if ((threadIdx.x & 31) == 0) { pairmask[w] = pm; singlemask[w] = sm; }
	}
	nshort = __reduce_add_sync(0xffffffffu, nshort);
	slow = __any_sync(0xffffffffu, slow);
	if ((threadIdx.x & 31) == 0) { if (nshort) atomicAdd(hdr + 4, nshort); if (slow) atomicOr(hdr + 5, DBT_RAW_SLOWPATH); }
}

__global__ void __launch_bounds__(256) pair_fill_kernel(const uint32_t* __restrict__ nl, uint32_t* __restrict__ hdr, uint32_t L, const uint32_t* __restrict__ pairlist,
                                                       const uint32_t* __restrict__ singlelist, const uint64_t* __restrict__ hash, uint32_t max_pairs, uint32_t max_single,
                                                       uint64_t* __restrict__ rstart, uint64_t* __restrict__ rend, uint64_t* __restrict__ rqual, uint32_t* __restrict__ pairoff,
                                                       uint64_t* __restrict__ pairhash, uint32_t* __restrict__ singleoff, uint32_t check_qual) {
	const uint32_t np = hdr[2], ns = hdr[3];
	if (blockIdx.x == 0 && threadIdx.x == 0 && (np > max_pairs || ns > max_single)) atomicOr(hdr + 5, DBT_RAW_OVERFLOW);
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < min(np, max_pairs); i += gridDim.x * blockDim.x) {
		const uint32_t r = pairlist[i];                              // the late mate; its partner is record r - 1
		rstart[2 * i] = (uint64_t)nl[L * r] + 1;           rend[2 * i] = nl[L * r + 1];
		rstart[2 * i + 1] = (uint64_t)nl[L * (r - 1)] + 1; rend[2 * i + 1] = nl[L * (r - 1) + 1];
		if (L == 4) {
			rqual[2 * i] = (uint64_t)nl[L * r + 2] + 1; rqual[2 * i + 1] = (uint64_t)nl[L * (r - 1) + 2] + 1;
			if (check_qual && (nl[L * r + 3] - nl[L * r + 2] != nl[L * r + 1] - nl[L * r] || nl[L * (r - 1) + 3] - nl[L * (r - 1) + 2] != nl[L * (r - 1) + 1] - nl[L * (r - 1)]))
				atomicOr(hdr + 5, DBT_RAW_SLOWPATH);                     // one phred score per base or the chunk is the host's
		}
		pairoff[i] = r >= 2 ? nl[L * (r - 1) - 1] + 1 : 0;
		pairhash[i] = hash[r];
	}
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < min(ns, max_single); i += gridDim.x * blockDim.x) {
		const uint32_t r = singlelist[i];
		singleoff[i] = r ? nl[L * r - 1] + 1 : 0;
	}
}

cudaError_t launch_fastx_split(cudaStream_t st, int nsm, const RawView& v, uint64_t nbytes, uint32_t L, uint32_t min_len) {
	const unsigned g = (unsigned)nsm * 8;
	cudaError_t e = cudaMemsetAsync(v.hdr, 0, 8 * sizeof(uint32_t), st);
	if (e != cudaSuccess) return e;
	nl_mask_kernel<<<g, 256, 0, st>>>(reinterpret_cast<const uint4*>(v.raw), nbytes, v.nlmask);
	e = launch_compact_bits(st, v.nlmask, nbytes, v.blocksum, v.nl, v.hdr + 0, v.max_lines);
	if (e != cudaSuccess) return e;
	rec_hash_kernel<<<g, 256, 0, st>>>(v.raw, v.nl, v.hdr, L, v.max_lines, v.max_recs, v.hash, v.tlen, v.slen);
	rec_eq_kernel<<<g, 256, 0, st>>>(v.raw, v.nl, v.hdr, L, v.hash, v.tlen, v.eq);
	rec_class_kernel<<<g, 256, 0, st>>>(v.hdr, v.eq, v.slen, min_len, v.max_recs, v.pairmask, v.singlemask);
	e = launch_compact_bits(st, v.pairmask, v.max_recs, v.blocksum, v.pairlist, v.hdr + 2, v.max_recs);
	if (e != cudaSuccess) return e;
	e = launch_compact_bits(st, v.singlemask, v.max_recs, v.blocksum, v.singlelist, v.hdr + 3, v.max_recs);
	if (e != cudaSuccess) return e;
	pair_fill_kernel<<<g, 256, 0, st>>>(v.nl, v.hdr, L, v.pairlist, v.singlelist, v.hash, v.max_pairs, v.max_single, v.rstart, v.rend, v.rqual, v.pairoff, v.pairhash, v.singleoff, v.check_qual);
	return cudaGetLastError();
}

}  // namespace dbt
