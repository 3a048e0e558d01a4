// One read, ASCII -> packed record (kmer.cuh layout), 32-bit SWAR: the per-lane body of pack_kernel (kernels.cu).
// Host-compilable so that tests/native/pack_check.cpp can run exactly this code against the host packer.
// Alphabet = exactly the upper-case "ACGT" of the reference (h:52,297); code = encodeSeq's A0 C1 G2 T3 (h:126-132).
#pragma once
#include <stdint.h>
#include "kmer.cuh"

namespace dbt {

// four ASCII bytes (byte 0 = first base) -> v: 0x01 in every byte that is A/C/G/T; codes: the 2-bit code in every valid byte
DBT_HD void pack4_swar(uint32_t x, uint32_t& codes, uint32_t& v) {
	// y = x ^ '@': A 0x01, C 0x03, G 0x07, T 0x14.  Valid <=> bits 7,6,5,3 clear and (b4 b2 b1 b0) in {0001, 0011, 0111, 1100}
	const uint32_t y = x ^ 0x40404040u;
	const uint32_t b1 = y >> 1, b2 = y >> 2, b4 = y >> 4;
	const uint32_t hi = (y >> 7) | (y >> 6) | (y >> 5) | (y >> 3);
	const uint32_t acg = ~b4 & y & (b1 | ~b2), t = b4 & b2 & ~b1 & ~y;
	v = (acg | t) & ~hi & 0x01010101u;
	codes = ((x >> 1) ^ (x >> 2)) & (v * 3u);
}

DBT_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {
#ifdef __CUDA_ARCH__
	return __funnelshift_r(lo, hi, sh);
#else
	return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}

// sw: 4-byte-aligned words holding the read from byte sh/8 of sw[0] on (words up to the one after the read's last byte are
// readable); orec: W64 output words.  Returns the meta word (length | "no valid k-mer window").
DBT_HD uint16_t pack_read_swar(const uint32_t* sw, uint32_t sh, uint32_t len, uint32_t k, ReadLayout lay, uint64_t* orec) {
	const uint32_t WB = lay.WB, WM = lay.WM;
	uint32_t lo = sw[0];
	uint64_t mw = 0; bool clean = true;
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
	for (uint32_t w = 0; w < 2 * WM; ++w) {                                        // 32 bases per round: one base word, half a mask word
		uint64_t bw = 0; uint32_t vb = 0;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
		for (uint32_t j = 0; j < 8; ++j) {
			const uint32_t i0 = w * 32 + j * 4;
			uint32_t x = 0;
			if (i0 < len) {
				const uint32_t hi = sw[w * 8 + j + 1];
				x = funnel_r(lo, hi, sh); lo = hi;
				if (len - i0 < 4) x &= 0xFFFFFFFFu >> (8 * (4 - (len - i0)));
			}
			uint32_t codes, v; pack4_swar(x, codes, v);
			bw |= (uint64_t)((codes * 0x01041040u) >> 24) << (8 * j);              // 4 x 2-bit codes -> one byte
			vb |= ((v * 0x01020408u) >> 24) << (4 * j);                            // 4 valid bits -> one nibble
		}
		if (w < WB) orec[w] = bw;
		const uint32_t nin = len > w * 32 ? (len - w * 32 < 32u ? len - w * 32 : 32u) : 0u;   // bases of this round inside the read
		if (nin && (vb | (nin == 32 ? 0u : 0xFFFFFFFFu << nin)) != 0xFFFFFFFFu) clean = false;
		mw |= (uint64_t)(~vb) << (32 * (w & 1));
		if (w & 1) { orec[WB + (w >> 1)] = mw; mw = 0; }
	}
	for (uint32_t w = WB + WM; w < lay.W64; ++w) orec[w] = 0;
	// is there a window of k consecutive valid bases?  (read2kmers_edges leaves the vectors empty otherwise, h:282)
	bool any = len >= k;
	if (any && !clean) {
		any = false;
		uint64_t prev = 0;                                                         // valid bits of the previous 64 bases
		for (uint32_t m = 0; m < WM && !any; ++m) {
			const uint64_t cur = ~orec[WB + m];
			// runs of k ones starting anywhere in (cur:prev): AND of k shifted copies of the 128-bit value, log steps (k <= 32)
			uint64_t rl = prev, rh = cur; uint32_t have = 1;
			while (have * 2 <= k) { rl &= (rl >> have) | (rh << (64 - have)); rh &= rh >> have; have *= 2; }
			if (have < k) { const uint32_t d = k - have; rl &= (rl >> d) | (rh << (64 - d)); rh &= rh >> d; }
			any = (rh | rl) != 0;
			prev = cur;
		}
	}
	return (uint16_t)(len | (any ? 0 : META_NOVALID));
}

}  // namespace dbt
