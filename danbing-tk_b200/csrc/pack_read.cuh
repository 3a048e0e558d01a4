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

// ---- fast path: a read that holds nothing but upper-case ACGT (almost every read) ----
// Four bytes per step, ~half the instructions of the exact path: the codes are computed unconditionally, validity is only
// ACCUMULATED (any byte outside the alphabet sets a bit of `bad`), and the mask words follow from the length alone.
//   byte in {A 0x41, C 0x43, G 0x47, T 0x54}  <=>  bits 7,5,3 clear, bit 6 set, and (b4 ? b2 & ~b1 & ~b0 : b0 & (b1 | ~b2))
DBT_HD uint32_t pack4_fast(uint32_t x, uint32_t& bad) {
	const uint32_t t1 = x >> 1, t2 = x >> 2, t4 = x >> 4;
	const uint32_t q = x & (t1 | ~t2), r = t2 & ~t1 & ~x;
	const uint32_t g = (t4 & r) | (~t4 & q);                                       // bit 0 of every byte: the relation holds
	bad |= ((x & 0xE8E8E8E8u) ^ 0x40404040u) | (~g & 0x01010101u);
	return (((t1 ^ t2) & 0x03030303u) * 0x01041040u);                              // 4 x 2-bit codes in the TOP byte
}

// byte 3 of `prod` into byte j of `acc` (acc's byte j is still zero): PRMT on the device; the host build emulates the same
// selectors so that tests/native/pack_check.cpp covers them
DBT_HD uint32_t byte_perm(uint32_t x, uint32_t y, uint32_t sel) {
#ifdef __CUDA_ARCH__
	return __byte_perm(x, y, sel);
#else
	const uint64_t v = ((uint64_t)y << 32) | x;
	uint32_t r = 0;
	for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
	return r;
#endif
}
template <int J>
DBT_HD uint32_t put_top_byte(uint32_t acc, uint32_t prod) {
	return byte_perm(acc, prod, J == 0 ? 0x3217u : J == 1 ? 0x3270u : J == 2 ? 0x3710u : 0x7210u);
}

// Same arguments as pack_read_swar.  Returns false -- with orec in an unspecified state -- when the read holds a byte outside the
// alphabet (the caller then runs the exact path); true with the finished record and meta word otherwise.
DBT_HD bool pack_read_fast(const uint32_t* sw, uint32_t sh, uint32_t len, uint32_t k, ReadLayout lay, uint64_t* orec, uint16_t* meta) {
	const uint32_t WB = lay.WB, WM = lay.WM;
	const uint32_t nfull = len >> 2, rem = len & 3u, nwords = nfull + (rem ? 1u : 0u);
	const uint32_t keep = rem ? (0xFFFFFFFFu >> (8 * (4 - rem))) : 0xFFFFFFFFu;     // bytes of the last, partial word that belong to the read
	uint32_t lo = sw[0], bad = 0;
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
	for (uint32_t w = 0; w < WB; ++w) {                                            // 32 bases per round = one base word
		uint32_t a0 = 0, a1 = 0;
		const uint32_t w0 = w * 8;
		if (w0 + 8 <= nfull) {                                                     // the round lies inside the read: no guards
#define DBT_STEP(J, ACC) { const uint32_t hi = sw[w0 + J + 1]; const uint32_t x = funnel_r(lo, hi, sh); lo = hi; ACC = put_top_byte<(J) & 3>(ACC, pack4_fast(x, bad)); }
			DBT_STEP(0, a0) DBT_STEP(1, a0) DBT_STEP(2, a0) DBT_STEP(3, a0) DBT_STEP(4, a1) DBT_STEP(5, a1) DBT_STEP(6, a1) DBT_STEP(7, a1)
#undef DBT_STEP
		} else if (w0 < nwords) {                                                  // the round the read ends in
#define DBT_STEP(J, ACC) if (w0 + J < nwords) { const uint32_t hi = sw[w0 + J + 1]; uint32_t x = funnel_r(lo, hi, sh); lo = hi; \
				if (w0 + J == nfull) x = (x & keep) | (0x41414141u & ~keep);               /* past the end: 'A' = code 0, valid */ \
				ACC = put_top_byte<(J) & 3>(ACC, pack4_fast(x, bad)); }
			DBT_STEP(0, a0) DBT_STEP(1, a0) DBT_STEP(2, a0) DBT_STEP(3, a0) DBT_STEP(4, a1) DBT_STEP(5, a1) DBT_STEP(6, a1) DBT_STEP(7, a1)
#undef DBT_STEP
		}
		orec[w] = (uint64_t)a0 | ((uint64_t)a1 << 32);
	}
	if (bad) return false;
	for (uint32_t m = 0; m < WM; ++m)                                              // invalid mask = the positions past the end
		orec[WB + m] = len >= 64 * (m + 1) ? 0ull : (len <= 64 * m ? ~0ull : ~0ull << (len - 64 * m));
	for (uint32_t w = WB + WM; w < lay.W64; ++w) orec[w] = 0;
	*meta = (uint16_t)(len | (len >= k ? 0 : META_NOVALID));
	return true;
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
