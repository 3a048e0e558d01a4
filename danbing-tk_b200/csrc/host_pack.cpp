// Host side of the read batcher (BASELINE.json north_star item 1): ASCII reads -> the device's packed read records
// (kmer.cuh: WB words of 2-bit codes in little-endian base order, WM words of invalid mask, 16-byte aligned) so that
// the H2D copy moves 64 B per 150 bp read instead of 150 B + offsets.  Semantics = encodeSeq / the alphabet test of
// read2kmers_edges (reference src/aQueryFasta_thread.h:52,126-132,297): only upper-case ACGT is a base, anything else
// (and every position >= len) is flagged invalid.  Compiled by g++ (AVX2 path chosen at run time, SWAR otherwise).
#include <stdint.h>
#include <string.h>
#include <immintrin.h>
#include "kmer.cuh"
#include "host_pack.h"

namespace dbt {

// 8 ASCII bytes -> 16 bits of codes + 8 invalid bits, without SIMD
static inline void pack8_swar(uint64_t x, uint32_t& codes, uint32_t& inval) {
	// code = ((c >> 1) ^ (c >> 2)) & 3 maps A C G T -> 0 1 2 3
	uint64_t c = ((x >> 1) ^ (x >> 2)) & 0x0303030303030303ull;
	c = (c | (c >> 6)) & 0x000F000F000F000Full;
	c = (c | (c >> 12)) & 0x000000FF000000FFull;
	codes = (uint32_t)((c | (c >> 24)) & 0xFFFF);
	uint32_t bad = 0;
	for (int i = 0; i < 8; ++i) { const uint8_t b = (uint8_t)(x >> (8 * i)); bad |= (uint32_t)(b != 'A' && b != 'C' && b != 'G' && b != 'T') << i; }
	inval = bad;
}

static void pack_swar(const uint8_t* s, uint32_t len, const ReadLayout& lay, uint64_t* out) {
	for (uint32_t w = 0; w < lay.W64; ++w) out[w] = 0;
	const uint32_t nbits = lay.WM * 64;
	for (uint32_t i = 0; i < nbits; i += 8) {
		uint64_t x = 0;
		if (i + 8 <= len) memcpy(&x, s + i, 8); else if (i < len) memcpy(&x, s + i, len - i);
		uint32_t codes, inval; pack8_swar(x, codes, inval);
		if (i + 8 > len) { const uint32_t keep = i < len ? len - i : 0; inval |= 0xFFu << keep; inval &= 0xFF; codes &= (1u << 2 * keep) - 1; }
		// invalid positions carry code 0 (the device packer writes nothing there)
		uint32_t spread = 0; for (int b = 0; b < 8; ++b) if (!((inval >> b) & 1u)) spread |= 3u << (2 * b);
		codes &= spread;
		if ((i >> 5) < lay.WB) out[i >> 5] |= (uint64_t)codes << (2 * (i & 31));
		out[lay.WB + (i >> 6)] |= (uint64_t)inval << (i & 63);
	}
}

// 32 ASCII bytes -> 64 bits of codes + 32 invalid bits
__attribute__((target("avx2"))) static inline void pack32_avx2(__m256i x, uint64_t& codes, uint32_t& inval) {
	// validity: table[low nibble] holds the only upper-case base with that nibble (A 0x41, C 0x43, T 0x54, G 0x47)
	const __m256i tbl = _mm256_setr_epi8((char)0xFF, 0x41, (char)0xFF, 0x43, 0x54, (char)0xFF, (char)0xFF, 0x47, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, 0x00,
	                                     (char)0xFF, 0x41, (char)0xFF, 0x43, 0x54, (char)0xFF, (char)0xFF, 0x47, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, (char)0xFF, 0x00);
	const __m256i lo = _mm256_and_si256(x, _mm256_set1_epi8(0x0F));
	const __m256i ok = _mm256_cmpeq_epi8(_mm256_shuffle_epi8(tbl, lo), x);
	inval = ~(uint32_t)_mm256_movemask_epi8(ok);
	// code = ((c >> 1) ^ (c >> 2)) & 3, zeroed where invalid
	const __m256i s1 = _mm256_srli_epi16(x, 1), s2 = _mm256_srli_epi16(x, 2);
	__m256i c = _mm256_and_si256(_mm256_xor_si256(s1, s2), _mm256_set1_epi8(3));
	c = _mm256_and_si256(c, ok);
	// 4 codes per byte: (c0 + 4 c1) per 16-bit lane, then (lo + 16 hi) per 32-bit lane
	const __m256i p16 = _mm256_maddubs_epi16(c, _mm256_set1_epi16(0x0401));
	const __m256i p32 = _mm256_madd_epi16(p16, _mm256_set1_epi32(0x00100001));
	// gather the low byte of each 32-bit lane: 4 bytes per 128-bit half
	const __m256i sh = _mm256_setr_epi8(0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, 0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
	const __m256i g = _mm256_shuffle_epi8(p32, sh);
	const uint32_t a = (uint32_t)_mm256_extract_epi32(g, 0), b = (uint32_t)_mm256_extract_epi32(g, 4);
	codes = (uint64_t)a | ((uint64_t)b << 32);
}

// avail = bytes that may be read starting at s (the rest of the caller's buffer): full 32-byte loads run past the read
// into its neighbour wherever that is safe, the bytes >= len are masked off afterwards
__attribute__((target("avx2"))) static void pack_avx2(const uint8_t* s, uint32_t len, uint64_t avail, const ReadLayout& lay, uint64_t* out) {
	const uint32_t nbits = lay.WM * 64;
	uint64_t mask = 0;
	for (uint32_t i = 0; i < nbits; i += 32) {
		uint64_t codes = 0; uint32_t inval = 0xFFFFFFFFu;
		if (i < len) {
			__m256i x;
			if (i + 32 <= avail) x = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(s + i));
			else { alignas(32) uint8_t tmp[32] = {0}; memcpy(tmp, s + i, len - i); x = _mm256_load_si256(reinterpret_cast<const __m256i*>(tmp)); }
			pack32_avx2(x, codes, inval);
			if (i + 32 > len) { const uint32_t keep = len - i; inval |= 0xFFFFFFFFu << keep; codes &= (1ull << 2 * keep) - 1; }
		}
		if ((i >> 5) < lay.WB) out[i >> 5] = codes;
		mask |= (uint64_t)inval << (i & 63);
		if ((i & 63) == 32 || i + 32 >= nbits) { out[lay.WB + (i >> 6)] = mask; mask = 0; }
	}
	for (uint32_t w = lay.WB + lay.WM; w < lay.W64; ++w) out[w] = 0;
}

// is there a window of k consecutive valid bases?  (read2kmers_edges leaves the vectors empty otherwise, h:282)
static bool any_valid_window(const uint64_t* mask, uint32_t WM, uint32_t len, uint32_t k) {
	if (len < k) return false;
	bool clean = true;
	for (uint32_t w = 0; w < WM && clean; ++w) {
		const uint32_t lo = w * 64;
		if (lo >= len) break;
		const uint64_t in = len - lo >= 64 ? ~0ull : ((1ull << (len - lo)) - 1);
		if (mask[w] & in) clean = false;
	}
	if (clean) return true;
	uint32_t run = 0;
	for (uint32_t i = 0; i < len; ++i) {
		if ((mask[i >> 6] >> (i & 63)) & 1ull) run = 0;
		else if (++run >= k) return true;
	}
	return false;
}

static int g_have_avx2 = -1;

void host_pack_read(const uint8_t* s, uint32_t len, uint64_t avail, uint32_t k, const ReadLayout& lay, uint64_t* out, uint16_t* meta, int impl) {
	if (g_have_avx2 < 0) g_have_avx2 = __builtin_cpu_supports("avx2") ? 1 : 0;
	if (impl != HOST_PACK_SWAR && g_have_avx2) pack_avx2(s, len, avail, lay, out); else pack_swar(s, len, lay, out);
	*meta = (uint16_t)(len | (any_valid_window(out + lay.WB, lay.WM, len, k) ? 0 : META_NOVALID));
}

}  // namespace dbt
