// k-mer primitives on 2-bit packed reads.
//
// Reference semantics being reproduced (src/aQueryFasta_thread.h):
//   encodeSeq  h:126-132   k-mer = sum(code[i] << 2*(k-1-i)), A=0 C=1 G=2 T=3, first base most significant
//   getNuRC    h:165-178   reverse complement through the byteRC table
//   read2kmers_edges h:274-311  canonical = min(kmer, rc); windows touching a byte outside
//                              upper-case "ACGT" (h:52,297) are invalid (-1)
//
// Device layout of one read ("packed record", W64 u64 words, 16-byte aligned):
//   words [0, WB)        2-bit codes, base i in bits 2*(i%32) of word i/32 (LITTLE-endian base order)
//   words [WB, WB+WM)    invalid mask, bit (i%64) of word i/64 = 1 if base i is not ACGT or i >= len
// With little-endian packing the window w = bits [2i, 2i+2k) gives
//   reverse complement = (~w) & mask            (free), and
//   forward k-mer      = rev2(w) >> (64-2k)     (one brev + one pair swap),
// because rc = sum((3-c[i+m]) << 2m) and fwd = sum(c[i+m] << 2(k-1-m)).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define DBT_HD __host__ __device__ __forceinline__
#else
#define DBT_HD inline
#endif

namespace dbt {

constexpr uint64_t KMER_INVALID = 0xFFFFFFFFFFFFFFFFull;   // the reference's -1 slot (h:277,283)

struct ReadLayout {
	uint32_t WB;    // u64 words of bases   = ceil(max_read_len/32)
	uint32_t WM;    // u64 words of mask    = ceil(max_read_len/64)
	uint32_t W64;   // record words, (WB+WM) rounded up to even
};

DBT_HD ReadLayout make_layout(uint32_t max_read_len) {
	ReadLayout l;
	l.WB = (max_read_len + 31) / 32;
	l.WM = (max_read_len + 63) / 64;
	l.W64 = (l.WB + l.WM + 1) & ~1u;
	return l;
}

// meta word per read: low 15 bits length, bit 15 = "no valid k-mer window" (read2kmers_edges left
// the vectors empty, h:282 -> the pair is skipped as short, cpp:2037)
constexpr uint16_t META_NOVALID = 0x8000;
constexpr uint16_t META_LENMASK = 0x7FFF;

// reverse the order of the 32 2-bit groups of x
DBT_HD uint64_t rev2(uint64_t x) {
#ifdef __CUDA_ARCH__
	x = __brevll(x);
#else
	x = ((x >> 1) & 0x5555555555555555ull) | ((x & 0x5555555555555555ull) << 1);
	x = ((x >> 2) & 0x3333333333333333ull) | ((x & 0x3333333333333333ull) << 2);
	x = ((x >> 4) & 0x0F0F0F0F0F0F0F0Full) | ((x & 0x0F0F0F0F0F0F0F0Full) << 4);
	x = __builtin_bswap64(x);
#endif
	return ((x >> 1) & 0x5555555555555555ull) | ((x & 0x5555555555555555ull) << 1);
}

// bits [bit, bit+nbits) of a little-endian u64 array (nbits <= 64; caller guarantees the array
// holds word (bit>>6)+1 whenever the range straddles a word)
template <typename P>
DBT_HD uint64_t extract_bits(const P* words, uint32_t bit, uint32_t nbits) {
	uint32_t q = bit >> 6, s = bit & 63;
	uint64_t v = words[q] >> s;
	if (s + nbits > 64) v |= words[q + 1] << (64 - s);
	return nbits == 64 ? v : (v & ((1ull << nbits) - 1));
}

// canonical k-mer of window i of a packed record, or KMER_INVALID.  rec = WB base words then WM mask words.
template <typename P>
DBT_HD uint64_t canonical_kmer_at(const P* rec, uint32_t WB, uint32_t i, uint32_t k) {
	if (extract_bits(rec + WB, i, k) != 0) return KMER_INVALID;
	uint64_t w = extract_bits(rec, 2 * i, 2 * k);
	uint64_t mask = (1ull << (2 * k)) - 1;
	uint64_t rc = (~w) & mask;
	uint64_t fw = rev2(w) >> (64 - 2 * k);
	return fw < rc ? fw : rc;
}

// ASCII -> code (0..3) or 4 for anything that is not upper-case ACGT
DBT_HD uint32_t base_code(uint8_t c) {
	return c == 'A' ? 0u : c == 'C' ? 1u : c == 'G' ? 2u : c == 'T' ? 3u : 4u;
}

}  // namespace dbt
