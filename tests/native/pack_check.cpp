// Host build of the per-lane body of pack_kernel (csrc/pack_read.cuh) against the host packer (csrc/host_pack.cpp, itself pinned
// to a numpy model by tests/test_abi.py): random reads of every length 0..max with N / lower-case / arbitrary bytes, at every
// byte alignment inside a staged tile, k = 17/21/25, max_read_len = 150 / 250.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "../../danbing-tk_b200/csrc/pack_read.cuh"
#include "../../danbing-tk_b200/csrc/host_pack.h"
using namespace dbt;
int main() {
	std::mt19937_64 rng(12345);
	uint64_t ncase = 0, nnovalid = 0, nfast = 0;
	for (uint32_t maxlen : {150u, 250u, 37u}) for (uint32_t k : {17u, 21u, 25u}) {
		const ReadLayout lay = make_layout(maxlen);
		std::vector<uint64_t> a(lay.W64), b(lay.W64);
		for (int it = 0; it < 6000; ++it) {
			const uint32_t len = it < 2 * (int)maxlen ? (uint32_t)it / 2 : (uint32_t)(rng() % (maxlen + 1));
			const uint32_t so = (uint32_t)(rng() % 64);
			alignas(16) uint8_t tile[1024];
			for (auto& c : tile) c = (uint8_t)rng();                      // neighbours / uninitialised shared memory
			const int mode = (int)(rng() % 4);
			for (uint32_t i = 0; i < len; ++i) {
				uint8_t c = "ACGT"[rng() % 4];
				const uint32_t r = (uint32_t)(rng() % 1000);
				if (mode == 1 && r < 20) c = 'N';
				if (mode == 2 && r < 100) c = "acgtN\n@U"[rng() % 8];
				if (mode == 3 && r < 300) c = (uint8_t)rng();
				tile[so + i] = c;
			}
			uint16_t ma = pack_read_swar(reinterpret_cast<const uint32_t*>(tile + (so & ~3u)), (so & 3u) * 8, len, k, lay, a.data());
			uint16_t mb = 0;
			host_pack_read(tile + so, len, sizeof(tile) - so, k, lay, b.data(), &mb, it & 1 ? HOST_PACK_SWAR : HOST_PACK_AUTO);
			if (ma != mb || memcmp(a.data(), b.data(), lay.W64 * 8)) {
				printf("MISMATCH maxlen=%u k=%u len=%u so=%u mode=%d meta %04x vs %04x\n", maxlen, k, len, so, mode, ma, mb);
				for (uint32_t w = 0; w < lay.W64; ++w) printf("  %u: %016llx %016llx\n", w, (unsigned long long)a[w], (unsigned long long)b[w]);
				return 1;
			}
			// the fast path (clean reads): refuses exactly the reads that hold a byte outside ACGT, else the same record
			bool clean = true;
			for (uint32_t i = 0; i < len; ++i) clean = clean && (tile[so + i] == 'A' || tile[so + i] == 'C' || tile[so + i] == 'G' || tile[so + i] == 'T');
			std::vector<uint64_t> f(lay.W64, 0x5555555555555555ull);
			uint16_t mf = 0xFFFF;
			const bool took = pack_read_fast(reinterpret_cast<const uint32_t*>(tile + (so & ~3u)), (so & 3u) * 8, len, k, lay, f.data(), &mf);
			if (took != clean || (took && (mf != mb || memcmp(f.data(), b.data(), lay.W64 * 8)))) {
				printf("FAST MISMATCH maxlen=%u k=%u len=%u so=%u mode=%d took=%d clean=%d meta %04x vs %04x\n", maxlen, k, len, so, mode, (int)took, (int)clean, mf, mb);
				for (uint32_t w = 0; w < lay.W64; ++w) printf("  %u: %016llx %016llx\n", w, (unsigned long long)f[w], (unsigned long long)b[w]);
				return 1;
			}
			nfast += took;
			++ncase; nnovalid += (ma & META_NOVALID) != 0;
		}
	}
	printf("OK %llu %llu %llu\n", (unsigned long long)ncase, (unsigned long long)nnovalid, (unsigned long long)nfast);
	return 0;
}
