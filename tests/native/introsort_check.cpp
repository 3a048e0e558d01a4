// Differential test: danbing-tk_b200/csrc/introsort.cuh (the code the GPU lane runs, compiled for the
// host) and oracle/dbt_oracle.c's restatement, both against this toolchain's std::sort with the
// comparator of getSortedIndex (reference src/aQueryFasta_thread.cpp:247-250).
// Prints "OK <ncases> <heapsort_cases>" or the first mismatch.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <numeric>
#include <random>
#include <vector>
static long nheap;
#define DBT_SORT_ON_HEAPSORT() (++nheap)
#include "../../danbing-tk_b200/csrc/introsort.cuh"
extern "C" void dbo_introsort_by_key(uint64_t* idx, uint32_t n, const uint64_t* key);

static long ncmp;
static bool run(const std::vector<uint64_t>& key) {
	int n = (int)key.size();
	std::vector<uint64_t> ref(n);
	std::iota(ref.begin(), ref.end(), 0);
	std::sort(ref.begin(), ref.end(), [&key](uint64_t a, uint64_t b) { ++ncmp; return key[a] < key[b]; });
	std::vector<uint16_t> idx(n);
	std::iota(idx.begin(), idx.end(), 0);
	std::vector<uint32_t> k32(key.begin(), key.end());
	dbt::IntroSort<uint16_t, uint32_t> s{idx.data(), k32.data()};
	s.sort(n);
	std::vector<uint64_t> o(n);
	std::iota(o.begin(), o.end(), 0);
	dbo_introsort_by_key(o.data(), n, key.data());
	for (int i = 0; i < n; ++i) if (ref[i] != idx[i] || ref[i] != o[i]) { printf("MISMATCH n=%d at %d: std=%lu dev=%u oracle=%lu\n", n, i, ref[i], idx[i], o[i]); return false; }
	return true;
}

int main() {
	std::mt19937_64 rng(12345);
	long cases = 0, heavy = 0;
	for (int n = 1; n <= 700; ++n) {
		for (int rep = 0; rep < 12; ++rep) {
			std::vector<uint64_t> key(n);
			switch (rep % 6) {
			case 0: for (auto& k : key) k = 1; break;                                     // all ties (the common case)
			case 1: for (auto& k : key) k = (rng() % 100 < 5) ? 2 + rng() % 40 : 1; break; // mostly 1, a few multi-locus
			case 2: for (auto& k : key) k = rng() % 4; break;
			case 3: for (auto& k : key) k = rng() & 0xFFFFFFFFull; break;
			case 4: for (int i = 0; i < n; ++i) key[i] = i; break;
			case 5: for (int i = 0; i < n; ++i) key[i] = n - i; break;
			}
			if (!run(key)) return 1;
			++cases;
		}
	}
	// McIlroy's "antiquicksort" adversary builds, for THIS std::sort, a key array that makes the
	// quicksort phase quadratic and so drives introsort into its heap-sort fallback
	for (int n = 40; n <= 700; n += 3) {
		std::vector<long> val(n);
		const long gas = n;
		long nsolid = 0, candidate = 0;
		std::fill(val.begin(), val.end(), gas);
		std::vector<uint64_t> ord(n);
		std::iota(ord.begin(), ord.end(), 0);
		std::sort(ord.begin(), ord.end(), [&](uint64_t x, uint64_t y) {
			if (val[x] == gas && val[y] == gas) { if ((long)x == candidate) val[x] = nsolid++; else val[y] = nsolid++; }
			if (val[x] == gas) candidate = x; else if (val[y] == gas) candidate = y;
			return val[x] < val[y];
		});
		std::vector<uint64_t> key(n);
		for (int i = 0; i < n; ++i) key[i] = (uint64_t)val[i];
		long before = nheap;
		if (!run(key)) return 1;
		if (nheap > before) ++heavy;
		++cases;
	}
	printf("OK %ld %ld\n", cases, heavy);
	return 0;
}
