// `-b` / `-tb` on the host: the bait file loader and the replay that `-tb` needs.
//
//   PREF.bt.kmdb / PREF.qc.bt.kmdb   readBinaryBaitDB (src/aQueryFasta_thread.h:542-547) -> deserializeKmapDB (binaryKmerIO.hpp:70-95):
//                                    u64 nloci | u64 n[nloci] | u64 nk | u64 sizeof(T) = 2 | u64 keys[nk] | u16 vals[nk]  (val = min << 8 | max)
//   -tb                              bfilter_FPSv1 (cpp:1377-1396) records the FIRST bait k-mer whose count is out of bounds in the iteration order
//                                    of the mate's histogram (kc8_t = unordered_map<uint64_t, uint8_t> filled in k-mer order).  The GPU only
//                                    says which mates it removed (mate flag DBT_MF_BF); for those the same histogram is built here from the
//                                    read's text and walked in the same order.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>
#include "rpgg_files.hpp"

namespace dbt_host {

struct BaitDB {
	std::vector<uint64_t> n, kmers; std::vector<uint16_t> vals;            // flat, for dbt_index
	std::vector<std::unordered_map<uint64_t, uint16_t>> per_locus;        // for the -tb replay
};

inline bool read_bait_kmdb(const std::string& fn, uint64_t nloci, bool want_maps, BaitDB& db, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	auto bad = [&]() { err = fn + ": truncated or malformed .bt.kmdb"; return false; };
	if (b.size() < 8) return bad();
	uint64_t nl; memcpy(&nl, b.data(), 8);
	if (nl != nloci) { err = fn + ": " + std::to_string(nl) + " loci, the RPGG has " + std::to_string(nloci); return false; }
	if (b.size() < 8 + nl * 8 + 16) return bad();
	db.n.resize(nl); memcpy(db.n.data(), b.data() + 8, nl * 8);
	uint64_t nk, szv; memcpy(&nk, b.data() + 8 + nl * 8, 8); memcpy(&szv, b.data() + 16 + nl * 8, 8);
	uint64_t sum = 0; for (uint64_t x : db.n) sum += x;
	if (szv != 2 || sum != nk || b.size() != 24 + nl * 8 + nk * 10) return bad();
	db.kmers.resize(nk); db.vals.resize(nk);
	memcpy(db.kmers.data(), b.data() + 24 + nl * 8, nk * 8);
	memcpy(db.vals.data(), b.data() + 24 + nl * 8 + nk * 8, nk * 2);
	if (want_maps) {
		db.per_locus.assign(nl, {});
		uint64_t o = 0;
		for (uint64_t l = 0; l < nl; ++l) for (uint64_t i = 0; i < db.n[l]; ++i, ++o) db.per_locus[l][db.kmers[o]] = db.vals[o];
	}
	return true;
}

// canonical k-mers of a read as read2kmers_edges leaves them (src/aQueryFasta_thread.h:274-311): one per window, -1 where the window
// holds a byte that is not upper-case ACGT
inline void canonical_kmers(const char* s, size_t len, uint32_t k, std::vector<uint64_t>& out) {
	out.clear();
	if (len < k) return;
	out.assign(len - k + 1, ~0ull);
	const uint64_t mask = k < 32 ? (1ull << 2 * k) - 1 : ~0ull;
	uint64_t fw = 0, rc = 0; uint32_t run = 0;
	for (size_t i = 0; i < len; ++i) {
		uint64_t c;
		switch (s[i]) { case 'A': c = 0; break; case 'C': c = 1; break; case 'G': c = 2; break; case 'T': c = 3; break; default: run = 0; fw = rc = 0; continue; }
		fw = ((fw << 2) | c) & mask;
		rc = (rc >> 2) | ((3 - c) << 2 * (k - 1));
		if (++run >= k) out[i + 1 - k] = fw < rc ? fw : rc;
	}
}

// qString2qMask (src/aQueryFasta_thread.h:1038-1071): which k-mers of a FASTQ read enter the bait histogram -- those whose k bases all
// have phred >= qth, as far as the reference's scan gets (its base index is bounded by the NUMBER OF K-MERS, so it stops k-1 bases
// before the end of the read).  Same state machine as quality_mask in csrc/bait.cuh.
inline void quality_mask(const char* q, size_t nq_, int k, int qth, std::vector<bool>& m) {
	const int nq = (int)nq_, nk = nq - k + 1;
	m.assign(nk > 0 ? nk : 0, false);
	if (nk < 1) return;
	auto lowq = [&](int i) { return (int)q[i] - 33 < qth; };
	int qi = 0, ki = 0;
	while (lowq(qi)) { ++qi; ++ki; if (qi >= nk) return; }
	while (qi < nk) {
		bool pass = true;
		for (const int qj = qi; qi < qj + k; ++qi) {
			if (lowq(qi)) {
				pass = false; ki = qi;
				while (lowq(qi)) { ++qi; ++ki; if (qi >= nk) return; }
				break;
			}
		}
		if (pass) {
			m[ki] = true; ++ki;
			if (qi >= nk) return;
			while (!lowq(qi)) { m[ki] = true; ++qi; ++ki; if (qi >= nk) return; }
			ki = qi;
			while (lowq(qi)) { ++qi; ++ki; if (qi >= nk) return; }
		}
	}
}

// the k-mer bfilter_FPSv1 stops at (cpp:1386-1393; with a quality string, the masked form cpp:1398-1419); false when none is out of bounds
inline bool first_bait_violation(const std::unordered_map<uint64_t, uint16_t>& baitdb, const char* seq, size_t len, uint32_t k, uint64_t& km_out,
                                 const char* qual = nullptr, int qth = 20) {
	std::vector<uint64_t> ks;
	canonical_kmers(seq, len, k, ks);
	if (ks.empty()) return false;
	std::unordered_map<uint64_t, uint8_t> kc;                              // kc8_t
	if (qual) {
		std::vector<bool> m;
		quality_mask(qual, len, (int)k, qth, m);
		for (size_t i = 0; i < ks.size(); ++i) if (m[i]) ++kc[ks[i]];
	} else
		for (uint64_t km : ks) ++kc[km];
	for (auto& p : kc) {
		auto it = baitdb.find(p.first);
		if (it == baitdb.end()) continue;
		const uint8_t mi = (uint8_t)(it->second >> 8), ma = (uint8_t)(it->second & 0xff);
		if (p.second < mi || p.second > ma) { km_out = p.first; return true; }
	}
	return false;
}

}  // namespace dbt_host
