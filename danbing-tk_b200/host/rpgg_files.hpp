// Host loaders and writers for the RPGG files of the reference, feeding dbt_load_index.
//
//   PREF.kmers.dbi   readBinaryIndex        src/aQueryFasta_thread.h:654-673   u64 nk | u64 keys[nk] | u32 vals[nk] | u64 nvv | u32 vv[nvv]
//   PREF.fl.kdb      readBinaryKmerSetDB    src/aQueryFasta_thread.h:675-698   u64 nloci | u64 n[nloci] | u64 nk | u64 kmers[nk]
//   PREF.tr.kmers    countLoci              src/kmerIO.hpp:33-45               '>' lines
//                    readKmersWithZeroCount src/aQueryFasta_thread.h:469-480   decimal k-mer = first field of every other line
//   -qc FILE         readQCFile             src/kmerIO.hpp:111-120             one ASCII '0'/'1' per locus
//   OUT.trkmc.ar     dumpTRKmers            src/aQueryFasta_thread.h:968-973 -> binaryKmerIO.hpp:32-51, 179-188
//   OUT.tr.summary.txt writeTRKmerSummary   src/aQueryFasta_thread.h:975-983
//   OUT.tr.kmers     writeKmersWithName     src/aQueryFasta_thread.h:926-937
//
// Output order: the reference dumps each locus' unordered_map<size_t, atomic<size_t>> in ITERATION order
// after inserting the keys in file order.  That order is a property of libstdc++'s hashtable (identity
// hash, prime bucket counts, insertion at the head of the singly linked node list); it is reproduced
// here by building the same std::unordered_map<size_t, ...> per locus with the same insertion sequence
// -- this file must be compiled against the libstdc++ the reference would be built with.  The device
// count array is laid out in that order, so the final dump is a flat copy.
//
// Unlike the reference (which reads garbage when PREF.fl.kdb is missing, h:681), every short or
// missing file is an error.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace dbt_host {

struct Rpgg {
	uint64_t nloci = 0;
	std::vector<uint64_t> dbi_keys; std::vector<uint32_t> dbi_vals, vv;
	std::vector<uint64_t> tr_n, tr_kmers;      // per-locus sizes + keys in OUTPUT order
	std::vector<uint64_t> fl_n, fl_kmers;
	std::vector<uint8_t> qc;
	std::vector<uint64_t> g_n, g_kmers;        // PREF.graph.kmers: per-locus node counts + strand-specific k-mers (sorted, unique)
	std::vector<uint8_t> g_masks;              // 4-bit out-edge masks
	std::vector<uint64_t> tre_n, tre_kmers;    // PREF.tre.kdb (`-bu`): per-locus TR (k+1)-mer edges
};

inline bool read_whole_file(const std::string& fn, std::vector<char>& buf, std::string& err) {
	FILE* f = fopen(fn.c_str(), "rb");
	if (!f) { err = "cannot open " + fn; return false; }
	fseek(f, 0, SEEK_END);
	long sz = ftell(f);
	fseek(f, 0, SEEK_SET);
	buf.resize(sz > 0 ? (size_t)sz : 0);
	if (sz > 0 && fread(buf.data(), 1, (size_t)sz, f) != (size_t)sz) { fclose(f); err = "short read on " + fn; return false; }
	fclose(f);
	return true;
}

inline bool read_dbi(const std::string& fn, Rpgg& g, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	auto bad = [&]() { err = fn + ": truncated or malformed .kmers.dbi"; return false; };
	if (b.size() < 16) return bad();
	uint64_t nk; memcpy(&nk, b.data(), 8);
	if (nk > (b.size() - 16) / 12) return bad();
	g.dbi_keys.resize(nk); g.dbi_vals.resize(nk);
	memcpy(g.dbi_keys.data(), b.data() + 8, nk * 8);
	memcpy(g.dbi_vals.data(), b.data() + 8 + nk * 8, nk * 4);
	size_t o = 8 + nk * 12;
	uint64_t nvv; memcpy(&nvv, b.data() + o, 8);
	if (b.size() != o + 8 + nvv * 4) return bad();
	g.vv.resize(nvv);
	memcpy(g.vv.data(), b.data() + o + 8, nvv * 4);
	return true;
}

inline bool read_kdb(const std::string& fn, std::vector<uint64_t>& n, std::vector<uint64_t>& ks, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	auto bad = [&]() { err = fn + ": truncated or malformed .kdb"; return false; };
	if (b.size() < 16) return bad();
	uint64_t nloci; memcpy(&nloci, b.data(), 8);
	if (nloci > (b.size() - 16) / 8) return bad();
	n.resize(nloci);
	memcpy(n.data(), b.data() + 8, nloci * 8);
	uint64_t nk; memcpy(&nk, b.data() + 8 + nloci * 8, 8);
	if (b.size() != 16 + nloci * 8 + nk * 8) return bad();
	ks.resize(nk);
	memcpy(ks.data(), b.data() + 16 + nloci * 8, nk * 8);
	uint64_t s = 0;
	for (uint64_t x : n) s += x;
	if (s != nk) return bad();
	return true;
}

// PREF.tr.kmers -> nloci, per-locus k-mers in the reference's output order.  Parsing and the per-locus
// hashtable replay run on nthreads host threads (loci are independent).
inline bool read_tr_kmers_ordered(const std::string& fn, Rpgg& g, int nthreads, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	// pass 1: locus boundaries (countLoci: a line whose first character is '>')
	std::vector<size_t> start;                       // offset of the first k-mer line of each locus
	const char* p = b.data(); const char* e = p + b.size();
	const char* q = p;
	while (q < e) {
		const char* nl = (const char*)memchr(q, '\n', e - q);
		const char* next = nl ? nl + 1 : e;
		if (*q == '>') start.push_back(next - p);
		else if (start.empty()) { err = fn + ": k-mer line before the first '>' header"; return false; }   // reference: kmerDB[-1], undefined
		q = next;
	}
	const size_t nloci = start.size();
	g.nloci = nloci;
	g.tr_n.assign(nloci, 0);
	std::vector<std::vector<uint64_t>> per(nloci);
	std::atomic<size_t> nextl{0};
	std::atomic<bool> ok{true};
	auto work = [&]() {
		for (;;) {
			const size_t l = nextl.fetch_add(1);
			if (l >= nloci) break;
			std::unordered_map<size_t, uint32_t> m;       // fresh table per locus: same bucket-count trajectory as the reference's
			const char* s = p + start[l];
			const char* lim = (l + 1 < nloci) ? p + start[l + 1] : e;
			while (s < lim) {
				const char* nl = (const char*)memchr(s, '\n', lim - s);
				const char* next = nl ? nl + 1 : lim;
				if (*s == '>') { s = next; continue; }        // the next locus' header line
				if (s == (nl ? nl : lim)) { ok = false; s = next; continue; }   // empty line: stoul would throw in the reference
				// stoul(line): leading whitespace, optional sign, decimal digits; stops at the first other char
				char* endp = nullptr;
				const unsigned long v = strtoul(s, &endp, 10);
				if (endp == s) { ok = false; s = next; continue; }
				m[(size_t)v] = 0;
				s = next;
			}
			std::vector<uint64_t>& out = per[l];
			out.reserve(m.size());
			for (auto& kv : m) out.push_back(kv.first);
		}
	};
	if (nthreads < 1) nthreads = 1;
	std::vector<std::thread> th;
	for (int t = 1; t < nthreads; ++t) th.emplace_back(work);
	work();
	for (auto& t : th) t.join();
	if (!ok) { err = fn + ": malformed k-mer line"; return false; }
	size_t tot = 0;
	for (size_t l = 0; l < nloci; ++l) { g.tr_n[l] = per[l].size(); tot += per[l].size(); }
	g.tr_kmers.resize(tot);
	size_t o = 0;
	for (size_t l = 0; l < nloci; ++l) { if (!per[l].empty()) memcpy(g.tr_kmers.data() + o, per[l].data(), per[l].size() * 8); o += per[l].size(); }
	return true;
}

// PREF.graph.kmers (text; readGraphKmers, src/aQueryFasta_thread.h:549-575: `kmerDB[idx][kmer] |= c`) -> per-locus nodes,
// duplicate lines merged with OR like the reference's operator|=.  Loci are parsed on nthreads host threads.
inline bool read_graph_kmers(const std::string& fn, uint64_t nloci, Rpgg& g, int nthreads, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	std::vector<size_t> start;
	const char* p = b.data(); const char* e = p + b.size();
	for (const char* q = p; q < e; ) {
		const char* nl = (const char*)memchr(q, '\n', e - q);
		const char* next = nl ? nl + 1 : e;
		if (*q == '>') start.push_back(next - p);
		else if (start.empty()) { err = fn + ": node line before the first '>' header"; return false; }
		q = next;
	}
	if (start.size() != nloci) { err = fn + ": " + std::to_string(start.size()) + " loci, the RPGG has " + std::to_string(nloci); return false; }
	std::vector<std::vector<std::pair<uint64_t, uint8_t>>> per(nloci);
	std::atomic<size_t> nextl{0};
	std::atomic<bool> ok{true};
	auto work = [&]() {
		for (;;) {
			const size_t l = nextl.fetch_add(1);
			if (l >= nloci) break;
			auto& v = per[l];
			const char* s = p + start[l];
			const char* lim = (l + 1 < nloci) ? p + start[l + 1] : e;
			while (s < lim) {
				const char* nl = (const char*)memchr(s, '\n', lim - s);
				const char* next = nl ? nl + 1 : lim;
				if (*s == '>') { s = next; continue; }
				char* e1 = nullptr; char* e2 = nullptr;
				const unsigned long km = strtoul(s, &e1, 10);
				if (e1 == s || *e1 != '\t') { ok = false; s = next; continue; }
				const unsigned long mk = strtoul(e1 + 1, &e2, 10);
				if (e2 == e1 + 1) { ok = false; s = next; continue; }
				v.emplace_back((uint64_t)km, (uint8_t)mk);
				s = next;
			}
			std::sort(v.begin(), v.end());
			size_t w = 0;
			for (size_t i = 0; i < v.size(); ++i) {
				if (w && v[w - 1].first == v[i].first) v[w - 1].second |= v[i].second;
				else v[w++] = v[i];
			}
			v.resize(w);
		}
	};
	if (nthreads < 1) nthreads = 1;
	std::vector<std::thread> th;
	for (int t = 1; t < nthreads; ++t) th.emplace_back(work);
	work();
	for (auto& t : th) t.join();
	if (!ok) { err = fn + ": malformed node line (expected `kmer<TAB>mask`)"; return false; }
	size_t tot = 0;
	g.g_n.assign(nloci, 0);
	for (size_t l = 0; l < nloci; ++l) { g.g_n[l] = per[l].size(); tot += per[l].size(); }
	g.g_kmers.resize(tot); g.g_masks.resize(tot);
	size_t o = 0;
	for (size_t l = 0; l < nloci; ++l) for (auto& kv : per[l]) { g.g_kmers[o] = kv.first; g.g_masks[o] = kv.second; ++o; }
	return true;
}

// LEGACY v1.3 PREF.graph.umap: u64 nloci | per locus { u64 n | n x { u64 key, u8 mask } }
inline bool read_graph_umap(const std::string& fn, uint64_t nloci, Rpgg& g, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	auto bad = [&]() { err = fn + ": truncated or malformed .graph.umap"; return false; };
	size_t o = 0;
	auto rd64 = [&](uint64_t& v) { if (o + 8 > b.size()) return false; memcpy(&v, b.data() + o, 8); o += 8; return true; };
	uint64_t nl;
	if (!rd64(nl)) return bad();
	if (nl != nloci) { err = fn + ": " + std::to_string(nl) + " loci, the RPGG has " + std::to_string(nloci); return false; }
	g.g_n.assign(nloci, 0); g.g_kmers.clear(); g.g_masks.clear();
	for (uint64_t l = 0; l < nloci; ++l) {
		uint64_t n;
		if (!rd64(n) || n > (b.size() - o) / 9) return bad();
		std::vector<std::pair<uint64_t, uint8_t>> v(n);
		for (uint64_t i = 0; i < n; ++i) { memcpy(&v[i].first, b.data() + o, 8); v[i].second = (uint8_t)b[o + 8]; o += 9; }
		std::sort(v.begin(), v.end());
		size_t w = 0;
		for (size_t i = 0; i < v.size(); ++i) { if (w && v[w - 1].first == v[i].first) v[w - 1].second |= v[i].second; else v[w++] = v[i]; }
		g.g_n[l] = w;
		for (size_t i = 0; i < w; ++i) { g.g_kmers.push_back(v[i].first); g.g_masks.push_back(v[i].second); }
	}
	if (o != b.size()) return bad();
	return true;
}

inline bool file_exists(const std::string& fn) { FILE* f = fopen(fn.c_str(), "rb"); if (!f) return false; fclose(f); return true; }

inline bool read_qc(const std::string& fn, uint64_t nloci, std::vector<uint8_t>& qc, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	if (b.size() < nloci) { err = fn + ": fewer QC flags than loci"; return false; }
	qc.resize(nloci);
	for (uint64_t i = 0; i < nloci; ++i) qc[i] = (uint8_t)(b[i] - 48);
	return true;
}

// ---- writers ----
inline bool write_trkmc_ar(const std::string& pref, const std::vector<uint64_t>& counts, std::string& err) {
	const std::string fn = pref + ".trkmc.ar";
	FILE* f = fopen(fn.c_str(), "wb");
	if (!f) { err = "cannot create " + fn; return false; }
	const uint64_t nk = counts.size();
	bool ok = fwrite(&nk, 8, 1, f) == 1 && (nk == 0 || fwrite(counts.data(), 8, nk, f) == nk);
	ok = (fclose(f) == 0) && ok;
	if (!ok) err = "short write on " + fn;
	return ok;
}

inline bool write_tr_summary(const std::string& fn, const std::vector<uint64_t>& kmc, const std::vector<uint32_t>& nmapread, std::string& err) {
	FILE* f = fopen(fn.c_str(), "w");
	if (!f) { err = "cannot create " + fn; return false; }
	for (size_t i = 0; i < kmc.size(); ++i) fprintf(f, "%u\t%llu\n", nmapread[i], (unsigned long long)kmc[i]);
	if (fclose(f) != 0) { err = "short write on " + fn; return false; }
	return true;
}

inline bool write_kmers_with_name(const std::string& fn, const Rpgg& g, const std::vector<uint64_t>& counts, std::string& err) {
	FILE* f = fopen(fn.c_str(), "w");
	if (!f) { err = "cannot create " + fn; return false; }
	size_t j = 0;
	for (uint64_t l = 0; l < g.nloci; ++l) {
		fprintf(f, ">%llu\n", (unsigned long long)l);
		for (uint64_t i = 0; i < g.tr_n[l]; ++i, ++j) fprintf(f, "%llu\t%llu\n", (unsigned long long)g.tr_kmers[j], (unsigned long long)counts[j]);
	}
	if (fclose(f) != 0) { err = "short write on " + fn; return false; }
	return true;
}

}  // namespace dbt_host
