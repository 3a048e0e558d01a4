// Index construction + legacy loaders.
//
//   serialize_rpgg(PREF)   = `ktools serialize PREF` (src/kmertools.cpp:221-345): PREF.tr.kmers + PREF.fl.kmers ->
//                            PREF.kmers.dbi, PREF.fl.kmers -> PREF.fl.kdb, PREF.tre.kmers -> PREF.tre.kdb, byte for byte.
//   read_legacy_dbi(PREF)  = the v1.3 release containers PREF.kmerDBi.umap / PREF.kmerDBi.vv (no reader
//                            is left in the reference at HEAD): u64 n | n x { u64 key, u64 val }  and  u64 n | u32 vv[n].
//   read_text_kset(fn)     = PREF.fl.kmers / the v1.3 PREF.ntr.kmers as a per-locus k-mer set when no PREF.fl.kdb exists.
//
// Byte equality with ktools rests on two libstdc++ behaviours (SURVEY §0 trap 6): `.kmers.dbi` lists the keys of
// unordered_map<size_t,size_t> in iteration order after the insertions of readKmerIndex (src/kmerIO.hpp:47-78), and the
// `.kdb` files list every locus' unordered_set<uint64_t> in iteration order after the insertions of readKmers_ksetDB
// (src/kmerIO.hpp:80-90).  The same containers are built here with the same insertion sequence, so this file must be
// compiled against the libstdc++ the reference would be built with (as host/rpgg_files.hpp already requires).
// Parsing is this file's own (one pass over the file image, strtoul per line).
#pragma once
#include <unordered_set>
#include "rpgg_files.hpp"

namespace dbt_host {

// calls fn(locus, kmer) for every k-mer line, locus = number of '>' lines seen so far - 1 (a k-mer line before the first
// header is an error: the reference indexes [-1]); returns the number of '>' lines (countLoci, src/kmerIO.hpp:33-45)
template <typename F>
inline bool for_each_kmer_line(const std::string& fn, F&& f, uint64_t& nloci, std::string& err) {
	std::vector<char> b;
	if (!read_whole_file(fn, b, err)) return false;
	const char* s = b.data(); const char* e = s + b.size();
	int64_t locus = -1;
	while (s < e) {
		const char* nl = (const char*)memchr(s, '\n', e - s);
		const char* le = nl ? nl : e;
		if (*s == '>') ++locus;
		else {
			if (locus < 0) { err = fn + ": k-mer line before the first '>' header"; return false; }
			char* endp = nullptr;
			const unsigned long v = strtoul(s, &endp, 10);           // stoul: first field only
			if (endp == s || endp > le) { err = fn + ": malformed k-mer line"; return false; }
			f((uint32_t)locus, (uint64_t)v);
		}
		s = nl ? nl + 1 : e;
	}
	nloci = (uint64_t)(locus + 1);
	return true;
}

struct DbiBuilder {                                  // readKmerIndex, src/kmerIO.hpp:47-78
	std::unordered_map<size_t, size_t> m;            // kmerIndex_uint32_umap
	std::vector<std::vector<uint32_t>> lists;        // kmerDBi_vec
	bool add_file(const std::string& fn, std::string& err) {
		uint32_t vsize = (uint32_t)lists.size();     // sampled once per file, like the reference's local
		uint64_t nl = 0;
		return for_each_kmer_line(fn, [&](uint32_t idx, uint64_t kmer) {
			auto it = m.find(kmer);
			if (it == m.end()) { m[kmer] = (size_t)(idx << 1); return; }
			const uint32_t vi = (uint32_t)it->second;
			if (vi & 1) {                             // already shared: append the locus once
				auto& v = lists[vi >> 1];
				if (std::find(v.begin(), v.end(), idx) == v.end()) v.push_back(idx);
			} else if ((vi >> 1) != idx) {            // second locus of a so-far unique k-mer
				lists.push_back(std::vector<uint32_t>{vi >> 1, idx});
				it->second = (size_t)(((vsize++) << 1) + 1);
			}
		}, nl, err);
	}
};

inline bool write_file(const std::string& fn, const std::vector<std::pair<const void*, size_t>>& parts, std::string& err) {
	FILE* f = fopen(fn.c_str(), "wb");
	if (!f) { err = "cannot create " + fn; return false; }
	bool ok = true;
	for (auto& p : parts) if (p.second && fwrite(p.first, 1, p.second, f) != p.second) ok = false;
	ok = (fclose(f) == 0) && ok;
	if (!ok) err = "short write on " + fn;
	return ok;
}

// PREF.<tp>.kmers -> PREF.<tp>.kdb (readKmers_ksetDB + flattenKsetDB + serializeKsetDB, src/binaryKmerIO.hpp:116-139)
inline bool serialize_kset(const std::string& text_fn, const std::string& kdb_fn, uint64_t nloci, std::string& err) {
	std::vector<std::unordered_set<uint64_t>> db(nloci);
	uint64_t nl = 0; bool range_ok = true;
	if (!for_each_kmer_line(text_fn, [&](uint32_t l, uint64_t km) { if (l < nloci) db[l].insert(km); else range_ok = false; }, nl, err)) return false;
	if (!range_ok) { err = text_fn + ": more loci than PREF.tr.kmers"; return false; }
	std::vector<uint64_t> index(nloci), ks;
	for (uint64_t l = 0; l < nloci; ++l) { index[l] = db[l].size(); for (uint64_t km : db[l]) ks.push_back(km); }
	const uint64_t nk = ks.size();
	return write_file(kdb_fn, {{&nloci, 8}, {index.data(), 8 * nloci}, {&nk, 8}, {ks.data(), 8 * nk}}, err);
}

inline bool count_loci(const std::string& fn, uint64_t& nloci, std::string& err) {
	return for_each_kmer_line(fn, [](uint32_t, uint64_t) {}, nloci, err);
}

inline bool serialize_rpgg(const std::string& pref, FILE* log, std::string& err) {
	uint64_t nloci = 0;
	if (!count_loci(pref + ".tr.kmers", nloci, err)) return false;
	{	// kmerDBi + vv, src/kmertools.cpp:232-279
		DbiBuilder b;
		if (log) fprintf(log, "reading kmers from %s.tr.kmers\n", pref.c_str());
		if (!b.add_file(pref + ".tr.kmers", err)) return false;
		if (log) fprintf(log, "reading kmers from %s.fl.kmers\n", pref.c_str());
		if (!b.add_file(pref + ".fl.kmers", err)) return false;
		std::vector<uint32_t> vv, vvi;
		for (auto& v : b.lists) { vvi.push_back((uint32_t)vv.size()); vv.push_back((uint32_t)v.size()); vv.insert(vv.end(), v.begin(), v.end()); }
		const uint64_t nvv = vv.size(), nk = b.m.size();
		std::vector<uint64_t> keys(nk); std::vector<uint32_t> vals(nk);
		uint64_t i = 0;
		for (auto& p : b.m) {
			keys[i] = p.first;
			vals[i] = (p.second & 1) ? (uint32_t)((vvi[p.second >> 1] << 1) + 1) : (uint32_t)p.second;
			++i;
		}
		if (log) fprintf(log, "serializing kmerDBi as *.kmers.dbi\n");
		if (!write_file(pref + ".kmers.dbi", {{&nk, 8}, {keys.data(), 8 * nk}, {vals.data(), 4 * nk}, {&nvv, 8}, {vv.data(), 4 * nvv}}, err)) return false;
	}
	if (log) fprintf(log, "Generating flank binary kmers fl.kdb\n");
	if (!serialize_kset(pref + ".fl.kmers", pref + ".fl.kdb", nloci, err)) return false;
	if (log) fprintf(log, "Generating TR edge (k+1) binary kmers tre.kdb\n");
	if (!serialize_kset(pref + ".tre.kmers", pref + ".tre.kdb", nloci, err)) return false;
	if (log) fprintf(log, "done\n");
	return true;
}

// ---- legacy v1.3 containers ----
inline bool read_legacy_dbi(const std::string& pref, Rpgg& g, std::string& err) {
	std::vector<char> b;
	const std::string fu = pref + ".kmerDBi.umap", fv = pref + ".kmerDBi.vv";
	if (!read_whole_file(fu, b, err)) return false;
	if (b.size() < 8) { err = fu + ": truncated"; return false; }
	uint64_t n; memcpy(&n, b.data(), 8);
	if (n > (b.size() - 8) / 16 || b.size() != 8 + n * 16) { err = fu + ": truncated or malformed .kmerDBi.umap"; return false; }
	g.dbi_keys.resize(n); g.dbi_vals.resize(n);
	for (uint64_t i = 0; i < n; ++i) {
		uint64_t k, v;
		memcpy(&k, b.data() + 8 + i * 16, 8); memcpy(&v, b.data() + 16 + i * 16, 8);
		if (v >> 32) { err = fu + ": value does not fit 32 bits"; return false; }
		g.dbi_keys[i] = k; g.dbi_vals[i] = (uint32_t)v;
	}
	if (!read_whole_file(fv, b, err)) return false;
	if (b.size() < 8) { err = fv + ": truncated"; return false; }
	memcpy(&n, b.data(), 8);
	if (n > (b.size() - 8) / 4 || b.size() != 8 + n * 4) { err = fv + ": truncated or malformed .kmerDBi.vv"; return false; }
	g.vv.resize(n);
	if (n) memcpy(g.vv.data(), b.data() + 8, n * 4);
	return true;
}

// per-locus k-mer set from a text file (unique per locus, file order)
inline bool read_text_kset(const std::string& fn, uint64_t nloci, std::vector<uint64_t>& n, std::vector<uint64_t>& ks, std::string& err) {
	std::vector<std::vector<uint64_t>> per(nloci);
	uint64_t nl = 0; bool range_ok = true;
	if (!for_each_kmer_line(fn, [&](uint32_t l, uint64_t km) { if (l < nloci) per[l].push_back(km); else range_ok = false; }, nl, err)) return false;
	if (!range_ok || nl != nloci) { err = fn + ": " + std::to_string(nl) + " loci, the RPGG has " + std::to_string(nloci); return false; }
	n.assign(nloci, 0); ks.clear();
	for (uint64_t l = 0; l < nloci; ++l) {
		auto& v = per[l];
		std::sort(v.begin(), v.end());
		v.erase(std::unique(v.begin(), v.end()), v.end());
		n[l] = v.size(); ks.insert(ks.end(), v.begin(), v.end());
	}
	return true;
}

// The RPGG as align loads it (cpp:2459-2504), plus the fallbacks that let a v1.3 release directory or an un-serialized
// text RPGG work unchanged: PREF.kmers.dbi, else PREF.kmerDBi.umap/.vv; PREF.fl.kdb, else PREF.fl.kmers, else PREF.ntr.kmers.
inline bool load_rpgg(const std::string& pref, const std::string& trFname, const std::string& qcFn, int nthreads, Rpgg& g, std::string& err) {
	if (!read_tr_kmers_ordered(trFname, g, nthreads, err)) return false;
	if (file_exists(pref + ".kmers.dbi") || !file_exists(pref + ".kmerDBi.umap")) { if (!read_dbi(pref + ".kmers.dbi", g, err)) return false; }
	else if (!read_legacy_dbi(pref, g, err)) return false;
	if (file_exists(pref + ".fl.kdb")) { if (!read_kdb(pref + ".fl.kdb", g.fl_n, g.fl_kmers, err)) return false; }
	else if (file_exists(pref + ".fl.kmers")) { if (!read_text_kset(pref + ".fl.kmers", g.nloci, g.fl_n, g.fl_kmers, err)) return false; }
	else if (file_exists(pref + ".ntr.kmers")) { if (!read_text_kset(pref + ".ntr.kmers", g.nloci, g.fl_n, g.fl_kmers, err)) return false; }
	else { err = "cannot open " + pref + ".fl.kdb"; return false; }
	if (g.fl_n.size() != g.nloci) { err = pref + ".fl.kdb: " + std::to_string(g.fl_n.size()) + " loci, " + trFname + " has " + std::to_string(g.nloci); return false; }
	if (!qcFn.empty() && !read_qc(qcFn, g.nloci, g.qc, err)) return false;
	return true;
}

}  // namespace dbt_host
