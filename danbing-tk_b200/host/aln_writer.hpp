// Legacy `.aln` lines on stdout (`-a` / `-ae` with graph threading): writeAlignments, src/aQueryFasta_thread.cpp:1742-1759,
// with writeCigar (cpp:1701-1740; a D next to an I prints as one X) and writeAnnot (cpp:1683-1699) restated literally,
// quirks included (a trailing single edit prints as "1X"/"1D"/"1I" through the final `cout << ct << e0.t`).
// Columns: src dst title seq(read 2) seq(read 1) ops2 annot2 ops1 annot1; the title keeps its leading '>' / '@'
// (titles[] is printed without the substr(1) the kam writer applies, cpp:1747 vs 1664).
#pragma once
#include <cstdint>
#include <string>
#include "../../include/danbing_b200.h"
#include "fastx_reader.hpp"
#include "kam_writer.hpp"

namespace dbt_host {

inline void append_cigar(std::string& o, const uint8_t* es, uint32_t n) {
	static const char T[8] = {'*', '=', 'X', 'D', 'I', '?', '?', '?'};
	static const char G[8] = {'\0', 'A', 'C', 'G', 'T', '*', '?', '?'};
	if (!n) { o.push_back('*'); return; }
	auto t = [&](uint32_t i) { return T[DBT_ES_TYPE(es[i])]; };
	auto g = [&](uint32_t i) { return G[DBT_ES_BASE(es[i])]; };
	auto putg = [&](char c) { if (c) o.push_back(c); else o.push_back('\0'); };
	long long ct = 1;
	uint32_t i0 = 0;                                   // index of e0
	for (uint32_t i = 1; i < n; ++i) {
		const char t0 = t(i0);
		if (t0 == '=' || t0 == '*') {
			while (t(i) == t0) { ++ct; ++i; if (i == n) break; }
			append_int(o, ct); o.push_back(t0);
		}
		else if (t0 == 'X') { o.push_back('X'); putg(g(i0)); }
		else if (t0 == 'D') {
			if (t(i) == 'I') { o.push_back('X'); putg(g(i0)); ++i; }
			else { o.push_back('D'); putg(g(i0)); }
		}
		else if (t0 == 'I') {
			if (t(i) == 'D') { o.push_back('X'); putg(g(i)); ++i; }
			else o.push_back('I');
		}
		else o.push_back(t0);
		if (i == n) return;
		ct = 1;
		i0 = i;
	}
	append_int(o, ct); o.push_back(t(i0));
}

inline void append_annot(std::string& o, const uint8_t* tr, uint32_t n) {
	static const char C[4] = {'*', '.', '=', '?'};
	if (!n) { o.push_back('*'); return; }
	long long ct = 1;
	char c0 = C[tr[0] & 3];
	for (uint32_t i = 1; i < n; ++i) {
		while (C[tr[i] & 3] == c0) { ++ct; ++i; if (i == n) break; }
		append_int(o, ct); o.push_back(c0);
		if (i == n) return;
		ct = 1;
		c0 = C[tr[i] & 3];
	}
	append_int(o, ct); o.push_back(c0);
}

// rec: one record of dbt_record_stride bytes in the threading layout (header + two dbt_aln_mate blobs)
inline void append_aln_line(std::string& o, const uint8_t* rec, uint32_t aln_stride, uint32_t es_cap, const PairText& t) {
	const dbt_pair_record* h = reinterpret_cast<const dbt_pair_record*>(rec);
	o.append(".\t");                                   // src: -1 without -s (cpp:1744)
	append_int(o, (int)h->dst); o.push_back('\t');
	o.append(t.title, t.title_len); o.push_back('\t');
	o.append(t.seq[1], t.seq_len[1]); o.push_back('\t');     // read 2
	o.append(t.seq[0], t.seq_len[0]); o.push_back('\t');     // read 1
	for (int m = 1; m >= 0; --m) {                      // r2 first (cpp:1750-1756)
		const uint8_t* b = rec + sizeof(dbt_pair_record) + (size_t)m * aln_stride;
		const dbt_aln_mate* a = reinterpret_cast<const dbt_aln_mate*>(b);
		append_cigar(o, b + sizeof(dbt_aln_mate), a->nes); o.push_back('\t');
		append_annot(o, b + sizeof(dbt_aln_mate) + es_cap, a->ntr);
		o.push_back(m ? '\t' : '\n');
	}
}
inline void append_aln_line(std::string& o, const uint8_t* rec, uint32_t aln_stride, uint32_t es_cap, const HostBatch& hb) {
	const uint64_t i = reinterpret_cast<const dbt_pair_record*>(rec)->pair_index - hb.first_pair_index;
	append_aln_line(o, rec, aln_stride, es_cap, pair_text(hb, i, false));
}

}  // namespace dbt_host
