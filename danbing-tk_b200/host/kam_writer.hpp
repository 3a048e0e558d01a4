// kam ("k-mer assignment") lines on stdout: writeKmerAssignments, src/aQueryFasta_thread.cpp:1646-1681,
// with the run-length annotation of km_asgn_t::annot2str_ (cpp:121-138).  One line per dbt_pair_record
// returned by dbt_wait, in pair order (= the order in which the reference pushes kams, cpp:2169-2175).
// Columns: src dst dst0 span2 span1 header fields2 fields1 rle(as2) rle(as1) title seq2 qual2|. seq1 qual1|.
// where "1"/"2" are the reference's internal slots: read 1 = the mate that arrived second (cpp:1972-1974).
#pragma once
#include <charconv>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include "../../include/danbing_b200.h"
#include "fastx_reader.hpp"

namespace dbt_host {

inline void append_int(std::string& o, long long v) {
	char tmp[24];
	const auto r = std::to_chars(tmp, tmp + sizeof tmp, v);      // (snprintf was two thirds of the formatting time of a kam line)
	o.append(tmp, (size_t)(r.ptr - tmp));
}
inline void append_na(std::string& o, int v) { if (v == -1) o.push_back('.'); else append_int(o, v); }

inline void append_rle(std::string& o, const uint8_t* as2bit, uint32_t nk) {
	static const char chs[3] = {'*', '.', '='};
	if (!nk) { o.push_back('*'); return; }            // `string as1{'*'}` stays when the annotation is empty (cpp:1652)
	auto st = [&](uint32_t i) { return (as2bit[i >> 2] >> (2 * (i & 3))) & 3u; };
	uint32_t ct = 1, a0 = st(0);
	for (uint32_t i = 1; i < nk; ++i) {
		const uint32_t a1 = st(i);
		if (a1 != a0) { append_int(o, ct); o.push_back(chs[a0]); ct = 1; } else ++ct;
		a0 = a1;
	}
	append_int(o, ct); o.push_back(chs[a0]);
}

inline void append_fields(std::string& o, const dbt_mate_record& r) {
	auto flag = [&](int bit) { o.push_back((r.flags & bit) ? '1' : '0'); o.push_back(':'); };
	flag(DBT_MF_KF); flag(DBT_MF_HF); flag(DBT_MF_BF); flag(DBT_MF_QF); flag(DBT_MF_AF); flag(DBT_MF_RM);
	o.append("0:0:");                                  // qn, qm: qfilter is disabled at HEAD (cpp:2129-2135)
	append_na(o, r.si); o.push_back(':'); append_na(o, r.nt); o.push_back(':'); append_na(o, r.bs); o.push_back(':'); append_na(o, r.ti);
}

// rec: one record of dbt_record_stride bytes; t: the text of its pair (HostBatch, or the raw chunk the GPU split)
// `-s 1|2`: the source locus a simulated read carries in its (pruned) title.  simmode 1: ">LOCUS.rest" (parseReadName,
// cpp:477-489); simmode 2: ">CHR:START-END:LOCUS", LOCUS "." = not from a locus = nloci (cpp:492-506).  -2: malformed.
inline long long sim_source_locus(const PairText& t, int simmode, uint64_t nloci) {
	const std::string title(t.title, t.title_len);
	char* e = nullptr;
	if (simmode == 1) {
		if (title.size() < 2) return -2;
		const unsigned long v = strtoul(title.c_str() + 1, &e, 10);
		return e == title.c_str() + 1 ? -2 : (long long)v;
	}
	const size_t p1 = title.find(':');
	const size_t p2 = p1 == std::string::npos ? p1 : title.find(':', p1 + 1);
	if (p2 == std::string::npos || p2 + 1 >= title.size()) return -2;
	if (title[p2 + 1] == '.') return (long long)nloci;
	const unsigned long long v = strtoull(title.c_str() + p2 + 1, &e, 10);
	return e == title.c_str() + p2 + 1 ? -2 : (long long)(int)v;          // meta holds it as int (cpp:492)
}

// src < 0: no `-s` (the column reads "."); with `-s` the caller drops a record unless src != nloci or dst != nloci (cpp:2169)
inline void append_kam_line(std::string& o, const uint8_t* rec, uint32_t as_stride, const PairText& t, bool isFastq, long long src = -1) {
	const dbt_pair_record* h = reinterpret_cast<const dbt_pair_record*>(rec);
	const dbt_mate_record& r1 = h->mate[0];
	const dbt_mate_record& r2 = h->mate[1];
	if (src < 0) o.push_back('.'); else append_int(o, src);   // cpp:1653
	o.push_back('\t');
	append_int(o, h->dst); o.push_back('\t');
	append_int(o, (h->dst0 != h->dst) ? (long long)(int)h->dst0 : -1); o.push_back('\t');   // km_asgn_t::assign (cpp:115-119), int fields
	append_int(o, (int)r2.ei - (int)r2.si); o.push_back('\t');
	append_int(o, (int)r1.ei - (int)r1.si); o.push_back('\t');
	o.append("kf:hf:bf:qf:af:rm:qn:qm:si:nt:bs:ti\t");
	append_fields(o, r2); o.push_back('\t');
	append_fields(o, r1); o.push_back('\t');
	append_rle(o, rec + sizeof(dbt_pair_record) + as_stride, r2.nk); o.push_back('\t');
	append_rle(o, rec + sizeof(dbt_pair_record), r1.nk); o.push_back('\t');
	if (t.title_len > 1) o.append(t.title + 1, t.title_len - 1);                          // titles[..].substr(1)
	o.push_back('\t');
	o.append(t.seq[1], t.seq_len[1]); o.push_back('\t');                                  // read 2
	if (isFastq) o.append(t.qual[1], t.qual_len[1]); else o.push_back('.');
	o.push_back('\t');
	o.append(t.seq[0], t.seq_len[0]); o.push_back('\t');                                  // read 1
	if (isFastq) o.append(t.qual[0], t.qual_len[0]); else o.push_back('.');
	o.push_back('\n');
}
// `-e 1|2`: writeExtractedReads, cpp:1618-1644 -- read 2 first, then read 1, both under the pair's title; `-e 2` appends ":<locus>"
inline void append_extracted(std::string& o, const uint8_t* rec, const PairText& t, bool isFastq, int mode) {
	const dbt_pair_record* h = reinterpret_cast<const dbt_pair_record*>(rec);
	for (int m = 1; m >= 0; --m) {
		o.append(t.title, t.title_len);
		if (mode == 2) { o.push_back(':'); append_int(o, h->dst); }
		o.push_back('\n');
		o.append(t.seq[m], t.seq_len[m]); o.push_back('\n');
		if (isFastq) { o.append("+\n"); o.append(t.qual[m], t.qual_len[m]); o.push_back('\n'); }
	}
}

inline void append_kam_line(std::string& o, const uint8_t* rec, uint32_t as_stride, const HostBatch& hb, bool isFastq) {
	const uint64_t i = reinterpret_cast<const dbt_pair_record*>(rec)->pair_index - hb.first_pair_index;
	append_kam_line(o, rec, as_stride, pair_text(hb, i, isFastq), isFastq);
}

}  // namespace dbt_host
