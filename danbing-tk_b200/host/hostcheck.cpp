// dbt-hostcheck: exercises the host-side pieces of the CLI without a GPU (used by tests/test_host.py).
//   tr-order TR.kmers OUT.u64                      loader + libstdc++ iteration-order replay -> u64 nloci | n[nloci] | kmers[]
//   index PREF [QCFILE]                            loads PREF.{tr.kmers,kmers.dbi,fl.kdb}; prints sizes
//   pair FASTX isFastq minReadSize readsPerRefBatch pairsCap    reader/pairer -> "title\tseq1\tseq2[\tqual1\tqual2]" per pair
//   kam RECORDS.bin stride as_stride FASTX isFastq minReadSize   formats records (e.g. from the CPU oracle) as kam lines
//   args ...                                        parses a command line, prints the resolved options
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "options.hpp"
#include "bait.hpp"
#include "serialize.hpp"
#include "fastx_reader.hpp"
#include "kam_writer.hpp"
#include "raw_ingest.hpp"

using namespace dbt_host;

static bool alloc_batch(HostBatch& hb, uint64_t pairs, uint64_t bytes) {
	hb.bases = (uint8_t*)malloc(bytes); hb.bases_cap = bytes;
	hb.offsets = (uint64_t*)malloc((2 * pairs + 1) * 8); hb.pairs_cap = pairs;
	return hb.bases && hb.offsets;
}

static int cmd_tr_order(const char* in, const char* out) {
	Rpgg g; std::string err;
	if (!read_tr_kmers_ordered(in, g, 4, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
	FILE* f = fopen(out, "wb");
	if (!f) return 1;
	fwrite(&g.nloci, 8, 1, f); fwrite(g.tr_n.data(), 8, g.tr_n.size(), f); fwrite(g.tr_kmers.data(), 8, g.tr_kmers.size(), f);
	fclose(f);
	return 0;
}

static int cmd_index(const std::string& pref, const std::string& qcfn) {
	Rpgg g; std::string err;
	if (!load_rpgg(pref, pref + ".tr.kmers", qcfn, 4, g, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
	printf("nloci %llu nk_dbi %zu nvv %zu ntr %zu nfl %zu nqc %zu\n", (unsigned long long)g.nloci, g.dbi_keys.size(), g.vv.size(),
	       g.tr_kmers.size(), g.fl_kmers.size(), g.qc.size());
	return 0;
}

// order-independent digest of the loaded index (legacy-container tests: the same index must come out of either file set)
static int cmd_index_sum(const std::string& pref) {
	Rpgg g; std::string err;
	if (!load_rpgg(pref, pref + ".tr.kmers", "", 4, g, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
	uint64_t keysum = 0, kvsum = 0, vvsum = 0, flsum = 0;
	for (size_t i = 0; i < g.dbi_keys.size(); ++i) { keysum += g.dbi_keys[i]; kvsum += g.dbi_keys[i] * (uint64_t)(g.dbi_vals[i] + 1u); }
	for (size_t i = 0; i < g.vv.size(); ++i) vvsum = vvsum * 1000003u + g.vv[i];
	size_t o = 0;
	for (size_t l = 0; l < g.fl_n.size(); ++l) for (uint64_t i = 0; i < g.fl_n[l]; ++i) flsum += g.fl_kmers[o++] * (uint64_t)(l + 1);
	printf("nloci %llu nk_dbi %zu nvv %zu ntr %zu nfl %zu keysum %llu kvsum %llu vvsum %llu flsum %llu\n", (unsigned long long)g.nloci, g.dbi_keys.size(),
	       g.vv.size(), g.tr_kmers.size(), g.fl_kmers.size(), (unsigned long long)keysum, (unsigned long long)kvsum, (unsigned long long)vvsum, (unsigned long long)flsum);
	return 0;
}

// RawReader + StreamPairer with the GPU's record split (csrc/fastx.cu) restated on the host for the test: neighbours with
// equal pruned titles pair up, runs of three raise the slow-path flag, the rest are singletons.  Prints the pairs in output order.
static int cmd_pair_raw(const char* fn, bool fq, uint64_t minlen, uint64_t refbatch, uint64_t target) {
	RawReader rd(fq, target >= (4u << 20) ? 4 : 1);
	if (!rd.open(fn)) { fprintf(stderr, "cannot open %s\n", fn); return 1; }
	RawChunk c;
	c.cap = RawReader::HEADROOM + target + 4096;
	std::vector<uint8_t> buf(c.cap + 64);
	c.buf = buf.data();
	StreamPairer pr(fq, minlen, refbatch);
	StreamPairer::Extra ex;
	const uint32_t L = fq ? 4 : 2;
	uint64_t nfast = 0, nslow = 0;
	while (rd.fill(c, target)) {
		// ---- the device split, on the host ----
		const uint8_t* t = c.text(); const uint64_t n = c.nbytes();
		std::vector<uint32_t> rec;                              // record start offsets
		{ uint64_t p = 0; while (p < n) { rec.push_back((uint32_t)p); for (uint32_t i = 0; i < L; ++i) p = (uint64_t)((const uint8_t*)rawmemchr(t + p, '\n') - t) + 1; } }
		auto title = [&](uint32_t r) { const uint8_t* a = t + rec[r]; const uint8_t* e = (const uint8_t*)rawmemchr(a, '\n'); return std::string((const char*)a, pruned_len(a, (size_t)(e - a))); };
		auto slen = [&](uint32_t r) { const uint8_t* a = (const uint8_t*)rawmemchr(t + rec[r], '\n') + 1; return (uint64_t)((const uint8_t*)rawmemchr(a, '\n') - a); };
		const uint32_t nr = (uint32_t)rec.size();
		std::vector<uint8_t> eq(nr + 2, 0);
		for (uint32_t r = 1; r < nr; ++r) eq[r] = title(r) == title(r - 1);
		std::vector<uint32_t> po, so; std::vector<uint64_t> ph; uint32_t flags = 0, nshort = 0;
		for (uint32_t r = 0; r < nr; ++r) {
			const bool e0 = r ? eq[r - 1] : false, e1 = eq[r], e2 = eq[r + 1];
			if (e1 && (e0 || e2)) flags |= DBT_RAW_SLOWPATH;
			else if (e1) { if (slen(r) >= minlen && slen(r - 1) >= minlen) { po.push_back(rec[r - 1]); const std::string tt = title(r); uint64_t h = 0xcbf29ce484222325ull; for (unsigned char ch : tt) { h ^= ch; h *= 0x100000001b3ull; } ph.push_back(h); } else ++nshort; }
			else if (!e2) so.push_back(rec[r]);
		}
		c.pair_offset = po.data(); c.pair_hash = ph.data(); c.single_offset = so.data();
		const bool fast = pr.chunk(c, true, flags, (uint32_t)po.size(), (uint32_t)so.size(), nshort, ex);
		(fast ? nfast : nslow) += 1;
		// ---- merged output: GPU pairs (key = pair_offset) and host pairs (key) ascending ----
		size_t i = 0, j = 0; const size_t ng = fast ? po.size() : 0, nh = ex.key.size();
		while (i < ng || j < nh) {
			PairText p;
			if (j >= nh || (i < ng && (uint64_t)po[i] < ex.key[j])) p = pair_text_raw(t, po[i++], fq); else p = pair_text(ex.hb, j++, fq);
			fwrite(p.title, 1, p.title_len, stdout); putchar('\t'); fwrite(p.seq[0], 1, p.seq_len[0], stdout); putchar('\t'); fwrite(p.seq[1], 1, p.seq_len[1], stdout);
			if (fq) { putchar('\t'); fwrite(p.qual[0], 1, p.qual_len[0], stdout); putchar('\t'); fwrite(p.qual[1], 1, p.qual_len[1], stdout); }
			putchar('\n');
		}
	}
	if (!rd.error().empty()) { fprintf(stderr, "%s\n", rd.error().c_str()); return 1; }
	fprintf(stderr, "reads %llu unpaired %llu fast %llu slow %llu\n", (unsigned long long)(2 * pr.total_pairs()), (unsigned long long)pr.unpaired(), (unsigned long long)nfast, (unsigned long long)nslow);
	return 0;
}

static int cmd_pair(const char* fn, bool fq, uint64_t minlen, uint64_t refbatch, uint64_t cap) {
	FastxReader rd(fq, minlen, refbatch);
	if (!rd.open(fn)) { fprintf(stderr, "cannot open %s\n", fn); return 1; }
	HostBatch hb;
	if (!alloc_batch(hb, cap, cap * 2 * 400)) return 1;
	uint64_t first = 0;
	for (;;) {
		hb.reset(first);
		if (!rd.fill(hb)) break;
		for (uint64_t i = 0; i < hb.npairs; ++i) {
			fwrite(hb.names.data() + hb.name_off[i], 1, hb.name_off[i + 1] - hb.name_off[i], stdout); fputc('\t', stdout);
			fwrite(hb.bases + hb.offsets[2 * i], 1, hb.offsets[2 * i + 1] - hb.offsets[2 * i], stdout); fputc('\t', stdout);
			fwrite(hb.bases + hb.offsets[2 * i + 1], 1, hb.offsets[2 * i + 2] - hb.offsets[2 * i + 1], stdout);
			if (fq) {
				fputc('\t', stdout); fwrite(hb.quals.data() + hb.qual_off[2 * i], 1, hb.qual_off[2 * i + 1] - hb.qual_off[2 * i], stdout);
				fputc('\t', stdout); fwrite(hb.quals.data() + hb.qual_off[2 * i + 1], 1, hb.qual_off[2 * i + 2] - hb.qual_off[2 * i + 1], stdout);
			}
			fputc('\n', stdout);
		}
		first += hb.npairs;
	}
	fprintf(stderr, "reads %llu unpaired %llu\n", (unsigned long long)rd.total_reads(), (unsigned long long)rd.unpaired());
	return 0;
}

static int cmd_kam(const char* recfn, uint32_t stride, uint32_t as_stride, const char* fx, bool fq, uint64_t minlen) {
	std::vector<char> recs; std::string err;
	if (!read_whole_file(recfn, recs, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
	FastxReader rd(fq, minlen, 300000);
	if (!rd.open(fx)) { fprintf(stderr, "cannot open %s\n", fx); return 1; }
	HostBatch hb;
	if (!alloc_batch(hb, 1 << 16, (uint64_t)(1 << 16) * 2 * 400)) return 1;
	uint64_t first = 0; size_t ri = 0; const size_t nrec = recs.size() / stride;
	std::string out;
	for (;;) {
		hb.reset(first);
		if (!rd.fill(hb)) break;
		out.clear();
		while (ri < nrec) {
			const dbt_pair_record* h = reinterpret_cast<const dbt_pair_record*>(recs.data() + ri * stride);
			if (h->pair_index >= first + hb.npairs) break;
			append_kam_line(out, reinterpret_cast<const uint8_t*>(h), as_stride, hb, fq);
			++ri;
		}
		fwrite(out.data(), 1, out.size(), stdout);
		first += hb.npairs;
	}
	return ri == nrec ? 0 : 3;
}

static int cmd_args(int argc, char** argv) {
	Options o;
	std::string err = parse_args(argc, argv, o);
	if (!err.empty()) { printf("error: %s\n", err.c_str()); return 1; }
	printf("k=%llu kf=%llu,%llu cth=%llu c=%llu p=%llu r=%g threading=%d thread_cth=%llu maxnc=%llu tc=%d aln=%d alnmin=%d okam=%d qc=%d fq=%d named=%d tr=%s fx=%s out=%s qcfn=%s gpus=%d\n",
	       (unsigned long long)o.ksize, (unsigned long long)o.N_FILTER, (unsigned long long)o.NM_FILTER, (unsigned long long)o.Cthreshold,
	       (unsigned long long)o.NM_TR, (unsigned long long)o.nproc, o.readsPerBatchFactor, o.threading, (unsigned long long)o.thread_cth,
	       (unsigned long long)o.maxncorrection, o.tc, o.aln, o.aln_minimal, o.okam, o.qc, o.isFastq, o.writeKmerName, o.trFname.c_str(),
	       o.fastxFname.c_str(), o.outPrefix.c_str(), o.qcFn.c_str(), o.gpus);
	report_parameters(o, stderr);
	return 0;
}

int main(int argc, char** argv) {
	if (argc < 2) { fprintf(stderr, "usage: dbt-hostcheck tr-order|index|pair|kam|args ...\n"); return 2; }
	const std::string cmd = argv[1];
	if (cmd == "tr-order" && argc == 4) return cmd_tr_order(argv[2], argv[3]);
	if (cmd == "index" && argc >= 3) return cmd_index(argv[2], argc > 3 ? argv[3] : "");
	if (cmd == "index-sum" && argc == 3) return cmd_index_sum(argv[2]);
	if (cmd == "pair-raw" && argc == 7) return cmd_pair_raw(argv[2], atoi(argv[3]) != 0, strtoull(argv[4], nullptr, 10), strtoull(argv[5], nullptr, 10), strtoull(argv[6], nullptr, 10));
	if (cmd == "pair" && argc == 7) return cmd_pair(argv[2], atoi(argv[3]) != 0, strtoull(argv[4], nullptr, 10), strtoull(argv[5], nullptr, 10), strtoull(argv[6], nullptr, 10));
	if (cmd == "kam" && argc == 8) return cmd_kam(argv[2], (uint32_t)atoi(argv[3]), (uint32_t)atoi(argv[4]), argv[5], atoi(argv[6]) != 0, strtoull(argv[7], nullptr, 10));
	if (cmd == "qmask" && argc == 4) {                      // stdin: one quality string per line -> the bait histogram's k-mer mask as 0/1
		std::string line; std::vector<bool> m;
		char buf[1 << 16];
		while (fgets(buf, sizeof buf, stdin)) {
			size_t n = strlen(buf); while (n && (buf[n - 1] == '\n' || buf[n - 1] == '\r')) --n;
			quality_mask(buf, n, atoi(argv[2]), atoi(argv[3]), m);
			line.clear(); for (bool x : m) line.push_back(x ? '1' : '0');
			puts(line.c_str());
		}
		return 0;
	}
	if (cmd == "args") return cmd_args(argc - 1, argv + 1);
	fprintf(stderr, "bad arguments\n");
	return 2;
}
