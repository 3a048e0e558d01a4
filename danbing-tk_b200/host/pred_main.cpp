// danbing-tk-pred, B200 build: drop-in for the reference's bias-correction tool (src/pred.cpp), SURVEY §8(f) row 4.
//
//   danbing-tk-pred [--device N] <trkmers.meta.txt> <ikmer.meta> <raw.gt> <corrected.gt> <bias.tsv>
//
// Same inputs (metadata file "OUT.trkmc.ar <TAB> read depth" per sample, pred.h:41-50; binary invariant k-mer table,
// pred.h:69-123; each OUT.trkmc.ar = u64 nk | u64[nk], pred.h:141-164) and the same three outputs, byte for byte: the two
// matrices as save_matrix writes them (u32 rows, u32 cols, column-major float32, pred.h:197-210) and the bias matrix as Eigen
// prints it with IOFormat(StreamPrecision, DontAlignCols, "\t", "\n") (pred.cpp:49, pred.h:212-219).  The arithmetic runs on
// the GPU (dbt_pred, csrc/pred.cu); there is no CPU fallback.  One deviation: the bias column of a locus the reference skips
// (no k-mers / no invariant k-mers) is 0 here, uninitialised memory there.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <algorithm>
#include <memory>
#include <thread>
#include <string>
#include <vector>
#include "danbing_b200.h"

namespace {

struct Mapped { void* p = MAP_FAILED; size_t n = 0; };

bool map_file(const std::string& fn, Mapped& m, std::string& err) {
	int fd = open(fn.c_str(), O_RDONLY);
	if (fd < 0) { err = "cannot open " + fn; return false; }
	struct stat st;
	if (fstat(fd, &st) != 0) { close(fd); err = "cannot stat " + fn; return false; }
	m.n = (size_t)st.st_size;
	if (m.n) m.p = mmap(nullptr, m.n, PROT_READ, MAP_PRIVATE, fd, 0);
	close(fd);
	if (m.n && m.p == MAP_FAILED) { err = "cannot map " + fn; return false; }
	return true;
}

bool write_matrix(const std::string& fn, const float* d, uint64_t nrow, uint64_t ncol) {
	FILE* f = fopen(fn.c_str(), "wb");
	if (!f) return false;
	const uint32_t r = (uint32_t)nrow, c = (uint32_t)ncol;          // the reference writes the low 4 bytes of its size_t's
	bool ok = fwrite(&r, 4, 1, f) == 1 && fwrite(&c, 4, 1, f) == 1;
	const size_t n = (size_t)nrow * ncol;
	ok = ok && (n == 0 || fwrite(d, 4, n, f) == n);
	return fclose(f) == 0 && ok;
}

}  // namespace

int main(int argc, char** argv) {
	if (argc < 2) {
		fprintf(stderr, "\nUsage: danbing-tk-pred [--device N] <INPUT1> <INPUT2> <OUTPUT1> <OUTPUT2> <OUTPUT3>\n"
		                "INPUT1      metadata of *.trkmc.ar files, consisting of 2 columns.\n"
		                " col1       *.trkmc.ar file name\n"
		                " col2       read depth\n"
		                "INPUT2      invariant kmers of an RPGG build\n"
		                "OUTPUT1     raw genotype matrix. Row: sample. Column: kmer.\n"
		                "OUTPUT2     bias-corrected genotype matrix. Row: sample. Column: kmer.\n"
		                "OUTPUT3     bias matrix. Row: sample. Column: TR locus.\n\n");
		return 0;
	}
	int argi = 1, device = 0;
	while (argi < argc && argv[argi][0] == '-') {
		if (!strcmp(argv[argi], "--device") && argi + 1 < argc) { device = atoi(argv[argi + 1]); argi += 2; }
		else { fprintf(stderr, "danbing-tk-pred: unsupported option %s (the reference's developer switch -f does nothing there either)\n", argv[argi]); return 1; }
	}
	if (argc - argi != 5) { fprintf(stderr, "danbing-tk-pred: expected 5 file arguments\n"); return 1; }
	// the CUDA context (about half a second) comes up on a helper thread while the inputs are parsed and mapped
	const auto tstart = std::chrono::steady_clock::now();
	auto since = [&](std::chrono::steady_clock::time_point a) { return std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count(); };
	double t_parse = 0, t_initwait = 0, t_alloc = 0, t_pred = 0, t_write = 0;
	int init_rc = 0;
	std::thread warm([&] { init_rc = dbt_init_device(device); });
	struct Joiner { std::thread& t; ~Joiner() { if (t.joinable()) t.join(); } } warm_guard{warm};
	const std::string finGtMeta = argv[argi], finIkMeta = argv[argi + 1], foutRaw = argv[argi + 2], fout = argv[argi + 3], foutBias = argv[argi + 4];
	printf("metadata of *.trkmc.ar: %s\ninvariant kmers: %s\nraw genotype matrix will be written to: %s\n"
	       "bias-corrected genotype matrix will be written to: %s\nbias matrix will be written to: %s\n",
	       finGtMeta.c_str(), finIkMeta.c_str(), foutRaw.c_str(), fout.c_str(), foutBias.c_str());

	// ---- read_gt_meta, pred.h:41-50 ----
	std::vector<std::string> fns; std::vector<float> rds;
	{
		std::ifstream fin(finGtMeta);
		if (!fin) { fprintf(stderr, "danbing-tk-pred: cannot open %s\n", finGtMeta.c_str()); return 1; }
		std::string f1, f2;
		while (std::getline(fin, f1, '\t') && std::getline(fin, f2)) {
			fns.push_back(f1);
			try { rds.push_back(std::stof(f2)); } catch (...) { fprintf(stderr, "danbing-tk-pred: %s: bad read depth '%s'\n", finGtMeta.c_str(), f2.c_str()); return 1; }
		}
	}
	const uint64_t ns = fns.size();
	if (!ns) { fprintf(stderr, "danbing-tk-pred: %s lists no sample\n", finGtMeta.c_str()); return 1; }

	// ---- read_ikmer, pred.h:69-123 (little-endian: u64 nk, n_ik, n_tr | u32 nk_cum[n_tr] | u32 nik_cum[n_tr] | n_ik x {u32, u8}) ----
	std::string err;
	Mapped ik;
	if (!map_file(finIkMeta, ik, err)) { fprintf(stderr, "danbing-tk-pred: %s\n", err.c_str()); return 1; }
	if (ik.n < 24) { fprintf(stderr, "danbing-tk-pred: %s: truncated\n", finIkMeta.c_str()); return 1; }
	const uint8_t* ib = static_cast<const uint8_t*>(ik.p);
	uint64_t nk, n_ik, n_tr;
	memcpy(&nk, ib, 8); memcpy(&n_ik, ib + 8, 8); memcpy(&n_tr, ib + 16, 8);
	if (ik.n != 24 + n_tr * 8 + n_ik * 5) { fprintf(stderr, "danbing-tk-pred: %s: %zu bytes, expected %llu\n", finIkMeta.c_str(), ik.n, (unsigned long long)(24 + n_tr * 8 + n_ik * 5)); return 1; }
	printf("%llu loci in total.\n%llu/%llu kmers are invariant.\n", (unsigned long long)n_tr, (unsigned long long)n_ik, (unsigned long long)nk);
	std::vector<uint32_t> nk_cum(n_tr), nik_cum(n_tr), iki(n_ik); std::vector<uint8_t> ikmc(n_ik);
	memcpy(nk_cum.data(), ib + 24, n_tr * 4);
	memcpy(nik_cum.data(), ib + 24 + n_tr * 4, n_tr * 4);
	for (uint64_t i = 0; i < n_ik; ++i) { memcpy(&iki[i], ib + 24 + n_tr * 8 + i * 5, 4); ikmc[i] = ib[24 + n_tr * 8 + i * 5 + 4]; }

	// ---- load_eachBinGT, pred.h:141-164: the count arrays stay where the page cache has them ----
	const auto t0 = std::chrono::steady_clock::now();
	printf("reading %llu gt files\n", (unsigned long long)ns);
	std::vector<Mapped> maps(ns); std::vector<const uint64_t*> kmc(ns);
	for (uint64_t s = 0; s < ns; ++s) {
		if (!map_file(fns[s], maps[s], err)) { fprintf(stderr, "danbing-tk-pred: %s\n", err.c_str()); return 1; }
		uint64_t n = 0;
		if (maps[s].n >= 8) memcpy(&n, maps[s].p, 8);
		if (maps[s].n < 8 || n != nk) { fprintf(stderr, "nk %llu != nk_ %llu\n", (unsigned long long)n, (unsigned long long)nk); return 1; }   // pred.h:154
		if (maps[s].n != 8 + nk * 8) { fprintf(stderr, "danbing-tk-pred: %s: %zu bytes, expected %llu\n", fns[s].c_str(), maps[s].n, (unsigned long long)(8 + nk * 8)); return 1; }
		kmc[s] = reinterpret_cast<const uint64_t*>(static_cast<const uint8_t*>(maps[s].p) + 8);
	}

	// result buffers: pageable, but touched by a few threads while the CUDA context is still coming up, so that the device-to-host
	// copies land in pages that exist (pinning 2 x ns*nk*4 bytes instead costs 0.3 s per 0.8 GB, more than the copies themselves)
	const size_t nelem = (size_t)ns * nk;
	std::unique_ptr<float[]> raw_mem(new float[nelem]), cor_mem(new float[nelem]);
	{
		std::vector<std::thread> th;
		const unsigned nt = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
		for (unsigned i = 0; i < nt; ++i) th.emplace_back([&, i] {
			for (float* base : {raw_mem.get(), cor_mem.get()}) {
				char* b = reinterpret_cast<char*>(base);
				const size_t bytes = nelem * 4, lo = bytes * i / nt, hi = bytes * (i + 1) / nt;
				for (size_t o = lo; o < hi; o += 4096) b[o] = 0;
			}
		});
		for (auto& t : th) t.join();
	}
	t_parse = since(tstart);
	auto tp = std::chrono::steady_clock::now();
	warm.join();
	t_initwait = since(tp); tp = std::chrono::steady_clock::now();
	if (init_rc) { fprintf(stderr, "danbing-tk-pred: %s: %s\n", dbt_strerror(init_rc), dbt_last_error()); return 1; }
	struct { float* p; float* data() { return p; } } raw{raw_mem.get()}, cor{cor_mem.get()};
	std::vector<float> bias((size_t)ns * n_tr, 0.f);
	t_alloc = since(tp); tp = std::chrono::steady_clock::now();
	dbt_pred_input in{};
	in.ns = ns; in.nk = nk; in.n_tr = n_tr; in.n_ik = n_ik; in.kmc = kmc.data(); in.rd = rds.data();
	in.nk_cum = nk_cum.data(); in.nik_cum = nik_cum.data(); in.iki = iki.data(); in.ikmc = ikmc.data();
	float kms[2] = {0.f, 0.f};
	printf("normalizaing read depth\ncomputing/correcting bias\n");
	const int rc = dbt_pred(device, &in, raw.data(), cor.data(), bias.data(), kms);
	if (rc) { fprintf(stderr, "danbing-tk-pred: %s: %s\n", dbt_strerror(rc), dbt_last_error()); return 1; }
	t_pred = since(tp); tp = std::chrono::steady_clock::now();
	const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	fprintf(stderr, "[danbing-tk-pred] GPU %d: %llu samples x %llu k-mers, %llu loci: normalise %.3f ms, bias correction %.3f ms (kernels); read + compute %.3f s\n",
	        device, (unsigned long long)ns, (unsigned long long)nk, (unsigned long long)n_tr, kms[0], kms[1], sec);

	// the three outputs are written side by side: two threads stream the matrices out while this one formats the bias table
	printf("saving matrix to %s\nsaving matrix to %s\nsaving matrix to %s\n", foutRaw.c_str(), fout.c_str(), foutBias.c_str());
	bool ok_raw = false, ok_cor = false;
	std::thread w_raw([&] { ok_raw = write_matrix(foutRaw, raw.data(), ns, nk); });
	std::thread w_cor([&] { ok_cor = write_matrix(fout, cor.data(), ns, nk); });
	struct Join2 { std::thread &a, &b; ~Join2() { if (a.joinable()) a.join(); if (b.joinable()) b.join(); } } wguard{w_raw, w_cor};
	{	// Eigen's operator<< with StreamPrecision: every coefficient through an ostream at its default precision (6 significant digits, %g)
		std::string out;
		out.reserve((size_t)ns * n_tr * 10);
		char buf[64];
		for (uint64_t s = 0; s < ns; ++s) {
			if (s) out.push_back('\n');
			for (uint64_t t = 0; t < n_tr; ++t) {
				if (t) out.push_back('\t');
				out.append(buf, (size_t)snprintf(buf, sizeof buf, "%g", (double)bias[s + t * ns]));
			}
		}
		FILE* f = fopen(foutBias.c_str(), "wb");
		if (!f || fwrite(out.data(), 1, out.size(), f) != out.size() || fclose(f) != 0) { fprintf(stderr, "danbing-tk-pred: cannot write %s\n", foutBias.c_str()); return 1; }
	}
	w_raw.join(); w_cor.join();
	if (!ok_raw) { fprintf(stderr, "danbing-tk-pred: cannot write %s\n", foutRaw.c_str()); return 1; }
	if (!ok_cor) { fprintf(stderr, "danbing-tk-pred: cannot write %s\n", fout.c_str()); return 1; }
	t_write = since(tp);
	fprintf(stderr, "[danbing-tk-pred] phases: parse+map+touch buffers %.3f s, wait for the CUDA context %.3f s, dbt_pred %.3f s, write %.3f s, total %.3f s\n",
	        t_parse, t_initwait, t_alloc + t_pred, t_write, since(tstart));
	return 0;
}
