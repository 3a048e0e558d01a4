// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// Function-level oracle for the `danbing-tk align` hot path.
// This translation unit #includes the UNMODIFIED reference source where it lies under
// /root/reference (nothing is copied into this repo) with its main() renamed, and drives the
// reference's own functions once per read pair, dumping the per-pair internals that the
// reference binary itself never prints (probe counts, top.fc/top.rc at the early stop, ...).
//
// The only code here that is not the reference's is the glue that sequences the calls; it
// follows the live (non-commented) statements of CountWords, aQueryFasta_thread.cpp:2002-2176,
// restricted to the default "asgn" mode (no -b, -bu, -s, -e, -g*). tests/test_oracle_ref.py
// cross-checks that glue against the reference binary's own kam lines, counters and .trkmc.ar.
//
// Build (oracle/Makefile):  g++ -std=c++11 -O2 -w -pthread -I/root/reference/src ref_harness.cpp
//
// Threading mode (P1; SURVEY.md §8(c) decision D1): the reference's isThreadFeasible and helpers
// (aQueryFasta_thread.cpp:526-1260) compile but their only caller is commented out at HEAD
// (cpp:2072-2088). With `thread=T` this harness re-creates that caller as GLUE (below, marked) --
// the statements of the commented blocks cpp:2072-2088 / 2195-2209 / 2232-2248 with the variable
// renames seq->seq1, seq1->seq2 -- and loads PREF.graph.kmers with the reference's readGraphKmers
// (aQueryFasta_thread.h:549-575). Every algorithmic line that runs is still the reference's.
//
// Usage: ref_harness PREF PAIRS.tsv OUT K CTH NFILTER NMFILTER NMTR OKAM [QCFILE] [key=value ...]
//   thread=T     -g T   (graph threading, thread_cth = T)
//   corr=0|1     0 = -g (no error correction), 1 = -gc (default 1)
//   maxnc=E      -gc T E (maxncorrection, default 4)
//   aln=0|1|2    1 = -a, 2 = -ae: legacy .aln lines on stdout through the reference's writeAlignments
//   trace=0      do not write OUT.trace.tsv (bench.py's timed runs: the dump is harness work, not the reference's)
//   PAIRS.tsv : one pair per line "seq1<TAB>seq2" in the reference's INTERNAL mate order
//               (seq1 = seqs[2i] = the mate that arrived second, aQueryFasta_thread.cpp:1972-1974)
//   writes OUT.trace.tsv, OUT.trkmc.ar, OUT.tr.summary.txt, OUT.counters.txt
#define main ref_main
#include "aQueryFasta_thread.cpp"
#undef main

static string rle(vector<int>& as) {
	km_asgn_t t;
	string s{'*'};
	t.annot2str_(as, s);
	return s;
}

int main(int argc, char* argv[]) {
	if (argc < 10) { cerr << "usage: ref_harness PREF PAIRS.tsv OUT K CTH NF NMF NMTR OKAM [QCFILE]\n"; return 2; }
	string pref = argv[1], pairsfn = argv[2], out = argv[3];
	ksize = stoul(argv[4]);
	rmask = (1ULL << 2*(ksize-1)) - 1;
	uint16_t Cthreshold = stoi(argv[5]);
	N_FILTER = stoul(argv[6]);
	NM_FILTER = stoul(argv[7]);
	NM_TR = stoul(argv[8]);
	bool okam = stoi(argv[9]);
	bool threading = false, correction = true;
	int aln = 0;
	bool dotrace = true;
	uint64_t thread_cth = 100;
	string qcfile;
	for (int i = 10; i < argc; ++i) {
		string a = argv[i];
		size_t eq = a.find('=');
		if (eq == string::npos) { qcfile = a; continue; }
		string key = a.substr(0, eq), val = a.substr(eq+1);
		if      (key == "thread") { threading = true; thread_cth = stoul(val); }
		else if (key == "corr")   { correction = stoi(val); }
		else if (key == "maxnc")  { maxncorrection = stoul(val); }
		else if (key == "aln")    { aln = stoi(val); }
		else if (key == "trace")  { dotrace = stoi(val); }
		else { cerr << "unknown option " << a << '\n'; return 2; }
	}
	bool qc = qcfile.size();

	string trFname = pref + ".tr.kmers";
	uint64_t nloci = countLoci(trFname);
	vector<kmer_aCount_umap> trKmerDB(nloci);
	kset_db_t flankDB;
	kmerIndex_uint32_umap kmerDBi;
	vector<uint32_t> kmerDBi_vv;
	vector<uint8_t> qcFilter(nloci);
	vector<atomic_uint32_t> nmapread(nloci);
	vector<atomic_uint64_t> kmc(nloci);
	readBinaryIndex(kmerDBi, kmerDBi_vv, pref);
	readBinaryKmerSetDB(flankDB, pref+".fl");
	readKmersWithZeroCount(trKmerDB, trFname);
	if (qc) { readQCFile(qcFilter, qcfile); }
	vector<GraphType> graphDB(nloci);
	if (threading) { readGraphKmers(graphDB, pref + ".graph.kmers"); }

	vector<uint32_t> hits1(nloci+1,0), hits2(nloci+1,0);
	vector<string> seqs, titles;          // for writeAlignments (cpp:1742-1759)
	vector<uint64_t> alnindices;
	vector<sam_t> sams;
	uint64_t nReads = 0, nShort = 0, nSub = 0, nKf = 0, nLaf = 0, nQC = 0, nThr = 0, nFea = 0, nAsgn = 0, nh0 = 0, nh1 = 0;

	ifstream fin(pairsfn);
	assert(fin);
	ofstream tr(out + ".trace.tsv");
	string line;
	uint64_t idx = 0;
	while (getline(fin, line)) {
		size_t tab = line.find('\t');
		string s1 = line.substr(0, tab), s2 = line.substr(tab+1);
		nReads += 2;
		// ---- glue: aQueryFasta_thread.cpp:2004-2062 ----
		vector<uint64_t> caks1, caks2, caes1, caes2;
		vector<kmerIndex_uint32_umap::iterator> its1, its2;
		vector<PE_KMC> dup;
		log_t log;
		int rm1 = 0, rm2 = 0, kf1 = 0, kf2 = 0, hf1 = 0, hf2 = 0, af1 = 0, af2 = 0;
		int nm1 = 0, nm2 = 0;
		uint64_t nhash0 = 0, nhash1 = 0, destLocus = nloci, destLocus0 = NAN32;
		int status = -1;
		km_asgn_t kam;

		read2kmers_edges(caks1, caes1, s1, ksize);
		read2kmers_edges(caks2, caes2, s2, ksize);
		if (not caks1.size() or not caks2.size()) { ++nShort; status = 0; }
		if (status < 0 and N_FILTER and NM_FILTER) {
			if (subfilter(caks1, caks2, kmerDBi, nhash0)) { nSub += 2; status = 1; }
		}
		if (status < 0) {
			kfilter(caks1, caks2, its1, its2, kmerDBi, Cthreshold, nhash1, kf1, kf2, rm1, rm2);
			nKf += kf1 + kf2;
			if (rm1 and rm2) { status = 2; }
		}
		if (status < 0) {
			destLocus = countHit(kmerDBi_vv, its1, its2, hits1, hits2, dup, nloci, Cthreshold, log, destLocus0, nm1, nm2, hf1, hf2, rm1, rm2);
			nLaf += hf1 + hf2;
			if (destLocus == nloci) { status = 3; }
		}
		if (status < 0 and qc and not qcFilter[destLocus]) { nQC += 2 - rm1 - rm2; status = 4; }
		string thr_dump;
		if (status < 0 and threading) {
			// ---- GLUE (decision D1): the commented caller cpp:2072-2088, counting cpp:2195-2209
			//      without the raw-k-mer double count, .aln selection cpp:2232-2239 (no simmode) ----
			nThr += 2;
			sam_t sam;
			vector<uint64_t> nonca[2], ak[2];
			int alned[2] = {0, 0};
			string* ss[2] = { &s1, &s2 };
			cigar_t* cgs[2] = { &sam.r1, &sam.r2 };
			for (int m = 0; m < 2; ++m) {
				cgs[m]->init(*ss[m]);
				// a read without any valid k-mer leaves `kmers` empty and the reference then reads kmers[0]
				// (cpp:1124): undefined behaviour. D1 defines such a mate as not threadable, k-mer vector empty.
				bool anyvalid = false;
				{ vector<uint64_t> t; read2kmers(t, *ss[m], ksize, 0, 0, false, true); anyvalid = t.size(); }
				if (anyvalid) { alned[m] = isThreadFeasible(graphDB[destLocus], *ss[m], nonca[m], ak[m], thread_cth, correction, *cgs[m], trKmerDB[destLocus], log); }
			}
			uint64_t dst_before = destLocus;
			if (alned[0] or alned[1]) {
				nFea += 2;
				kmer_aCount_umap& trKmers = trKmerDB[destLocus];
				for (int m = 0; m < 2; ++m) {
					for (uint64_t km : ak[m]) {
						if (km == -1ULL) { continue; }
						auto it = trKmers.find(toCaKmer(km, ksize));
						if (it != trKmers.end()) { ++(it->second); }
					}
				}
				status = 6;
			}
			else { destLocus = nloci; status = 5; }
			if (aln == 1 or (aln == 2 and destLocus != nloci)) {
				seqs.push_back(s1); seqs.push_back(s2);
				titles.resize(seqs.size()/2, ">p" + to_string(idx));   // titles[] keep the leading '>' (cpp:1747 prints them whole)
				alnindices.push_back(seqs.size());
				sam.src = -1;
				sam.dst = destLocus;
				sams.push_back(sam);
			}
			ostringstream td;
			if (dotrace) {
			td << '\t' << dst_before;
			for (int m = 0; m < 2; ++m) {
				td << '\t' << alned[m] << '\t';
				for (auto& e : cgs[m]->es) { td << (char)(e.t) << (char)(e.g ? e.g : '_'); }
				td << '\t';
				for (char c : cgs[m]->tr) { td << c; }
				td << '\t';
				for (size_t i = 0; i < ak[m].size(); ++i) { if (i) { td << ','; } td << (int64_t)ak[m][i]; }
				if (not ak[m].size()) { td << '.'; }
			}
			}
			thr_dump = td.str();
		}
		if (status < 0) {
			// ---- glue: aQueryFasta_thread.cpp:2070-2175 (threading == false, no bait) ----
			nThr += 2;
			nFea += 2;
			vector<kmer_aCount_umap::iterator> kits1, kits2;
			if (okam or not rm1 or not rm2) {
				kmer_aCount_umap &trKmers = trKmerDB[destLocus0];
				unordered_set<uint64_t>& flKmers = flankDB[destLocus0];
				assignTRkmc(caks1, trKmers, flKmers, kits1, kam.r1, af1, rm1, okam);
				assignTRkmc(caks2, trKmers, flKmers, kits2, kam.r2, af2, rm2, okam);
			}
			if (rm1 and rm2) { destLocus = nloci; status = 5; }
			else {
				int n = 2 - rm1 - rm2;
				nmapread[destLocus] += n;
				nAsgn += n;
				kmc[destLocus] += (kam.r1.ei - kam.r1.si) + (kam.r2.ei - kam.r2.si);
				auto& as1 = kam.r1.as;
				auto& as2 = kam.r2.as;
				if (not rm1) { for (int i = 0; i < as1.size(); ++i) { if (as1[i] == 2) { ++(kits1[i]->second); } } }
				if (not rm2) { for (int i = 0; i < as2.size(); ++i) { if (as2[i] == 2) { ++(kits2[i]->second); } } }
				status = 6;
			}
		}
		nh0 += nhash0;
		nh1 += nhash1;
		// ---- dump ----
		if (not dotrace) { ++idx; continue; }
		tr << idx << '\t' << status << '\t' << nhash0 << '\t' << nhash1 << '\t'
		   << (status >= 3 ? (int64_t)destLocus : -1) << '\t' << (status >= 3 ? (int64_t)destLocus0 : -1) << '\t'
		   << (status >= 3 ? nm1 : 0) << '\t' << (status >= 3 ? nm2 : 0);
		km_asgn_read_t* rs[2] = { &kam.r1, &kam.r2 };
		int kf[2] = {kf1, kf2}, hf[2] = {hf1, hf2}, af[2] = {af1, af2}, rm[2] = {rm1, rm2};
		for (int m = 0; m < 2; ++m) {
			km_asgn_read_t& r = *rs[m];
			tr << '\t' << kf[m] << ':' << hf[m] << ':' << af[m] << ':' << rm[m] << ':'
			   << r.si << ':' << r.ei << ':' << r.si_ << ':' << r.ei_ << ':' << r.nt << ':' << r.bs << ':' << r.ti
			   << '\t' << rle(r.as);
		}
		tr << thr_dump << '\n';
		++idx;
	}
	tr.close();
	if (aln) { writeAlignments(seqs, titles, alnindices, sams); }

	dumpTRKmers(out, trKmerDB);
	writeTRKmerSummary(out+".tr.summary.txt", kmc, nmapread);
	ofstream fc(out + ".counters.txt");
	fc << "nReads\t" << nReads << "\nnShort\t" << nShort << "\nnSubFiltered\t" << nSub << "\nnKmerFiltered\t" << nKf
	   << "\nnLocusAssignFiltered\t" << nLaf << "\nnQCFiltered\t" << nQC << "\nnThreadingReads\t" << nThr
	   << "\nnFeasibleReads\t" << nFea << "\nnAsgnReads\t" << nAsgn << "\nnhash0\t" << nh0 << "\nnhash1\t" << nh1 << '\n';
	return 0;
}
