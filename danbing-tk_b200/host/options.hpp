// Command line of the drop-in `danbing-tk` (align).  Same hand-rolled, order-sensitive scan as the
// reference's main() (src/aQueryFasta_thread.cpp:2344-2429): same flag names, same defaults
// (cpp:25-34, 2336-2339), same look-ahead rules (`-b [STR]` cpp:2345-2348, `-gc INT1 [INT2]`
// cpp:2360-2365), `-qs` resolving file names from the flags seen BEFORE it (cpp:2387-2409), `-o/-on`
// truncating OUT.trkmc.ar while parsing (cpp:2414-2420).  Where the reference `assert`s or `throw`s
// (= abort) this returns an error message and the CLI exits with status 1.
// Long options (`--gpus`, `--batch-pairs`, `--max-read-len`, `--table-load`, `--bucket-slots`) are
// additions that cannot collide with the reference's single-dash flags.
#pragma once
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>

namespace dbt_host {

struct Options {
	// reference flags
	bool bait = false, aug = false, threading = false, correction = true, tc = false, aln = false, aln_minimal = false,
	     okam = true, g2pan = false, writeKmerName = false, outputBubbles = false, invkmer = false, isFastq = false,
	     trackBait = false, qc = false;
	int simmode = 0, extractFastX = 0, verbosity = 0;
	uint64_t trim = 0, thread_cth = 100, Cthreshold = 10, nproc = 1;
	uint64_t ksize = 21, N_FILTER = 4, NM_FILTER = 1, NM_TR = 40, maxncorrection = 4, qth = 20;
	float readsPerBatchFactor = 1;
	std::string trPrefix, trFname, fastxFname, outPrefix, qcFn, baitFname, mapFname;
	// additions
	int gpus = 1;
	uint64_t batch_pairs = 0;          // pairs per GPU submission; 0 = auto
	uint32_t max_read_len = 0;         // 0 = longest read of the first batch (the context is rebuilt if a longer one shows up)
	float table_load = 0.f;
	uint32_t bucket_slots = 0;
	bool host_reader = false;          // --host-reader: parse and pair every record on the host (the line reader) instead of on the GPU
	uint64_t chunk_mb = 0;             // --chunk-mb: raw text per GPU submission; 0 = auto
	bool timing = false;               // --timing: per-phase wall times on stderr
};

inline bool file_readable(const std::string& fn) { std::ifstream f(fn); return (bool)f; }

// returns "" on success, else the error text
inline std::string parse_args(int argc, char** argv, Options& o) {
	std::vector<std::string> args(argv, argv + argc);
	size_t argi = 1;
	const size_t n = args.size();
	auto need = [&](size_t k) { return argi + k < n; };
	try {
		while (argi < n) {
			const std::string& a = args[argi];
			if (a == "-b") {
				o.bait = true;
				if (need(1) && args[argi + 1][0] != '-') o.baitFname = args[++argi];
			}
			else if (a == "-v") { if (!need(1)) return "missing value for -v"; o.verbosity = std::stoi(args[++argi]); }
			else if (a == "-e") { if (!need(1)) return "missing value for -e"; o.extractFastX = std::stoi(args[++argi]); }
			else if (a == "-bu") o.outputBubbles = true;
			else if (a == "-t") { if (!need(1)) return "missing value for -t"; o.trim = std::stoi(args[++argi]); }
			else if (a == "-s") { if (!need(1)) return "missing value for -s"; o.simmode = std::stoi(args[++argi]); }
			else if (a == "-m") {
				if (!need(1)) return "missing value for -m";
				o.g2pan = true; o.mapFname = args[++argi];
				if (!file_readable(o.mapFname)) return "cannot open " + o.mapFname;
			}
			else if (a == "-au") o.aug = true;
			else if (a == "-g" || a == "-gc" || a == "-gcc") {
				o.threading = true;
				if (a == "-gcc") o.tc = true;
				if (!need(1)) return "missing value for " + a;
				o.thread_cth = std::stoi(args[++argi]);
				if (need(1) && args[argi + 1][0] != '-') o.maxncorrection = std::stoi(args[++argi]);
			}
			else if (a == "-a") o.aln = true;
			else if (a == "-ae") { o.aln = true; o.aln_minimal = true; }
			else if (a == "-ka") o.okam = false;
			else if (a == "-kf") {
				if (!need(2)) return "missing values for -kf";
				o.N_FILTER = std::stoi(args[++argi]);
				o.NM_FILTER = std::stoi(args[++argi]);
			}
			else if (a == "-r") { if (!need(1)) return "missing value for -r"; o.readsPerBatchFactor = std::stof(args[++argi]); }
			else if (a == "-c") { if (!need(1)) return "missing value for -c"; o.NM_TR = std::stoul(args[++argi]); }
			else if (a == "-ik") o.invkmer = true;
			else if (a == "-k") { if (!need(1)) return "missing value for -k"; o.ksize = std::stoi(args[++argi]); }
			else if (a == "-tb") o.trackBait = true;
			else if (a == "-qc") {
				if (!need(1)) return "missing value for -qc";
				o.qc = true; o.qcFn = args[++argi];
				if (!file_readable(o.qcFn)) return "cannot open " + o.qcFn;
			}
			else if (a == "-qs") {
				if (!need(1)) return "missing value for -qs";
				o.trPrefix = args[++argi];
				o.trFname = o.trim ? o.trPrefix + ".tr.trim" + std::to_string(o.trim) + ".kmers" : o.trPrefix + ".tr.kmers";
				if (!file_readable(o.trFname)) return "cannot open " + o.trFname;
				if (o.aug && !file_readable(o.trPrefix + ".tr.aug.kmers")) return "cannot open " + o.trPrefix + ".tr.aug.kmers";
				if (o.bait) {
					if (o.baitFname.empty()) o.baitFname = o.trPrefix + (o.qc ? ".qc.bt.kmdb" : ".bt.kmdb");
					if (!file_readable(o.baitFname)) return "cannot open " + o.baitFname;
				}
			}
			else if (a == "-fa" || a == "-fq") {
				if (!need(1)) return "missing value for " + a;
				o.isFastq = a == "-fq";
				o.fastxFname = args[++argi];
				if (!file_readable(o.fastxFname)) return "cannot open " + o.fastxFname;
			}
			else if (a == "-o" || a == "-on") {
				if (!need(1)) return "missing value for " + a;
				o.writeKmerName = a == "-on";
				o.outPrefix = args[++argi];
				std::ofstream f(o.outPrefix + ".trkmc.ar");          // created (truncated) during parsing, cpp:2417-2419
				if (!f) return "cannot create " + o.outPrefix + ".trkmc.ar";
			}
			else if (a == "-p") { if (!need(1)) return "missing value for -p"; o.nproc = std::stoi(args[++argi]); }
			else if (a == "-cth") { if (!need(1)) return "missing value for -cth"; o.Cthreshold = std::stoi(args[++argi]); }
			else if (a == "-qth") { if (!need(1)) return "missing value for -qth"; o.qth = std::stoi(args[++argi]); }
			else if (a == "--gpus") { if (!need(1)) return "missing value for --gpus"; o.gpus = std::stoi(args[++argi]); }
			else if (a == "--batch-pairs") { if (!need(1)) return "missing value for --batch-pairs"; o.batch_pairs = std::stoull(args[++argi]); }
			else if (a == "--max-read-len") { if (!need(1)) return "missing value for --max-read-len"; o.max_read_len = (uint32_t)std::stoul(args[++argi]); }
			else if (a == "--table-load") { if (!need(1)) return "missing value for --table-load"; o.table_load = std::stof(args[++argi]); }
			else if (a == "--bucket-slots") { if (!need(1)) return "missing value for --bucket-slots"; o.bucket_slots = (uint32_t)std::stoul(args[++argi]); }
			else if (a == "--host-reader") o.host_reader = true;
			else if (a == "--timing") o.timing = true;
			else if (a == "--chunk-mb") { if (!need(1)) return "missing value for --chunk-mb"; o.chunk_mb = std::stoull(args[++argi]); }
			else return "invalid option: " + a;
			++argi;
		}
	} catch (const std::exception&) {
		return "malformed value after " + args[argi < n ? argi - 1 : n - 1];
	}
	return "";
}

// the parameter echo of cpp:2432-2460 (same text, same order)
inline void report_parameters(const Options& o, FILE* err) {
	const int MAX_NT = 2;
	fprintf(err, "use baitDB: %d\nextract fastX: %d\noutput bubbles: %d\nis Fastq: %d\nsim mode: %d\ntrim mode: %llu\n"
	             "augmentation mode: %d\ngraph threading mode: %d\ntrack bait kmer hitting rate: %d\noutput alignment: %d\n"
	             "output successfully aligned reads only: %d\noutput kmer assignment (kam): %d\nwrite invariant kmer counts: %d\n"
	             "k: %llu\n# of subsampled kmers in pre-filtering: %llu\nminimal # of matches in pre-filtering: %llu\n"
	             "Cthreshold: %llu\nthreading Cthreshold: %llu\nmax # of read corrections in threading: %llu\n"
	             "max # of TR-flank transitions: %d\nmin # of kmer matches for TR spanning read: %s\n"
	             "step1 kmer-based filtering: on\nstep2 threading: %s\nfastx: %s\nquery: %s.(tr/ntr).kmers\n\n",
	        o.bait, o.extractFastX, o.outputBubbles, o.isFastq, o.simmode, (unsigned long long)o.trim, o.aug, o.threading, o.trackBait,
	        o.aln, o.aln_minimal, o.okam, o.invkmer, (unsigned long long)o.ksize, (unsigned long long)o.N_FILTER,
	        (unsigned long long)o.NM_FILTER, (unsigned long long)o.Cthreshold, (unsigned long long)o.thread_cth,
	        (unsigned long long)o.maxncorrection, MAX_NT, std::to_string(o.NM_TR).c_str(), o.threading ? "on" : "off",
	        o.fastxFname.c_str(), o.trPrefix.c_str());
}

inline const char* usage_text() {
	return "\nUsage: danbing-tk [-bu] [-ka] [-qc] [-k] [-kf] [-cth] [-qth] [-b] [-c] [-r] [-p] <-fa|-fq> -qs <-o|-on>\n"
	       "Options:\n-Input:\n"
	       "  -fa <STR>             Fasta file e.g. generated by samtools fasta -n; Reads are paired on the fly.\n"
	       "  -fq <STR>             Fastq file e.g. generated by samtools fastq -n; Reads are paired on the fly.\n"
	       "  -qs <STR>             Prefix for graph/index/bait files\n"
	       "-Output:\n"
	       "  -o <STR>              Output prefix\n"
	       "  -on <STR>             Same as the -o option, but write locus and kmer name as well\n"
	       "  -ka                   Turn off kmer assignment output. [on]\n"
	       "-Algorithm:\n"
	       "  -qc <STR>             Locus QC. Skip genotyping loci failing QC\n"
	       "  -k <INT>              Kmer size [21]\n"
	       "  -kf <INT1> <INT2>     Kmer-based pre-filtering: # of sub-sampled kmers [4], minimal # of matches [1]\n"
	       "  -cth <INT>            Discard both pe reads if maxhit of one pe read is below this threshold. [10]\n"
	       "  -c <INT>              Minimal number of exact TR kmer matches for TR-spanning reads [40]\n"
	       "  -g <INT> | -gc <INT1> [INT2] | -gcc <INT1> [INT2]   graph threading (w/o, w/ error correction, + self check)\n"
	       "  -a | -ae              Output alignments for all / successfully threaded reads (with -g/-gc)\n"
	       "-Multiprocessing:\n"
	       "  -r <FLOAT>            scaling factor for readsPerBatch [1]\n"
	       "  -p <INT>              host threads for parsing / formatting [1]\n"
	       "-B200:\n"
	       "  --gpus <INT>          GPUs to shard the reads over (index replicated, NCCL sum at the end) [1]\n"
	       "  --batch-pairs <INT>   read pairs per GPU submission [auto]\n"
	       "  --max-read-len <INT>  longest read to expect [auto]\n"
	       "  --chunk-mb <INT>      raw FASTA/FASTQ text per GPU submission, MiB [128]\n"
	       "  --host-reader         parse and pair the reads on the host instead of on the GPU\n"
	       "  --timing              per-phase wall times on stderr\n\n";
}

}  // namespace dbt_host
