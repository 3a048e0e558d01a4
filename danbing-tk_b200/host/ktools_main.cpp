// dbt-ktools: the index-construction step of the reference's `ktools` (src/kmertools.cpp:221-345) for the B200 drop-in.
//   dbt-ktools serialize PREF     PREF.{tr,fl,tre}.kmers -> PREF.kmers.dbi, PREF.fl.kdb, PREF.tre.kdb (byte-identical to ktools)
// Host-only (no CUDA): these files are what danbing-tk loads into the GPU tables (host/serialize.hpp).
#include <cstdio>
#include <string>
#include "serialize.hpp"

int main(int argc, char** argv) {
	const std::string cmd = argc > 1 ? argv[1] : "";
	if (cmd != "serialize" || argc != 3) {
		fputs("Usage: dbt-ktools serialize <pref>\n\n  PREF     prefix of *.(fl|tr|tre).kmers\n", stderr);
		return cmd == "serialize" && argc == 2 ? 0 : 1;
	}
	std::string err;
	if (!dbt_host::serialize_rpgg(argv[2], stderr, err)) { fprintf(stderr, "dbt-ktools: %s\n", err.c_str()); return 1; }
	return 0;
}
