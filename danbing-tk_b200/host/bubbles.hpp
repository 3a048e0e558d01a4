// `-bu`: OUT.bub.kmdb from the novel-edge events the GPU hands back (csrc/bubbles.cu, dbt_bubble_events).
//
// The reference counts novel edges in a per-batch unordered_map<locus, unordered_map<edge, count>> (bubbles_t, declared anew
// for every batch of 300000*r reads, src/aQueryFasta_thread.cpp:1992), merges it into bubbleDB[locus] at the end of the batch
// (accumBubbles, cpp:1599-1606) and finally writes, per locus, the entries counted at least 5 times in the ITERATION order of
// bubbleDB[locus] (dumpBubbles -> flattenKmapDB(th = 5) -> serializeKmapDB, src/aQueryFasta_thread.h:1006, binaryKmerIO.hpp:32-68).
// That order is a property of libstdc++'s hashtable and of the insertion sequence, so the same containers are filled here with
// the same sequence: events sorted by (pair, mate, position), batch borders every ceil(300000*r / 2) pairs.  Deterministic in
// the reference only with one thread (-p 1); with more, its batches are merged in completion order.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <string>
#include <unordered_map>
#include <vector>

namespace dbt_host {

struct BubbleEvent { uint64_t key; uint64_t edge; uint32_t locus; };   // key = global pair ordinal << 10 | mate << 9 | position

// The same shape serves `-tb` (bait k-mer hit tracking): per batch tkr[locus][k-mer] (cpp:1993), merged by accumBaitKmerHits
// (cpp:1608-1616) into btTK[locus], written as OUT.btk.kmdb without a threshold (dumpBaitKmerHits, h:1010-1012).
class BubbleAcc {
public:
	BubbleAcc(uint64_t nloci, uint64_t readsPerRefBatch, uint64_t min_count = 5, const char* suffix = ".bub.kmdb")
		: db_(nloci), ppb_((std::max<uint64_t>(readsPerRefBatch, 1) + 1) / 2), th_(min_count), suffix_(suffix) {}
	// events of one submission, ascending keys; submissions arrive in stream order
	void replay(const std::vector<BubbleEvent>& ev) {
		for (const BubbleEvent& e : ev) {
			const uint64_t b = (e.key >> 10) / ppb_;
			if (b != cur_batch_) { flush(); cur_batch_ = b; }
			if (e.locus < db_.size()) ++cur_[e.locus][e.edge];
		}
	}
	bool write(const std::string& pref, uint64_t& nskipped, std::string& err) {
		flush();
		const uint64_t nloci = db_.size();
		std::vector<uint64_t> index(nloci), ks, vs;
		nskipped = 0;
		for (uint64_t l = 0; l < nloci; ++l) {
			uint64_t nskip = 0;
			for (auto& p : db_[l]) { if (p.second >= th_) { ks.push_back(p.first); vs.push_back(p.second); } else ++nskip; }
			index[l] = db_[l].size() - nskip;
			nskipped += nskip;
		}
		const uint64_t nk = ks.size(), sizeofval = 8;
		const std::string fn = pref + suffix_;
		FILE* f = fopen(fn.c_str(), "wb");
		if (!f) { err = "cannot create " + fn; return false; }
		bool ok = fwrite(&nloci, 8, 1, f) == 1 && (!nloci || fwrite(index.data(), 8, nloci, f) == nloci) && fwrite(&nk, 8, 1, f) == 1 && fwrite(&sizeofval, 8, 1, f) == 1
		          && (!nk || (fwrite(ks.data(), 8, nk, f) == nk && fwrite(vs.data(), 8, nk, f) == nk));
		ok = (fclose(f) == 0) && ok;
		if (!ok) err = "short write on " + fn;
		return ok;
	}
private:
	std::vector<std::unordered_map<size_t, uint32_t>> db_;                       // bubble_db_t
	std::unordered_map<uint64_t, std::unordered_map<size_t, uint32_t>> cur_;    // bubbles_t of the batch being replayed
	uint64_t ppb_, cur_batch_ = 0, th_; std::string suffix_;
	void flush() {
		for (auto& p : cur_) { auto& o = db_[p.first]; for (auto& q : p.second) o[q.first] += q.second; }
		cur_ = std::unordered_map<uint64_t, std::unordered_map<size_t, uint32_t>>();   // a NEW table per batch, like the reference's local
	}
};

}  // namespace dbt_host
