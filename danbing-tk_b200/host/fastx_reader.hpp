// FASTA/FASTQ reader + on-the-fly mate pairing, host side of the batcher.
//
// Restates the read loop of CountWords (src/aQueryFasta_thread.cpp:1869-1981) including its quirks:
//   * titles lose a trailing "/1" or "/2" (prunePEinfo, cpp:455-462); mates are paired by title through a hash
//     map whatever their distance in the stream (cpp:1955-1964 FASTA, cpp:1928-1939 FASTQ);
//   * the mate that arrives SECOND becomes "read 1" (seqs[2i]), the first one "read 2" (cpp:1972-1974);
//   * a pair with a mate shorter than Cthreshold + k - 1 is dropped silently (cpp:1957);
//   * state is per reference batch of 300000*r reads (cpp:1835): in FASTQ mode a trailing unpaired record at
//     end of file is emitted as a pair with the STALE second mate of the previous pair of that batch (the
//     `if (se and in->peek() == EOF) break;` of the FASTA branch, cpp:1967, is missing from the FASTQ branch).
// Output goes straight into a HostBatch: concatenated ASCII bases + offsets in (pinned) memory that
// dbt_upload copies to the GPU as is, plus titles/qualities for the kam writer.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>
#include <fcntl.h>
#include <unistd.h>

namespace dbt_host {

struct HostBatch {
	uint64_t first_pair_index = 0, npairs = 0;
	uint64_t seq = 0;                                         // ordinal of the batch in the stream
	uint8_t* bases = nullptr; uint64_t bases_cap = 0;       // caller-provided (pinned) buffer
	uint64_t* offsets = nullptr; uint64_t pairs_cap = 0;    // caller-provided (pinned), 2*pairs_cap+1 entries
	std::vector<char> names; std::vector<uint64_t> name_off;    // title of pair i = names[name_off[i], name_off[i+1]) incl. leading '>'/'@'
	std::vector<char> quals; std::vector<uint64_t> qual_off;    // FASTQ only: quality of read r = quals[qual_off[r], qual_off[r+1])
	uint32_t max_len = 0;                                     // longest read in the batch
	uint64_t nreads_total_after = 0, unpaired_after = 0;      // for the "Buffered reading" stderr line
	void reset(uint64_t first) {
		first_pair_index = first; npairs = 0; max_len = 0;
		names.clear(); name_off.assign(1, 0); quals.clear(); qual_off.assign(1, 0);
		if (offsets) offsets[0] = 0;
	}
	bool room_for(uint64_t l1, uint64_t l2) const { return npairs < pairs_cap && offsets[2 * npairs] + l1 + l2 <= bases_cap; }
};

// ---- a pair's text, wherever it lives ----
struct PairText {
	const char* title = nullptr; size_t title_len = 0;       // pruned, with its leading '>' / '@'
	const char* seq[2] = {nullptr, nullptr}; size_t seq_len[2] = {0, 0};    // [0] = read 1 (late mate), [1] = read 2
	const char* qual[2] = {nullptr, nullptr}; size_t qual_len[2] = {0, 0};
};

inline PairText pair_text(const HostBatch& hb, uint64_t i, bool fq) {
	PairText t;
	t.title = hb.names.data() + hb.name_off[i]; t.title_len = hb.name_off[i + 1] - hb.name_off[i];
	for (int m = 0; m < 2; ++m) {
		t.seq[m] = (const char*)hb.bases + hb.offsets[2 * i + m]; t.seq_len[m] = hb.offsets[2 * i + m + 1] - hb.offsets[2 * i + m];
		if (fq) { t.qual[m] = hb.quals.data() + hb.qual_off[2 * i + m]; t.qual_len[m] = hb.qual_off[2 * i + m + 1] - hb.qual_off[2 * i + m]; }
	}
	return t;
}

class FastxReader {
public:
	FastxReader(bool isFastq, uint64_t minReadSize, uint64_t readsPerRefBatch)
		: fq_(isFastq), minReadSize_(minReadSize), refBatch_(readsPerRefBatch ? readsPerRefBatch : 2) { buf_.resize(1 << 24); }
	~FastxReader() { if (fd_ >= 0 && fd_ != 0) close(fd_); }
	bool open(const std::string& fn) { fd_ = ::open(fn.c_str(), O_RDONLY); return fd_ >= 0; }
	uint64_t total_reads() const { return nReads_; }
	uint64_t unpaired() const { return db_.size() + (has_pending_ ? 1 : 0); }

	// Appends whole reference batches to hb until it cannot take another full one (or EOF).
	// Returns false when the input is exhausted and nothing was appended.
	bool fill(HostBatch& hb) {
		bool any = false;
		while (!eof_at_batch_start()) {
			// a reference batch holds up to refBatch_ reads; make sure the worst case fits: stop early otherwise
			if (hb.npairs + refBatch_ / 2 > hb.pairs_cap && hb.npairs) break;
			if (!read_ref_batch(hb)) break;          // host buffer full in the middle of a reference batch: resume next time
			any = true;
			if (pending_resume_) break;
		}
		hb.nreads_total_after = nReads_; hb.unpaired_after = unpaired();
		return any || hb.npairs;
	}

private:
	struct Rec { std::string seq, qual; };
	bool fq_; uint64_t minReadSize_, refBatch_;
	int fd_ = -1;
	std::vector<char> buf_; size_t pos_ = 0, end_ = 0; bool src_eof_ = false;
	std::unordered_map<std::string, Rec> db_;        // readDB / fqDB ...
	std::string pending_title_; Rec pending_; bool has_pending_ = false;   // ... whose most recent entry is kept outside the map (adjacent mates never touch it)
	uint64_t nReads_ = 0;
	// per reference batch state (cpp:1871: `string title, seq1, seq2, qtitle, qual1, qual2;` inside while(true))
	std::string title_, seq1_, seq2_, qtitle_, qual1_, qual2_;
	uint64_t nReadsBatch_ = 0; bool in_batch_ = false, pending_resume_ = false;

	bool refill() {
		if (src_eof_) return false;
		if (pos_ > 0) { memmove(buf_.data(), buf_.data() + pos_, end_ - pos_); end_ -= pos_; pos_ = 0; }
		if (end_ == buf_.size()) buf_.resize(buf_.size() * 2);
		ssize_t r = ::read(fd_, buf_.data() + end_, buf_.size() - end_);
		if (r <= 0) { src_eof_ = true; return false; }
		end_ += (size_t)r;
		return true;
	}
	bool peek_eof() { while (pos_ == end_) { if (!refill()) return true; } return false; }      // in->peek() == EOF
	// std::getline semantics: at EOF the string is left untouched and false is returned
	bool getline(std::string& s) {
		if (peek_eof()) return false;
		size_t scanned = 0;                          // bytes after pos_ already searched for a newline
		for (;;) {
			const char* base = buf_.data() + pos_;
			const size_t avail = end_ - pos_;
			const char* nl = (const char*)memchr(base + scanned, '\n', avail - scanned);
			if (nl) { s.assign(base, nl - base); pos_ += (size_t)(nl - base) + 1; return true; }
			scanned = avail;
			if (!refill()) { s.assign(buf_.data() + pos_, end_ - pos_); pos_ = end_; return true; }   // last line without '\n'
		}
	}
	static void prune(std::string& t) {              // prunePEinfo; titles shorter than 2 are left alone (reference: out-of-bounds read)
		const size_t len = t.size();
		if (len >= 2 && t[len - 2] == '/' && (t[len - 1] == '1' || t[len - 1] == '2')) t.resize(len - 2);
	}
	bool eof_at_batch_start() { return !in_batch_ && peek_eof(); }

	void emit(HostBatch& hb) {
		const uint64_t r = 2 * hb.npairs;
		uint64_t o = hb.offsets[r];
		memcpy(hb.bases + o, seq1_.data(), seq1_.size()); o += seq1_.size(); hb.offsets[r + 1] = o;
		memcpy(hb.bases + o, seq2_.data(), seq2_.size()); o += seq2_.size(); hb.offsets[r + 2] = o;
		hb.names.insert(hb.names.end(), title_.begin(), title_.end()); hb.name_off.push_back(hb.names.size());
		if (fq_) {
			hb.quals.insert(hb.quals.end(), qual1_.begin(), qual1_.end()); hb.qual_off.push_back(hb.quals.size());
			hb.quals.insert(hb.quals.end(), qual2_.begin(), qual2_.end()); hb.qual_off.push_back(hb.quals.size());
		}
		if (seq1_.size() > hb.max_len) hb.max_len = (uint32_t)seq1_.size();
		if (seq2_.size() > hb.max_len) hb.max_len = (uint32_t)seq2_.size();
		++hb.npairs;
	}

	// One reference batch (cpp:1869-1981).  Returns false if hb ran out of room before the batch ended
	// (the batch is resumed by the next call, with its state kept).
	bool read_ref_batch(HostBatch& hb) {
		if (!in_batch_) {
			title_.clear(); seq1_.clear(); seq2_.clear(); qtitle_.clear(); qual1_.clear(); qual2_.clear();
			nReadsBatch_ = 0; in_batch_ = true;
		}
		if (pending_resume_) {                       // the pair formed last time did not fit: emit it first
			if (!hb.room_for(seq1_.size(), seq2_.size())) return false;
			emit(hb); nReadsBatch_ += 2; nReads_ += 2; pending_resume_ = false;
		}
		while (nReadsBatch_ < refBatch_ && !peek_eof()) {
			bool se = true;
			while (se) {
				getline(title_); getline(seq1_);
				if (fq_) { getline(qtitle_); getline(qual1_); }
				prune(title_);
				Rec* mate = nullptr;
				auto it = db_.end();
				if (has_pending_ && pending_title_ == title_) mate = &pending_;
				else { it = db_.find(title_); if (it != db_.end()) mate = &it->second; }
				if (mate) {
					const bool is_short = seq1_.size() < minReadSize_ || mate->seq.size() < minReadSize_;
					if (!is_short) { seq2_.swap(mate->seq); if (fq_) qual2_.swap(mate->qual); }
					if (mate == &pending_) has_pending_ = false; else db_.erase(it);
					if (is_short) continue;                  // cpp:1957: both mates forgotten, keep reading (even at EOF)
					se = false;
					break;
				}
				if (has_pending_) db_.emplace(std::move(pending_title_), std::move(pending_));
				pending_title_ = title_; pending_.seq = seq1_; if (fq_) pending_.qual = qual1_;
				has_pending_ = true;
				if (peek_eof()) break;
			}
			if (!fq_ && se && peek_eof()) break;      // cpp:1967 (absent from the FASTQ branch)
			if (!hb.room_for(seq1_.size(), seq2_.size())) { pending_resume_ = true; return false; }
			emit(hb); nReadsBatch_ += 2; nReads_ += 2;
		}
		in_batch_ = false;
		return true;
	}
};

}  // namespace dbt_host
