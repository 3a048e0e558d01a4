// Raw-text ingestion for the CLI: the host half of the reader/pairer when the GPU splits and pairs
// the records (csrc/fastx.cu, dbt_upload_raw).
//
//   RawReader     cuts the input stream into chunks of WHOLE records in pinned memory without parsing them: it only counts
//                 newlines (AVX2 when the CPU has it) -- a record is L lines, as in the reference's read loop
//                 (src/aQueryFasta_thread.cpp:1918-1976: two / four getline calls, no look at '>' or '@').  The last complete
//                 record of a chunk is carried over to the next one, together with every record before it that has the
//                 same pruned title, so that mates written next to each other are never separated by a chunk border.
//   StreamPairer  the reference's map (readDB / fqDB, cpp:1926-1964) for what the GPU leaves over: the records it could not
//                 pair with a neighbour ("singletons": mates far apart in the file), whole chunks the GPU hands back
//                 (three equal titles in a row; a neighbour pair whose title a straggler is waiting for), the end-of-file
//                 quirks (FASTQ: a trailing unpaired record is emitted with a stale mate, cpp:1940-1946 lacks the break of
//                 cpp:1967; a truncated last record reuses the previous record's lines: std::getline leaves its string alone
//                 at EOF).  Pairs formed here go to the GPU as an ordinary ASCII batch and are merged into the chunk's
//                 output by the position of their late mate.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>
#include <fcntl.h>
#include <unistd.h>
#include <sys/stat.h>
#include <zlib.h>
#include <algorithm>
#include <thread>
#if defined(__x86_64__)
#include <immintrin.h>
#endif
#include "fastx_reader.hpp"

namespace dbt_host {

// ---- newline counting ----
#if defined(__x86_64__)
__attribute__((target("avx2"))) inline uint64_t count_nl_avx2(const uint8_t* p, size_t n) {
	uint64_t c = 0; size_t i = 0;
	const __m256i nl = _mm256_set1_epi8('\n');
	for (; i + 32 <= n; i += 32) c += (uint64_t)__builtin_popcount((unsigned)_mm256_movemask_epi8(_mm256_cmpeq_epi8(_mm256_loadu_si256((const __m256i*)(p + i)), nl)));
	for (; i < n; ++i) c += p[i] == '\n';
	return c;
}
#endif
inline uint64_t count_nl(const uint8_t* p, size_t n) {
#if defined(__x86_64__)
	static const bool avx2 = __builtin_cpu_supports("avx2");
	if (avx2) return count_nl_avx2(p, n);
#endif
	uint64_t c = 0;
	for (const uint8_t* q = p; (q = (const uint8_t*)memchr(q, '\n', (size_t)(p + n - q))) != nullptr; ++q) ++c;
	return c;
}

inline size_t pruned_len(const uint8_t* t, size_t len) {             // prunePEinfo, cpp:455-462
	return (len >= 2 && t[len - 2] == '/' && (t[len - 1] == '1' || t[len - 1] == '2')) ? len - 2 : len;
}

struct RawChunk {
	uint8_t* buf = nullptr; uint64_t cap = 0;        // pinned
	uint64_t begin = 0, end = 0;                     // the chunk's records: buf[begin, end), buf[end-1] == '\n'
	uint64_t seq = 0;                                // chunk ordinal
	bool eof = false;                                // the stream ended with this chunk
	std::string tail;                                // eof only: the lines of a truncated last record
	uint32_t* pair_offset = nullptr; uint64_t* pair_hash = nullptr; uint32_t* single_offset = nullptr;   // pinned outputs of dbt_upload_raw
	uint32_t pairs_cap = 0, single_cap = 0;
	const uint8_t* text() const { return buf + begin; }
	uint64_t nbytes() const { return end - begin; }
};

class RawReader {
public:
	static constexpr uint64_t HEADROOM = 1 << 20;     // carried records are copied in front of the freshly read bytes
	explicit RawReader(bool isFastq, int nthreads = 1) : L_(isFastq ? 4 : 2), nthreads_(std::max(1, nthreads)) {}
	~RawReader() { if (gz_) gzclose(gz_); else if (fd_ > 0) close(fd_); }
	bool open(const std::string& fn) {
		fd_ = ::open(fn.c_str(), O_RDONLY);
		if (fd_ < 0) return false;
		struct stat st;
		if (fstat(fd_, &st) == 0 && S_ISREG(st.st_mode)) { regular_ = true; size_ = (uint64_t)st.st_size; }   // pipes / stdin: one sequential read
		// gzip input (a convenience the reference does not have: its pipelines pipe `zcat` into /dev/stdin): a regular file that starts
		// with the gzip magic, or any name ending in .gz, is inflated on the fly (zlib, one thread) and then treated like a pipe
		unsigned char magic[2] = {0, 0};
		const bool named_gz = fn.size() > 3 && fn.compare(fn.size() - 3, 3, ".gz") == 0;
		if ((regular_ && pread(fd_, magic, 2, 0) == 2 && magic[0] == 0x1f && magic[1] == 0x8b) || named_gz) {
			gz_ = gzdopen(fd_, "rb");
			if (!gz_) return false;
			gzbuffer(gz_, 1u << 20);
			regular_ = false;
		}
		return true;
	}
	const std::string& error() const { return err_; }

	// Fills c with the next chunk (about `target` fresh bytes).  false: nothing more (or error(): non-empty).
	// The last chunk has c.eof set and may be empty; it is always delivered so that the end-of-stream rules run.
	bool fill(RawChunk& c, uint64_t target) {
		if (done_) return false;
		if (carry_.size() > HEADROOM) { err_ = "a record is longer than 1 MiB"; done_ = true; return false; }
		target = std::min<uint64_t>(target, c.cap - HEADROOM - 1);
		const uint64_t start = HEADROOM - carry_.size();
		memcpy(c.buf + start, carry_.data(), carry_.size());
		uint64_t have = HEADROOM;                     // bytes [start, have) are valid
		uint64_t nl = 0;
		if (regular_ && nthreads_ > 1) {
			// a regular file: the chunk's byte range is read by nthreads_ threads at once (pread), each counting its newlines
			const uint64_t want = std::min<uint64_t>(target, size_ - std::min(size_, off_));
			const int T = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)nthreads_, want >> 20));
			const uint64_t part = (want + T - 1) / T;
			std::vector<uint64_t> cnt(T, 0); std::vector<int> bad(T, 0);
			auto work = [&](int i) {
				uint64_t a = std::min(want, part * i), b = std::min(want, part * (i + 1));
				while (a < b) {
					const ssize_t r = ::pread(fd_, c.buf + HEADROOM + a, (size_t)std::min<uint64_t>(b - a, 1u << 30), (off_t)(off_ + a));
					if (r <= 0) { bad[i] = 1; return; }
					cnt[i] += count_nl(c.buf + HEADROOM + a, (size_t)r);
					a += (uint64_t)r;
				}
			};
			std::vector<std::thread> th;
			for (int i = 1; i < T; ++i) th.emplace_back(work, i);
			work(0);
			for (auto& t : th) t.join();
			for (int i = 0; i < T; ++i) { if (bad[i]) { err_ = "read error (file truncated while reading?)"; done_ = true; return false; } nl += cnt[i]; }
			have += want; off_ += want;
			if (off_ >= size_) src_eof_ = true;
			nl += count_nl(c.buf + start, HEADROOM - start);
		} else {
			while (have - HEADROOM < target && !src_eof_) {
				const size_t want = (size_t)std::min<uint64_t>(target - (have - HEADROOM), 1u << 30);
				const ssize_t r = gz_ ? (ssize_t)gzread(gz_, c.buf + have, (unsigned)want) : ::read(fd_, c.buf + have, want);
				if (r < 0) { err_ = gz_ ? "gzip read error (corrupt or truncated .gz?)" : "read error"; done_ = true; return false; }
				if (r == 0) { src_eof_ = true; break; }
				have += (uint64_t)r;
			}
			nl = count_nl(c.buf + start, have - start);
		}
		if (src_eof_ && have > start && c.buf[have - 1] != '\n') { c.buf[have++] = '\n'; ++nl; }   // std::getline hands out an unterminated last line too
		c.seq = seq_++; c.eof = false; c.tail.clear(); c.begin = start;
		const uint64_t nrec = nl / L_;
		const uint64_t rec_end = back_lines(c.buf, start, have, nl % L_);   // end of the last complete record
		if (src_eof_) {
			c.end = rec_end; c.eof = true; done_ = true;
			c.tail.assign((const char*)c.buf + rec_end, have - rec_end);    // lines of a truncated last record
			carry_.clear();
			return true;
		}
		if (nrec == 0) { err_ = "a record does not fit the read buffer"; done_ = true; return false; }
		// keep the last complete record, and every record before it that carries the same title, for the next chunk
		uint64_t keep_from = back_lines(c.buf, start, rec_end, L_), kept = 1;
		while (kept < nrec) {
			const uint64_t ps = back_lines(c.buf, start, keep_from, L_);
			if (!same_title(c.buf, ps, keep_from)) break;
			keep_from = ps; ++kept;
		}
		if (kept == nrec) keep_from = rec_end;           // one title throughout: no border is better than another
		c.end = keep_from;
		carry_.assign((const char*)c.buf + keep_from, have - keep_from);
		return true;
	}

private:
	uint32_t L_; int nthreads_;
	bool regular_ = false; uint64_t size_ = 0, off_ = 0;
	int fd_ = -1; gzFile gz_ = nullptr; bool src_eof_ = false, done_ = false;
	std::string carry_; uint64_t seq_ = 0;
	std::string err_;

	// [lo, hi) minus its unterminated last fragment and minus its last n whole lines: the position where what is left ends
	static uint64_t back_lines(const uint8_t* b, uint64_t lo, uint64_t hi, uint64_t n) {
		uint64_t p = hi;
		while (p > lo && b[p - 1] != '\n') --p;
		for (uint64_t i = 0; i < n; ++i) {
			if (p <= lo) return lo;
			--p;                                       // the newline that ends the line being dropped
			while (p > lo && b[p - 1] != '\n') --p;
		}
		return p;
	}
	// records starting at a and b (a < b, both complete): equal pruned titles?
	static bool same_title(const uint8_t* buf, uint64_t a, uint64_t b) {
		const uint8_t* ea = (const uint8_t*)memchr(buf + a, '\n', b - a);
		const uint8_t* eb = (const uint8_t*)rawmemchr(buf + b, '\n');
		if (!ea) return false;
		const size_t la = pruned_len(buf + a, (size_t)(ea - (buf + a))), lb = pruned_len(buf + b, (size_t)(eb - (buf + b)));
		return la == lb && memcmp(buf + a, buf + b, la) == 0;
	}
};

// lines of the record that starts at p (the text holds whole records)
inline void record_lines(const char* p, uint32_t L, const char* ls[4], size_t ln[4], const char** next) {
	for (uint32_t i = 0; i < L; ++i) { const char* e = (const char*)rawmemchr(p, '\n'); ls[i] = p; ln[i] = (size_t)(e - p); p = e + 1; }
	*next = p;
}

// pair formed by the GPU: two neighbouring records starting at text + off, the second one is read 1
inline PairText pair_text_raw(const uint8_t* text, uint32_t off, bool fq) {
	const uint32_t L = fq ? 4 : 2;
	const char* ls[4]; size_t ln[4]; const char* nx;
	PairText t;
	record_lines((const char*)text + off, L, ls, ln, &nx);
	t.seq[1] = ls[1]; t.seq_len[1] = ln[1];
	if (fq) { t.qual[1] = ls[3]; t.qual_len[1] = ln[3]; }
	record_lines(nx, L, ls, ln, &nx);
	t.title = ls[0]; t.title_len = pruned_len((const uint8_t*)ls[0], ln[0]);
	t.seq[0] = ls[1]; t.seq_len[0] = ln[1];
	if (fq) { t.qual[0] = ls[3]; t.qual_len[0] = ln[3]; }
	return t;
}

// ---- the reference's map for everything the GPU does not pair ----
class StreamPairer {
public:
	struct Extra {                                    // pairs formed on the host for one chunk: an ordinary ASCII batch + merge keys
		HostBatch hb;
		std::vector<uint8_t> bases; std::vector<uint64_t> offsets;
		std::vector<uint64_t> key;                    // position of the late mate in the chunk (byte offset; end-of-stream pairs: ~0 - 1, ~0)
		void clear() { hb.bases = nullptr; hb.offsets = nullptr; hb.reset(0); bases.clear(); offsets.assign(1, 0); key.clear(); }
		void seal() { bases.resize(bases.size() + 16); hb.bases = bases.data(); hb.offsets = offsets.data(); hb.bases_cap = bases.size(); hb.pairs_cap = key.size(); }
	};
	StreamPairer(bool fq, uint64_t minlen, uint64_t readsPerRefBatch) : fq_(fq), L_(fq ? 4 : 2), minlen_(minlen), ppb_((std::max<uint64_t>(readsPerRefBatch, 1) + 1) / 2) {}
	uint64_t total_pairs() const { return npairs_; }
	uint64_t unpaired() const { return db_.size(); }

	// One chunk, in stream order.  flags / np / ns: dbt_raw_info of the GPU's split (flags != 0, or use_gpu false: no split).
	// Returns true when the GPU's np pairs stand (ex = the pairs of the singletons, ascending keys); false when the whole
	// chunk was paired here (ex = all its pairs; the GPU's, if any, are void).
	bool chunk(const RawChunk& c, bool use_gpu, uint32_t flags, uint32_t np, uint32_t ns, uint32_t nshort_pairs, Extra& ex) {
		ex.clear();
		const uint8_t* text = c.text();
		bool fast = use_gpu && flags == 0;
		if (fast && nshort_pairs && (ns || !db_.empty())) fast = false;   // a dropped short pair may have met a straggler first: not worth a hash list
		if (fast && (ns || !db_.empty())) {
			// A neighbour pair whose title a straggler carries -- one the map holds, or a singleton of this chunk -- is a
			// third record of that title: the reference pairs in order of arrival.  Rare: the host takes the chunk.
			std::unordered_set<uint64_t> live;
			for (auto& kv : db_) live.insert(fnv((const uint8_t*)kv.first.data(), kv.first.size()));
			for (uint32_t j = 0; j < ns; ++j) {
				const uint8_t* t = text + c.single_offset[j];
				const uint8_t* e = (const uint8_t*)rawmemchr(t, '\n');
				live.insert(fnv(t, pruned_len(t, (size_t)(e - t))));
			}
			for (uint32_t i = 0; fast && i < np; ++i) if (live.count(c.pair_hash[i])) fast = false;
		}
		if (fast) {
			uint32_t gpu_done = 0;                        // GPU pairs that precede the record being handled
			for (uint32_t j = 0; j < ns; ++j) {
				const uint32_t g = (uint32_t)(std::lower_bound(c.pair_offset, c.pair_offset + np, c.single_offset[j]) - c.pair_offset);
				npairs_ += g - gpu_done; gpu_done = g;
				const char* ls[4]; size_t ln[4]; const char* nx;
				record_lines((const char*)text + c.single_offset[j], L_, ls, ln, &nx);
				record(ls, ln, c.single_offset[j], ex);
			}
			npairs_ += np - gpu_done;
			if (c.nbytes()) note_chunk_end(c, np, ns, ex);
		} else {
			const char* p = (const char*)text; const char* e = p + c.nbytes();
			while (p < e) { const char* q = p; const char* ls[4]; size_t ln[4]; record_lines(p, L_, ls, ln, &p); record(ls, ln, (uint64_t)(q - (const char*)text), ex); }
		}
		if (c.eof) finish(c, ex);
		ex.seal();
		return fast;
	}

private:
	struct Rec { std::string seq, qual; };
	bool fq_; uint32_t L_; uint64_t minlen_, ppb_;       // ppb_: pairs per reference batch (300000*r reads, cpp:1835, 1874)
	std::unordered_map<std::string, Rec> db_;
	uint64_t npairs_ = 0;                             // pairs emitted so far (GPU + host)
	// what the end-of-stream rules read: the reference's per-batch strings (cpp:1871) = the lines of the last record read in
	// the current reference batch and the early mate of the last pair emitted in it
	std::string last_line_[4]; bool last_valid_ = false;
	std::string stale_seq2_, stale_qual2_;
	bool last_inserted_ = false; std::string last_title_;   // the last record of the stream so far went into the map unmatched
	bool last_short_drop_ = false;                    // ... or completed a pair that was forgotten for a short mate

	static uint64_t fnv(const uint8_t* t, size_t len) {  // = dbt_title_hash on an already pruned title
		uint64_t h = 0xcbf29ce484222325ull;
		for (size_t i = 0; i < len; ++i) { h ^= t[i]; h *= 0x100000001b3ull; }
		return h;
	}
	void emit(Extra& ex, const std::string& title, const char* s1, size_t n1, const char* q1, size_t nq1, const std::string& s2, const std::string& q2, uint64_t key) {
		ex.bases.insert(ex.bases.end(), s1, s1 + n1); ex.offsets.push_back(ex.bases.size());
		ex.bases.insert(ex.bases.end(), s2.begin(), s2.end()); ex.offsets.push_back(ex.bases.size());
		ex.hb.names.insert(ex.hb.names.end(), title.begin(), title.end()); ex.hb.name_off.push_back(ex.hb.names.size());
		if (fq_) {
			ex.hb.quals.insert(ex.hb.quals.end(), q1, q1 + nq1); ex.hb.qual_off.push_back(ex.hb.quals.size());
			ex.hb.quals.insert(ex.hb.quals.end(), q2.begin(), q2.end()); ex.hb.qual_off.push_back(ex.hb.quals.size());
		}
		ex.hb.max_len = std::max<uint32_t>(ex.hb.max_len, (uint32_t)std::max(n1, s2.size()));
		ex.key.push_back(key);
		++ex.hb.npairs; ++npairs_;
		stale_seq2_ = s2; stale_qual2_ = q2;
	}
	// one record through the map (cpp:1926-1939 / 1951-1964)
	void record(const char* const ls[4], const size_t ln[4], uint64_t key, Extra& ex) {
		for (uint32_t i = 0; i < L_; ++i) last_line_[i].assign(ls[i], ln[i]);
		last_valid_ = true; last_inserted_ = false; last_short_drop_ = false;
		std::string title(ls[0], pruned_len((const uint8_t*)ls[0], ln[0]));
		auto it = db_.find(title);
		if (it != db_.end()) {
			Rec mate = std::move(it->second);
			db_.erase(it);
			if (ln[1] < minlen_ || mate.seq.size() < minlen_) { last_short_drop_ = true; return; }   // cpp:1957: both mates forgotten
			emit(ex, title, ls[1], ln[1], fq_ ? ls[3] : nullptr, fq_ ? ln[3] : 0, mate.seq, mate.qual, key);
			if (npairs_ % ppb_ == 0) last_valid_ = false;                             // the reference batch is full: the next one re-declares the strings
			return;
		}
		Rec r; r.seq.assign(ls[1], ln[1]); if (fq_) r.qual.assign(ls[3], ln[3]);
		db_.emplace(title, std::move(r));
		last_inserted_ = true; last_title_ = std::move(title);
	}
	// After a chunk whose pairs the GPU formed: bring the end-of-stream state to what a record-by-record pass would have left.
	void note_chunk_end(const RawChunk& c, uint32_t np, uint32_t ns, Extra& ex) {
		const uint8_t* text = c.text();
		uint64_t rs = c.nbytes();                          // start of the chunk's last record
		for (uint32_t i = 0; i < L_; ++i) { if (rs) --rs; while (rs && text[rs - 1] != '\n') --rs; }
		uint64_t late_start = ~0ull;                       // start of the late mate of the GPU's last pair
		if (np) {
			const char* ls[4]; size_t ln[4]; const char* nx;
			record_lines((const char*)text + c.pair_offset[np - 1], L_, ls, ln, &nx);
			late_start = (uint64_t)(nx - (const char*)text);
			if (ex.key.empty() || ex.key.back() < c.pair_offset[np - 1]) {           // the chunk's last pair is the GPU's
				stale_seq2_.assign(ls[1], ln[1]);
				if (fq_) stale_qual2_.assign(ls[3], ln[3]);
			}
		}
		if (ns && c.single_offset[ns - 1] == rs) return;   // the last record went through record(): the state is its
		const char* ls[4]; size_t ln[4]; const char* nx;
		record_lines((const char*)text + rs, L_, ls, ln, &nx);
		for (uint32_t i = 0; i < L_; ++i) last_line_[i].assign(ls[i], ln[i]);
		last_inserted_ = false; last_short_drop_ = false;
		last_valid_ = !(late_start == rs && npairs_ % ppb_ == 0);
		if (rs > 0 && late_start != rs) {                  // the late mate of a pair the GPU forgot for a short mate?
			uint64_t ps = rs;
			for (uint32_t i = 0; i < L_; ++i) { if (ps) --ps; while (ps && text[ps - 1] != '\n') --ps; }
			const char* pl[4]; size_t pn[4];
			record_lines((const char*)text + ps, L_, pl, pn, &nx);
			const size_t a = pruned_len((const uint8_t*)ls[0], ln[0]), b = pruned_len((const uint8_t*)pl[0], pn[0]);
			if (a == b && memcmp(ls[0], pl[0], a) == 0 && (ln[1] < minlen_ || pn[1] < minlen_)) last_short_drop_ = true;
		}
	}
	// end of the stream (cpp:1940-1946, 1965-1968)
	void finish(const RawChunk& c, Extra& ex) {
		if (!c.tail.empty()) {
			// A truncated last record: std::getline leaves its string alone at end of file, so the missing lines keep what the
			// previous record of the reference batch left there (nothing, at the start of a batch).
			const char* ls[4]; size_t ln[4];
			const char* p = c.tail.data(); const char* e = p + c.tail.size();
			std::string keep[4];
			for (uint32_t i = 0; i < L_; ++i) {
				if (p < e) { const char* q = (const char*)memchr(p, '\n', (size_t)(e - p)); if (!q) q = e; ls[i] = p; ln[i] = (size_t)(q - p); p = q < e ? q + 1 : e; }
				else { keep[i] = last_valid_ ? last_line_[i] : std::string(); ls[i] = keep[i].data(); ln[i] = keep[i].size(); }
			}
			record(ls, ln, ~0ull - 2, ex);
		}
		if (last_short_drop_) {
			// The stream ended on a pair that was forgotten for a short mate: the reference's loop goes round once more
			// (cpp:1957 `continue`), every getline fails, and the strings of the last record -- its title pruned a second
			// time -- go into the map as a record of their own.
			std::string keep[4]; const char* ls[4]; size_t ln[4];
			for (uint32_t i = 0; i < L_; ++i) keep[i] = last_line_[i];
			keep[0].resize(pruned_len((const uint8_t*)keep[0].data(), keep[0].size()));
			for (uint32_t i = 0; i < L_; ++i) { ls[i] = keep[i].data(); ln[i] = keep[i].size(); }
			record(ls, ln, ~0ull - 1, ex);
		}
		if (fq_ && last_inserted_) {
			// FASTQ lacks the `se and EOF` break of the FASTA branch (cpp:1967): the trailing unpaired record is emitted with the
			// early mate of the previous pair of its reference batch (an empty read at the start of a batch); it stays in the map
			const bool fresh = npairs_ % ppb_ == 0;
			auto it = db_.find(last_title_);
			if (it != db_.end()) {
				const std::string s1 = it->second.seq, q1 = it->second.qual;
				emit(ex, last_title_, s1.data(), s1.size(), q1.data(), q1.size(), fresh ? std::string() : stale_seq2_, fresh ? std::string() : stale_qual2_, ~0ull);
			}
		}
	}
};

}  // namespace dbt_host
