// danbing-tk (B200): drop-in for the reference's `danbing-tk` align CLI (src/aQueryFasta_thread.cpp:2286-2660).
//
// Same flags, same input files (PREF.tr.kmers / .kmers.dbi / .fl.kdb, -qc FILE, FASTA/FASTQ incl. /dev/stdin),
// same outputs (OUT.trkmc.ar + OUT.tr.summary.txt, or OUT.tr.kmers with -on; kam lines on stdout; counters on
// stderr).  What CountWords did per pair on pthreads (cpp:2002-2249) runs on the GPU(s) behind the C ABI of
// include/danbing_b200.h; this file is the L3 driver around it:
//
//   reader thread      FASTA/FASTQ -> paired, pinned HostBatch (fastx_reader.hpp)          [cpp:1885-1986]
//   one worker per GPU dbt_submit / dbt_wait on two slots, batch b -> GPU b mod G            [cpp:2002-2249]
//                      + formatting of the returned records into kam lines                   [cpp:1646-1681]
//   writer (main)      stdout in batch order, per-batch stderr line                          [cpp:2253-2279]
//   end of run         NCCL sum of counts/kmc/nmapread/counters over the GPUs, outputs       [cpp:2617-2656]
//
// There is no CPU fallback: without a CUDA device the program fails.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/danbing_b200.h"
#include "options.hpp"
#include "serialize.hpp"
#include "fastx_reader.hpp"
#include "raw_ingest.hpp"
#include "kam_writer.hpp"
#include "aln_writer.hpp"
#include "bubbles.hpp"
#include "bait.hpp"

using namespace dbt_host;

namespace {

template <typename T>
class Channel {
public:
	void push(T v) { { std::lock_guard<std::mutex> l(m_); q_.push_back(std::move(v)); } cv_.notify_all(); }
	bool pop(T& v) {
		std::unique_lock<std::mutex> l(m_);
		cv_.wait(l, [&] { return !q_.empty() || closed_; });
		if (q_.empty()) return false;
		v = std::move(q_.front()); q_.pop_front();
		return true;
	}
	void close() { { std::lock_guard<std::mutex> l(m_); closed_ = true; } cv_.notify_all(); }
private:
	std::mutex m_; std::condition_variable cv_; std::deque<T> q_; bool closed_ = false;
};

struct BatchOut {                       // what a worker hands to the writer
	uint64_t index = 0;                 // ordinal of the batch / chunk in the stream: stdout and the per-batch stderr lines go out in this order
	std::string text;                   // kam lines
	std::vector<BubbleEvent> bait_events;   // `-tb`: the bait k-mer that removed a mate (key = pair ordinal << 10 | mate << 9)
	std::vector<BubbleEvent> events;    // `-bu`: novel edges of the batch, ascending (pair, mate, position)
	std::vector<std::string> parts;     // ... or several pieces in order (formatted by several threads); the strings come from and return to Shared::strings
	uint64_t counters[DBT_N_COUNTERS] = {};
	uint64_t npairs = 0, nreads_total = 0, unpaired = 0;
	double seconds = 0;
};

struct PairTurn {                       // raw ingestion: chunks meet the host pairer (one map, like the reference's) in stream order
	std::mutex m; std::condition_variable cv; uint64_t next = 0;
};

struct StringPool {                     // output buffers circulate between the formatting threads and the writer: no fresh pages per chunk
	std::mutex m; std::vector<std::string> free_;
	std::string get() { std::lock_guard<std::mutex> l(m); if (free_.empty()) return std::string(); std::string s = std::move(free_.back()); free_.pop_back(); return s; }
	void put(std::string&& s) { s.clear(); std::lock_guard<std::mutex> l(m); if (free_.size() < 64) free_.push_back(std::move(s)); }
};

struct Shared {
	Options opt;
	StringPool strings;
	std::unique_ptr<StreamPairer> pairer; PairTurn turn;
	BaitDB baits; std::unique_ptr<BubbleAcc> bait_hits;                          // `-b` / `-tb`
	std::unique_ptr<BubbleAcc> bubbles;                                          // `-bu`: replayed by the writer thread in stream order
	uint64_t chunk_target = 0;          // raw text bytes per GPU submission
	std::atomic<uint64_t> t_split_us{0}, t_pair_us{0}, t_wait_us{0}, t_format_us{0}, t_read_us{0}, n_slow_chunks{0}, n_chunks{0};
	std::atomic<uint64_t> t_launch_us{0}, t_write_us{0}, t_reader_blocked_us{0}, t_worker_idle_us{0}, t_turn_us{0};
	Rpgg rpgg;
	dbt_index index{};
	uint64_t batch_pairs = 0, bases_cap = 0;
	bool thr_graph = false;             // -g/-gc with PREF.graph.kmers loaded: restored graph threading
	bool want_records = false;          // kam lines (no threading) or .aln lines (-a/-ae with threading)
	std::atomic<bool> failed{false};
	std::mutex err_mutex; std::string err;
	void fail(const std::string& m) { std::lock_guard<std::mutex> l(err_mutex); if (!failed.exchange(true)) err = m; }
};

struct Worker {
	int gpu = 0;
	dbt_ctx* ctx = nullptr;
	uint32_t ctx_len = 0;
	Channel<HostBatch*> in;
	Channel<RawChunk*> rin;
	uint8_t* recs[DBT_N_SLOTS] = {};                                             // raw path: pinned record buffers of the two slots
	std::vector<uint64_t> acc_tr, acc_kmc; std::vector<uint32_t> acc_nmap;      // counts of contexts retired by a rebuild
	uint64_t acc_cnt[DBT_N_COUNTERS] = {};
	std::thread th;
};

// one kam line, or none: with `-s` a record is printed when its read name carries a source locus or the pair was assigned (cpp:2169)
bool append_kam_checked(Shared& S, std::string& o, const uint8_t* rec, uint32_t as_stride, const PairText& t) {
	long long src = -1;
	const dbt_pair_record* h0 = reinterpret_cast<const dbt_pair_record*>(rec);
	if (!S.opt.okam) return true;                                                // (records only travel for `-tb`)
	if (!S.opt.simmode && h0->dst == S.rpgg.nloci) return true;                  // removed pairs print only with a known source locus (cpp:2169)
	if (S.opt.simmode) {
		src = sim_source_locus(t, S.opt.simmode, S.rpgg.nloci);
		if (src == -2) { S.fail("-s " + std::to_string(S.opt.simmode) + ": cannot read the source locus from read name " + std::string(t.title, t.title_len)); return false; }
		const dbt_pair_record* h = reinterpret_cast<const dbt_pair_record*>(rec);
		if (!(src != (long long)S.rpgg.nloci || h->dst != S.rpgg.nloci)) return true;
	}
	append_kam_line(o, rec, as_stride, t, S.opt.isFastq, src);
	return true;
}

std::string dbt_err(const char* what, int rc) { return std::string(what) + ": " + dbt_strerror(rc) + ": " + dbt_last_error(); }

// `-bu`: the novel-edge events of the submission that just finished on `slot`; ordinal(p) = global ordinal of pair p of the submission
template <typename F>
bool fetch_bubbles(Shared& S, Worker& w, int slot, uint64_t npairs, F&& ordinal, std::vector<BubbleEvent>& out) {
	if (!S.opt.outputBubbles || !npairs) return true;
	const uint64_t cap = 4 * npairs + (1u << 20);
	std::vector<uint64_t> key(cap), edge(cap); std::vector<uint32_t> loc(cap);
	uint64_t n = 0;
	const int rc = dbt_bubble_events(w.ctx, slot, key.data(), edge.data(), loc.data(), cap, &n);
	if (rc) { S.fail(dbt_err("dbt_bubble_events", rc)); return false; }
	out.reserve(out.size() + n);
	for (uint64_t i = 0; i < n; ++i) out.push_back(BubbleEvent{(ordinal(key[i] >> 10) << 10) | (key[i] & 1023), edge[i], loc[i]});
	return true;
}

// `-b` with FASTQ: the batch's quality strings ride along, laid out like the bases (one phred score per base)
bool batch_quals(const HostBatch& hb, dbt_batch& b) {
	if (hb.qual_off.size() != 2 * hb.npairs + 1 || memcmp(hb.qual_off.data(), hb.offsets, (2 * hb.npairs + 1) * 8) != 0) return false;
	b.quals = reinterpret_cast<const uint8_t*>(hb.quals.data());
	return true;
}

// `-tb`: for every mate of this record that the bait filter removed, the bait k-mer bfilter_FPSv1 stopped at (host/bait.hpp)
void bait_track(Shared& S, const uint8_t* rec, const PairText& t, uint64_t ordinal, std::vector<BubbleEvent>& out) {
	const dbt_pair_record* h = reinterpret_cast<const dbt_pair_record*>(rec);
	if (h->dst0 >= S.baits.per_locus.size()) return;
	for (int m = 0; m < 2; ++m) {
		if (!(h->mate[m].flags & DBT_MF_BF)) continue;
		uint64_t km = 0;
		if (first_bait_violation(S.baits.per_locus[h->dst0], t.seq[m], t.seq_len[m], (uint32_t)S.opt.ksize, km, S.opt.isFastq ? t.qual[m] : nullptr, (int)S.opt.qth))
			out.push_back(BubbleEvent{(ordinal << 10) | ((uint64_t)m << 9), km, h->dst0});
	}
}

bool make_ctx(Shared& S, Worker& w, uint32_t max_len) {
	dbt_params p;
	dbt_default_params(&p);
	const Options& o = S.opt;
	p.device = w.gpu; p.k = (uint32_t)o.ksize; p.n_filter = (uint32_t)o.N_FILTER; p.nm_filter = (uint32_t)o.NM_FILTER;
	p.cth = (uint32_t)o.Cthreshold; p.nm_tr = (uint32_t)o.NM_TR; p.thread_cth = (uint32_t)o.thread_cth; p.max_ncorrection = (uint32_t)o.maxncorrection;
	p.flags = (o.extractFastX ? DBT_F_EXTRACT : ((o.okam || o.trackBait) ? DBT_F_KAM : 0)) | ((o.simmode || o.trackBait) ? DBT_F_SIM : 0) | (o.outputBubbles ? DBT_F_BUBBLES : 0) |
	          (o.bait ? DBT_F_BAIT : 0) | (o.qc ? DBT_F_QC : 0) | (o.threading ? DBT_F_THREADING : 0) | (o.threading && o.correction ? DBT_F_CORRECTION : 0) |
	          (o.aln ? DBT_F_ALN : 0) | (o.aln_minimal ? DBT_F_ALN_MIN : 0);
	p.qth = (uint32_t)o.qth; p.max_read_len = max_len; p.max_batch_pairs = (uint32_t)S.batch_pairs; p.max_records = (uint32_t)S.batch_pairs;
	p.table_load = o.table_load; p.bucket_slots = o.bucket_slots;
	int rc = dbt_create(&w.ctx, &p);
	if (rc) { S.fail(dbt_err("dbt_create", rc)); return false; }
	rc = dbt_load_index(w.ctx, &S.index);
	if (rc) { S.fail(dbt_err("dbt_load_index", rc)); return false; }
	rc = dbt_prepare(w.ctx, S.chunk_target ? S.chunk_target + RawReader::HEADROOM : 0);     // slot and splitter buffers now, not inside the first batches
	if (rc) { S.fail(dbt_err("dbt_prepare", rc)); return false; }
	w.ctx_len = max_len;
	return true;
}

// download the counts of the current context into the worker's host accumulators
bool retire_ctx(Shared& S, Worker& w) {
	if (!w.ctx) return true;
	const uint64_t nk = dbt_nk_tr(w.ctx), nl = dbt_nloci(w.ctx);
	std::vector<uint64_t> tr(std::max<uint64_t>(nk, 1)), kmc(std::max<uint64_t>(nl, 1)); std::vector<uint32_t> nm(std::max<uint64_t>(nl, 1));
	uint64_t cnt[DBT_N_COUNTERS];
	int rc = dbt_download(w.ctx, tr.data(), kmc.data(), nm.data(), cnt);
	if (rc) { S.fail(dbt_err("dbt_download", rc)); return false; }
	if (w.acc_tr.empty()) { w.acc_tr.assign(tr.size(), 0); w.acc_kmc.assign(kmc.size(), 0); w.acc_nmap.assign(nm.size(), 0); }
	for (size_t i = 0; i < tr.size(); ++i) w.acc_tr[i] += tr[i];
	for (size_t i = 0; i < kmc.size(); ++i) { w.acc_kmc[i] += kmc[i]; w.acc_nmap[i] += nm[i]; }
	for (int i = 0; i < DBT_N_COUNTERS; ++i) w.acc_cnt[i] += cnt[i];
	dbt_destroy(w.ctx); w.ctx = nullptr;
	return true;
}

void worker_main(Shared& S, Worker& w, Channel<BatchOut>& out, Channel<HostBatch*>& recycle) {
	struct Inflight { HostBatch* hb = nullptr; dbt_batch_result res{}; std::vector<uint8_t> recs; std::chrono::steady_clock::time_point t0; };
	Inflight fl[DBT_N_SLOTS];
	auto finish = [&](int s) -> bool {
		Inflight& f = fl[s];
		if (!f.hb) return true;
		int rc = dbt_wait(w.ctx, s);
		if (rc) { S.fail(dbt_err("dbt_wait", rc)); return false; }
		BatchOut bo;
		bo.index = f.hb->seq;
		bo.npairs = f.hb->npairs; bo.nreads_total = f.hb->nreads_total_after; bo.unpaired = f.hb->unpaired_after;
		memcpy(bo.counters, f.res.counters, sizeof bo.counters);
		if (!fetch_bubbles(S, w, s, f.hb->npairs, [&](uint64_t p) { return f.hb->first_pair_index + p; }, bo.events)) return false;
		std::sort(bo.events.begin(), bo.events.end(), [](const BubbleEvent& a, const BubbleEvent& b) { return a.key < b.key; });
		std::sort(bo.bait_events.begin(), bo.bait_events.end(), [](const BubbleEvent& a, const BubbleEvent& b) { return a.key < b.key; });
		if (S.bait_hits) {
			const uint32_t stride = dbt_record_stride(w.ctx);
			for (uint64_t i = 0; i < f.res.n_records; ++i) {
				const uint8_t* rec = f.recs.data() + i * stride;
				const uint64_t pi = reinterpret_cast<const dbt_pair_record*>(rec)->pair_index - f.hb->first_pair_index;
				bait_track(S, rec, pair_text(*f.hb, pi, S.opt.isFastq), f.hb->first_pair_index + pi, bo.bait_events);
			}
		}
		if (S.want_records && f.res.n_records) {
			const uint32_t stride = dbt_record_stride(w.ctx), as_stride = dbt_record_as_stride(w.ctx);
			bo.text.reserve((size_t)f.res.n_records * 520);
			if (S.opt.extractFastX) {
				for (uint64_t i = 0; i < f.res.n_records; ++i) {
					const uint8_t* rec = f.recs.data() + i * stride;
					append_extracted(bo.text, rec, pair_text(*f.hb, reinterpret_cast<const dbt_pair_record*>(rec)->pair_index - f.hb->first_pair_index, S.opt.isFastq), S.opt.isFastq, S.opt.extractFastX);
				}
			} else if (S.thr_graph) {
				const uint32_t aln_stride = dbt_record_aln_stride(w.ctx), es_cap = dbt_record_es_cap(w.ctx);
				for (uint64_t i = 0; i < f.res.n_records; ++i) append_aln_line(bo.text, f.recs.data() + i * stride, aln_stride, es_cap, *f.hb);
			} else
				for (uint64_t i = 0; i < f.res.n_records; ++i) {
					const uint8_t* rec = f.recs.data() + i * stride;
					if (!append_kam_checked(S, bo.text, rec, as_stride, pair_text(*f.hb, reinterpret_cast<const dbt_pair_record*>(rec)->pair_index - f.hb->first_pair_index, S.opt.isFastq))) return false;
				}
		}
		bo.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - f.t0).count();
		recycle.push(f.hb);
		f.hb = nullptr;
		out.push(std::move(bo));
		return true;
	};
	HostBatch* hb = nullptr;
	int k = 0;
	while (!S.failed && w.in.pop(hb)) {
		if (hb->max_len > w.ctx_len || !w.ctx) {              // first batch, or a read longer than the context was built for
			for (int s = 0; s < DBT_N_SLOTS; ++s) if (!finish((k + s) & 1)) break;      // older slot first
			if (S.failed || !retire_ctx(S, w)) break;
			uint32_t L = std::max<uint32_t>({hb->max_len, S.opt.max_read_len, (uint32_t)S.opt.ksize, 150u});
			if (w.ctx_len) L = std::max<uint32_t>(L, w.ctx_len + w.ctx_len / 4);
			if (!make_ctx(S, w, L)) break;
		}
		const int s = k++ & 1;
		if (!finish(s)) break;
		Inflight& f = fl[s];
		f.hb = hb; f.t0 = std::chrono::steady_clock::now();
		const uint32_t stride = dbt_record_stride(w.ctx);
		if (S.want_records) f.recs.resize((size_t)S.batch_pairs * stride);
		f.res = dbt_batch_result{};
		f.res.capacity = S.want_records ? S.batch_pairs : 0; f.res.records = S.want_records ? f.recs.data() : nullptr;
		dbt_batch b{};
		b.first_pair_index = hb->first_pair_index; b.npairs = hb->npairs; b.layout = DBT_LAYOUT_ASCII; b.bases = hb->bases; b.offsets = hb->offsets;
		if (S.opt.bait && S.opt.isFastq && !batch_quals(*hb, b)) { S.fail("a FASTQ record whose quality string is not as long as its sequence (-b needs one phred score per base)"); break; }
		int rc = dbt_submit(w.ctx, &b, &f.res, s);
		if (rc) { S.fail(dbt_err("dbt_submit", rc)); break; }
	}
	if (!S.failed) for (int s = 0; s < DBT_N_SLOTS; ++s) { const int ss = (k + s) & 1; if (!finish(ss)) break; }
	if (S.failed) recycle.close();                           // unblock the reader
}

using Clock = std::chrono::steady_clock;
static uint64_t us_since(Clock::time_point t0) { return (uint64_t)std::chrono::duration_cast<std::chrono::microseconds>(Clock::now() - t0).count(); }

// One GPU's share of the raw chunks: dbt_upload_raw (H2D + record split + neighbour pairing + 2-bit pack on the device),
// the chunk's turn at the host pairer (singletons, end-of-stream rules; whole chunks the device hands back), the pair
// loop on the device, and -- when the pairer formed pairs of its own -- one more ordinary ASCII submission whose lines
// are merged into the chunk's output by the position of the late mate.
void raw_worker_main(Shared& S, Worker& w, Channel<BatchOut>& out, Channel<RawChunk*>& recycle) {
	struct Inflight {
		RawChunk* ch = nullptr; bool gpu = false; dbt_raw_info info{}; dbt_batch_result res{};
		StreamPairer::Extra ex;
		uint64_t nreads_total = 0, unpaired = 0, first_pair = 0; Clock::time_point t0;
	};
	Inflight fl[DBT_N_SLOTS];
	const bool fq = S.opt.isFastq;
	std::vector<uint8_t> recs2;
	std::function<bool(int)> finish;
	auto rebuild = [&](uint32_t need_len, int keep_slot) -> bool {       // a read longer than the context was built for
		for (int s = 0; s < DBT_N_SLOTS; ++s) if (s != keep_slot && !finish(s)) return false;
		if (!retire_ctx(S, w)) return false;
		uint32_t L = std::max<uint32_t>({need_len, S.opt.max_read_len, (uint32_t)S.opt.ksize, 150u});
		if (w.ctx_len) L = std::max<uint32_t>(L, w.ctx_len + w.ctx_len / 4);
		for (auto& r : w.recs) if (r) { dbt_host_free(r); r = nullptr; }             // sized for the old record stride
		return make_ctx(S, w, L);
	};
	auto ensure_recs = [&](int s) -> bool {
		if (!S.want_records || w.recs[s]) return true;
		w.recs[s] = (uint8_t*)dbt_host_alloc((size_t)S.batch_pairs * dbt_record_stride(w.ctx));
		if (!w.recs[s]) { S.fail("cannot allocate pinned record buffer"); return false; }
		return true;
	};
	finish = [&](int s) -> bool {
		Inflight& f = fl[s];
		if (!f.ch) return true;
		RawChunk* ch = f.ch; f.ch = nullptr;
		BatchOut bo;
		bo.index = ch->seq; bo.nreads_total = f.nreads_total; bo.unpaired = f.unpaired;
		uint64_t ngpu = 0;
		if (f.gpu) {
			auto tw = Clock::now();
			int rc = dbt_wait(w.ctx, s);
			S.t_wait_us += us_since(tw);
			if (rc) { S.fail(dbt_err("dbt_wait", rc)); return false; }
			for (int i = 0; i < DBT_N_COUNTERS; ++i) bo.counters[i] += f.res.counters[i];
			bo.npairs += f.info.npairs; ngpu = f.res.n_records;
			// global ordinal of device pair p = pairs before the chunk + p + the host-formed pairs of the chunk that come before it
			if (!fetch_bubbles(S, w, s, f.info.npairs, [&](uint64_t p) {
				return f.first_pair + p + (uint64_t)(std::lower_bound(f.ex.key.begin(), f.ex.key.end(), (uint64_t)ch->pair_offset[p]) - f.ex.key.begin()); }, bo.events)) return false;
		}
		const uint32_t np_gpu = f.gpu ? f.info.npairs : 0;
		const HostBatch& xb = f.ex.hb;
		auto format = [&](std::string& o, const uint8_t* rec, bool from_gpu) {
			const uint64_t pi = reinterpret_cast<const dbt_pair_record*>(rec)->pair_index;
			const PairText t = from_gpu ? pair_text_raw(ch->text(), ch->pair_offset[pi], fq) : pair_text(xb, pi, fq);
			if (S.opt.extractFastX) append_extracted(o, rec, t, fq, S.opt.extractFastX);
			else if (S.thr_graph) append_aln_line(o, rec, dbt_record_aln_stride(w.ctx), dbt_record_es_cap(w.ctx), t);
			else append_kam_checked(S, o, rec, dbt_record_as_stride(w.ctx), t);
		};
		auto tf = Clock::now();
		if (S.bait_hits && ngpu) {
			const uint32_t stride = dbt_record_stride(w.ctx);
			for (uint64_t i = 0; i < ngpu; ++i) {
				const uint8_t* rec = w.recs[s] + i * stride;
				const uint64_t pi = reinterpret_cast<const dbt_pair_record*>(rec)->pair_index;
				const uint64_t ord = f.first_pair + pi + (uint64_t)(std::lower_bound(f.ex.key.begin(), f.ex.key.end(), (uint64_t)ch->pair_offset[pi]) - f.ex.key.begin());
				bait_track(S, rec, pair_text_raw(ch->text(), ch->pair_offset[pi], fq), ord, bo.bait_events);
			}
		}
		if (!xb.npairs) {
			if (S.want_records && ngpu) {
				const uint32_t stride = dbt_record_stride(w.ctx);
				const int T = (int)std::min<uint64_t>(std::min<uint64_t>(std::max<uint64_t>(S.opt.nproc, 1), 16), ngpu / 1024 + 1);   // -p host threads format the lines
				if (T <= 1) {
					bo.parts.resize(1); bo.parts[0] = S.strings.get();
					for (uint64_t i = 0; i < ngpu; ++i) format(bo.parts[0], w.recs[s] + i * stride, true);
				} else {
					bo.parts.resize(T);
					for (auto& x : bo.parts) x = S.strings.get();
					std::vector<std::thread> th;
					for (int t = 0; t < T; ++t) th.emplace_back([&, t] {
						const uint64_t a = ngpu * t / T, b = ngpu * (t + 1) / T;
						for (uint64_t i = a; i < b; ++i) format(bo.parts[t], w.recs[s] + i * stride, true);
					});
					for (auto& x : th) x.join();
				}
			}
		} else {
			// Pairs the host pairer formed for this chunk: an ordinary ASCII batch on the same slot, its lines merged into the
			// chunk's by the position of the late mate.  The device pairs are formatted first: the context may have to be
			// rebuilt for a longer read, which changes the record stride.
			std::vector<std::pair<uint64_t, std::string>> lines;
			if (S.want_records && ngpu) {
				const uint32_t stride = dbt_record_stride(w.ctx);
				lines.reserve(ngpu + xb.npairs);
				for (uint64_t i = 0; i < ngpu; ++i) {
					const uint8_t* rec = w.recs[s] + i * stride;
					lines.emplace_back((uint64_t)ch->pair_offset[reinterpret_cast<const dbt_pair_record*>(rec)->pair_index], std::string());
					format(lines.back().second, rec, true);
				}
			}
			if (!w.ctx || xb.max_len > w.ctx_len) { if (!rebuild(xb.max_len, s)) return false; }
			const uint32_t stride = dbt_record_stride(w.ctx);
			dbt_batch_result res2{};
			if (S.want_records) recs2.resize((size_t)xb.npairs * stride);
			res2.capacity = S.want_records ? xb.npairs : 0; res2.records = S.want_records ? recs2.data() : nullptr;
			dbt_batch b{};
			b.first_pair_index = 0; b.npairs = xb.npairs; b.layout = DBT_LAYOUT_ASCII; b.bases = xb.bases; b.offsets = xb.offsets;
			if (S.opt.bait && fq && !batch_quals(xb, b)) { S.fail("a FASTQ record whose quality string is not as long as its sequence (-b needs one phred score per base)"); return false; }
			int rc = dbt_submit(w.ctx, &b, &res2, s);
			if (!rc) rc = dbt_wait(w.ctx, s);
			if (rc) { S.fail(dbt_err("dbt_submit (host-paired reads)", rc)); return false; }
			for (int i = 0; i < DBT_N_COUNTERS; ++i) bo.counters[i] += res2.counters[i];
			bo.npairs += xb.npairs;
			if (!fetch_bubbles(S, w, s, xb.npairs, [&](uint64_t j) {
				const uint64_t kj = f.ex.key[j];
				return f.first_pair + j + (uint64_t)(kj > 0xFFFFFFFFull ? np_gpu : std::lower_bound(ch->pair_offset, ch->pair_offset + np_gpu, (uint32_t)kj) - ch->pair_offset); }, bo.events)) return false;
			for (uint64_t j = 0; S.bait_hits && j < res2.n_records; ++j) {
				const uint8_t* rec = recs2.data() + j * stride;
				const uint64_t pj = reinterpret_cast<const dbt_pair_record*>(rec)->pair_index, kj = f.ex.key[pj];
				const uint64_t ord = f.first_pair + pj + (uint64_t)(kj > 0xFFFFFFFFull ? np_gpu : std::lower_bound(ch->pair_offset, ch->pair_offset + np_gpu, (uint32_t)kj) - ch->pair_offset);
				bait_track(S, rec, pair_text(xb, pj, fq), ord, bo.bait_events);
			}
			for (uint64_t j = 0; j < res2.n_records; ++j) {
				const uint8_t* rec = recs2.data() + j * stride;
				lines.emplace_back(f.ex.key[reinterpret_cast<const dbt_pair_record*>(rec)->pair_index], std::string());
				format(lines.back().second, rec, false);
			}
			std::stable_sort(lines.begin(), lines.end(), [](const auto& a, const auto& b2) { return a.first < b2.first; });
			for (auto& l : lines) bo.text += l.second;
		}
		S.t_format_us += us_since(tf);
		std::sort(bo.events.begin(), bo.events.end(), [](const BubbleEvent& a, const BubbleEvent& b) { return a.key < b.key; });
		bo.seconds = std::chrono::duration<double>(Clock::now() - f.t0).count();
		recycle.push(ch);
		out.push(std::move(bo));
		return true;
	};
	RawChunk* ch = nullptr;
	int k = 0;
	const uint32_t minlen = (uint32_t)(S.opt.Cthreshold + S.opt.ksize - 1);
	for (;;) {
		auto ti = Clock::now();
		if (S.failed || !w.rin.pop(ch)) break;
		S.t_worker_idle_us += us_since(ti);
		const int s = k++ & 1;
		if (!finish(s)) break;
		Inflight& f = fl[s];
		f.t0 = Clock::now(); f.gpu = false; f.info = dbt_raw_info{};
		bool use_gpu = ch->nbytes() > 0, ok = true;
		if (use_gpu) {
			if (!w.ctx && !make_ctx(S, w, std::max<uint32_t>({S.opt.max_read_len, (uint32_t)S.opt.ksize, 150u}))) break;
			if (!ensure_recs(s)) break;
			auto ts = Clock::now();
			for (int attempt = 0; attempt < 2; ++attempt) {
				dbt_raw_batch rb{};
				rb.first_pair_index = 0; rb.text = ch->text(); rb.nbytes = ch->nbytes(); rb.is_fastq = fq ? 1 : 0; rb.min_read_len = minlen;
				rb.pair_offset = ch->pair_offset; rb.pair_hash = ch->pair_hash; rb.single_offset = ch->single_offset; rb.single_cap = ch->single_cap;
				const int rc = dbt_upload_raw(w.ctx, &rb, &f.info, s);
				if (rc == DBT_OK) break;
				if (rc == DBT_E_CAPACITY && f.info.max_read_len > w.ctx_len && attempt == 0) {
					if (!rebuild(f.info.max_read_len, s) || !ensure_recs(s)) { ok = false; break; }
					continue;
				}
				S.fail(dbt_err("dbt_upload_raw", rc)); ok = false; break;
			}
			S.t_split_us += us_since(ts);
		}
		// the chunk's turn at the pairer (stream order over all GPUs)
		{
			auto tt = Clock::now();
			std::unique_lock<std::mutex> l(S.turn.m);
			S.turn.cv.wait(l, [&] { return S.turn.next == ch->seq || S.failed; });
			S.t_turn_us += us_since(tt);
			if (ok && !S.failed) {
				auto tp = Clock::now();
				const bool fast = S.pairer->chunk(*ch, use_gpu, f.info.flags, f.info.npairs, f.info.nsingle, f.info.nshort_pairs, f.ex);
				f.gpu = fast && f.info.npairs > 0;
				if (!fast) f.info.npairs = 0;
				if (use_gpu && !fast) ++S.n_slow_chunks;
				++S.n_chunks;
				f.nreads_total = 2 * S.pairer->total_pairs(); f.unpaired = S.pairer->unpaired();
				f.first_pair = S.pairer->total_pairs() - (f.gpu ? f.info.npairs : 0) - f.ex.hb.npairs;
				S.t_pair_us += us_since(tp);
			}
			++S.turn.next;
			S.turn.cv.notify_all();
		}
		if (!ok || S.failed) { recycle.push(ch); break; }
		f.ch = ch;
		if (f.gpu) {
			if (!ensure_recs(s)) break;
			f.res = dbt_batch_result{};
			f.res.capacity = S.want_records ? S.batch_pairs : 0; f.res.records = S.want_records ? w.recs[s] : nullptr;
			auto tl = Clock::now();
			int rc = dbt_process(w.ctx, s);
			if (!rc) rc = dbt_fetch(w.ctx, &f.res, s);
			S.t_launch_us += us_since(tl);
			if (rc) { S.fail(dbt_err("dbt_process", rc)); break; }
		}
	}
	if (!S.failed) for (int s = 0; s < DBT_N_SLOTS; ++s) { const int ss = (k + s) & 1; if (!finish(ss)) break; }
	// (w.recs are released by main() after the last line is out: unpinning 1.4 GB takes longer than a chunk)
	if (S.failed) { recycle.close(); S.turn.cv.notify_all(); }
}

}  // namespace

int main(int argc, char** argv) {
	if (argc < 2) { fputs(usage_text(), stderr); return 0; }
	Shared S;
	Options& o = S.opt;
	std::string err = parse_args(argc, argv, o);
	if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
	if (o.trPrefix.empty() || o.fastxFname.empty() || o.outPrefix.empty()) { fprintf(stderr, "-qs, -fa|-fq and -o|-on are required\n"); return 1; }
	for (const char* unsupported : {
	                                o.g2pan ? "-m" : "", o.aug ? "-au" : "", o.trim ? "-t" : "", o.invkmer ? "-ik" : ""})
		if (*unsupported) { fprintf(stderr, "option %s is not part of the B200 align path (SURVEY.md 8(f)); use the reference binary for it\n", unsupported); return 1; }
	if (o.ksize < 1 || o.ksize > 27) { fprintf(stderr, "-k %llu unsupported (1..27)\n", (unsigned long long)o.ksize); return 1; }
	report_parameters(o, stderr);
	fprintf(stderr, "total number of loci in %s: ", o.trFname.c_str());

	const int ndev = dbt_device_count();
	if (ndev <= 0) { fprintf(stderr, "\nno CUDA device: %s (this build has no CPU path)\n", dbt_last_error()); return 1; }
	const int G = std::max(1, o.gpus);
	if (G > ndev) { fprintf(stderr, "\n--gpus %d but only %d CUDA device(s)\n", G, ndev); return 1; }

	// ---- load the RPGG (cpp:2459-2504) ----
	auto t_load = std::chrono::steady_clock::now();
	const int hw = (int)std::max(1u, std::thread::hardware_concurrency());
	if (!load_rpgg(o.trPrefix, o.trFname, o.qc ? o.qcFn : std::string(), std::max<int>((int)o.nproc, std::min(hw, 16)), S.rpgg, err)) { fprintf(stderr, "\n%s\n", err.c_str()); return 1; }
	const Rpgg& g = S.rpgg;
	fprintf(stderr, "%llu\n", (unsigned long long)g.nloci);
	if (!g.nloci) { fprintf(stderr, "no loci in %s\n", o.trFname.c_str()); return 1; }
	S.index.nloci = g.nloci; S.index.nk_dbi = g.dbi_keys.size(); S.index.dbi_keys = g.dbi_keys.data(); S.index.dbi_vals = g.dbi_vals.data();
	S.index.nvv = g.vv.size(); S.index.vv = g.vv.empty() ? nullptr : g.vv.data();
	S.index.tr_n = g.tr_n.data(); S.index.tr_kmers = g.tr_kmers.data(); S.index.fl_n = g.fl_n.data(); S.index.fl_kmers = g.fl_kmers.data();
	S.index.qc = o.qc ? g.qc.data() : nullptr;
	if (o.threading) {
		// The reference's HEAD never loads a graph and its threading step passes nothing (cpp:2072-2090, 2468).  When the RPGG
		// ships its graph (PREF.graph.kmers from `fa2kmers -g`, or a v1.3 PREF.graph.umap) the restored threading runs instead.
		const std::string gk = o.trPrefix + ".graph.kmers", gu = o.trPrefix + ".graph.umap";
		bool okg = true;
		if (file_exists(gk)) okg = read_graph_kmers(gk, g.nloci, S.rpgg, std::max<int>((int)o.nproc, std::min(hw, 16)), err);
		else if (file_exists(gu)) okg = read_graph_umap(gu, g.nloci, S.rpgg, err);
		else fprintf(stderr, "[danbing-tk b200] no %s: -g/-gc behaves like the reference at HEAD (no read passes threading)\n", gk.c_str());
		if (!okg) { fprintf(stderr, "\n%s\n", err.c_str()); return 1; }
		if (!g.g_n.empty()) {
			S.index.g_n = g.g_n.data(); S.index.g_kmers = g.g_kmers.data(); S.index.g_masks = g.g_masks.data();
			S.thr_graph = true;
			fprintf(stderr, "[danbing-tk b200] graph threading restored: %llu nodes\n", (unsigned long long)g.g_kmers.size());
		}
	}
	if (o.extractFastX && S.thr_graph) { fprintf(stderr, "\n-e with a graph in the RPGG (restored threading) is outside the reference's behaviour; drop -g/-gc or the graph file\n"); return 1; }
	if (o.extractFastX && o.extractFastX != 1 && o.extractFastX != 2) { fprintf(stderr, "\n-e takes 1 or 2\n"); return 1; }
	if (o.trackBait && !o.bait) { fprintf(stderr, "\n-tb needs -b\n"); return 1; }
	if (o.bait) {
		if (S.thr_graph) { fprintf(stderr, "\n-b with a graph in the RPGG (restored threading): the reference filters baits only without threading\n"); return 1; }
		if (o.extractFastX) { fprintf(stderr, "\n-b with -e: extraction stops before the bait filter\n"); return 1; }
		if (!read_bait_kmdb(o.trPrefix + (o.qc ? ".qc" : "") + ".bt.kmdb", g.nloci, o.trackBait, S.baits, err)) { fprintf(stderr, "\n%s\n", err.c_str()); return 1; }
		S.index.bt_n = S.baits.n.data(); S.index.bt_kmers = S.baits.kmers.data(); S.index.bt_vals = S.baits.vals.data();
		if (o.trackBait) S.bait_hits.reset(new BubbleAcc(g.nloci, (uint64_t)(300000 * o.readsPerBatchFactor), 0, ".btk.kmdb"));
	}
	if (o.outputBubbles) {
		if (S.thr_graph) { fprintf(stderr, "\n-bu with a graph in the RPGG (restored threading): the reference counts bubbles only without threading\n"); return 1; }
		if (o.extractFastX) { fprintf(stderr, "\n-bu with -e: extraction counts nothing\n"); return 1; }
		if (!read_kdb(o.trPrefix + ".tre.kdb", S.rpgg.tre_n, S.rpgg.tre_kmers, err) || S.rpgg.tre_n.size() != g.nloci) { fprintf(stderr, "\n%s\n", err.empty() ? (o.trPrefix + ".tre.kdb: locus count differs").c_str() : err.c_str()); return 1; }
		S.index.tre_n = S.rpgg.tre_n.data(); S.index.tre_kmers = S.rpgg.tre_kmers.data();
		S.bubbles.reset(new BubbleAcc(g.nloci, (uint64_t)(300000 * o.readsPerBatchFactor)));
	}
	S.want_records = o.extractFastX ? true : (S.thr_graph ? o.aln : (o.okam || o.trackBait));
	fprintf(stderr, "%llu bait loci in baitDB\ndeserialized graph/index and read tr.kmers in %lld sec.\n# unique kmers in kmerDBi: %llu\n", (unsigned long long)(o.bait ? g.nloci : 0),
	        (long long)std::chrono::duration_cast<std::chrono::seconds>(std::chrono::steady_clock::now() - t_load).count(), (unsigned long long)g.dbi_keys.size());

	// ---- batches ----
	const uint64_t readsPerRefBatch = (uint64_t)(300000 * o.readsPerBatchFactor);          // cpp:1835
	const uint64_t minlen = o.Cthreshold + o.ksize - 1;
	const bool raw = !o.host_reader;
	const int NB = 2 * G + 4;
	std::vector<std::unique_ptr<HostBatch>> pool;
	std::vector<std::unique_ptr<RawChunk>> rpool;
	Channel<HostBatch*> recycle;
	Channel<RawChunk*> rrecycle;
	if (raw) {
		// a submission = a chunk of raw text; it cannot hold more pairs than bytes / (two reads of the minimum length + their lines' overhead)
		const uint64_t per_pair = 2 * (minlen + 4);
		if (o.batch_pairs) { S.batch_pairs = o.batch_pairs; S.chunk_target = std::max<uint64_t>(o.batch_pairs * per_pair, 4096); }
		else { S.chunk_target = (o.chunk_mb ? o.chunk_mb : 128) << 20; S.batch_pairs = S.chunk_target / per_pair + 16; }
		if (o.batch_pairs && o.chunk_mb) S.chunk_target = o.chunk_mb << 20;
		S.batch_pairs = std::max<uint64_t>(S.batch_pairs, (S.chunk_target + RawReader::HEADROOM) / per_pair + 16);
		if (S.chunk_target >= (3ull << 30) || S.batch_pairs >= (1ull << 31)) { fprintf(stderr, "--chunk-mb / --batch-pairs too large (a submission stays below 3 GiB)\n"); return 1; }
		S.pairer.reset(new StreamPairer(o.isFastq, minlen, readsPerRefBatch));
		const uint32_t single_cap = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(S.batch_pairs / 8, 1024), 1u << 22);
		for (int i = 0; i < NB; ++i) {
			std::unique_ptr<RawChunk> c(new RawChunk());
			c->cap = RawReader::HEADROOM + S.chunk_target + 4096;
			c->buf = (uint8_t*)dbt_host_alloc(c->cap + 64);
			c->pair_offset = (uint32_t*)dbt_host_alloc(S.batch_pairs * 4); c->pair_hash = (uint64_t*)dbt_host_alloc(S.batch_pairs * 8);
			c->single_offset = (uint32_t*)dbt_host_alloc((size_t)single_cap * 4); c->pairs_cap = (uint32_t)S.batch_pairs; c->single_cap = single_cap;
			if (!c->buf || !c->pair_offset || !c->pair_hash || !c->single_offset) { fprintf(stderr, "cannot allocate pinned chunk buffers\n"); return 1; }
			rrecycle.push(c.get());
			rpool.push_back(std::move(c));
		}
	} else {
		S.batch_pairs = o.batch_pairs ? o.batch_pairs : (1ull << 20);
		S.batch_pairs = std::max<uint64_t>(S.batch_pairs, readsPerRefBatch / 2 + 1);
		if (S.batch_pairs >= (1ull << 31)) { fprintf(stderr, "--batch-pairs too large\n"); return 1; }
		S.bases_cap = S.batch_pairs * 2 * 160;
		for (int i = 0; i < NB; ++i) {
			std::unique_ptr<HostBatch> hb(new HostBatch());
			hb->bases = (uint8_t*)dbt_host_alloc(S.bases_cap); hb->bases_cap = S.bases_cap;
			hb->offsets = (uint64_t*)dbt_host_alloc((2 * S.batch_pairs + 1) * 8); hb->pairs_cap = S.batch_pairs;
			if (!hb->bases || !hb->offsets) { fprintf(stderr, "cannot allocate pinned batch buffers\n"); return 1; }
			recycle.push(hb.get());
			pool.push_back(std::move(hb));
		}
	}

	fprintf(stderr, "creating data for each process...\n");
	std::vector<std::unique_ptr<Worker>> workers;
	Channel<BatchOut> outq;
	for (int i = 0; i < G; ++i) { workers.emplace_back(new Worker()); workers.back()->gpu = i; }
	{	// the device-resident index of every GPU is part of loading the RPGG (cpp:2484-2504), not of the query phase
		auto t_ctx = std::chrono::steady_clock::now();
		const uint32_t L0 = std::max<uint32_t>({o.max_read_len, (uint32_t)o.ksize, 150u});
		std::vector<std::thread> th;
		for (auto& w : workers) th.emplace_back([&S, L0, raw, wp = w.get()] {
			if (!make_ctx(S, *wp, L0) || !raw || !S.want_records) return;
			for (auto& r : wp->recs) { r = (uint8_t*)dbt_host_alloc((size_t)S.batch_pairs * dbt_record_stride(wp->ctx)); if (!r) S.fail("cannot allocate pinned record buffer"); }
		});
		for (auto& t : th) t.join();
		if (S.failed) { fprintf(stderr, "danbing-tk: %s\n", S.err.c_str()); return 1; }
		if (o.timing) fprintf(stderr, "[danbing-tk b200 timing] device index built on %d GPU(s) in %.3f s\n", G, std::chrono::duration<double>(std::chrono::steady_clock::now() - t_ctx).count());
	}
	auto t_query = std::chrono::steady_clock::now();
	for (auto& w : workers) {
		if (raw) w->th = std::thread([&S, &outq, &rrecycle, wp = w.get()] { raw_worker_main(S, *wp, outq, rrecycle); });
		else w->th = std::thread([&S, &outq, &recycle, wp = w.get()] { worker_main(S, *wp, outq, recycle); });
	}

	// ---- reader thread (the serial section of CountWords, cpp:1885-1986) ----
	std::thread reader([&] {
		if (raw) {
			// raw text in whole records; the GPU splits and pairs it (dbt_upload_raw), the StreamPairer takes what is left over
			RawReader rd(o.isFastq, (int)std::min<uint64_t>(std::max<uint64_t>(o.nproc, 1), 16));
			if (!rd.open(o.fastxFname)) { S.fail("cannot open " + o.fastxFname); }
			RawChunk* ch = nullptr;
			for (;;) {
				auto tb = Clock::now();
				if (S.failed || !rrecycle.pop(ch)) break;
				S.t_reader_blocked_us += us_since(tb);
				auto tr = Clock::now();
				const bool got = rd.fill(*ch, S.chunk_target);
				S.t_read_us += us_since(tr);
				if (!got) { if (!rd.error().empty()) S.fail(o.fastxFname + ": " + rd.error()); rrecycle.push(ch); break; }
				workers[ch->seq % G]->rin.push(ch);
			}
			for (auto& w : workers) w->rin.close();
			return;
		}
		FastxReader rd(o.isFastq, minlen, readsPerRefBatch);
		if (o.fastxFname.size() > 3 && o.fastxFname.compare(o.fastxFname.size() - 3, 3, ".gz") == 0) S.fail("--host-reader reads plain text only (the default reader inflates .gz)");
		else if (!rd.open(o.fastxFname)) { S.fail("cannot open " + o.fastxFname); }
		uint64_t first = 0, b = 0;
		HostBatch* hb = nullptr;
		while (!S.failed && recycle.pop(hb)) {
			hb->reset(first);
			if (!rd.fill(*hb)) { recycle.push(hb); break; }
			if (hb->npairs == 0) {                            // fill() made progress without emitting: the pending pair does not fit an EMPTY batch
				S.fail("a read pair does not fit the batch buffer (raise --batch-pairs)"); recycle.push(hb); break;
			}
			hb->seq = b;
			first += hb->npairs;
			workers[b % G]->in.push(hb);
			++b;
		}
		for (auto& w : workers) w->in.close();
	});

	// ---- writer: kam lines in batch order (cpp:2253-2279) ----
	std::thread closer([&] { reader.join(); for (auto& w : workers) w->th.join(); outq.close(); });
	std::map<uint64_t, BatchOut> held;
	uint64_t next_seq = 0;
	std::vector<char> obuf(1 << 22);
	setvbuf(stdout, obuf.data(), _IOFBF, obuf.size());
	uint64_t tot[DBT_N_COUNTERS] = {};
	BatchOut bo;
	auto emit = [&](BatchOut& x) {
		auto tw = Clock::now();
		if (!x.text.empty()) fwrite(x.text.data(), 1, x.text.size(), stdout);
		for (auto& part : x.parts) { if (!part.empty()) fwrite(part.data(), 1, part.size(), stdout); S.strings.put(std::move(part)); }
		S.t_write_us += us_since(tw);
		const uint64_t* c = x.counters;
		const double nr = (double)(2 * x.npairs);
		fprintf(stderr, "Buffered reading %llu\t%llu\t%llu\n", (unsigned long long)(2 * x.npairs), (unsigned long long)x.nreads_total, (unsigned long long)x.unpaired);
		fprintf(stderr, "Batch query in %lld sec. %llu/%g/%g/%llu/%llu/%llu/%llu/%llu/%llu/%llu/%llu\n", (long long)x.seconds,
		        (unsigned long long)c[DBT_C_NSHORT], (float)c[DBT_C_NHASH0] / nr, (float)c[DBT_C_NHASH1] / (nr - (double)c[DBT_C_NSUBFILTERED]),
		        (unsigned long long)c[DBT_C_NSUBFILTERED], (unsigned long long)c[DBT_C_NKMERFILTERED], (unsigned long long)c[DBT_C_NLOCUSASSIGNFILTERED],
		        (unsigned long long)c[DBT_C_NQCFILTERED], (unsigned long long)c[DBT_C_NTHREADINGREADS], (unsigned long long)c[DBT_C_NFEASIBLEREADS],
		        (unsigned long long)c[DBT_C_NBAITFILTERED], (unsigned long long)c[DBT_C_NASGNREADS]);
		if (S.bubbles) S.bubbles->replay(x.events);
		if (S.bait_hits) S.bait_hits->replay(x.bait_events);
		for (int i = 0; i < DBT_N_COUNTERS; ++i) tot[i] += c[i];
		next_seq = x.index + 1;
	};
	while (outq.pop(bo)) {
		if (bo.index != next_seq) { held.emplace(bo.index, std::move(bo)); continue; }
		emit(bo);
		for (auto it = held.find(next_seq); it != held.end(); it = held.find(next_seq)) { emit(it->second); held.erase(it); }
	}
	closer.join();
	fflush(stdout);
	const double query_wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_query).count();
	if (S.failed) { fprintf(stderr, "danbing-tk: %s\n", S.err.c_str()); return 1; }

	// ---- reduce over GPUs and collect ----
	std::vector<uint64_t> counts(g.tr_kmers.size(), 0), kmc(g.nloci, 0); std::vector<uint32_t> nmap(g.nloci, 0);
	{
		std::vector<Worker*> live;
		for (auto& w : workers) if (w->ctx) live.push_back(w.get());
		if (live.size() > 1) {
			uint8_t id[DBT_NCCL_ID_BYTES];
			int rc = dbt_nccl_unique_id(id);
			if (rc) { fprintf(stderr, "danbing-tk: %s\n", dbt_err("dbt_nccl_unique_id", rc).c_str()); return 1; }
			std::vector<std::thread> th; std::atomic<int> bad{0};
			for (size_t r = 0; r < live.size(); ++r) th.emplace_back([&, r] {
				int rc2 = dbt_comm_init(live[r]->ctx, (int)live.size(), (int)r, id);
				if (!rc2) rc2 = dbt_allreduce(live[r]->ctx);
				if (rc2) { bad = rc2; S.fail(dbt_err("NCCL reduce", rc2)); }
			});
			for (auto& t : th) t.join();
			if (bad) { fprintf(stderr, "danbing-tk: %s\n", S.err.c_str()); return 1; }
		}
		if (!live.empty()) {
			std::vector<uint64_t> tr(std::max<size_t>(counts.size(), 1)), km(std::max<size_t>(kmc.size(), 1)); std::vector<uint32_t> nm(std::max<size_t>(nmap.size(), 1));
			int rc = dbt_download(live[0]->ctx, tr.data(), km.data(), nm.data(), nullptr);
			if (rc) { fprintf(stderr, "danbing-tk: %s\n", dbt_err("dbt_download", rc).c_str()); return 1; }
			for (size_t i = 0; i < counts.size(); ++i) counts[i] += tr[i];
			for (size_t i = 0; i < kmc.size(); ++i) { kmc[i] += km[i]; nmap[i] += nm[i]; }
		}
		for (auto& w : workers) {
			for (size_t i = 0; i < w->acc_tr.size() && i < counts.size(); ++i) counts[i] += w->acc_tr[i];
			for (size_t i = 0; i < w->acc_kmc.size() && i < kmc.size(); ++i) { kmc[i] += w->acc_kmc[i]; nmap[i] += w->acc_nmap[i]; }
			if (w->ctx) { dbt_destroy(w->ctx); w->ctx = nullptr; }
		}
	}
	for (auto& w : workers) for (auto& r : w->recs) if (r) { dbt_host_free(r); r = nullptr; }
	for (auto& hb : pool) { dbt_host_free(hb->bases); dbt_host_free(hb->offsets); }
	for (auto& c : rpool) { dbt_host_free(c->buf); dbt_host_free(c->pair_offset); dbt_host_free(c->pair_hash); dbt_host_free(c->single_offset); }
	if (o.timing) {
		fprintf(stderr, "[danbing-tk b200 timing] query wall %.3f s; reader %.3f s; upload+split (sync) %.3f s; pairer %.3f s; wait %.3f s; format %.3f s; chunks %llu (%llu paired on the host); launch %.3f s; pairer turn wait %.3f s; worker idle %.3f s; reader blocked %.3f s; stdout write %.3f s\n",
		        query_wall, S.t_read_us / 1e6, S.t_split_us / 1e6, S.t_pair_us / 1e6, S.t_wait_us / 1e6, S.t_format_us / 1e6,
		        (unsigned long long)S.n_chunks.load(), (unsigned long long)S.n_slow_chunks.load(), S.t_launch_us / 1e6, S.t_turn_us / 1e6, S.t_worker_idle_us / 1e6,
		        S.t_reader_blocked_us / 1e6, S.t_write_us / 1e6);
	}

	// ---- final counters (cpp:2617-2627) and outputs (cpp:2631-2656) ----
	fprintf(stderr, "%llu reads processed in total.\n%llu reads removed by subsampled kmer-filter.\n%llu reads removed by kmer-filter.\n"
	                "%llu reads removed by bait locus.\n%llu reads removed by qual filter.\n%llu reads removed during locus assignment.\n"
	                "%llu reads removed by QC filter.\n%llu reads entered threading step.\n%llu reads passsed threading.\n"
	                "%llu reads assigned to TR region.\nparallel query completed in %lld sec.\n",
	        (unsigned long long)tot[DBT_C_NREADS], (unsigned long long)tot[DBT_C_NSUBFILTERED], (unsigned long long)tot[DBT_C_NKMERFILTERED],
	        (unsigned long long)tot[DBT_C_NBAITFILTERED], (unsigned long long)tot[DBT_C_NQUALFILTERED], (unsigned long long)tot[DBT_C_NLOCUSASSIGNFILTERED],
	        (unsigned long long)tot[DBT_C_NQCFILTERED], (unsigned long long)tot[DBT_C_NTHREADINGREADS], (unsigned long long)tot[DBT_C_NFEASIBLEREADS],
	        (unsigned long long)tot[DBT_C_NASGNREADS],
	        (long long)std::chrono::duration_cast<std::chrono::seconds>(std::chrono::steady_clock::now() - t_query).count());
	if (o.extractFastX) { fprintf(stderr, "all done!\n"); return 0; }      // cpp:2631: extraction writes no count files
	fprintf(stderr, "writing kmers...\n");
	bool ok;
	if (o.writeKmerName) ok = write_kmers_with_name(o.outPrefix + ".tr.kmers", g, counts, err);
	else {
		fprintf(stderr, "# kmer skipped: 0\ndone\nserializing %s.trkmc.ar\n", o.outPrefix.c_str());
		ok = write_trkmc_ar(o.outPrefix, counts, err) && write_tr_summary(o.outPrefix + ".tr.summary.txt", kmc, nmap, err);
	}
	if (ok && S.bubbles) {                                                       // cpp:2648-2651
		fprintf(stderr, "writing bubbles...\n");
		uint64_t nskip = 0;
		ok = S.bubbles->write(o.outPrefix, nskip, err);
		fprintf(stderr, "# kmer skipped: %llu\nserializing %s.bub.kmdb\n", (unsigned long long)nskip, o.outPrefix.c_str());
	}
	if (ok && S.bait_hits) {                                                     // cpp:2652-2655
		fprintf(stderr, "writing bait kmer hit statistics...\n");
		uint64_t nskip = 0;
		ok = S.bait_hits->write(o.outPrefix, nskip, err);
		fprintf(stderr, "# kmer skipped: %llu\nserializing %s.btk.kmdb\n", (unsigned long long)nskip, o.outPrefix.c_str());
	}
	if (!ok) { fprintf(stderr, "danbing-tk: %s\n", err.c_str()); return 1; }
	fprintf(stderr, "all done!\n");
	return 0;
}
