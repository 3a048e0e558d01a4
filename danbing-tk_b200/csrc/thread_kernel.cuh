// thread_kernel: restored graph threading (decision D1 of §8(c)) for the mates that carry an
// error.  thread_fast_kernel (thread_fast.cuh) has already finished every error-free mate of the pairs that passed the locus
// vote and the QC gate; its work list sv.thr2 names, per pair, the mates left over and whether the other mate threaded.
// One warp per work item: the listed mates are threaded through the de Bruijn graph of the assigned locus
// (isThreadFeasible, cpp:1114-1260), the pair counts if either mate threads, and then every k-mer of the edited k-mer
// vectors that is a TR k-mer of the locus is incremented (commented caller cpp:2072-2088 / 2195-2209; the clean mate's
// share was counted by thread_fast_kernel).  "cpp" = src/aQueryFasta_thread.cpp of the reference.
//
// How the sequential walk is made parallel.  The reference advances one k-mer at a time, looking the CURRENT node's
// out-edge mask up and comparing its successors with the next read k-mer (cpp:1167-1180).  Whenever that step is
// reached the current node equals the previous read k-mer (it was just matched, anchored or corrected), so the
// outcome of every step is a function of per-position data: the lanes probe the graph table once for ALL k-mers of
// the read (one 32 B sector each, MAXJ independent probes in flight per lane), classify every position by ballot --
// M match, N invalid window, H repeat of the previous k-mer (cpp:1152), A first valid k-mer after an N (cpp:1159),
// E no matching edge -- and the warp then jumps from event to event with find-next-clear scans over the M mask,
// filling `tr` / `es` for whole runs at once.  Only E events (sequencing errors) enter the correction code:
//   * errorCorrection_forward/backward (cpp:898-1106): the 3-level graph expansion is 4 + 16 probes taken by
//     lanes 0-3 / 0-15, the 4x4x4 reachability cube is one 64-bit mask, and the up to 62 extension walks of the
//     eight edit families run one per lane (two rounds); thread_ext_t::get_edit's scan order is replayed on the
//     score table (first maximum wins, cpp:628-647);
//   * edit_kmers_forward/backward (cpp:649-862) -- vector inserts/erases and the CIGAR merge rules -- are literal
//     and run on lane 0 over the shared-memory copies of kmers / cg.es / cg.tr.
// After an edit the k-mer vector has changed (up to k rewritten k-mers, +-1 or 2 entries), so all positions are
// probed and classified again; errors are a few per read, a refresh costs as much as the initial probe round.
#pragma once
#include "device_types.cuh"
#include "graph_table.cuh"

namespace dbt {

#ifndef THR_WARPS_N
#define THR_WARPS_N 4
#endif
#ifndef THR_MINBLOCKS
#define THR_MINBLOCKS 7                            // 72 registers; with THR_STAGE_GROUPS 1 the 6.4 KB of shared memory per warp admit 7 blocks per SM
#endif
#ifndef THR_STAGE_GROUPS
#define THR_STAGE_GROUPS 1                         // 32-position groups probed per round of thr_refresh (capped at MAXJ)
#endif
// inlining policy of the correction code, each measured on the GPU (profiles/r1/s17_thread_by_function.txt): out-of-line
// edit_kmers_backward 8.80 -> 9.21 ms, + edit_kmers_forward 10.55 ms, out-of-line vector primitives 9.56 ms (the calls spill
// through local memory, which misses an L1 squeezed by the shared-memory carve-out) -- all three stay inlined
#ifdef THR_NI_BACK
#define THR_BACK_INLINE __noinline__
#else
#define THR_BACK_INLINE __forceinline__
#endif
#ifdef THR_NI_FWD
#define THR_FWD_INLINE __noinline__
#else
#define THR_FWD_INLINE __forceinline__
#endif
#ifdef THR_NI_TS
#define THR_TS_INLINE __noinline__
#else
#define THR_TS_INLINE __forceinline__
#endif
#ifndef THR_REFRESH_INLINE
#define THR_REFRESH_INLINE __noinline__
#endif
constexpr int THR_WARPS = THR_WARPS_N;
constexpr uint64_t THR_MSC = 5;                    // min score for thread extension, cpp:1120

// CIGAR entry codes (include/danbing_b200.h, dbt_aln_mate): type 0 '*' 1 '=' 2 'X' 3 'D' 4 'I'; base 0 none 1-4 ACGT 5 '*'
enum : uint8_t { ET_STAR = 0, ET_EQ = 1, ET_X = 2, ET_D = 3, ET_I = 4 };
constexpr uint8_t EG_STAR = 5, ER_INVALID = 6;     // read base that is not ACGT: equals no graph base
__device__ __forceinline__ bool et_is_edit(uint8_t t) { return t >= ET_X; }
__device__ __forceinline__ uint8_t eg_comp(uint8_t g) { return (g >= 1 && g <= 4) ? (uint8_t)(5 - g) : (uint8_t)127; }   // baseComplement on characters, h:71-87

struct Txt {                                       // thread_ext_t, cpp:596-625 (the part that outlives the scorer)
	uint32_t score, nedits;
	uint8_t et[2], eg[2];
	uint32_t nm, nd, nins;
	uint64_t nrk;
	int a_lo, a_n;                                 // annotation range to fill from the refreshed TR bits
};

struct MateResult { int alned, err, nkm, nes, ntr, clean; };   // clean: error-free mate, es / tr not materialised (all '=', tr from the TR ballots)

// Mutable state of one mate's walk (warp-uniform).  It lives in SHARED memory (ThreadWarp::S): find_anchor / score / apply
// are single out-of-line functions working on it by reference (the kernel is instruction-fetch sensitive: every inlined
// copy of them cost more than the calls, profiles/r1 s5-s7), and as a per-thread local struct it took 32 x the footprint in
// an L1 that the 200 KB shared-memory carve-out leaves at ~56 KB -- ncu (s17): 55 % of the 87 M local loads/stores missed
// L1, i.e. the walk's own loop variables came back from L2.
struct MS {
	int nkm, nes, ntr, ni, err, ki;
	uint64_t nskip, ncorr;
	Txt x;
};

template <int MAXJ, int SB>
struct ThreadWarp {
	static constexpr int CAP = MAXJ * 32;          // k-mer vector / annotation capacity
	static constexpr int ECAP = MAXJ * 32 + 32;    // CIGAR capacity (read length + deletions)
	static constexpr int SG = MAXJ < THR_STAGE_GROUPS ? MAXJ : THR_STAGE_GROUPS;   // groups whose buckets are in flight together
	uint64_t km[CAP];                              // edited k-mer vector of the mate being threaded (strand-specific, KMER_INVALID = -1)
	uint32_t trx[2][CAP];                          // count-array position + 1 of each k-mer of both edited vectors (0: not a TR k-mer of the locus)
	uint64_t rck[CAP];                             // reverse complement of km[m][i] of the mate being threaded (free at extraction, kmer.cuh)
	uint32_t cls[6][MAXJ];                         // ballots over positions: M, N, H, A, present, TR
	uint4    stage[SG][SB][32];                    // home buckets of SG*32 positions of the mate being probed (cp.async); between
	                                               // refreshes its first 320 bytes are the k-mer window handed to the edit scorer
	__device__ __forceinline__ uint64_t* win() { return reinterpret_cast<uint64_t*>(&stage[0][0][0]); }
	uint32_t info[CAP];                            // GNode.bits of every position of the mate being threaded (thr_refresh)
	uint8_t  es_t[ECAP], es_r[ECAP], es_g[ECAP];   // cigar_t.es (cpp:53-60)
	uint8_t  tr[CAP];                              // cigar_t.tr: 0 '*' 1 '.' 2 '='
	MS       S;                                    // walk state of the mate being threaded
	MateResult res[2];                             // outcome of both mates
	int      kk;                                   // k-mer index handed back by lane 0's edit (ms_apply)
	uint32_t cnt[3];                               // this warp's probe counters: walk, ec, refresh (lane 0 adds; measurement)
};
enum { CL_M = 0, CL_N, CL_H, CL_A, CL_P, CL_T };

// lane-0 view of one read's threading state (thr_t of the oracle restatement)
struct TState {
	uint64_t* km; uint8_t *es_t, *es_r, *es_g, *tr;
	int nkm, nes, ntr, ni, err;
	int cap, ecap;
	uint32_t k; uint64_t rmask;
};

// ---- lane-0 vector primitives (bounds violations = undefined behaviour in the reference -> err) ----
__device__ THR_TS_INLINE void ts_km_insert(TState& T, int pos, uint64_t v) {
	if (T.nkm + 1 > T.cap || pos < 0 || pos > T.nkm) { T.err = 1; return; }
	for (int i = T.nkm; i > pos; --i) T.km[i] = T.km[i - 1];
	T.km[pos] = v; ++T.nkm;
}
__device__ THR_TS_INLINE void ts_km_erase(TState& T, int a, int b) {
	if (a < 0 || b > T.nkm || a > b) { T.err = 1; return; }
	for (int i = b; i < T.nkm; ++i) T.km[i - (b - a)] = T.km[i];
	T.nkm -= (b - a);
}
__device__ THR_TS_INLINE void ts_tr_insert(TState& T, int pos, uint8_t v) {
	if (T.ntr + 1 > T.cap || pos < 0 || pos > T.ntr) { T.err = 1; return; }
	for (int i = T.ntr; i > pos; --i) T.tr[i] = T.tr[i - 1];
	T.tr[pos] = v; ++T.ntr;
}
__device__ THR_TS_INLINE void ts_tr_erase(TState& T, int a, int b) {
	if (a < 0 || b > T.ntr || a > b) { T.err = 1; return; }
	for (int i = b; i < T.ntr; ++i) T.tr[i - (b - a)] = T.tr[i];
	T.ntr -= (b - a);
}
__device__ THR_TS_INLINE void ts_es_insert(TState& T, int pos, uint8_t t, uint8_t r, uint8_t g) {
	if (T.nes + 1 > T.ecap || pos < 0 || pos > T.nes) { T.err = 1; return; }
	for (int i = T.nes; i > pos; --i) { T.es_t[i] = T.es_t[i - 1]; T.es_r[i] = T.es_r[i - 1]; T.es_g[i] = T.es_g[i - 1]; }
	T.es_t[pos] = t; T.es_r[pos] = r; T.es_g[pos] = g; ++T.nes;
}
__device__ THR_TS_INLINE void ts_es_erase(TState& T, int pos) {
	if (pos < 0 || pos >= T.nes) { T.err = 1; return; }
	for (int i = pos + 1; i < T.nes; ++i) { T.es_t[i - 1] = T.es_t[i]; T.es_r[i - 1] = T.es_r[i]; T.es_g[i - 1] = T.es_g[i]; }
	--T.nes;
}
__device__ __forceinline__ bool ts_es_ok(TState& T, int i) { if (i < 0 || i >= T.nes) { T.err = 1; return false; } return true; }
__device__ __forceinline__ uint8_t ts_es_type(TState& T, int i) { return ts_es_ok(T, i) ? T.es_t[i] : (uint8_t)ET_EQ; }
__device__ __forceinline__ uint64_t ts_km(TState& T, int i) { if (i < 0 || i >= T.nkm) { T.err = 1; return KMER_INVALID; } return T.km[i]; }

// thread_ext_t::edit_kmers_forward, cpp:828-862 (lane 0).  The reference snapshots good[]/nts[] of the tail before
// editing (cpp:829-832); element `dt_nti` of those snapshots is exactly the not-yet-rewritten k-mer that sits at
// position i when it is consumed (cpp:842-845), so the tail is read in place.
__device__ THR_FWD_INLINE void edit_kmers_forward(TState& T, Txt& x, int& ki_io, uint64_t& ncorrection) {
	const int ksize = (int)T.k;
	int ki = ki_io;
	x.nm = x.nd = x.nins = 0;
	for (uint32_t e = 0; e < x.nedits; ++e) {
		const uint8_t t = x.et[e];
		if (t == ET_X)      {                             if (ki < 1 || ki >= T.nkm) { T.err = 1; return; } T.km[ki] = ((T.km[ki-1] & T.rmask) << 2) + (uint64_t)(x.eg[e] - 1); ++ki; ++x.nm; }
		else if (t == ET_D) { ts_km_insert(T, ki, 0);     if (T.err || ki < 1) { T.err = 1; return; }       T.km[ki] = ((T.km[ki-1] & T.rmask) << 2) + (uint64_t)(x.eg[e] - 1); ++ki; ++x.nd; }
		else if (t == ET_I) { ts_km_erase(T, ki, ki + 1); if (T.err) return;                                                                                               ++x.nins; }
	}
	const int dt_ki = (int)(x.nm + x.nd), dt_km = (int)x.nd - (int)x.nins;
	for (int i = ki; i < min(T.nkm, ki + ksize); ++i) {
		if (T.km[i] == KMER_INVALID) break;
		T.km[i] = ((T.km[i-1] & T.rmask) << 2) + (T.km[i] & 3);
	}
	if (dt_km > 0) { for (int i = 0; i < dt_km; ++i) ts_tr_insert(T, T.ntr, 0); }          // cg.tr.resize(size + dt_km, '*')
	else if (dt_km < 0) { ts_tr_erase(T, T.ntr + dt_km, T.ntr); }
	for (uint32_t i = 0; i < x.nd; ++i) ts_es_insert(T, T.ni + ksize - 1 + (int)x.nm, ET_D, 0, EG_STAR);
	x.a_lo = ki - dt_ki; x.a_n = dt_ki + (int)x.score;
	if (x.a_lo < 0 || x.a_lo + x.a_n > T.ntr || x.a_lo + x.a_n > T.nkm) { T.err = 1; return; }
	for (uint32_t i = 0; i < x.nedits; ++i, ++T.ni) {
		const int q = T.ni + ksize - 1;
		if (!ts_es_ok(T, q)) return;
		T.es_t[q] = x.et[i]; T.es_g[q] = x.eg[i];
	}
	for (uint32_t i = 0; i < x.score; ++i, ++T.ni) {
		const int q = T.ni + ksize - 1;
		if (!ts_es_ok(T, q)) return;
		T.es_t[q] = ET_EQ;
	}
	--T.ni;
	ki += (int)x.score - 1;
	ncorrection += x.nedits;
	ki_io = ki;
}

// thread_ext_t::edit_kmers_backward, cpp:649-826 (lane 0).  good[]/nts[] (cpp:651-656) are read in place for the
// same reason as above: position i-1 is still the original prefix k-mer when cpp:683-686 rewrites it.
__device__ THR_BACK_INLINE void edit_kmers_backward(TState& T, Txt& x, int& ki_io, uint64_t& ncorrection, uint64_t& nskip) {
	const int ksize = (int)T.k;
	int ki = ki_io;
	const uint64_t lmask = 3ull << 2 * (ksize - 1), lbase = 1ull << 2 * (ksize - 1);
	x.nm = x.nd = x.nins = 0; x.nrk = 0;
	for (uint32_t e = 0; e < x.nedits; ++e) { const uint8_t t = x.et[e]; if (t == ET_X) ++x.nm; else if (t == ET_D) ++x.nd; else if (t == ET_I) ++x.nins; }
	const int dt_km = (int)x.nd - (int)x.nins;
	T.ni += (int)x.nd;
	if (dt_km > 0) { for (int i = 0; i < dt_km; ++i) { ts_km_insert(T, ki, 0); ts_tr_insert(T, ki, 0); } }
	else if (dt_km < 0) { ts_km_erase(T, ki + dt_km, ki); ts_tr_erase(T, ki + dt_km, ki); }
	if (T.err) return;
	ki += dt_km;
	int ki_ = ki;
	for (uint32_t e = 0; e < x.nedits; ++e) {                       // corrected k-mers, cpp:679-681
		if (x.et[e] == ET_X || x.et[e] == ET_D) {
			if (ki_ - 1 < 0 || ki_ >= T.nkm) { T.err = 1; return; }
			T.km[ki_-1] = (T.km[ki_] >> 2) + (uint64_t)(3 - (x.eg[e] - 1)) * lbase;   // baseComplement[baseNumConversion[e.g]]
			--ki_;
		}
	}
	for (int i = ki_; i > max(0, ki_ - ksize); --i) {               // extended k-mers, cpp:683-686
		if (T.km[i-1] == KMER_INVALID) break;
		T.km[i-1] = (T.km[i] >> 2) + (T.km[i-1] & lmask);
	}
	const int lb = ki - (int)x.nm - (int)x.nd - (int)x.score;
	if (lb < 0 || ki > T.ntr) { T.err = 1; return; }
	for (int i = ki - 1; i >= lb; --i) if (T.tr[i] == 0) ++x.nrk;   // tr values themselves come from the refreshed TR bits
	x.a_lo = lb; x.a_n = ki - lb;
	x.nrk -= (x.nm + x.nd);
	nskip -= x.nrk;
	ncorrection += x.nedits;
	{	// CIGAR, cpp:696-823
		int cni = 0;
		const int nti_ = ki - dt_km;
		for (int i = 0; i < nti_ + cni; ++i) if (ts_es_type(T, i) == ET_I) ++cni;
		int nti = nti_ + cni - 1;
		for (uint32_t i = 0; i < x.nedits; ++i, --nti) {
			const uint8_t t1 = x.et[i], g1 = x.eg[i];
			if (t1 == ET_D) { ++nti; ts_es_insert(T, nti, ET_D, 0, EG_STAR); }
			if (!ts_es_ok(T, nti)) return;
			if (T.es_t[nti] == ET_D) {
				if (t1 == ET_I) { ts_es_erase(T, nti); --T.ni; }
				else T.es_g[nti] = eg_comp(g1);
			} else {
				while (ts_es_type(T, nti) == ET_I && !T.err) --nti;
				if (!ts_es_ok(T, nti)) return;
				T.es_t[nti] = t1;
				T.es_g[nti] = g1 ? eg_comp(g1) : (uint8_t)0;
			}
		}
		int e0 = nti + 1, e1 = e0;
		for (uint32_t i = 0; i < x.score; ++i, --nti) {
			const uint8_t t = ts_es_type(T, nti);
			if (T.err) return;
			if (t == ET_EQ) { }
			else if (t == ET_STAR) T.es_t[nti] = ET_EQ;
			else break;
		}
		{	// find the edit tract
			uint8_t t = ts_es_type(T, e1);
			while (et_is_edit(t) && !T.err) { ++e1; t = ts_es_type(T, e1); }
			t = ts_es_type(T, e0 - 1);
			while (et_is_edit(t) && !T.err) { --e0; t = ts_es_type(T, e0 - 1); }
			if (T.err) return;
		}
		// merge edits if possible: rnts / gnts = the non-empty read / graph bases of the tract, in order
		int nets = e1 - e0, nr = 0, ng = 0;
		for (int i = e0; i < e1; ++i) { if (T.es_r[i]) ++nr; if (T.es_g[i]) ++ng; }
		auto nth = [&](const uint8_t* a, int n) -> uint8_t {         // n-th non-empty base of the tract (as it was BEFORE the rewrite below)
			for (int i = e0; i < e1; ++i) if (a[i]) { if (n == 0) return a[i]; --n; }
			return 0;
		};
		if (nr == ng) {
			bool no_edit = true;
			{ int i = e0, j = e0;
			  for (int q = 0; q < nr; ++q) { while (!T.es_r[i]) ++i; while (!T.es_g[j]) ++j; if (T.es_r[i] != T.es_g[j]) { no_edit = false; break; } ++i; ++j; } }
			if (no_edit) {                                            // edits cancelled out, cpp:755-770
				int dt_es = 0;
				for (int i = e0; i < e1; ++i) {
					const int q = i + dt_es;
					if (!ts_es_ok(T, q)) return;
					if (T.es_t[q] == ET_D) { ts_es_erase(T, q); --dt_es; }
					else { T.es_t[q] = ET_EQ; T.es_g[q] = 0; }
				}
				T.ni += dt_es;
				ncorrection -= (uint64_t)(e1 - e0);
				nskip -= (uint64_t)(e1 - e0);
			} else if (nets != nr) {                                  // D+I at one position -> X, cpp:772-798
				// snapshot the base lists: the loop below rewrites the tract while consuming them
				uint8_t rn[48], gn[48];
				if (nr > 48) { T.err = 1; return; }
				for (int q = 0; q < nr; ++q) { rn[q] = nth(T.es_r, q); gn[q] = nth(T.es_g, q); }
				int dt_es = 0;
				const int dt_es_ = nr - nets;
				int j = 0, kk = 0;
				for (int i = e0; i < e1; ++i) {
					const int q = i + dt_es;
					if (!ts_es_ok(T, q)) return;
					if (T.es_t[q] == ET_D && dt_es != dt_es_) { ts_es_erase(T, q); --dt_es; }
					else {
						if (kk >= nr || j >= ng) { T.err = 1; return; }
						if (rn[kk] == gn[kk]) { T.es_t[q] = ET_EQ; T.es_g[q] = 0; }
						else { T.es_t[q] = ET_X; T.es_g[q] = gn[j]; }
						++j; ++kk;
					}
				}
				if (dt_es != dt_es_) { T.err = 1; return; }               // assert, cpp:794
				T.ni += dt_es;
				ncorrection += (uint64_t)(int64_t)dt_es;
				nskip += (uint64_t)(int64_t)dt_es;
			} else {                                                  // match / mismatch only, cpp:799-809: every entry has both bases
				for (int i = 0; i < nr; ++i) {
					const int q = e0 + i;
					if (T.es_r[q] == T.es_g[q]) { T.es_t[q] = ET_EQ; T.es_g[q] = 0; --ncorrection; --nskip; }
				}
			}
		} else {                                                      // cpp:812-822
			for (int i = 0; i < nets; ++i) {
				const int q = e0 + i;
				if (T.es_r[q] == T.es_g[q]) { T.es_t[q] = ET_EQ; T.es_g[q] = 0; --ncorrection; --nskip; }
			}
		}
	}
	ki_io = ki;
}

template <int B>
__device__ __noinline__ GNode g_lookup_nl(const Table gt, uint64_t fw, uint32_t locus, uint32_t k) { return g_lookup<B>(gt, fw, locus, k); }
#ifdef THR_INLINE_LOOKUP
#define EC_LOOKUP g_lookup
#else
#define EC_LOOKUP g_lookup_nl                      // one out-of-line copy for the three look-up sites of the edit scorer: 8.80 -> 8.61 ms
#endif

// ---- warp-parallel edit scorer: errorCorrection_forward, cpp:898-1089 ----
// W.win()[j+1] holds kmers[ki+j] for j = -1 .. k+1 (KMER_INVALID beyond the vector), nrem = nkmers - ki.
// Returns (warp-uniform) true when an edit was found; x.et/eg/score/nedits describe it.  err |= graph not closed.
template <int B, int MAXJ>
__device__ __noinline__ bool ec_score(const Table gt, uint32_t locus, uint32_t K, uint64_t rmask, ThreadWarp<MAXJ, B>& W,
                                      uint32_t nrem, uint32_t mes, int lane, Txt& x, int& err, uint32_t g0bits) {
	const uint64_t node = W.win()[0];
	uint32_t n_ec = lane == 0 ? 1u : 0u;
	// nts0 / nnds: getOutNodes(g, node), cpp:1170 / 1097 -- the node is a k-mer of the read (or its reverse complement)
	// that thr_refresh has already probed: g0bits = its GNode bits in the walk's orientation (W.info)
	if (!(g0bits & GN_PRESENT)) err = 1;
	const uint32_t nts0 = gn_mask(g0bits);
	// level 1 (lanes 16-19) and level 2 (lanes 0-15) of the expansion, cpp:907-920, probed in ONE round: the level-2 node
	// ((n1_i & rmask) << 2) + j is known without the level-1 answer, only whether the reference would visit it (edge j of
	// n1_i set) is not -- so it is probed speculatively and counted / checked / used only if that edge exists
	uint32_t myb = 0; bool pres = false, probe = false; uint64_t nd = 0;
	{
		const uint32_t i = lane < 16 ? (uint32_t)(lane >> 2) : (uint32_t)(lane - 16);
		if (lane < 20 && ((nts0 >> i) & 1u)) {
			nd = ((node & rmask) << 2) + (uint64_t)i;
			if (lane < 16) nd = ((nd & rmask) << 2) + (uint64_t)(lane & 3);
			probe = true;
		}
	}
	if (probe) { const GNode g = EC_LOOKUP<B>(gt, nd, locus, K); pres = (g.bits & GN_PRESENT) != 0; myb = gn_mask(g.bits); }
	const uint32_t m1i = __shfl_sync(0xffffffffu, myb, 16 + ((lane >> 2) & 3));     // out-edges of n1_i, i = (lane >> 2) & 3
	const uint32_t nts1 = __reduce_or_sync(0xffffffffu, (lane >= 16 && lane < 20) ? myb : 0u);
	const bool real2 = lane < 16 && probe && ((m1i >> (lane & 3)) & 1u);
	const uint32_t m2 = real2 ? myb : 0u;
	if ((lane >= 16 && probe) || real2) ++n_ec;
	const bool bad = ((lane >= 16 && probe) || real2) && !pres;
	if (__any_sync(0xffffffffu, bad)) err = 1;
	const uint32_t nts2 = __reduce_or_sync(0xffffffffu, m2);
	const uint32_t mlo = __reduce_or_sync(0xffffffffu, lane < 8 ? (m2 << (4 * lane)) : 0u);
	const uint32_t mhi = __reduce_or_sync(0xffffffffu, (lane >= 8 && lane < 16) ? (m2 << (4 * (lane - 8))) : 0u);
	const uint64_t mat = ((uint64_t)mhi << 32) | mlo;                 // bit nt0*16 + nt1*4 + nt2, graph_triplet_t cpp:865-875
	auto nnts2 = [&](uint32_t i, uint32_t j) -> uint32_t { return (uint32_t)(mat >> (i * 16 + j * 4)) & 0xFu; };
	auto nnts1 = [&](uint32_t i) -> uint32_t { uint32_t r = 0; for (uint32_t j = 0; j < 4; ++j) if (nnts2(i, j)) r |= 1u << j; return r; };

	const uint32_t gl = nrem < K + 2 ? nrem : K + 2;                  // good[j], cpp:922-923
	auto good = [&](uint32_t j) -> bool { return j < gl && W.win()[j + 1] != KMER_INVALID; };
	const uint64_t w1 = W.win()[1], w2 = W.win()[2], w3 = W.win()[3];
	const uint32_t oldnt = (uint32_t)(w1 & 3);
	const bool g0k = good(0), g1k = good(1), g2k = good(2);
	// family pre-conditions; the reference reads kmers[ki+j] % 4 before good[j], device code guards first
	const bool c_m1 = g1k && ((nts1 >> (w2 & 3)) & 1u);
	const bool c_m2 = !c_m1 && g2k && ((nts2 >> (w3 & 3)) & 1u) && mes >= 2;       // else-if, cpp:944
	const bool c_mi = g2k && ((nts1 >> (w3 & 3)) & 1u) && mes >= 2;
	const bool c_md = g1k && ((nts2 >> (w2 & 3)) & 1u) && mes >= 2;
	const bool c_i1 = g1k && ((nts0 >> (w2 & 3)) & 1u);
	const bool c_d1 = g0k && ((nts1 >> (w1 & 3)) & 1u);
	const bool c_i2 = g2k && ((nts0 >> (w3 & 3)) & 1u) && mes >= 2;
	const bool c_d2 = g0k && ((nts2 >> (w1 & 3)) & 1u) && mes >= 2;
	const uint32_t e1 = nrem < K + 1 ? nrem : K + 1, e2 = nrem < K + 2 ? nrem : K + 2, e0 = nrem < K ? nrem : K;

	// parallel-walk set-up (k <= 29: the k+3 window entries fit one ballot): validity of win[q] and the last base of
	// win[q] packed at bits 2*(31-q) of PB, q = 1..k+2
	if (K + 3 > 32) { err = 1; x.score = 0; x.nedits = 0; return false; }   // dbt_create admits k <= 27
	const uint64_t kmask = K < 32 ? (1ull << 2 * K) - 1 : ~0ull;
	uint32_t vmask = 0; uint64_t PB = 0;
	{
		const bool inq = (uint32_t)lane < K + 3 && (uint32_t)lane < gl + 1;       // good(j): j = q-1 < gl
		const uint64_t wq = inq ? W.win()[lane] : KMER_INVALID;
		const bool vq = wq != KMER_INVALID;
		vmask = __ballot_sync(0xffffffffu, vq);
		const uint64_t pb = (vq && lane >= 1) ? ((wq & 3) << (2 * (31 - lane))) : 0ull;
		PB = ((uint64_t)__reduce_or_sync(0xffffffffu, (uint32_t)(pb >> 32)) << 32) | __reduce_or_sync(0xffffffffu, (uint32_t)pb);
	}
	const uint32_t ms1 = (uint32_t)THR_MSC, ms2 = 2 * (uint32_t)THR_MSC;
	uint32_t mykey = 0;
	// score slots in get_edit's scan order families: [0,4) nem1, [4,8) ned1, 8 nei1, [9,25) nem2, [25,41) nemd, [41,57) ned2, [57,61) nemi, 61 nei2
#pragma unroll 1
	for (int round = 0; round < 2; ++round) {
		const uint32_t w = (uint32_t)lane + 32u * round;
		bool act = false; uint64_t cr = 0; uint32_t nn = 0, jb = 0, je = 0;
		if (w < 4)        { const uint32_t a = w;      act = c_m1 && ((nts0 >> a) & 1u); cr = w1 - oldnt + a; nn = nnts1(a); jb = 1; je = e1; }
		else if (w < 8)   { const uint32_t a = w - 4;  act = c_d1 && ((nts0 >> a) & 1u); cr = w1 - oldnt + a; nn = nnts1(a); jb = 0; je = e0; }
		else if (w == 8)  { act = c_i1; cr = node; nn = nts0; jb = 1; je = e1; }
		else if (w < 57)  {
			const uint32_t f = (w - 9) >> 4, q = (w - 9) & 15, a = q >> 2, b = q & 3;
			const bool c = f == 0 ? c_m2 : (f == 1 ? c_md : c_d2);
			act = c && ((nts0 >> a) & 1u) && ((nnts1(a) >> b) & 1u);
			cr = (((w1 - oldnt + a) & rmask) << 2) + b; nn = nnts2(a, b);
			jb = f == 0 ? 2 : (f == 1 ? 1 : 0); je = f == 0 ? e2 : (f == 1 ? e1 : e0);
		}
		else if (w < 61)  { const uint32_t a = w - 57; act = c_mi && ((nts0 >> a) & 1u); cr = w1 - oldnt + a; nn = nnts1(a); jb = 2; je = e2; }
		else if (w == 61) { act = c_i2; cr = node; nn = nts0; jb = 2; je = e2; }
		uint32_t n = 0;
		{
			// One candidate at a time, its walk spread over the lanes: step t of the walk visits the k-mer made of the
			// candidate's start k-mer followed by t+1 read bases, which does not depend on the earlier steps, so all
			// steps are probed at once (one round of latency instead of up to k+2) and the walk length is read off a
			// ballot.  The chain "mask of the previous node" is exact while every visited node exists; otherwise the
			// reference keeps the stale mask (cpp:553) and the scan below replays that literally.
			uint32_t amask = __ballot_sync(0xffffffffu, act);
			while (amask) {
				const int src = __ffs(amask) - 1; amask &= amask - 1;
				const uint64_t cr0 = __shfl_sync(0xffffffffu, cr, src);
				const uint32_t nn0 = __shfl_sync(0xffffffffu, nn, src), jb0 = __shfl_sync(0xffffffffu, jb, src), je0 = __shfl_sync(0xffffffffu, je, src);
				const uint32_t T = je0 > jb0 ? je0 - jb0 : 0u, t = (uint32_t)lane;
				const uint32_t need = (2u << t) - 1u;                                   // t+1 low bits
				const bool in = t < T && ((~(vmask >> (jb0 + 1)) & need) == 0u);       // win[jb0+1 .. jb0+t+1] all valid (good(j), cpp:922)
				uint32_t msk = 0; bool pres = false; uint32_t b = 0;
				if (in) {
					const uint32_t nb = t + 1;                                            // bases appended so far
					const uint64_t low = (PB >> (2 * (31 - (jb0 + nb)))) & ((1ull << (2 * nb)) - 1);
					const uint64_t c = ((cr0 << (2 * nb)) | low) & kmask;
					b = (uint32_t)(c & 3);
					const GNode g = EC_LOOKUP<B>(gt, c, locus, K);
					pres = (g.bits & GN_PRESENT) != 0; msk = gn_mask(g.bits);
				}
				const uint32_t pm = __shfl_up_sync(0xffffffffu, msk, 1);
				const uint32_t nnt = t == 0 ? nn0 : pm;
				const uint32_t okb = __ballot_sync(0xffffffffu, in && ((nnt >> b) & 1u));
				const uint32_t prb = __ballot_sync(0xffffffffu, pres);
				uint32_t nw = (uint32_t)__ffs(~okb) - 1u;                               // leading ok steps (T <= 29 < 32: ~okb != 0)
				if ((prb & ((1u << nw) - 1u)) != ((1u << nw) - 1u)) {                   // a visited node is missing: literal scan
					uint32_t cur = nn0; nw = 0;
					for (uint32_t q = 0; q < T; ++q) {
						const bool inq = (__shfl_sync(0xffffffffu, (uint32_t)in, q)) != 0;
						const uint32_t bq = __shfl_sync(0xffffffffu, b, q), mq = __shfl_sync(0xffffffffu, msk, q);
						const bool pq = (__shfl_sync(0xffffffffu, (uint32_t)pres, q)) != 0;
						if (!inq || !((cur >> bq) & 1u)) break;
						++nw; if (pq) cur = mq;
					}
				}
				if (lane == src) n = nw;
				if (lane == 0) n_ec += nw;
			}
		}
		// thread_ext_t::get_edit, cpp:628-647: a sequential scan with strict '>' = the FIRST maximum in scan order among the
		// entries that reach their threshold.  key = score << 16 | (255 - scan rank) << 8 | slot, reduced with max.
		if (w < 62 && n >= (w < 9 ? ms1 : ms2) && (w < 9 || mes > 1)) {
			uint32_t rank = w;
			if (w >= 9 && w < 57)  { const uint32_t f = (w - 9) >> 4, q = (w - 9) & 15; rank = 9 + (q >> 2) * 13 + (q & 3) * 3 + f; }
			else if (w >= 57 && w < 61) rank = 9 + (w - 57) * 13 + 12;
			const uint32_t key = (n << 16) | ((255u - rank) << 8) | w;
			if (key > mykey) mykey = key;
		}
	}
	const uint32_t bkey = __reduce_max_sync(0xffffffffu, mykey);
	const uint32_t best = bkey >> 16, bw = bkey ? (bkey & 0xFFu) : 0xFFu;
	__syncwarp();
	n_ec = __reduce_add_sync(0xffffffffu, n_ec);
	if (lane == 0) W.cnt[1] += n_ec;
	x.score = best; x.nedits = 0;
	if (best == 0) return false;
	if (bw < 4)        { x.nedits = 1; x.et[0] = ET_X; x.eg[0] = (uint8_t)(bw + 1); }
	else if (bw < 8)   { x.nedits = 1; x.et[0] = ET_D; x.eg[0] = (uint8_t)(bw - 4 + 1); }
	else if (bw == 8)  { x.nedits = 1; x.et[0] = ET_I; x.eg[0] = 0; }
	else if (bw < 57)  {
		const uint32_t f = (bw - 9) >> 4, q = (bw - 9) & 15;
		x.nedits = 2;
		x.et[0] = f == 2 ? ET_D : ET_X; x.eg[0] = (uint8_t)((q >> 2) + 1);
		x.et[1] = f == 0 ? ET_X : ET_D; x.eg[1] = (uint8_t)((q & 3) + 1);
	}
	else if (bw < 61)  { x.nedits = 2; x.et[0] = ET_X; x.eg[0] = (uint8_t)(bw - 57 + 1); x.et[1] = ET_I; x.eg[1] = 0; }
	else               { x.nedits = 2; x.et[0] = ET_I; x.eg[0] = 0; x.et[1] = ET_I; x.eg[1] = 0; }
	return true;
}

// ---- per-position probes + classification (the parallel replacement of the node-by-node walk) ----
// Two ROLLED loops (the kernel is instruction-fetch sensitive: hot code must stay inside the 32 KB L1.5 I-cache,
// profiles/r1 s5-s7): loop A computes the home bucket of every position and starts its 32-byte copy into shared
// memory with cp.async -- all <= MAXJ*32 probes of the read are in flight together -- loop B matches the staged
// slots and classifies.  The class of position i needs the node of position i-1: shuffled in from the neighbour
// lane, lane 0 takes the carry of the previous iteration.
template <int B, int MAXJ>
__device__ THR_REFRESH_INLINE void thr_refresh(const Table gt, uint32_t locus, uint32_t K, uint64_t rmask, ThreadWarp<MAXJ, B>& W,
                                            int m, int nkm, int lo, int hi, int lane) {
	// [lo, hi) = positions whose k-mer changed since the last refresh (0, nkm: all).  Only they are probed; the ballots are
	// rebuilt for the 32-position groups that contain lo .. hi (position hi's match bit depends on node hi-1), from W.info
	// for the untouched positions.
	__syncwarp();
	const uint64_t lmix = (uint64_t)locus * 0xC2B2AE3D27D4EB4Full;         // g_hash_key
	constexpr int SG = ThreadWarp<MAXJ, B>::SG;
	const int j0 = lo >> 5, j1 = min(hi, nkm - 1) >> 5;                    // group range, inclusive
	uint64_t carry_f = KMER_INVALID; uint32_t carry_b = 0, nprobe = 0;
	if (j0 > 0) { carry_f = W.km[j0 * 32 - 1]; carry_b = W.info[j0 * 32 - 1]; }
#pragma unroll 1
	for (int jc = j0; jc <= j1; jc += SG) {                                // SG groups per round (all of them when the stage holds MAXJ)
	const int je = min(j1, jc + SG - 1);
#ifndef THR_DIRECT_PROBE
#pragma unroll 1
	for (int j = jc; j <= je; ++j) {
		const int i = j * 32 + lane;
		const uint64_t f = (i >= lo && i < hi && i < nkm) ? W.km[i] : KMER_INVALID;
		if (f != KMER_INVALID) {
			const uint64_t rc = W.rck[i];
			const uint64_t canon = rc < f ? rc : f;
			const uint4* bp = gt.slots + home_bucket(gt, canon ^ lmix) * B;
#pragma unroll
			for (int q = 0; q < B; ++q) cp_async16(&W.stage[j - jc][q][lane], bp + q);
		}
	}
	cp_async_wait_all();
	__syncwarp();
#endif
#pragma unroll 1
	for (int j = jc; j <= je; ++j) {
		const int i = j * 32 + lane;
		const bool in = i < nkm;
		const uint64_t f = in ? W.km[i] : KMER_INVALID;
		const bool valid = f != KMER_INVALID;
		uint32_t bi = 0, trp = 0;
		if (valid && (i < lo || i >= hi)) { bi = W.info[i]; trp = W.trx[m][i]; }
		else if (valid) {
			const uint64_t rc = W.rck[i];
			const bool is_rc = rc < f;
			const uint64_t canon = is_rc ? rc : f;
			bool full = true, found = false;
			uint64_t k0 = 0;
#ifdef THR_DIRECT_PROBE
			const uint4* bkt = gt.slots + home_bucket(gt, canon ^ lmix) * B;
			uint4 slv[B];
#pragma unroll
			for (int q = 0; q < B; ++q) slv[q] = ldg128(bkt + q);
#endif
#pragma unroll
			for (int q = 0; q < B; ++q) {
#ifdef THR_DIRECT_PROBE
				const uint4 sl = slv[q];
#else
				const uint4 sl = W.stage[j - jc][q][lane];
#endif
				const uint64_t kk = slot_key(sl);
				if (q == 0) k0 = kk;
				if (kk == SLOT_EMPTY) full = false;
				else if ((kk & SLOT_KEYMASK) == canon && (sl.z & G_LOCUS_MASK) == locus) { const GNode n = g_decode(sl.z, sl.w, is_rc); bi = n.bits; trp = n.tr; found = true; }
			}
#ifdef THR_REFRESH_INLINE_STASH
			if (!found && full && (k0 & SLOT_OVF)) { const GNode n = g_lookup<B>(gt, f, locus, K); bi = n.bits; trp = n.tr; }   // continues in the stash: rare
#else
			// continues in the stash: rare, and out of line -- inlined, the look-up made one iteration of this loop 5.3 KB of SASS, most of the 6 KB L0
			if (!found && full && (k0 & SLOT_OVF)) { const GNode n = g_lookup_nl<B>(gt, f, locus, K); bi = n.bits; trp = n.tr; }
#endif
			++nprobe;
		}
		if (in) { W.trx[m][i] = trp; W.info[i] = bi; }
		uint64_t pv = __shfl_up_sync(0xffffffffu, f, 1); uint32_t bp = __shfl_up_sync(0xffffffffu, bi, 1);
		if (lane == 0) { pv = carry_f; bp = carry_b; }
		carry_f = __shfl_sync(0xffffffffu, f, 31); carry_b = __shfl_sync(0xffffffffu, bi, 31);
		const bool isN = in && !valid;
		const bool isH = valid && i > 0 && f == pv;                      // cpp:1152 (the N test comes first, cpp:1145)
		const bool isA = valid && i > 0 && pv == KMER_INVALID;           // cpp:1159
		const bool isM = valid && i > 0 && pv != KMER_INVALID && f != pv && (((pv & rmask) << 2) + (f & 3)) == f
		                 && (bp & GN_PRESENT) && ((gn_mask(bp) >> (f & 3)) & 1u);   // cpp:1170-1178 with node == kmers[ki-1]
		const uint32_t bm = __ballot_sync(0xffffffffu, isM), bn = __ballot_sync(0xffffffffu, isN), bh = __ballot_sync(0xffffffffu, isH);
		const uint32_t ba = __ballot_sync(0xffffffffu, isA), bpz = __ballot_sync(0xffffffffu, valid && (bi & GN_PRESENT));
		const uint32_t bt = __ballot_sync(0xffffffffu, valid && (bi & GN_TR));
		if (lane == 0) { W.cls[CL_M][j] = bm; W.cls[CL_N][j] = bn; W.cls[CL_H][j] = bh; W.cls[CL_A][j] = ba; W.cls[CL_P][j] = bpz; W.cls[CL_T][j] = bt; }
	}
	__syncwarp();                                                          // the stage is rewritten by the next round
	}
	if (lo == 0 && hi >= nkm) {                                            // full refresh: groups past the vector read as empty
#pragma unroll 1
		for (int j = j1 + 1 + lane; j < MAXJ; j += 32) { W.cls[CL_M][j] = 0; W.cls[CL_N][j] = 0; W.cls[CL_H][j] = 0; W.cls[CL_A][j] = 0; W.cls[CL_P][j] = 0; W.cls[CL_T][j] = 0; }
	}
	nprobe = __reduce_add_sync(0xffffffffu, nprobe);
	if (lane == 0) W.cnt[2] += nprobe;
	__syncwarp();
}

// first position >= from whose bit in mask words w[] equals `want`, or n
template <int MAXJ>
__device__ __forceinline__ int thr_next(const uint32_t* w, int from, int n, bool want) {
	if (from >= n) return n;
	int j = from >> 5;
	uint32_t v = (want ? w[j] : ~w[j]) & (0xFFFFFFFFu << (from & 31));
	while (true) {
		if (v) { const int p = j * 32 + __ffs(v) - 1; return p < n ? p : n; }
		if (++j >= MAXJ || j * 32 >= n) return n;
		v = want ? w[j] : ~w[j];
	}
}
template <int MAXJ>
__device__ __forceinline__ bool thr_bit(const uint32_t* w, int i) { return (w[i >> 5] >> (i & 31)) & 1u; }

// find_anchor, cpp:878-888
template <int B, int MAXJ>
__device__ __noinline__ bool ms_find_anchor(ThreadWarp<MAXJ, B>& W, int K, int lane) {
	__syncwarp();
	MS& S = W.S;
	const int ki = S.ki, nkm = S.nkm, ni0 = S.ni, nes = S.nes;
	const uint64_t nskip = S.nskip;
	const int a = thr_next<MAXJ>(W.cls[CL_P], ki, nkm, true);
	const int ni = ni0 + (a - ki);
	__syncwarp();
	S.nskip = nskip + (uint64_t)(a - ki); S.ni = ni; S.ki = a;
	if (a >= nkm) return false;
	if (lane == 0) W.tr[a] = thr_bit<MAXJ>(W.cls[CL_T], a) ? 2 : 1;
	if (ni < 0 || ni + K > nes) S.err = 1;
	else {
#pragma unroll 1
		for (int q = lane; q < K; q += 32) if (W.es_t[ni + q] == ET_STAR) W.es_t[ni + q] = ET_EQ;
	}
	__syncwarp();
	return true;
}

// errorCorrection_forward (backward = false) / errorCorrection_backward (cpp:1091-1106: the same scorer on kmers_rc)
template <int B, int MAXJ>
__device__ __noinline__ bool ms_score(const Table gt, uint32_t locus, uint32_t K, uint64_t rmask, ThreadWarp<MAXJ, B>& W, int m,
                                      bool backward, int kk, uint32_t mes, int lane) {
	__syncwarp();
	MS& S = W.S;
	const int nkm = S.nkm;
	uint32_t nrem;
	if (!backward) {
		nrem = (uint32_t)(nkm - kk);
#pragma unroll 1
		for (int q = lane; q < (int)K + 3; q += 32) { const int i = kk - 1 + q; W.win()[q] = (i >= 0 && i < nkm) ? W.km[i] : KMER_INVALID; }
	} else {                                                         // kmers_rc, cpp:1098-1103
		nrem = (uint32_t)kk;
#pragma unroll 1
		for (int q = lane; q < (int)K + 3; q += 32) {
			const int i = kk - q;
			uint64_t v = KMER_INVALID;
			if (i >= 0 && i < nkm && W.km[i] != KMER_INVALID) v = kmer_rc(W.km[i], K);
			W.win()[q] = v;
		}
	}
	__syncwarp();
	int err = S.err;
	// level-0 node of the expansion = kmers[kk-1] (forward) / reverse complement of kmers[kk] (backward): already probed
	uint32_t g0 = 0;
	if (!backward) { if (kk >= 1 && kk - 1 < nkm && W.km[kk - 1] != KMER_INVALID) g0 = W.info[kk - 1]; }
	else if (kk >= 0 && kk < nkm && W.km[kk] != KMER_INVALID) { const uint32_t b = W.info[kk]; g0 = ((b & GN_RCPRESENT) ? GN_PRESENT : 0u) | (gn_rcmask(b) << 1); }
	const bool ok = ec_score<B, MAXJ>(gt, locus, K, rmask, W, nrem, mes, lane, S.x, err, g0);
	__syncwarp();
	S.err = err;
	return ok;
}

// edit_kmers_forward / _backward on lane 0, then re-probe the rewritten positions and fill the annotation range.
// Returns the new k-mer index (ki / ki_ of the caller).
template <int B, int MAXJ>
__device__ __noinline__ int ms_apply(const Table gt, uint32_t locus, uint32_t K, uint64_t rmask, ThreadWarp<MAXJ, B>& W, int m,
                                     bool backward, int kk0, int lane) {
	constexpr int CAP = ThreadWarp<MAXJ, B>::CAP, ECAP = ThreadWarp<MAXJ, B>::ECAP;
	__syncwarp();
	MS& S = W.S;
	if (lane == 0) {
		int kk = kk0;
		uint64_t ncorr = S.ncorr, nskip = S.nskip;
		TState T; T.km = W.km; T.es_t = W.es_t; T.es_r = W.es_r; T.es_g = W.es_g; T.tr = W.tr;
		T.nkm = S.nkm; T.nes = S.nes; T.ntr = S.ntr; T.ni = S.ni; T.err = S.err; T.cap = CAP; T.ecap = ECAP; T.k = K; T.rmask = rmask;
		if (backward) edit_kmers_backward(T, S.x, kk, ncorr, nskip); else edit_kmers_forward(T, S.x, kk, ncorr);
		S.nkm = T.nkm; S.nes = T.nes; S.ntr = T.ntr; S.ni = T.ni; S.err = T.err; S.ncorr = ncorr; S.nskip = nskip;
		W.kk = kk;
	}
	__syncwarp();
	const int nkm = S.nkm, err = S.err, kk = W.kk;
	const int xnm = (int)S.x.nm, xnd = (int)S.x.nd, xnins = (int)S.x.nins, a_lo = S.x.a_lo, a_n = S.x.a_n;
	// substitutions only (no k-mer inserted or erased): the rewritten k-mers are the <= nm + k positions after (forward,
	// cpp:842-845) or before (backward, cpp:679-686) the edit point; anything else moved the tail -> everything is re-probed
	int lo = 0, hi = nkm;
	if (xnd == 0 && xnins == 0 && !err) {
		if (!backward) { lo = kk0; hi = min(nkm, kk0 + xnm + (int)K); }
		else           { lo = max(0, kk0 - xnm - (int)K); hi = kk0; }
		if (lo >= hi) { lo = 0; hi = nkm; }
	}
#pragma unroll 1
	for (int i = lo + lane; i < hi; i += 32) { const uint64_t f = W.km[i]; W.rck[i] = f == KMER_INVALID ? KMER_INVALID : kmer_rc(f, K); }
	__syncwarp();
	thr_refresh<B, MAXJ>(gt, locus, K, rmask, W, m, nkm, lo, hi, lane);
	if (!err) {                                                        // cpp:850 / 690
#pragma unroll 1
		for (int q = lane; q < a_n; q += 32) W.tr[a_lo + q] = thr_bit<MAXJ>(W.cls[CL_T], a_lo + q) ? 2 : 1;
	}
	__syncwarp();
	return kk;
}

// isThreadFeasible, cpp:1114-1260, for mate m of the pair held by warp state W.  Warp-uniform.
template <int B, int MAXJ>
__device__ __noinline__ void thread_mate(const Table gt, uint32_t locus, uint32_t K, uint32_t WB, uint32_t thread_cth, uint32_t maxnc32,
                                         bool correction, ThreadWarp<MAXJ, B>& W, int m,
                                         const uint64_t* __restrict__ rec, uint32_t len, int lane) {
	constexpr int CAP = ThreadWarp<MAXJ, B>::CAP, ECAP = ThreadWarp<MAXJ, B>::ECAP;
	const uint64_t rmask = (1ull << 2 * (K - 1)) - 1;
	const uint64_t maxnc = maxnc32;
	__syncwarp();
	MateResult& R = W.res[m];                                          // W.res[m] and W.S: warp-uniform values, every lane stores the same
	MS& S = W.S;
	R.alned = 0; R.err = 0; R.clean = 0;
	S.nkm = (int)len - (int)K + 1; S.nes = (int)len; S.ntr = S.nkm; S.ni = 0; S.err = 0; S.ki = 0; S.nskip = 0; S.ncorr = 0;
	S.x.score = 0; S.x.nedits = 0; S.x.nm = S.x.nd = S.x.nins = 0; S.x.nrk = 0; S.x.a_lo = 0; S.x.a_n = 0;
	if (S.nkm < 1 || S.nkm > CAP || S.nes > ECAP) { R.err = 1; R.nkm = 0; R.nes = 0; R.ntr = 0; return; }
	// read2kmers(.., canonical=false, keepN=true) (h:246-271)
	__syncwarp();
	bool anyvalid = false;
#pragma unroll 1
	for (int i = lane; i < S.nkm; i += 32) {
		uint64_t f = KMER_INVALID, rc = KMER_INVALID;
		if (extract_bits(rec + WB, (uint32_t)i, K) == 0) {
			const uint64_t w = extract_bits(rec, 2u * i, 2 * K);
			f = rev2(w) >> (64 - 2 * K); rc = (~w) & (K < 32 ? (1ull << 2 * K) - 1 : ~0ull); anyvalid = true;
		}
		W.km[i] = f; W.rck[i] = rc;
	}
	anyvalid = __any_sync(0xffffffffu, anyvalid);
	__syncwarp();
	const uint64_t nkm0 = (uint64_t)S.nkm;
	const uint64_t maxnskip = nkm0 >= thread_cth ? nkm0 - thread_cth : 0;
	if (anyvalid) {
		thr_refresh<B, MAXJ>(gt, locus, K, rmask, W, m, S.nkm, 0, S.nkm, lane);
		if (lane == 0) W.cnt[0] += (uint32_t)S.nkm;                       // DBT_C_THR2_WALK (the walk itself was counted by thread_fast_kernel)
		// Error-free mate (three quarters of them): the first k-mer is a node and every later k-mer matches an out-edge of its
		// predecessor.  The reference then anchors at 0 (cpp:1129, find_anchor marks es[0..k) '='), walks the whole read through
		// cpp:1167-1180 (tr[i] = '=' / '.', es[i+k-1] = '=') and returns 1 (cpp:1259: no skip, no correction): every CIGAR entry
		// is '=' and tr follows the TR ballots -- the kernel writes the record from those, es / tr are not materialised.
		const int nkm = S.nkm;
		bool ok = true;
		if (lane < MAXJ && lane * 32 < nkm) {
			const int n = min(32, nkm - lane * 32);
			const uint32_t want = n == 32 ? 0xFFFFFFFFu : ((1u << n) - 1u);
			const uint32_t mm = W.cls[CL_M][lane] | (lane == 0 ? (W.cls[CL_P][0] & 1u) : 0u);
			ok = mm == want;
		}
		if (__all_sync(0xffffffffu, ok)) {
			R.alned = 1; R.err = 0; R.nkm = nkm; R.nes = S.nes; R.ntr = S.ntr; R.clean = 1;
			__syncwarp();
			return;
		}
	}
	// cigar_t::init (cpp:62-67)
#pragma unroll 1
	for (int i = lane; i < S.nes; i += 32) {
		const uint32_t code = (uint32_t)(rec[i >> 5] >> (2 * (i & 31))) & 3u;
		const bool inv = (rec[WB + (i >> 6)] >> (i & 63)) & 1ull;
		W.es_t[i] = ET_STAR; W.es_r[i] = inv ? ER_INVALID : (uint8_t)(code + 1); W.es_g[i] = 0;
		if (i < S.nkm) W.tr[i] = 0;
	}
	__syncwarp();
	if (!anyvalid) { R.nkm = 0; R.nes = S.nes; R.ntr = S.ntr; return; }     // D1: a read without a valid k-mer is not threadable (cpp:1124 would be UB)
	int result = -1;                                                   // -1: fall through to the final test (cpp:1259)
#define FIND_ANCHOR() ms_find_anchor<B, MAXJ>(W, (int)K, lane)
#define SCORE_AT(bw, kk, mes) ms_score<B, MAXJ>(gt, locus, K, rmask, W, m, bw, kk, mes, lane)
#define APPLY(bw, kkvar) { const int kk_new = ms_apply<B, MAXJ>(gt, locus, K, rmask, W, m, bw, kkvar, lane); kkvar = kk_new; }

	if (!FIND_ANCHOR()) result = 0;                                     // cpp:1129
	else if (S.ki > 0 && correction && S.ncorr < maxnc && (uint64_t)S.ki >= THR_MSC + 1) {   // leading unaligned k-mers, cpp:1131-1141
		const uint32_t mes = ((uint64_t)S.ki >= 2 * THR_MSC + 2) ? 2 : 1;
		if (SCORE_AT(true, S.ki, mes)) APPLY(true, S.ki);
	}
	if (result < 0) { ++S.ki; ++S.ni; }
	// what to do at the end of an iteration: 0 = step (++ki, ++ni), 1 = find_anchor then the skip test and step, 2 = stop
	while (result < 0 && S.ki < S.nkm && !S.err) {                      // cpp:1144
		int tail = 0;
		{	// run of matching k-mers: cpp:1167-1180 for every position at once
			const int ki = S.ki;
			const int e = thr_next<MAXJ>(W.cls[CL_M], ki, S.nkm, false);
			if (e > ki) {
				const int d = S.ni - ki + (int)K - 1;
				if (d < 0 || e - 1 + d >= S.nes || e > S.ntr) { S.err = 1; break; }
#pragma unroll 1
				for (int i = ki + lane; i < e; i += 32) { W.tr[i] = thr_bit<MAXJ>(W.cls[CL_T], i) ? 2 : 1; W.es_t[i + d] = ET_EQ; }
				__syncwarp();
				S.ni += e - ki; S.ki = e;
			}
			if (S.ki >= S.nkm) break;
		}
		const int ki = S.ki;
		if (thr_bit<MAXJ>(W.cls[CL_N], ki) || thr_bit<MAXJ>(W.cls[CL_H], ki)) {    // "N" in read / homopolymer run, cpp:1145-1158
			int e = ki;
			while (e < S.nkm && (thr_bit<MAXJ>(W.cls[CL_N], e) || thr_bit<MAXJ>(W.cls[CL_H], e))) ++e;
			const uint64_t run = (uint64_t)(e - ki);
			const uint64_t take = S.nskip > maxnskip ? 1 : (run < maxnskip - S.nskip + 1 ? run : maxnskip - S.nskip + 1);
			const int d = S.ni - ki + (int)K - 1;
			if (d < 0 || ki + (int)take - 1 + d >= S.nes || ki + (int)take > S.ntr) { S.err = 1; break; }
#pragma unroll 1
			for (int i = ki + lane; i < ki + (int)take; i += 32) { W.tr[i] = 0; W.es_t[i + d] = ET_STAR; }
			__syncwarp();
			S.nskip += take;
			if (S.nskip > maxnskip) { result = 0; break; }
			S.ki += (int)take; S.ni += (int)take;
			continue;
		}
		if (thr_bit<MAXJ>(W.cls[CL_A], ki)) tail = 1;                       // after an N, cpp:1159-1165
		else if ((uint64_t)ki + THR_MSC >= nkm0) {                          // no matching node, not enough info: cpp:1182-1185
			S.nskip += (nkm0 - (uint64_t)ki);
			result = S.nskip <= maxnskip ? (S.ncorr ? 2 : 1) : 0;
			break;
		}
		else if (!(correction && S.ncorr < maxnc)) tail = 1;                // cpp:1250-1256
		else {                                                             // error correction, cpp:1186-1248
			uint32_t mes = ((uint64_t)(S.nkm - ki) >= 2 * THR_MSC + 2) ? 2 : 1;
			if (SCORE_AT(false, ki, mes)) {                                   // cpp:1192-1197
				S.nskip += S.x.nedits;
				if (S.nskip > maxnskip) { result = 0; break; }
				APPLY(false, S.ki);
			} else {                                                          // cpp:1198-1248
				if (!FIND_ANCHOR()) break;
				bool skip = !SCORE_AT(true, S.ki, 2);
				if (!skip) {
					APPLY(true, S.ki);
					++S.ncorr;
					int64_t gap = (int64_t)min((int64_t)K, (int64_t)S.ki - (int64_t)S.x.nm - (int64_t)S.x.nd) - (int64_t)S.x.score;
					int64_t ki1 = S.ki;
					if (gap < 0) S.err = 1;
					while (!skip && gap && !S.err) {                              // forward and backward threads not fully patched
						const int64_t ki0 = ki1;
						ki1 = ki0 - (int64_t)S.x.nm - (int64_t)S.x.nd - (int64_t)S.x.score;
						if (ki1 < 0) { S.err = 1; break; }
						mes = ((uint64_t)ki1 >= 2 * THR_MSC + 2) ? 2 : 1;
						if ((uint64_t)ki1 < THR_MSC + 1) break;
						if (ki1 >= S.nkm || !thr_bit<MAXJ>(W.cls[CL_P], (int)ki1)) { S.err = 1; break; }   // assert(g.count(node_)), cpp:1222
						skip = !SCORE_AT(true, (int)ki1, mes);
						if (!skip) {
							int k1 = (int)ki1;
							APPLY(true, k1);
							ki1 = k1;
							S.ki += (int)S.x.nd - (int)S.x.nins;
							gap = (int64_t)min((int64_t)K, ki1 - (int64_t)S.x.nm - (int64_t)S.x.nd) - (int64_t)S.x.score;
							if (gap < 0) S.err = 1;
						}
					}
					if (gap > 0 && !S.err) {                                        // annot_gap, cpp:1108-1111
						if (ki1 - gap < 0 || ki1 > S.ntr) S.err = 1;
						else { for (int q = lane; q < (int)gap; q += 32) W.tr[ki1 - 1 - q] = 0; __syncwarp(); }
						S.nskip -= (uint64_t)gap;
					}
					if (S.nskip > maxnskip) { result = 0; break; }
				}
				if (skip) tail = 1;                                               // cpp:1234-1243
			}
		}
		if (tail == 1) {
			if (!FIND_ANCHOR()) break;
			if (S.nskip > maxnskip) { result = 0; break; }
		}
		++S.ki; ++S.ni;
	}
#undef FIND_ANCHOR
#undef SCORE_AT
#undef APPLY
	if (result < 0) result = (S.nskip <= maxnskip && S.ncorr <= maxnc) ? (S.ncorr ? 2 : 1) : 0;   // cpp:1259
	if (S.err) result = 0;
	R.alned = result; R.err = S.err; R.nkm = S.nkm; R.nes = S.nes; R.ntr = S.ntr;
	__syncwarp();
}

constexpr uint32_t THR_DST_MASK = (1u << G_LOCUS_BITS) - 1;           // sv.thr[t].z = dst | kf bits << 28 | rm bits << 30

template <int B, int MAXJ>
__global__ void __launch_bounds__(THR_WARPS * 32, THR_MINBLOCKS) thread_kernel(DevIndex ix, RunParams rp, SlotView sv) {
	extern __shared__ __align__(16) uint8_t thr_smem[];
	ThreadWarp<MAXJ, B>& W = reinterpret_cast<ThreadWarp<MAXJ, B>*>(thr_smem)[threadIdx.x >> 5];
	const int lane = threadIdx.x & 31;
	const uint32_t gwarp = blockIdx.x * THR_WARPS + (threadIdx.x >> 5), nwarps = gridDim.x * THR_WARPS;
	const uint32_t nwork = sv.ncand[4];
	const bool trace = (rp.flags & DBT_F_TRACE) != 0;
	const bool aln_all = (rp.flags & DBT_F_ALN) && !(rp.flags & DBT_F_ALN_MIN), aln_min = (rp.flags & DBT_F_ALN_MIN) != 0;
	const bool want_rec = trace || aln_all || aln_min;
	const uint32_t W64 = rp.lay.W64;
	uint32_t c_fea = 0, c_inc = 0, c_cls = 0, c_mates = 0;
	if (lane < 3) W.cnt[lane] = 0;                                     // walk, ec, refresh probes (measurement)
	__syncwarp();

	uint4 item = gwarp < nwork ? sv.thr2[gwarp] : make_uint4(0, 0, 0, 0);
#pragma unroll 1
	for (uint32_t t = gwarp; t < nwork; t += nwarps) {
#ifdef THR_BLOCKSYNC
		__syncthreads();                                                   // experiment: the warps of a block start every work item together (instruction-cache locality)
#endif
		const uint32_t c = item.x, p = item.y, dw = item.z, mw = item.w;
		if (t + nwarps < nwork) item = sv.thr2[t + nwarps];             // next work item: in flight while this one is threaded
		const uint32_t dst = dw & THR_DST_MASK;
		const bool pre = (mw & THR2_ALNED) != 0;                           // the other mate threaded cleanly: the pair counts whatever happens here
		const uint64_t slot = trace ? (uint64_t)p : (uint64_t)c;
		const bool slot_ok = want_rec && (trace || slot < rp.max_records);
		uint8_t* rec = sv.records + slot * rp.rec_stride;
		MateResult* const r = W.res;
		uint32_t nvalid = 0;                                               // valid k-mers of the edited vectors threaded here (lane 0)
#pragma unroll 1
		for (int m = 0; m < 2; ++m) {
			if (!((mw >> m) & 1u)) { r[m].alned = 0; r[m].err = 0; r[m].nkm = 0; r[m].nes = 0; r[m].ntr = 0; r[m].clean = 0; __syncwarp(); continue; }
			if (lane == 0) ++c_mates;
			const uint32_t len = (m ? sv.meta[1] : sv.meta[0])[p] & META_LENMASK;       // (no dynamic index into the parameter struct)
			thread_mate<B, MAXJ>(ix.gtable, dst, rp.k, rp.lay.WB, rp.thread_cth, rp.max_ncorrection, rp.correction != 0, W, m, (m ? sv.packed[1] : sv.packed[0]) + (uint64_t)p * W64, len, lane);
			if (slot_ok) {                                              // sam_t.r1 / r2 -> dbt_aln_mate blob
				uint8_t* b = rec + sizeof(dbt_pair_record) + (size_t)m * rp.aln_stride;
				uint8_t* es = b + sizeof(dbt_aln_mate); uint8_t* trb = es + rp.es_cap;
				const int nes = min(r[m].nes, (int)rp.es_cap), ntr = min(r[m].ntr, (int)rp.tr_cap);
				if (r[m].nes > (int)rp.es_cap || r[m].ntr > (int)rp.tr_cap) r[m].err = 1;
#pragma unroll 1
				// four entries per lane and store (caps are multiples of 8, blobs 8-byte aligned); entries >= nes / ntr read 0
				for (int q = lane; q < (int)rp.es_cap / 4; q += 32) {
					const int nb = nes - 4 * q;
					const uint32_t keep = nb >= 4 ? 0xFFFFFFFFu : (nb <= 0 ? 0u : (1u << 8 * nb) - 1u);
					const uint32_t v = r[m].clean ? 0x01010101u * ET_EQ
					                              : (reinterpret_cast<const uint32_t*>(W.es_t)[q] | ((reinterpret_cast<const uint32_t*>(W.es_g)[q] & 0x07070707u) << 3));
					reinterpret_cast<uint32_t*>(es)[q] = v & keep;
				}
#pragma unroll 1
				for (int q = lane; q < (int)rp.tr_cap / 4; q += 32) {
					const int nb = ntr - 4 * q;
					const uint32_t keep = nb >= 4 ? 0xFFFFFFFFu : (nb <= 0 ? 0u : (1u << 8 * nb) - 1u);
					uint32_t v;
					if (r[m].clean) {                                       // tr[i] = '=' (2) for TR k-mers, '.' (1) otherwise: 4 TR ballot bits -> 4 bytes
						const uint32_t bits = 4 * q < MAXJ * 32 ? (W.cls[CL_T][(4 * q) >> 5] >> ((4 * q) & 31)) & 0xFu : 0u;
						v = 0x01010101u + ((bits & 1u) | ((bits & 2u) << 7) | ((bits & 4u) << 14) | ((bits & 8u) << 21));
					} else v = reinterpret_cast<const uint32_t*>(W.tr)[q];
					reinterpret_cast<uint32_t*>(trb)[q] = v & keep;
				}
				if (lane == 0) {
					dbt_aln_mate* h = reinterpret_cast<dbt_aln_mate*>(b);
					h->alned = (uint8_t)r[m].alned; h->err = (uint8_t)r[m].err;
					h->nes = (uint16_t)r[m].nes; h->ntr = (uint16_t)r[m].ntr; h->nkm = (uint16_t)r[m].nkm;
					uint64_t hash = 0;
					if (trace) {                                            // FNV-1a over the edited k-mer vector (parity tests only)
						hash = 0xcbf29ce484222325ull;
#pragma unroll 1
						for (int i = 0; i < r[m].nkm; ++i) { const uint64_t v = W.km[i]; for (int q = 0; q < 8; ++q) { hash ^= (v >> (8 * q)) & 0xFF; hash *= 0x100000001b3ull; } }
					}
					h->kmers_hash = hash;
				}
			}
			if (r[m].nkm > 0) {                                          // the N ballots of the last refresh cover exactly [0, nkm)
				uint32_t nn = 0;
#pragma unroll 1
				for (int j = lane; j < MAXJ; j += 32) nn += __popc(W.cls[CL_N][j]);
				nvalid += (uint32_t)r[m].nkm - __reduce_add_sync(0xffffffffu, nn);
			}
			__syncwarp();
		}
		const bool alned = pre || r[0].alned || r[1].alned;                // commented caller, cpp:2082-2087
		if (alned) {
			if (lane == 0 && !pre) c_fea += 2;                                // (a pair with a clean mate was counted as feasible by thread_fast_kernel)
			uint32_t inc_sum = 0;
#pragma unroll 1
			for (int m = 0; m < 2; ++m) {                                     // D1: every TR k-mer of the edited vectors, cpp:2203-2208
				if (!((mw >> m) & 1u)) continue;
#pragma unroll 1
				for (int j = 0; j * 32 < r[m].nkm; ++j) {
					const int i = j * 32 + lane;
					const uint32_t tx = i < r[m].nkm ? W.trx[m][i] : 0u;
					const uint32_t tm = __ballot_sync(0xffffffffu, tx != 0);
					if (!tm) continue;
					inc_sum += __popc(tm);
					if (tx) {
						const uint32_t peers = __match_any_sync(tm, tx);
						if ((uint32_t)lane == (uint32_t)(__ffs(peers) - 1))
							atomicAdd(reinterpret_cast<unsigned long long*>(ix.counts + (tx - 1)), (unsigned long long)__popc(peers));
					}
				}
			}
			if (lane == 0) { c_inc += inc_sum; c_cls += nvalid; }
		}
		if (pre) continue;                                                 // header, status and the overflow flag were thread_fast_kernel's
		const uint32_t status = alned ? DBT_ST_COUNTED : DBT_ST_THREAD_FAILED;
		const uint32_t fdst = alned ? dst : (uint32_t)ix.nloci;
		if (want_rec && !slot_ok) { if (lane == 0 && (aln_all || alned)) atomicOr(sv.ncand + 1, ERR_RECORD_OVERFLOW); }
		else if (slot_ok && lane == 0) {
			dbt_pair_record* h = reinterpret_cast<dbt_pair_record*>(rec);
			if (!trace) {                                                     // under DBT_F_TRACE the generic candidate kernel left the full header
				h->pair_index = sv.first_pair_index + p; h->dst0 = dst; h->fc = 0; h->rc = 0; h->nhash0 = 0; h->nhash1 = 0;
				h->pad[0] = h->pad[1] = h->pad[2] = 0;
#ifdef DBT_DEBUG_DIRTY
				h->pad[0] = (uint8_t)(0xC0u | (mw & 3u));
#endif
#pragma unroll 1
				for (int m = 0; m < 2; ++m) {
					dbt_mate_record& o = h->mate[m];
					o.si = o.ei = o.si_ = o.ei_ = o.ti = -1; o.nt = 0; o.bs = 0; o.pad = 0; o.nk = 0;
					o.flags = (uint8_t)((((dw >> (28 + m)) & 1u) ? DBT_MF_KF : 0) | (((dw >> (30 + m)) & 1u) ? DBT_MF_RM : 0));
				}
			}
			h->dst = fdst; h->status = (uint8_t)status;
		}
	}
	__syncwarp();
	if (lane == 0) {
		if (W.cnt[0]) atomicAdd(sv.counters + DBT_C_THR2_WALK, (unsigned long long)W.cnt[0]);
		if (W.cnt[1]) atomicAdd(sv.counters + DBT_C_THR_EC, (unsigned long long)W.cnt[1]);
		if (W.cnt[2]) { atomicAdd(sv.counters + DBT_C_THR_REFRESH, (unsigned long long)W.cnt[2]); atomicAdd(sv.counters + DBT_C_THR2_REFRESH, (unsigned long long)W.cnt[2]); }
		if (c_cls) { atomicAdd(sv.counters + DBT_C_THR_CLS, (unsigned long long)c_cls); atomicAdd(sv.counters + DBT_C_THR2_CLS, (unsigned long long)c_cls); }
		if (c_fea) atomicAdd(sv.counters + DBT_C_NFEASIBLEREADS, (unsigned long long)c_fea);
		if (c_inc) { atomicAdd(sv.counters + DBT_C_NINC, (unsigned long long)c_inc); atomicAdd(sv.counters + DBT_C_THR2_NINC, (unsigned long long)c_inc); }
		if (c_mates) atomicAdd(sv.counters + DBT_C_THR2_MATES, (unsigned long long)c_mates);
	}
}

// ---- graph table build (load time) ----
// pass 0: nodes in canonical orientation claim their entry; pass 1: reverse-complement orientation merges into it (or
// claims one when the partner is absent from the file); pass 2: TR k-mers attach their count-array position.
template <int B>
__global__ void graph_insert_kernel(Table t, const uint64_t* __restrict__ kmers, const uint8_t* __restrict__ masks,
                                    const uint32_t* __restrict__ locus_of, uint64_t n, uint32_t k, int pass, uint32_t* err) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint64_t fw = kmers[i]; const uint32_t loc = locus_of[i];
	if (k < 32 && (fw >> (2 * k))) { atomicOr(err, ERR_INDEX_MISSING); return; }
	const uint64_t rc = kmer_rc(fw, k);
	if (pass == 2) {
		if (rc < fw) return;                                            // not canonical: unreachable through toCaKmer (h:180-183)
		uint4* s = g_find_slot<B>(t, fw, loc);
		if (s) { uint32_t* w = reinterpret_cast<uint32_t*>(s); if (w[3]) atomicOr(err, ERR_INDEX_DUPTR); w[3] = (uint32_t)i + 1; }
		else if (!g_claim<B>(t, fw, loc, loc, (uint32_t)i + 1)) atomicOr(err, ERR_INDEX_FULL);
		return;
	}
	const uint32_t mk = masks[i] & 0xFu;
	const bool is_rc = rc < fw;
	if (pass == 0) {
		if (is_rc) return;
		uint32_t lw = loc | (mk << G_MASK_CA_SHIFT) | G_PRESENT_CA;
		if (rc == fw) lw |= (mk << G_MASK_RC_SHIFT) | G_PRESENT_RC;                    // palindrome (even k): both orientations are this node
		if (!g_claim<B>(t, fw, loc, lw, 0)) atomicOr(err, ERR_INDEX_FULL);
	} else {
		if (!is_rc) return;
		uint4* s = g_find_slot<B>(t, rc, loc);
		const uint32_t add = (mk << G_MASK_RC_SHIFT) | G_PRESENT_RC;
		if (s) atomicOr(reinterpret_cast<uint32_t*>(s) + 2, add);
		else if (!g_claim<B>(t, rc, loc, loc | add, 0)) atomicOr(err, ERR_INDEX_FULL);
	}
}

// verification: every node must read back with its own mask (catches duplicate nodes within a locus)
template <int B>
__global__ void graph_verify_kernel(Table t, const uint64_t* __restrict__ kmers, const uint8_t* __restrict__ masks,
                                    const uint32_t* __restrict__ locus_of, uint64_t n, uint32_t k, uint32_t* err) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const GNode g = g_lookup<B>(t, kmers[i], locus_of[i], k);
	if (!(g.bits & GN_PRESENT) || gn_mask(g.bits) != (masks[i] & 0xFu)) atomicOr(err, ERR_INDEX_DUPKEY);
}

}  // namespace dbt
