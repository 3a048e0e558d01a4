// candidate_fast_kernel: one warp per read pair that survived subfilter -- the common case, with no
// sort and no shared-memory work arrays.
//
// The expensive part of the reference's locus vote (countDupRemove's sort by key, fillstats' unstable
// std::sort, the sequential early-stopping walk; cpp:257-422) only decides
//   (1) top.idx and (2) whether (top.fc, top.rc) at the stop pass the thresholds of cpp:439-440;
// nm1/nm2 themselves are never used again (cpp:2016, 2056).  When EVERY index hit of the surviving
// mates is a k-mer unique to one and the same locus A (all values even and equal), the walk is
// degenerate: `second` stays {NAN32,0,0}, top = running sums, so top.idx = A, and with
//   F, R = number of hit positions of mate 1 / mate 2 (>= cth for a mate that survived kfilter, cpp:190-224)
// the exit state of the loops at cpp:383-417 is: counting continues while a threshold is still
// reachable, i.e. until fc >= cth and rc >= cth (both mates alive: always reached -> accept), or,
// with one mate removed (its side stays 0), until fc >= cth and remain < cth, where fc = F - remain:
//   F <  2*cth       -> fc + 0 < 2*cth           -> reject, whatever the visiting order
//   F >= 3*cth - 1   -> fc >= F - (cth-1) >= 2*cth -> accept, whatever the visiting order
//   otherwise the outcome depends on the multiplicities met last -> generic kernel (exact replay).
// The same holds with multi-locus k-mers present, as long as A is in the locus list of EVERY hit and the
// hit positions unique to A (S1) are at least as many as those on shared k-mers (S2 >= the largest single
// multiplicity): fillstats visits the unique k-mers first (ascending #loci, cpp:320), so when a shared
// k-mer raises hits[B] to at most S2_before + d <= S1 + S2_before = hits[A], updatetop2's strict `>`
// (cpp:332) never moves top away from A, and every later entry counts for top.idx (cpp:405-416).
// Pairs with unique hits in two loci, a hit outside A, S1 < S2, or in that undecided band are appended to
// sv.slow and left untouched for candidate_kernel.  Counts are accumulated per k-mer OCCURRENCE exactly as
// cpp:2155-2158 does (one increment per position with as == 2), aggregated per warp with
// __match_any_sync before the 64-bit atomic.
#pragma once
#include "candidate.cuh"
#include "thread_fast.cuh"      // thrf_finish_mate / thrf_finish_pair: threading from the index entries when the graphs are merged into them

namespace dbt {

#ifndef CANDF_WARPS_N
#define CANDF_WARPS_N 8
#endif
#ifndef CANDF_MINBLOCKS
#define CANDF_MINBLOCKS 3
#endif
constexpr int CAND_FAST_WARPS = CANDF_WARPS_N;

// Per-warp shared-memory state.  Everything indexed by mate or by 32-position group lives here rather
// than in registers so that the stages are ROLLED loops (2 mates x MAXJ groups): ncu showed the fully
// unrolled kernel (~4.4k SASS instructions = 70 KB, each executed once per pair) to be instruction-fetch
// bound (stall_no_instruction 5.0 per issue, issue slots 29 % busy).  Index buckets are staged through
// shared memory with cp.async so that the probe loop can be rolled and still keep MAXJ x B 16-byte loads
// per lane in flight.
// class word of a k-mer w.r.t. locus `loc` (classify2) from the RAW value word, plus its graph entry there
__device__ __forceinline__ uint32_t classify2g(const DevIndex& ix, uint32_t z, uint32_t aux, uint32_t loc, uint32_t& g) {
	g = 0;
	if (aux == AUX_MISS) return 0;
	if (!(z & 1u)) { if (((z & VAL_LOCUS_MASK) >> 1) != loc) return 0; g = z >> VAL_LOCUS_BITS; return aux; }
	const uint32_t off = z >> 1, n = __ldg(ix.vv + off);
	for (uint32_t j = 1; j <= n; ++j) if (__ldg(ix.vv + off + j) == loc) { if (ix.vv_g) g = __ldg(ix.vv_g + off + j); return __ldg(ix.vv_aux + off + j); }
	return 0;
}

template <int B, int MAXJ>
struct FastWarp {
	uint4    slot[MAXJ][B][32];            // bucket staging: slot q of the bucket probed for position j*32+lane
	uint64_t pk[2 * 16];                   // packed records of both mates (W64 <= 16 when nk_max <= 255)
	uint32_t val[2][MAXJ * 32];            // index value per position              (during the probe: k-mer, low half)
	uint32_t aux[2][MAXJ * 32];            // aux word per position, AUX_MISS = not in the index (during the probe: k-mer, high half);
	                                       // after classification: class word w.r.t. dst0
	uint16_t g[2][MAXJ * 32];              // graph entry of the k-mer in dst0 (ten bits, canonical orientation first; graphs merged into the index only)
	uint32_t hit[2][MAXJ], F[2][MAXJ], T[2][MAXJ];   // ballots: in the index / flank of dst0 / TR of dst0
	dbt_mate_record mate[2];
};


// COOP: the bucket staging of the probe round -- false: every lane fetches the B slots of its own bucket (tables in L2: fewest
// instructions), true: B adjacent lanes fetch one bucket as one coalesced request (tables in HBM: table.cuh, stage_buckets_coop).
// A template parameter, not a run-time switch: the uniform branch alone cost 4 % of this issue-bound kernel (profiles/r2/k7).
template <int B, int MAXJ, bool COOP>
__global__ void __launch_bounds__(CAND_FAST_WARPS * 32, CANDF_MINBLOCKS) candidate_fast_kernel(DevIndex ix, RunParams rp, SlotView sv) {
	extern __shared__ __align__(16) uint8_t fast_smem[];
	FastWarp<B, MAXJ>* wsm = reinterpret_cast<FastWarp<B, MAXJ>*>(fast_smem);
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t gwarp = blockIdx.x * CAND_FAST_WARPS + wib, nwarps = gridDim.x * CAND_FAST_WARPS;
	const uint32_t W64 = rp.lay.W64, WB = rp.lay.WB, K = rp.k;
	const bool okam = (rp.flags & (DBT_F_KAM | DBT_F_EXTRACT | DBT_F_BUBBLES)) != 0 && !rp.thr_on;   // kam lines are only written without threading (cpp:2110); `-e` records ride the same slots
	const bool wstat = okam || (rp.thr_on && (rp.flags & (DBT_F_ALN | DBT_F_ALN_MIN)));   // record slots are keyed by candidate ordinal: keep their status current
	const uint32_t nloci = (uint32_t)ix.nloci, cth = rp.cth;
	FastWarp<B, MAXJ>& W = wsm[wib];
	uint32_t c_kf = 0, c_laf = 0, c_qc = 0, c_thr = 0, c_asgn = 0, c_h1 = 0, c_inc = 0, c_thr0 = 0, c_bait = 0;   // lane-0 accumulators
	const bool fused = rp.thr_on && ix.gmerged && !(rp.flags & DBT_F_TRACE);   // thread the pairs here, from the index entries (graphs merged into the index)
	const bool want_aln = (rp.flags & (DBT_F_ALN | DBT_F_ALN_MIN)) != 0;
	uint32_t t_fea = 0, t_walk = 0, t_items = 0;
	ThrfCount TC;
	const uint32_t nwork = sv.ncand[0];

#pragma unroll 1
	for (uint32_t c = gwarp; c < nwork; c += nwarps) {
		const uint32_t p = sv.cand[c];

		// ---- packed records -> shared memory ----
		__syncwarp();
		for (uint32_t i = lane; i < 2 * W64; i += 32) {
			const uint32_t m = i >= W64;
			W.pk[i] = sv.packed[m][(uint64_t)p * W64 + (i - m * W64)];
		}
		const uint32_t nk0 = (sv.meta[0][p] & META_LENMASK) - K + 1, nk1 = (sv.meta[1][p] & META_LENMASK) - K + 1;
		__syncwarp();

		// ---- all canonical k-mers + ONE index probe each (A2 + the look-ups of kfilter, cpp:207/218) ----
		uint32_t nhit0 = 0, nhit1 = 0;                                     // index hits per mate (warp-uniform)
#pragma unroll 1
		for (int m = 0; m < 2; ++m) {
			const uint32_t nkm = m ? nk1 : nk0;
			const uint64_t* rec = W.pk + m * W64;
			const uint32_t* mk32 = reinterpret_cast<const uint32_t*>(rec + WB);
			const uint32_t sh = 2 * lane;
			const uint64_t kmask = (1ull << (2 * K)) - 1, vmask = (1ull << K) - 1;
			// round 1: k-mer of position j*32+lane, home bucket -> shared memory (asynchronously)
#pragma unroll 1
			for (int j = 0; j < MAXJ; ++j) {
				const uint32_t i = j * 32 + lane;
				// k-mers at positions lane+32j: same bit offset (2*lane) in consecutive base words, same bit
				// offset (lane) in consecutive 32-bit halves of the invalid mask -> funnel shifts
				const uint64_t wlo = rec[j], whi = (j + 1 < (int)WB) ? rec[j + 1] : 0ull;
				const uint32_t mlo = mk32[j], mnx = ((uint32_t)(j + 1) < 2 * rp.lay.WM) ? mk32[j + 1] : 0xFFFFFFFFu;
				const uint64_t wv = sh ? ((wlo >> sh) | (whi << (64 - sh))) : wlo;
				const uint64_t mm = (((uint64_t)mnx << 32) | mlo) >> lane;
				uint64_t kk = KMER_INVALID;
				if (i < nkm && (mm & vmask) == 0) {
					const uint64_t x = wv & kmask, rcx = (~wv) & kmask, fw = rev2(x) >> (64 - 2 * K);
					kk = fw < rcx ? fw : rcx;
				}
				W.val[m][i] = (uint32_t)kk; W.aux[m][i] = (uint32_t)(kk >> 32);
if constexpr (COOP) {                                 // tables in HBM: one coalesced request per bucket (table.cuh)
					const bool valid = kk != KMER_INVALID;
					stage_buckets_coop<B>(ix.table.slots, valid, valid ? (uint32_t)home_bucket(ix.table, kk) : 0u, W.slot[j], lane);
				} else if (kk != KMER_INVALID) {
					const uint4* bp = ix.table.slots + home_bucket(ix.table, kk) * B;
#pragma unroll
					for (int q = 0; q < B; ++q) cp_async16(&W.slot[j][q][lane], bp + q);
				}
			}
			cp_async_wait_all();
			// inspect; a full bucket with the overflow flag sends the lane to the stash (round 2, again batched)
			uint32_t need = 0;                                    // bit j: position j*32+lane continues in the stash
#pragma unroll 1
			for (int j = 0; j < MAXJ; ++j) {
				const uint32_t i = j * 32 + lane;
				const uint64_t key = ((uint64_t)W.aux[m][i] << 32) | W.val[m][i];
				uint32_t v = 0, a = AUX_MISS;
				if (key != KMER_INVALID) {
					bool full = true, found = false; uint64_t k0 = 0;
#pragma unroll
					for (int q = 0; q < B; ++q) {
						const uint4 e = W.slot[j][q][lane];
						const uint64_t kk = slot_key(e);
						if (q == 0) k0 = kk;
						if (kk == SLOT_EMPTY) full = false;
						else if ((kk & SLOT_KEYMASK) == key) { v = e.z; a = e.w; found = true; }
					}
					if (!found && full && (k0 & SLOT_OVF)) need |= 1u << j;
				}
				if (!((need >> j) & 1u)) { W.val[m][i] = v; W.aux[m][i] = a; }
			}
			if (__any_sync(0xffffffffu, need != 0)) {
				// Round 2, compacted: the ~7 % of positions that continue in the stash are a few lanes in each of the MAXJ groups.
				// They are gathered into a list (the staging area is free now) and walked one entry per lane -- one pass at full
				// width instead of two rolled loops over all groups at 2-3 active lanes (6 % of the kernel's instructions before).
				uint16_t* list = reinterpret_cast<uint16_t*>(&W.slot[0][0][0]);
				uint32_t n = 0;
#pragma unroll 1
				for (int j = 0; j < MAXJ; ++j) {
					const bool mine = (need >> j) & 1u;
					const uint32_t bm = __ballot_sync(0xffffffffu, mine);
					if (mine) list[n + __popc(bm & ((1u << lane) - 1u))] = (uint16_t)(j * 32 + lane);
					n += __popc(bm);
				}
				__syncwarp();
#pragma unroll 1
				for (uint32_t t = lane; t < n; t += 32) {
					const uint32_t i = list[t];
					const uint64_t key = ((uint64_t)W.aux[m][i] << 32) | W.val[m][i];
					const uint2 r2 = probe_walk<B, true>(ix.table, key, stash_bucket(ix.table, key));
					W.val[m][i] = r2.x; W.aux[m][i] = r2.y;
				}
				__syncwarp();
			}
			uint32_t hsum = 0;
#pragma unroll 1
			for (int j = 0; j < MAXJ; ++j) {
				const uint32_t hm = __ballot_sync(0xffffffffu, W.aux[m][j * 32 + lane] != AUX_MISS);
				if (lane == 0) W.hit[m][j] = hm;
				hsum += __popc(hm);
			}
			if (m) nhit1 = hsum; else nhit0 = hsum;
		}
		__syncwarp();

		// ---- kfilter (cpp:190-224): abort a mate at its (nk-cth+1)-th miss ----
		uint32_t kfm = (nk0 < cth ? 1u : 0u) | (nk1 < cth ? 2u : 0u);      // bit m = kf[m]
		uint32_t rmm = kfm;                                                // bit m = rm[m]
		uint32_t nhash1 = 0;
		uint32_t cnt0 = 0, cnt1 = 0;                                       // hit positions per mate
		if (rmm != 3u) {
#pragma unroll 1
			for (int m = 0; m < 2; ++m) {
				if ((rmm >> m) & 1u) continue;
				const uint32_t nkm = m ? nk1 : nk0;
				const uint32_t max_ns = nkm - cth;
				// the common case first: too few misses (invalid k-mers included) in the whole mate to abort it -- nk probes, nothing to locate
				if (nkm - (m ? nhit1 : nhit0) <= max_ns) { nhash1 += nkm; if (m) cnt1 = nhit1; else cnt0 = nhit0; continue; }
				uint32_t cum = 0, probes = nkm, aword = 0, abase = 0, arank = 0, hits = 0; bool aborted = false;
#pragma unroll 1
				for (int j = 0; j < MAXJ; ++j) {
					const uint32_t inrange = (nkm >= (uint32_t)(j + 1) * 32) ? 0xFFFFFFFFu : ((nkm > (uint32_t)j * 32) ? ((1u << (nkm - j * 32)) - 1u) : 0u);
					const uint32_t hm = W.hit[m][j];
					const uint32_t mm = ~hm & inrange;
					hits += __popc(hm);
					if (!aborted && cum + __popc(mm) > max_ns) { aborted = true; aword = mm; abase = j * 32; arank = max_ns - cum; }
					cum += __popc(mm);
				}
				if (aborted) {                                   // position of the (arank+1)-th miss of that word
					uint32_t pos = 0, nn = arank;
#pragma unroll
					for (int shf = 16; shf >= 1; shf >>= 1) {
						const uint32_t cn = __popc((aword >> pos) & ((1u << shf) - 1u));
						if (nn >= cn) { nn -= cn; pos += shf; }
					}
					probes = abase + pos + 1;
				}
				nhash1 += probes;
				if (aborted) { kfm |= 1u << m; rmm |= 1u << m; }
				else if (m) cnt1 = hits; else cnt0 = hits;
			}
		}
		if (rmm == 3u) {                                         // cpp:2054
			if (lane == 0) {
				c_kf += __popc(kfm); c_h1 += nhash1;
				if (wstat && c < rp.max_records) reinterpret_cast<dbt_pair_record*>(sv.records + (uint64_t)c * rp.rec_stride)->status = (uint8_t)DBT_ST_KFILTERED;
			}
			continue;
		}

		// ---- degenerate vote: one DOMINANT locus A? ----
		// A = locus of the first hit that is unique to one locus (even value, kio:59-74)
		uint32_t val0 = 0; bool have0 = false;
#pragma unroll 1
		for (int m = 0; m < 2 && !have0; ++m) {
			if ((rmm >> m) & 1u) continue;
#pragma unroll 1
			for (int j = 0; j < MAXJ && !have0; ++j) {
				const uint32_t v = W.val[m][j * 32 + lane];
				const uint32_t em = W.hit[m][j] & __ballot_sync(0xffffffffu, !(v & 1u));
				if (em) { val0 = __shfl_sync(0xffffffffu, v, __ffs(em) - 1); have0 = true; }
			}
		}
		const uint32_t dst0 = (val0 & VAL_LOCUS_MASK) >> 1;             // (the value words are kept raw here: their top bits carry the graph entry)
		// class of every k-mer w.r.t. A (= what assignTRkmc looks up, cpp:1464-1469); a hit belongs to A iff its class != 0
		bool differs = false; uint32_t S2 = 0;                   // S2 = hit positions of surviving mates on multi-locus k-mers
		if (have0) {
#pragma unroll 1
			for (int m = 0; m < 2; ++m) {
				const bool rmv = (rmm >> m) & 1u;
				if (rmv && !okam && !fused) {                       // cpp:1451: a removed mate is only annotated for the kam line (threading needs both mates)
					if (lane == 0) {
#pragma unroll
						for (int j = 0; j < MAXJ; ++j) { W.F[m][j] = 0; W.T[m][j] = 0; }
					}
					continue;
				}
				const uint32_t nkm = m ? nk1 : nk0;
#pragma unroll
				for (int j = 0; j < MAXJ; ++j) {
					const uint32_t i = j * 32 + lane;
					const uint32_t v = W.val[m][i];
					uint32_t cw = 0, gg = 0;
					if (i < nkm) cw = classify2g(ix, v, W.aux[m][i], dst0, gg);
					W.aux[m][i] = cw;                               // from here on: class word w.r.t. A
					if (fused) W.g[m][i] = (uint16_t)gg;
					const uint32_t fm = __ballot_sync(0xffffffffu, (cw >> AUX_CLASS_SHIFT) == 1);
					const uint32_t tm = __ballot_sync(0xffffffffu, (cw >> AUX_CLASS_SHIFT) == 2);
					const uint32_t om = __ballot_sync(0xffffffffu, v & 1u);
					const uint32_t hm = W.hit[m][j];
					if (!rmv) { if (hm != (fm | tm)) differs = true; S2 += __popc(hm & om); }
					if (lane == 0) { W.F[m][j] = fm; W.T[m][j] = tm; }
				}
			}
		}
		// Certain outcomes (see the header): every hit of the surviving mates lies in A, and the k-mers unique to A
		// (visited first: fillstats sorts by #loci, cpp:320) outweigh any single shared k-mer, so top.idx never leaves A.
		int verdict = -1;                                        // 1 accept, 0 reject, -1 order-dependent
		if (have0 && !differs && rp.nk_max <= 255) {
			const uint32_t Tall = ((rmm & 1u) ? 0u : cnt0) + ((rmm & 2u) ? 0u : cnt1);
			if (Tall - S2 >= S2) {
				if (rmm == 0u || cth == 0) verdict = 1;
				else if (Tall < 2 * cth) verdict = 0;
				else if (Tall + 1 >= 3 * cth) verdict = 1;
			}
		}
		if (verdict < 0) {                                       // generic path (nothing of this pair has been counted yet)
			if (lane == 0) sv.slow[atomicAdd(sv.ncand + 2, 1u)] = c;
			continue;
		}
		if (lane == 0) { c_kf += __popc(kfm); c_h1 += nhash1; }

		// ---- countHit tail (cpp:436-452), QC gate (cpp:2059-2062) ----
		uint32_t dst = dst0, hfm = 0;
		int status = -1;
		if (!verdict) {
			hfm = ~rmm & 3u; rmm = 3u; dst = nloci; status = DBT_ST_UNASSIGNED;
			if (lane == 0) c_laf += __popc(hfm);
		} else if ((rp.flags & DBT_F_QC) && ix.qc && !ix.qc[dst]) {
			status = DBT_ST_QCFILTERED;
			if (lane == 0) c_qc += 2 - __popc(rmm);
		}

		if (status < 0 && (rp.flags & DBT_F_THREADING)) {
			if (rp.thr_on && fused) {
				// restored threading, from what this warp already holds: the index entry of every position carries the graph entry
				// of (k-mer, dst) and the count position of a TR k-mer, i.e. everything thread_fast_kernel would probe the graph table
				// for.  The staging area of the probes is free now and takes the per-mate arrays thrf_finish_mate works on.
				if (lane == 0) c_thr0 += 2;
				const uint32_t dw = dst | (kfm << 28) | (rmm << 30);
				const bool slot_ok = want_aln && c < rp.max_records;
				uint8_t* recp = sv.records + (uint64_t)c * rp.rec_stride;
				if (slot_ok && lane == 0) reinterpret_cast<dbt_pair_record*>(recp)->status = (uint8_t)DBT_ST_THREAD_PENDING;
				uint8_t* ov = reinterpret_cast<uint8_t*>(&W.slot[0][0][0]);
				ThrfArrays A;
				A.key = reinterpret_cast<uint64_t*>(ov); A.trx = reinterpret_cast<uint32_t*>(ov + MAXJ * 32 * 8);
				A.nb = reinterpret_cast<uint16_t*>(ov + MAXJ * 32 * 12); A.T = reinterpret_cast<uint32_t*>(ov + MAXJ * 32 * 12 + (MAXJ * 32 + 32) * 2);
				A.ok = A.T + MAXJ; A.P = A.ok + MAXJ;
				uint32_t dirty = 0;
#pragma unroll 1
				for (int m = 0; m < 2; ++m) {
					const uint32_t nkm = m ? nk1 : nk0;
					const uint32_t len = nkm + K - 1;
					if ((int)nkm < 1 || nkm > (uint32_t)MAXJ * 32) { dirty |= 1u << m; continue; }
					const uint64_t* rec = W.pk + m * W64;
					const uint32_t* mk32 = reinterpret_cast<const uint32_t*>(rec + WB);
					const uint32_t sh = 2 * lane;
					const uint64_t kmask = (1ull << (2 * K)) - 1, vmask = (1ull << K) - 1;
					const int ng = ((int)nkm + 31) >> 5;
					bool plain = true, anyvalid = false, b2clean = true;
					uint32_t carry_nb = 0;                              // node bits of position j*32-1 (lane 31 of the previous group)
					__syncwarp();
#pragma unroll 1
					for (int j = 0; j < ng; ++j) {
						const uint32_t i = j * 32 + lane;
						uint32_t f3 = 0;
						const uint64_t wlo = rec[j], whi = (j + 1 < (int)WB) ? rec[j + 1] : 0ull;
						const uint32_t mlo = mk32[j], mnx = ((uint32_t)(j + 1) < 2 * rp.lay.WM) ? mk32[j + 1] : 0xFFFFFFFFu;
						const uint64_t wv = sh ? ((wlo >> sh) | (whi << (64 - sh))) : wlo;
						const uint64_t mm = (((uint64_t)mnx << 32) | mlo) >> lane;
						uint64_t kk = KMER_INVALID; uint32_t nbv = 0, trp = 0;
						if (i < nkm) {
							if ((mm & vmask) == 0) {
								const uint64_t x = wv & kmask, rcx = (~wv) & kmask, fw = rev2(x) >> (64 - 2 * K);
								const bool is_rc = rcx < fw;
								kk = (is_rc ? rcx : fw) | ((fw & 3) << THRF_F3_SHIFT) | (is_rc ? THRF_RCFLAG : 0ull);
								f3 = (uint32_t)fw & 3u;
								anyvalid = true;
								if (i + 1 < nkm && ((wv >> 2) & kmask) == x) plain = false;
								const uint32_t gg = W.g[m][i], cw = W.aux[m][i];
								nbv = is_rc ? (((gg & 31u) << 5) | (gg >> 5)) : gg;
								trp = (cw >> AUX_CLASS_SHIFT) == 2 ? (cw & AUX_IDX_MASK) + 1u
								    : (((cw >> AUX_CLASS_SHIFT) == 1 && (cw & AUX_FLANK_TR)) ? (cw & (AUX_FLANK_TR - 1)) + 1u : 0u);   // a TR k-mer classed flank still counts here
							} else plain = false;
						}
						A.key[i] = kk; A.nb[i + 1] = (uint16_t)nbv; A.trx[i] = trp;
						// the B2 pass of thrf_finish_mate, from the registers: every k-mer on an out-edge of its predecessor (the first one a node)?
						{
							const uint32_t up = __shfl_up_sync(0xffffffffu, nbv, 1);
							const uint32_t pb = lane ? up : carry_nb;
							carry_nb = __shfl_sync(0xffffffffu, nbv, 31);
							const bool ok = i >= nkm || (i == 0 ? (nbv & 1u) != 0 : ((pb & 1u) && ((pb >> (1 + f3)) & 1u)));
							const uint32_t bo = __ballot_sync(0xffffffffu, ok);
							const uint32_t bt = __ballot_sync(0xffffffffu, i < nkm && trp != 0);
							const uint32_t bp = __ballot_sync(0xffffffffu, i < nkm && (nbv & 1u));
							if (bo != 0xffffffffu) b2clean = false;
							if (lane == 0) { A.T[j] = bt; A.ok[j] = bo; A.P[j] = bp; }
						}
					}
					__syncwarp();
					anyvalid = __any_sync(0xffffffffu, anyvalid);
					plain = __all_sync(0xffffffffu, plain);
					if (lane == 0 && anyvalid) t_walk += nkm;
					if (!plain) { dirty |= 1u << m; continue; }
					if (!thrf_finish_mate<B, MAXJ, true>(ix, rp, sv, A, m, len, (int)nkm, dst, recp, slot_ok, false, lane, TC, b2clean)) dirty |= 1u << m;
				}
				thrf_finish_pair(rp, sv, c, p, dw, dst, dirty, recp, slot_ok, want_aln, false, lane, t_fea, t_items);
				__syncwarp();
				continue;
			}
			if (rp.thr_on) {                                     // restored threading: hand the pair to thread_kernel (cpp:2070-2088)
				status = DBT_ST_THREAD_PENDING;
				if (lane == 0) { c_thr0 += 2; sv.thr[atomicAdd(sv.ncand + 3, 1u)] = make_uint4(c, p, dst | (kfm << 28) | (rmm << 30), 0u); }
			} else { status = DBT_ST_NOT_THREADED; if (lane == 0) c_thr0 += 2; }   // cpp:2070-2090 at HEAD
		}
		if (status < 0 && (rp.flags & DBT_F_EXTRACT)) {           // `-e`: the pair is handed out as it is, nothing is counted (cpp:2094-2099)
			if (lane == 0) {
				c_thr += 2;
#pragma unroll 1
				for (int m = 0; m < 2; ++m) {
					dbt_mate_record& o = W.mate[m];
					o.si = o.ei = o.si_ = o.ei_ = o.ti = -1; o.nt = 0; o.bs = 0; o.pad = 0; o.nk = 0;
					o.flags = (uint8_t)((((kfm >> m) & 1u) ? DBT_MF_KF : 0) | (((rmm >> m) & 1u) ? DBT_MF_RM : 0));
				}
			}
			__syncwarp();
			status = DBT_ST_COUNTED;
		}
		uint32_t bfm = 0;                                        // bit m = bf[m]
		if (status < 0 && ix.btable.slots) {                     // `-b`: bait filter (cpp:2110-2125): either mate removes the pair
#pragma unroll 1
			for (int m = 0; m < 2; ++m) {
				const uint32_t nkm = m ? nk1 : nk0;
				uint32_t* qm = nullptr;
				if (sv.qual) {                                      // FASTQ: the quality mask of the mate into the (free) probe staging area
					qm = reinterpret_cast<uint32_t*>(&W.slot[0][0][0]);
					__syncwarp();
					if (lane == 0) quality_mask(sv.qual + sv.qoff[2 * (uint64_t)p + m], (int)(nkm + K - 1), (int)K, (int)rp.qth, qm, MAXJ);
					__syncwarp();
				}
				if (bait_filtered<B>(ix.btable, W.pk + m * W64, WB, nkm, K, dst, lane, qm)) bfm |= 1u << m;
			}
			if (bfm) { if (lane == 0) c_bait += __popc(bfm & ~rmm); rmm = 3u; }
		}
		if (status < 0) {
			if (lane == 0) c_thr += 2;
			int span = 0;
			__syncwarp();
			// ---- assignTRkmc state machine (cpp:1470-1556) on the flank/TR masks ----
#pragma unroll 1
			for (int m = 0; m < 2; ++m) {
				const bool rmv = (rmm >> m) & 1u;
				MateOut mo;
				mo.si = mo.ei = mo.si_ = mo.ei_ = mo.ti = -1; mo.nt = 0; mo.bs = 0; mo.af = 0; mo.rm = rmv;
				uint32_t nas = 0;
				if (okam || !rmv) {                                 // cpp:1451
					const uint32_t nkm = m ? nk1 : nk0;
					nas = nkm;
					if (rmv) { mo.nt = -1; mo.bs = -1; mo.ti = -1; }   // cpp:1470-1475
					else {
						MaskSet<MAXJ> Fs, Ts;
						uint32_t ntr = 0;
#pragma unroll
						for (int j = 0; j < MAXJ; ++j) { Fs.w[j] = W.F[m][j]; Ts.w[j] = W.T[m][j]; ntr += __popc(Ts.w[j]); }
						mo = assign_masks<MAXJ>(Fs, Ts, (int)nkm, ntr & 0xFF, rp.nm_tr, mo);   // uint8_t ntr (cpp:1454)
					}
				}
				if (mo.rm) rmm |= 1u << m;
				span += mo.ei - mo.si;
				if (lane == 0) {
					dbt_mate_record& o = W.mate[m];
					o.si = (int16_t)mo.si; o.ei = (int16_t)mo.ei; o.si_ = (int16_t)mo.si_; o.ei_ = (int16_t)mo.ei_;
					o.ti = (int16_t)mo.ti; o.nt = (int8_t)mo.nt; o.bs = (int8_t)mo.bs;
					o.flags = (uint8_t)((((kfm >> m) & 1u) ? DBT_MF_KF : 0) | (((bfm >> m) & 1u) ? DBT_MF_BF : 0) | (mo.af ? DBT_MF_AF : 0) | (mo.rm ? DBT_MF_RM : 0));
					o.pad = 0; o.nk = (uint16_t)nas;
				}
			}
			__syncwarp();
			// ---- accumulate (cpp:2145-2158) ----
			if (rmm == 3u) { dst = nloci; status = DBT_ST_ASGN_REMOVED; }
			else {
				status = DBT_ST_COUNTED;
				const int nr = 2 - __popc(rmm);
				if (lane == 0) {
					atomicAdd(ix.nmapread + dst, (uint32_t)nr);
					c_asgn += nr;
					atomicAdd(reinterpret_cast<unsigned long long*>(ix.kmc + dst), (unsigned long long)(long long)span);
				}
				uint32_t inc_sum = 0;
#pragma unroll 1
				for (int m = 0; m < 2; ++m) {
					if ((rmm >> m) & 1u) continue;
#pragma unroll 1
					for (int j = 0; j < MAXJ; ++j) {
						const uint32_t tm = W.T[m][j];
						if (!tm) continue;
						inc_sum += __popc(tm);
						if ((tm >> lane) & 1u) {                    // one atomic per distinct TR k-mer among these 32 positions
							const uint32_t idx = W.aux[m][j * 32 + lane] & AUX_IDX_MASK;
							const uint32_t peers = __match_any_sync(tm, idx);
							if ((uint32_t)lane == (uint32_t)(__ffs(peers) - 1))
								atomicAdd(reinterpret_cast<unsigned long long*>(ix.counts + idx), (unsigned long long)__popc(peers));
						}
					}
				}
				if (lane == 0) c_inc += inc_sum;
			}
		}

		// ---- kam record (cpp:2169-2175); slot = candidate ordinal, compacted by dbt_wait ----
		if (wstat) {
			if (c >= rp.max_records) { if (lane == 0 && (status == DBT_ST_COUNTED || ((rp.flags & DBT_F_SIM) && status == DBT_ST_ASGN_REMOVED))) atomicOr(sv.ncand + 1, ERR_RECORD_OVERFLOW); }
			else if (status != DBT_ST_COUNTED && !((rp.flags & DBT_F_SIM) && status == DBT_ST_ASGN_REMOVED)) {
				if (lane == 0) reinterpret_cast<dbt_pair_record*>(sv.records + (uint64_t)c * rp.rec_stride)->status = (uint8_t)status;
			} else {
				uint8_t* rec = sv.records + (uint64_t)c * rp.rec_stride;
				if (lane == 0) {
					dbt_pair_record* h = reinterpret_cast<dbt_pair_record*>(rec);
					h->pair_index = sv.first_pair_index + p;
					h->dst = dst; h->dst0 = dst0; h->fc = 0; h->rc = 0;
					h->nhash0 = 0; h->nhash1 = (uint16_t)nhash1; h->status = (uint8_t)status;
					h->pad[0] = h->pad[1] = h->pad[2] = 0;
					h->mate[0] = W.mate[0]; h->mate[1] = W.mate[1];
				}
				// 2-bit annotation (cpp:122): byte b = k-mers 4b..4b+3, from the flank/TR masks
#pragma unroll 1
				for (int m = 0; m < 2; ++m) {
					uint8_t* dsta = rec + sizeof(dbt_pair_record) + (size_t)m * rp.as_stride;
					for (uint32_t b = lane; b < rp.as_stride; b += 32) {
						const uint32_t jw = b >> 3;
						const uint32_t fw = jw < (uint32_t)MAXJ ? W.F[m][jw] : 0u, tw = jw < (uint32_t)MAXJ ? W.T[m][jw] : 0u;
						const uint32_t fn = (fw >> ((b & 7) * 4)) & 0xF, tn = (tw >> ((b & 7) * 4)) & 0xF;
						uint32_t v = 0;
#pragma unroll
						for (int q = 0; q < 4; ++q) v |= (((fn >> q) & 1u) | (((tn >> q) & 1u) << 1)) << (2 * q);
						dsta[b] = (uint8_t)v;
					}
				}
			}
		}
	}
	if (lane == 0) {
		counter_add(sv.counters, DBT_C_NKMERFILTERED, c_kf);
		counter_add(sv.counters, DBT_C_NLOCUSASSIGNFILTERED, c_laf);
		counter_add(sv.counters, DBT_C_NQCFILTERED, c_qc);
		counter_add(sv.counters, DBT_C_NBAITFILTERED, c_bait);
		counter_add(sv.counters, DBT_C_NTHREADINGREADS, c_thr + c_thr0);
		counter_add(sv.counters, DBT_C_NFEASIBLEREADS, c_thr);
		counter_add(sv.counters, DBT_C_NASGNREADS, c_asgn);
		counter_add(sv.counters, DBT_C_NHASH1, c_h1);
		counter_add(sv.counters, DBT_C_NINC, c_inc);
	}
	if (fused) {                                                 // what thread_fast_kernel reports for the pairs it threads
		const uint32_t t_probe = __reduce_add_sync(0xffffffffu, TC.probe);
		if (lane == 0) {
			counter_add(sv.counters, DBT_C_THR_WALK, t_walk); counter_add(sv.counters, DBT_C_THR_REFRESH, t_probe);
			counter_add(sv.counters, DBT_C_THR_CLS, TC.cls); counter_add(sv.counters, DBT_C_NFEASIBLEREADS, t_fea);
			counter_add(sv.counters, DBT_C_NINC, TC.inc); counter_add(sv.counters, DBT_C_THR2_PAIRS, t_items);
			counter_add(sv.counters, DBT_C_THR1_FIXED, TC.fix); counter_add(sv.counters, DBT_C_THR_EC, TC.ec);
		}
	}
}

}  // namespace dbt
