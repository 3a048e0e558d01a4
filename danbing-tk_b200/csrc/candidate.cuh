// candidate_kernel: one warp per read pair that survived subfilter.
//
//   1. packed records -> shared memory
//   2. every canonical k-mer of both mates (lanes own positions lane, lane+32, ..) and ONE index
//      probe each; home buckets of a mate are fetched before any is inspected (memory-level
//      parallelism instead of 2x130 dependent look-ups, cpp:204-221)
//   3. kfilter's abort position from ballots of the miss flags (cpp:190-224)
//   4. countDupRemove (cpp:257-296): hits -> (k-mer<<10 | mate<<9 | pos) -> register bitonic sort ->
//      unique k-mers with per-mate multiplicities
//   5. fillstats (cpp:308-329): visiting order = libstdc++ std::sort by #mapped loci -- a table
//      look-up when every k-mer ties (the usual case), the exact introsort replay otherwise
//   6. find_matching_locus (cpp:364-453): closed form with warp scans when all k-mers belong to one
//      locus, the literal sequential vote (one lane) otherwise
//   7. assignTRkmc (cpp:1450-1556): per-k-mer class from the SAME index entry (no second probe); the
//      flank/TR state machine evaluated on ballot bit masks by every lane redundantly
//   8. accumulation (cpp:2145-2158): one 64-bit atomic per UNIQUE TR k-mer (multiplicity added at once)
//   9. kam record (cpp:2169-2175)
#pragma once
#include "device_types.cuh"
#include "introsort.cuh"
#include "bait.cuh"

namespace dbt {

// ------------------------------------------------------------------------------------------------
// register bitonic sort: 32*E keys, lane holds elements [lane*E, lane*E+E) ("blocked"), ascending.
// All-ascending formulation (flip step + half cleaners), so +inf padding never moves.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ce(uint64_t& a, uint64_t& b) { uint64_t lo = a < b ? a : b, hi = a < b ? b : a; a = lo; b = hi; }

template <int E>
__device__ __forceinline__ void sort_inlane(uint64_t (&v)[E], const int s, const bool flip) {
	// one compare-exchange layer inside a lane: partner register r^s (half cleaner) or r^(2s-1) (flip)
#pragma unroll
	for (int t = 1; t < E; t <<= 1) {
		if (s == t) {
			if (flip) {
#pragma unroll
				for (int r = 0; r < E; ++r) { const int q = r ^ (2 * t - 1); if (r < q) ce(v[r], v[q]); }
			} else {
#pragma unroll
				for (int r = 0; r < E; ++r) if (!(r & t)) ce(v[r], v[r | t]);
			}
		}
	}
}

// N = 32*E is passed at run time so that the stage loops stay rolled (instruction-cache footprint)
template <int E>
__device__ __forceinline__ void warp_bitonic_sort(uint64_t (&v)[E], const int lane, const int N) {
#pragma unroll 1
	for (int size = 2; size <= N; size <<= 1) {
		// ---- flip: element e pairs with e ^ (size-1) ----
		if (size <= E) sort_inlane<E>(v, size >> 1, true);
		else {
			const int lm = size / E - 1;                       // partner lane = lane ^ lm, partner register = E-1-r
			const bool keep_min = !(lane & ((lm + 1) >> 1));
#pragma unroll
			for (int r = 0; r < E / 2; ++r) {                  // registers r and E-1-r trade places with the partner lane
				const uint64_t oa = __shfl_xor_sync(0xffffffffu, v[E - 1 - r], lm), ob = __shfl_xor_sync(0xffffffffu, v[r], lm);
				const bool la = v[r] < oa, lb = v[E - 1 - r] < ob;
				v[r] = (la == keep_min) ? v[r] : oa;
				v[E - 1 - r] = (lb == keep_min) ? v[E - 1 - r] : ob;
			}
		}
		// ---- half cleaners: stride s >= E across lanes, s < E inside a lane ----
#pragma unroll 1
		for (int s = size >> 2; s >= 1; s >>= 1) {
			if (s >= E) {
				const int lm = s / E;
				const bool keep_min = !(lane & lm);
#pragma unroll
				for (int r = 0; r < E; ++r) {
					const uint64_t o = __shfl_xor_sync(0xffffffffu, v[r], lm);
					const bool lt = v[r] < o;
					v[r] = (lt == keep_min) ? v[r] : o;
				}
			} else sort_inlane<E>(v, s, false);
		}
	}
}

// ------------------------------------------------------------------------------------------------
// vote (cpp:331-453)
// ------------------------------------------------------------------------------------------------
struct Asgn { uint32_t idx; uint64_t fc, rc; };               // asgn_t, cpp:146-149

__device__ __forceinline__ void updatetop2(uint64_t cf, uint32_t ind, uint64_t cr, Asgn& top, Asgn& second) {   // cpp:331-347
	if (cf + cr > top.fc + top.rc) {
		if (top.idx != ind) { second = top; top.idx = ind; }
		top.fc = cf; top.rc = cr;
	} else if (cf + cr > second.fc + second.rc) {
		second.idx = ind; second.fc = cf; second.rc = cr;
	}
}

// per-pair hits1/hits2 (cpp:432-434): small shared-memory table; a pair that touches more than ~VOTE_SMEM_SLOTS loci (k-mers
// listed in hundreds of loci) spills into one array of a small POOL of epoch-tagged global arrays [nloci+1], taken under a
// lock for the duration of this pair's vote.  The pool is shared by every warp of every in-flight kernel of the context
// (both slots): holding a lock never waits on anything else, so there is no deadlock, and the arrays cost
// VOTE_POOL x (nloci+1) x 8 B instead of one array per resident warp (1.5 GB at 80 000 loci).
struct VoteHits {
	uint32_t* lkey; uint32_t* lf; uint32_t* lr;    // shared memory, VOTE_SMEM_SLOTS entries, key = locus+1
	uint64_t* pool; uint32_t* lock; uint32_t* epochs; uint64_t nloci; uint32_t seed;
	uint32_t used; int held; uint32_t epoch; volatile uint64_t* scratch; uint32_t spill_at;
	__device__ __noinline__ void acquire() {
		uint32_t i = seed % VOTE_POOL;
		while (atomicCAS(lock + i, 0u, 1u) != 0u) i = (i + 1) % VOTE_POOL;
		__threadfence();
		held = (int)i;
		scratch = pool + (uint64_t)i * (nloci + 1);
		epoch = *(volatile uint32_t*)(epochs + i) + 1;
		if (epoch >= 0xFFFFF0u) {                          // 24-bit tag about to wrap: clear the array
			for (uint64_t q = 0; q < nloci + 1; ++q) scratch[q] = 0;
			epoch = 1;
		}
	}
	__device__ __forceinline__ void release() {
		if (held < 0) return;
		*(volatile uint32_t*)(epochs + held) = epoch;
		__threadfence();
		atomicExch(lock + held, 0u);
		held = -1;
	}
	__device__ __forceinline__ void add(uint32_t locus, uint32_t df, uint32_t dr, uint32_t& f, uint32_t& r) {
		uint32_t h = (locus * 2654435761u) >> 26;            // 6 bits
		for (int probe = 0; probe < VOTE_SMEM_SLOTS; ++probe) {
			uint32_t s = (h + probe) & (VOTE_SMEM_SLOTS - 1);
			uint32_t kk = lkey[s];
			if (kk == locus + 1) { f = (lf[s] += df); r = (lr[s] += dr); return; }
			if (kk == 0) {
				if (used >= spill_at) break;                // keep the table sparse; spill to global
				lkey[s] = locus + 1; lf[s] = df; lr[s] = dr; ++used; f = df; r = dr; return;
			}
		}
		if (held < 0) acquire();
		uint64_t e = scratch[locus];
		uint32_t ef = 0, er = 0;
		if ((uint32_t)(e >> 40) == epoch) { ef = (uint32_t)(e >> 20) & 0xFFFFF; er = (uint32_t)e & 0xFFFFF; }
		ef += df; er += dr;
		scratch[locus] = ((uint64_t)epoch << 40) | ((uint64_t)(ef & 0xFFFFF) << 20) | (er & 0xFFFFF);
		f = ef; r = er;
	}
};

// class/idx word of an index entry with respect to locus `loc` (0 = k-mer unknown in that locus)
__device__ __forceinline__ uint32_t classify(const DevIndex& ix, uint32_t val, uint32_t aux, uint32_t loc) {
	if (aux == AUX_MISS) return 0;
	if (!(val & 1)) return ((val >> 1) == loc) ? aux : 0;
	uint32_t off = val >> 1, n = __ldg(ix.vv + off);
	for (uint32_t j = 1; j <= n; ++j) if (__ldg(ix.vv + off + j) == loc) return __ldg(ix.vv_aux + off + j);
	return 0;
}

// ------------------------------------------------------------------------------------------------
// assignTRkmc state machine (cpp:1477-1555) on bit masks.  F/T: bit i of word i/32 set when k-mer i is
// a flank / TR k-mer of destLocus0 (as[i] == 1 / 2).  Evaluated identically by every lane.
// MAX_NT is 2 in the reference (cpp:31, not settable) and dbt_create enforces it.
// ------------------------------------------------------------------------------------------------
struct MateOut { int si, ei, si_, ei_, ti, nt, bs, af, rm; };

template <int MAXJ>
__device__ __forceinline__ uint32_t mw_bit(const uint32_t (&w)[MAXJ], int i) {
	uint32_t x = 0;
#pragma unroll
	for (int j = 0; j < MAXJ; ++j) if (j == (i >> 5)) x = w[j];
	return (x >> (i & 31)) & 1u;
}
// highest set bit strictly below position t, or -1
template <int MAXJ>
__device__ __forceinline__ int mw_prev_set(const uint32_t (&w)[MAXJ], int t) {
	int best = -1;
#pragma unroll
	for (int j = 0; j < MAXJ; ++j) {
		uint32_t x = w[j];
		if (j == (t >> 5)) x &= (1u << (t & 31)) - 1u;
		else if (j > (t >> 5)) x = 0;
		if (x) best = 32 * j + 31 - __clz(x);
	}
	return best;
}

template <int MAXJ> struct MaskSet { uint32_t w[MAXJ]; };

// lowest set bit of w at a position >= from, or -1
template <int MAXJ>
__device__ __forceinline__ int mw_next_set(const uint32_t (&w)[MAXJ], int from) {
	int best = -1;
#pragma unroll
	for (int j = MAXJ - 1; j >= 0; --j) {
		uint32_t x = w[j];
		if (j == (from >> 5)) x &= ~((1u << (from & 31)) - 1u);
		else if (j < (from >> 5)) x = 0;
		if (x) best = 32 * j + __ffs(x) - 1;
	}
	return best;
}

template <int MAXJ>
__device__ __noinline__ MateOut assign_masks(const MaskSet<MAXJ> Fs, const MaskSet<MAXJ> Ts, int nk, uint32_t ntr8, uint32_t nm_tr, MateOut r) {
	const uint32_t (&F)[MAXJ] = Fs.w;
	const uint32_t (&T)[MAXJ] = Ts.w;
	uint32_t NZ[MAXJ];
#pragma unroll
	for (int j = 0; j < MAXJ; ++j) NZ[j] = F[j] | T[j];
	// A transition (cpp:1479-1481) is a non-zero k-mer whose state differs from the previous non-zero one.
	// States alternate, so transition n+1 is the first k-mer after transition n of the other class:
	// p0 = first non-zero k-mer (state bs, cpp:1500), t1 = first k-mer of the other class after p0,
	// t2 = first of class bs after t1, t3 = first of the other class after t2.
	const int pF = mw_next_set<MAXJ>(F, 0), pT = mw_next_set<MAXJ>(T, 0);
	const int p0 = (pF < 0) ? pT : (pT < 0 ? pF : (pF < pT ? pF : pT));
	const bool bsT = p0 >= 0 && p0 == pT;
	r.bs = p0 < 0 ? 0 : (bsT ? 2 : 1);
	int t1 = -1, t2 = -1, ntrans = 0;
	if (p0 >= 0) {
		t1 = bsT ? mw_next_set<MAXJ>(F, p0) : mw_next_set<MAXJ>(T, p0);
		if (t1 >= 0) {
			ntrans = 1;
			t2 = bsT ? mw_next_set<MAXJ>(T, t1) : mw_next_set<MAXJ>(F, t1);
			if (t2 >= 0) {
				ntrans = 2;
				const int t3 = bsT ? mw_next_set<MAXJ>(F, t2) : mw_next_set<MAXJ>(T, t2);
				if (t3 >= 0) ntrans = 3;
			}
		}
	}
	r.ti = t1;                                                                          // ti1 aliases r.ti (cpp:1457)
	if (ntrans >= 2 && r.bs == 2) { r.nt = 2; r.af = 1; r.rm = 1; return r; }            // TR-flank-TR (cpp:1491)
	if (ntrans >= 3) { r.nt = 3; r.af = 1; r.rm = 1; return r; }                          // noisy (cpp:1482)
	r.nt = ntrans;
	// unknown tract right before a transition (cpp:1485-1488, 1501-1515): [previous non-zero + 1, t)
	int si1 = -1, ei1 = -1, si2 = -1, ei2 = -1;
	if (ntrans >= 1 && !mw_bit<MAXJ>(NZ, t1 - 1)) { si1 = mw_prev_set<MAXJ>(NZ, t1) + 1; ei1 = t1; }
	if (ntrans >= 2 && !mw_bit<MAXJ>(NZ, t2 - 1)) { si2 = mw_prev_set<MAXJ>(NZ, t2) + 1; ei2 = t2; }
	if (ntrans == 0) {
		if (r.bs != 2) { r.af = 1; r.rm = 1; return r; }
		r.si = 0; r.ei = nk; r.si_ = 0; r.ei_ = nk;
	} else if (ntrans == 1) {
		if (r.bs == 1) { r.si = si1 >= 0 ? (si1 + ei1) / 2 : t1; r.ei = nk; r.si_ = si1 >= 0 ? ei1 : t1; r.ei_ = nk; }
		else           { r.si = 0; r.ei = si1 >= 0 ? (si1 + ei1) / 2 : t1; r.si_ = 0; r.ei_ = si1 >= 0 ? si1 : t1; }
	} else {
		if (ntr8 < nm_tr) { r.af = 1; r.rm = 1; return r; }
		r.si = (si1 >= 0 ? (si1 + ei1) / 2 : t1);
		r.ei = (si2 >= 0 ? (si2 + ei2) / 2 : t2);
		r.si_ = ei1 >= 0 ? ei1 : t1;
		r.ei_ = si2 >= 0 ? si2 : t2;
	}
	return r;
}


// ------------------------------------------------------------------------------------------------
// out-of-line rare paths (kept out of the hot loop's instruction footprint)
// ------------------------------------------------------------------------------------------------
// continue a probe in the stash, starting with stash bucket b (home bucket full + overflow flag)
// (RAW: hand the value word out as stored, with the graph entry in its top bits -- candidate_fast_kernel; else masked)
template <int B, bool RAW = false>
__device__ __noinline__ uint2 probe_walk(const Table t, uint64_t key, uint64_t b) {
	for (;;) {
		bool full = true; uint64_t k0 = 0;
#pragma unroll
		for (int q = 0; q < B; ++q) {
			const uint4 e = ldg128(t.slots + b * B + q);
			const uint64_t kk = slot_key(e);
			if (q == 0) k0 = kk;
			if (kk == SLOT_EMPTY) full = false;
			else if ((kk & SLOT_KEYMASK) == key) return make_uint2(RAW ? e.z : slot_val(e.z), e.w);
		}
		if (!full || !(k0 & SLOT_OVF)) return make_uint2(0u, AUX_MISS);
		b = stash_next(t, b);
	}
}

// multi-locus k-mer: class word of locus `loc` from the vv list (cpp:369-376 layout)
__device__ __noinline__ uint32_t classify_list(const uint32_t* __restrict__ vv, const uint32_t* __restrict__ vv_aux, uint32_t off, uint32_t loc) {
	const uint32_t n = __ldg(vv + off);
	for (uint32_t j = 1; j <= n; ++j) if (__ldg(vv + off + j) == loc) return __ldg(vv_aux + off + j);
	return 0;
}
__device__ __forceinline__ uint32_t classify2(const DevIndex& ix, uint32_t val, uint32_t aux, uint32_t loc) {
	if (aux == AUX_MISS) return 0;
	if (!(val & 1)) return ((val >> 1) == loc) ? aux : 0;
	return classify_list(ix.vv, ix.vv_aux, val >> 1, loc);
}

__device__ __noinline__ void perm_introsort(uint16_t* perm, const uint32_t* unml, int U) {
	IntroSort<uint16_t, uint32_t> s{perm, unml};
	s.sort(U);
}

// find_matching_locus, literal (cpp:364-422); one lane.  Returns top.idx / top.fc / top.rc.
__device__ __noinline__ uint3 vote_serial(const uint32_t* __restrict__ vv, const uint16_t* perm, const uint32_t* uval, const uint16_t* utf,
                                          const uint16_t* utr, uint32_t* lkey, uint32_t U, uint32_t total, uint32_t cth32,
                                          const DevIndex& ix, uint32_t seed, uint32_t spill_at) {
	Asgn top = {NAN32, 0, 0}, second = {NAN32, 0, 0};
	VoteHits vh{lkey, lkey + VOTE_SMEM_SLOTS, lkey + 2 * VOTE_SMEM_SLOTS, ix.vote_pool, ix.vote_lock, ix.vote_epoch, ix.nloci, seed, 0, -1, 0, nullptr, spill_at};
	uint64_t rem = total;
	const uint64_t cth = cth32;
	for (uint32_t i = 0; i < U; ++i) {
		const uint32_t u = perm[i];
		const uint32_t vi = uval[u], df = utf[u] & 0xFF, dr = utr[u] & 0xFF;      // PE_KMC is uint8 (cpp:42)
		rem = (i + 1 < U) ? rem - df - dr : 0;                                    // remain[i] (cpp:302-305; last entry stays 0)
		if (vi & 1) {
			uint32_t j0 = (vi >> 1) + 1, j1 = j0 + __ldg(vv + (vi >> 1));
			for (; j0 < j1; ++j0) {
				uint32_t locus = __ldg(vv + j0), f, r;
				vh.add(locus, df, dr, f, r);
				updatetop2(f, locus, r, top, second);
			}
		} else {
			uint32_t locus = vi >> 1, f, r;
			vh.add(locus, df, dr, f, r);
			updatetop2(f, locus, r, top, second);
		}
		if (!((top.fc + top.rc - second.fc - second.rc) < rem)) {                 // not get_acm2 (cpp:383)
			uint32_t j = i; uint64_t remj = rem;
			while ((top.fc < cth && cth - top.fc <= remj) || (top.rc < cth && cth - top.rc <= remj)) {   // get_acm1
				if (++j >= U) break;
				const uint32_t uj = perm[j];
				const uint32_t vj = uval[uj], dfj = utf[uj] & 0xFF, drj = utr[uj] & 0xFF;
				remj = (j + 1 < U) ? remj - dfj - drj : 0;
				if (vj & 1) {
					uint32_t j0 = (vj >> 1) + 1, j1 = j0 + __ldg(vv + (vj >> 1));
					for (; j0 < j1; ++j0) if (__ldg(vv + j0) == top.idx) { top.fc += dfj; top.rc += drj; break; }
				} else if ((vj >> 1) == top.idx) { top.fc += dfj; top.rc += drj; }
			}
			break;
		}
	}
	vh.release();
	return make_uint3(top.idx, (uint32_t)top.fc, (uint32_t)top.rc);
}

// ------------------------------------------------------------------------------------------------
// candidate_kernel: the GENERIC path -- exact countDupRemove / introsort replay / sequential vote.
// It serves (a) the pairs candidate_fast_kernel (candidate_fast.cuh) defers to sv.slow: pairs whose
// hits are not all unique to one locus, or whose accept/reject decision depends on the visiting
// order, and (b) with all != 0 every candidate (DBT_F_TRACE: the records then carry the exact
// top.fc/top.rc of the early stop).  Warps are free-running, one pair at a time.
// ------------------------------------------------------------------------------------------------
template <int B, int MAXJ>
__global__ void __launch_bounds__(CAND_WARPS_SLOW * 32, 4) candidate_kernel(DevIndex ix, RunParams rp, SlotView sv, int all) {
	constexpr int WPB = CAND_WARPS_SLOW;
	constexpr int E = (2 * MAXJ <= 16) ? 16 : 32;            // sort slots per lane: 32*E >= 2*nk_max, E >= 2*MAXJ
	extern __shared__ __align__(16) uint8_t smem_raw[];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t gwarp = blockIdx.x * WPB + wib;
	const uint32_t nwarps = gridDim.x * WPB;
	const uint32_t W64 = rp.lay.W64, WB = rp.lay.WB, K = rp.k, NKP = rp.nkp, UMAX = rp.umax;
	const bool okam = (rp.flags & (DBT_F_KAM | DBT_F_EXTRACT | DBT_F_BUBBLES)) != 0 && !rp.thr_on, trace = (rp.flags & DBT_F_TRACE) != 0;   // no kam lines with threading (cpp:2110)
	const bool wstat = okam || (rp.thr_on && (rp.flags & (DBT_F_ALN | DBT_F_ALN_MIN)));

	// ---- shared-memory carve-up (per warp); dbt_create mirrors the sizes ----
	uint8_t* ws = smem_raw + (size_t)wib * rp.warp_smem;
	uint64_t* pk    = reinterpret_cast<uint64_t*>(ws);                  // [2][W64] packed records
	uint64_t* pkey  = pk + 2 * W64;                                     // [2][NKP] canonical k-mer per position
	uint32_t* pval  = reinterpret_cast<uint32_t*>(pkey + 2 * NKP);      // [2][NKP] index value per position
	uint32_t* paux  = pval + 2 * NKP;                                   // [2][NKP] aux word or AUX_MISS
	uint32_t* uval  = paux + 2 * NKP;                                   // [UMAX]   unique k-mers: value
	uint32_t* uaux  = uval + UMAX;
	uint32_t* unml  = uaux + UMAX;                                      // number of mapped loci (cpp:314-317)
	uint32_t* lkey  = unml + UMAX;                                      // vote table: lkey | lf | lr
	uint16_t* utf   = reinterpret_cast<uint16_t*>(lkey + 3 * VOTE_SMEM_SLOTS);   // true multiplicity in mate 0 / mate 1
	uint16_t* utr   = utf + UMAX;
	uint16_t* ustart = utr + UMAX;                                      // [UMAX+2] group start in the sorted order
	uint16_t* umid  = ustart + UMAX + 2;                                // first mate-1 element of the group
	uint16_t* perm  = umid + UMAX;                                      // visiting order (cpp:320-327)

	uint32_t c_kf = 0, c_laf = 0, c_qc = 0, c_thr = 0, c_asgn = 0, c_h1 = 0, c_inc = 0, c_pairs = 0, c_thr0 = 0, c_bait = 0;   // lane-0 accumulators
	const uint32_t nwork = all ? sv.ncand[0] : sv.ncand[2];
	const uint32_t stride = nwarps;

#pragma unroll 1
	for (uint32_t w = gwarp; w < nwork; w += stride) {
		const bool active = true;
		++c_pairs;
		const uint32_t c = active ? (all ? w : sv.slow[w]) : 0;
		const uint32_t p = active ? sv.cand[c] : 0;
		uint32_t nk[2] = {0, 0};
		uint32_t hitmask[2][MAXJ];            // ballot of "k-mer is in the index" per 32 positions
		int kf[2] = {0, 0}, rm[2] = {1, 1}, hf[2] = {0, 0}, bf[2] = {0, 0};
		uint32_t nhash1 = 0;
		int status = -1;
		bool deferred = false;
		uint32_t dst = (uint32_t)ix.nloci, dst0 = NAN32, fc = 0, rc = 0;
		uint32_t U = 0, total = 0, val0 = 0;
		bool any_odd = false, one_locus = false;
		MateOut mo[2];
#pragma unroll
		for (int m = 0; m < 2; ++m) { mo[m].si = mo[m].ei = mo[m].si_ = mo[m].ei_ = mo[m].ti = -1; mo[m].nt = 0; mo[m].bs = 0; mo[m].af = 0; mo[m].rm = 0; }
		uint32_t nas[2] = {0, 0};
		MaskSet<MAXJ> Fm[2], Tm[2];
#pragma unroll
		for (int m = 0; m < 2; ++m)
#pragma unroll
			for (int j = 0; j < MAXJ; ++j) { Fm[m].w[j] = 0; Tm[m].w[j] = 0; hitmask[m][j] = 0; }

		// =========================== phase 1: k-mers, probes, kfilter ===========================
		if (active) {
			__syncwarp();
			for (uint32_t i = lane; i < 2 * W64; i += 32) {
				uint32_t m = i >= W64;
				pk[i] = sv.packed[m][(uint64_t)p * W64 + (i - m * W64)];
			}
			nk[0] = (sv.meta[0][p] & META_LENMASK) - K + 1; nk[1] = (sv.meta[1][p] & META_LENMASK) - K + 1;
			for (uint32_t i = lane; i < 3 * VOTE_SMEM_SLOTS; i += 32) lkey[i] = 0;
			__syncwarp();
			// all canonical k-mers + ONE index probe each (A2 + the lookups of kfilter, cpp:207/218)
#pragma unroll 1
			for (int m = 0; m < 2; ++m) {
				const uint32_t nkm = m ? nk[1] : nk[0];
				uint4 s[MAXJ][B];
				uint64_t key[MAXJ], hb[MAXJ];
				// k-mers at positions lane+32j: same bit offset (2*lane) in consecutive base words, same bit
				// offset (lane) in consecutive 32-bit halves of the invalid mask -> funnel shifts
				const uint64_t* rec = pk + m * W64;
				const uint32_t* mk32 = reinterpret_cast<const uint32_t*>(rec + WB);
				const uint32_t sh = 2 * lane;
				const uint64_t kmask = (1ull << (2 * K)) - 1, vmask = (1ull << K) - 1;
				uint64_t wlo = rec[0];
				uint32_t mlo = mk32[0];
#pragma unroll
				for (int j = 0; j < MAXJ; ++j) {
					const uint32_t i = j * 32 + lane;
					const uint64_t whi = (j + 1 < (int)WB) ? rec[j + 1] : 0ull;
					const uint32_t mnx = ((uint32_t)(j + 1) < 2 * rp.lay.WM) ? mk32[j + 1] : 0xFFFFFFFFu;
					const uint64_t wv = sh ? ((wlo >> sh) | (whi << (64 - sh))) : wlo;
					const uint64_t mm = (((uint64_t)mnx << 32) | mlo) >> lane;         // lane + K <= 58 bits: two halves suffice
					uint64_t kk = KMER_INVALID;
					if (i < nkm && (mm & vmask) == 0) {
						const uint64_t x = wv & kmask, rcx = (~wv) & kmask, fw = rev2(x) >> (64 - 2 * K);
						kk = fw < rcx ? fw : rcx;
					}
					key[j] = kk;
					wlo = whi; mlo = mnx;
					hb[j] = 0;
					if (kk != KMER_INVALID) {
						hb[j] = home_bucket(ix.table, kk);
#pragma unroll
						for (int q = 0; q < B; ++q) s[j][q] = ldg128(ix.table.slots + hb[j] * B + q);
					}
				}
#pragma unroll
				for (int j = 0; j < MAXJ; ++j) {
					const uint32_t i = j * 32 + lane;
					uint32_t v = 0, a = AUX_MISS;
					if (key[j] != KMER_INVALID) {
						bool full = true, found = false;
#pragma unroll
						for (int q = 0; q < B; ++q) {
							const uint64_t kk = slot_key(s[j][q]);
							if (kk == SLOT_EMPTY) full = false;
							else if ((kk & SLOT_KEYMASK) == key[j]) { v = slot_val(s[j][q].z); a = s[j][q].w; found = true; }
						}
						if (!found && full && (slot_key(s[j][0]) & SLOT_OVF)) { const uint2 r2 = probe_walk<B>(ix.table, key[j], stash_bucket(ix.table, key[j])); v = r2.x; a = r2.y; }
					}
					pkey[m * NKP + i] = key[j]; pval[m * NKP + i] = v; paux[m * NKP + i] = a;
				}
			}
			__syncwarp();
#pragma unroll
			for (int m = 0; m < 2; ++m)
#pragma unroll
				for (int j = 0; j < MAXJ; ++j) hitmask[m][j] = __ballot_sync(0xffffffffu, paux[m * NKP + j * 32 + lane] != AUX_MISS);

			// kfilter (cpp:190-224): abort a mate at its (nk-cth+1)-th miss
			kf[0] = nk[0] < rp.cth; kf[1] = nk[1] < rp.cth;
			rm[0] = kf[0]; rm[1] = kf[1];
			if (!(rm[0] && rm[1])) {
#pragma unroll
				for (int m = 0; m < 2; ++m) {
					if (rm[m]) continue;
					const uint32_t max_ns = nk[m] - rp.cth;
					uint32_t cum = 0, probes = nk[m], aword = 0, abase = 0, arank = 0; bool aborted = false;
#pragma unroll
					for (int j = 0; j < MAXJ; ++j) {
						const uint32_t inrange = (nk[m] >= (uint32_t)(j + 1) * 32) ? 0xFFFFFFFFu : ((nk[m] > (uint32_t)j * 32) ? ((1u << (nk[m] - j * 32)) - 1u) : 0u);
						const uint32_t mm = ~hitmask[m][j] & inrange;
						if (!aborted && cum + __popc(mm) > max_ns) { aborted = true; aword = mm; abase = j * 32; arank = max_ns - cum; }
						cum += __popc(mm);
					}
					if (aborted) {                                   // position of the (arank+1)-th miss of that word
						uint32_t pos = 0, nn = arank;
#pragma unroll
						for (int shf = 16; shf >= 1; shf >>= 1) {
							const uint32_t cnt = __popc((aword >> pos) & ((1u << shf) - 1u));
							if (nn >= cnt) { nn -= cnt; pos += shf; }
						}
						probes = abase + pos + 1;
					}
					nhash1 += probes;
					kf[m] = aborted; rm[m] |= kf[m];
				}
			}
			if (rm[0] && rm[1]) status = DBT_ST_KFILTERED;
		}

		// =========================== phase 2: countDupRemove (cpp:257-296) ===========================
		if (active && status < 0) {
			uint64_t v[E];
			uint32_t n = 0;
#pragma unroll
			for (int r = 0; r < E; ++r) v[r] = ~0ull;
#pragma unroll
			for (int m = 0; m < 2; ++m)
#pragma unroll
				for (int j = 0; j < MAXJ; ++j) {
					const bool hit = !rm[m] && ((hitmask[m][j] >> lane) & 1u);
					if (hit) v[m * MAXJ + j] = (pkey[m * NKP + j * 32 + lane] << 10) | ((uint64_t)m << 9) | (uint32_t)(j * 32 + lane);
					if (!rm[m]) n += __popc(hitmask[m][j]);
				}
			warp_bitonic_sort<E>(v, lane, (int)rp.sortn);
			// collapse equal keys (cpp:278-295).  Lane holds sorted elements lane*E .. lane*E+E-1.
			const uint64_t prevlast = __shfl_up_sync(0xffffffffu, v[E - 1], 1);
			uint32_t headbits = 0, midbits = 0;
#pragma unroll
			for (int r = 0; r < E; ++r) {
				const uint64_t e = v[r], prev = r ? v[r - 1] : prevlast;
				const bool valid = e != ~0ull, first = (lane == 0 && r == 0);
				const bool head = valid && (first || ((prev ^ e) >> 10) != 0);
				const bool mid = valid && (((uint32_t)e >> 9) & 1u) && (head || !(((uint32_t)prev >> 9) & 1u));
				headbits |= (uint32_t)head << r; midbits |= (uint32_t)mid << r;
			}
			uint32_t hc = __popc(headbits), incl = hc;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
			U = __shfl_sync(0xffffffffu, incl, 31);
			int g = (int)(incl - hc) - 1;                       // group of the element just before this lane's first
			const int g0 = g;
#pragma unroll
			for (int r = 0; r < E; ++r) {
				if ((headbits >> r) & 1u) {
					++g;
					const uint64_t e = v[r];
					ustart[g] = (uint16_t)(lane * E + r); umid[g] = 0xFFFF;
					const uint32_t m = (uint32_t)(e >> 9) & 1, pos = (uint32_t)e & 511;
					uval[g] = pval[m * NKP + pos]; uaux[g] = paux[m * NKP + pos];
				}
			}
			if (lane == 0) ustart[U] = (uint16_t)n;
			__syncwarp();
			g = g0;
#pragma unroll
			for (int r = 0; r < E; ++r) {
				if ((headbits >> r) & 1u) ++g;
				if ((midbits >> r) & 1u) umid[g] = (uint16_t)(lane * E + r);
			}
			__syncwarp();
		}

		// =========================== phase 3: fillstats + vote (cpp:308-453) ===========================
		if (active && status < 0) {
			bool odd = false, differs = false; uint32_t wsum8 = 0;
			val0 = uval[0];
			for (uint32_t u = lane; u < U; u += 32) {
				const uint32_t s = ustart[u], e = ustart[u + 1], md = umid[u] == 0xFFFF ? e : umid[u];
				utf[u] = (uint16_t)(md - s); utr[u] = (uint16_t)(e - md);
				wsum8 += ((md - s) & 0xFF) + ((e - md) & 0xFF);              // PE_KMC is uint8 (cpp:42)
				const uint32_t v = uval[u];
				uint32_t nm = 1;
				if (v & 1) { nm = __ldg(ix.vv + (v >> 1)); odd = true; }
				if (v != val0) differs = true;
				unml[u] = nm;
			}
			any_odd = __any_sync(0xffffffffu, odd);
			one_locus = !any_odd && !__any_sync(0xffffffffu, differs) && rp.nk_max <= 255;
			total = warp_sum(wsum8);                                          // dupsum of countRemain (cpp:300)
		}
		if (active && status < 0 && !deferred) {
			// visiting order = std::sort by number of mapped loci (cpp:320-327)
			if (!any_odd && U <= ix.permtab_maxU) {
				const uint16_t* pt = ix.permtab + (size_t)U * (U - 1) / 2;
				for (uint32_t u = lane; u < U; u += 32) perm[u] = pt[u];
			} else {
				for (uint32_t u = lane; u < U; u += 32) perm[u] = (uint16_t)u;
				__syncwarp();
				if (lane == 0) perm_introsort(perm, unml, (int)U);
			}
			__syncwarp();
			// find_matching_locus (cpp:364-422)
			if (one_locus) {
				// Every k-mer belongs to locus val0>>1: `second` never leaves {NAN32,0,0}, top = running
				// sums in visiting order.  Stop index i* = first i with cumF+cumR >= remain[i] (cpp:383);
				// then counting goes on while get_acm1 holds (cpp:398): j* = first j >= i* where it fails.
				uint32_t carryF = 0, carryR = 0; bool stopped = false, done = false;
				const uint32_t cth = rp.cth;
#pragma unroll 1
				for (uint32_t b0 = 0; b0 < U && !done; b0 += 32) {
					const uint32_t i = b0 + lane;
					uint32_t f = 0, r = 0;
					if (i < U) { const uint32_t u = perm[i]; f = utf[u]; r = utr[u]; }
					uint32_t cfr = (f << 16) | r;                            // both running sums in one scan (each < 2^16)
#pragma unroll
					for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, cfr, o); if (lane >= o) cfr += y; }
					const uint32_t cf = (cfr >> 16) + carryF, cr = (cfr & 0xFFFF) + carryR;
					const uint32_t rem = (i + 1 < U) ? total - cf - cr : 0;
					const bool in = i < U;
					uint32_t lo = 0;                                          // first lane allowed to end the count
					if (!stopped) {
						const uint32_t sm = __ballot_sync(0xffffffffu, in && (cf + cr >= rem));
						if (sm) { stopped = true; lo = __ffs(sm) - 1; }
					}
					if (stopped) {
						const bool cond = (cf < cth && cth - cf <= rem) || (cr < cth && cth - cr <= rem);
						const uint32_t em = __ballot_sync(0xffffffffu, in && !cond && (uint32_t)lane >= lo);
						if (em) {
							const int jl = __ffs(em) - 1;
							fc = __shfl_sync(0xffffffffu, cf, jl); rc = __shfl_sync(0xffffffffu, cr, jl);
							done = true;
						}
					}
					carryF = __shfl_sync(0xffffffffu, cf, 31); carryR = __shfl_sync(0xffffffffu, cr, 31);
				}
				dst0 = val0 >> 1;
			} else {
				uint3 t3 = make_uint3(NAN32, 0, 0);
				if (lane == 0) t3 = vote_serial(ix.vv, perm, uval, utf, utr, lkey, U, total, rp.cth, ix, gwarp, rp.vote_spill_at);
				dst0 = __shfl_sync(0xffffffffu, t3.x, 0);
				fc = __shfl_sync(0xffffffffu, t3.y, 0);
				rc = __shfl_sync(0xffffffffu, t3.z, 0);
			}
			// countHit tail (cpp:436-452)
			const bool test1 = fc >= rp.cth && rc >= rp.cth, test2 = (fc + rc) >= 2 * rp.cth;
			if ((test1 || test2) && dst0 != NAN32) dst = dst0;
			else { hf[0] = 1 & !rm[0]; hf[1] = 1 & !rm[1]; rm[0] = rm[1] = 1; status = DBT_ST_UNASSIGNED; if (lane == 0) c_laf += hf[0] + hf[1]; }
			// QC gate (cpp:2059-2062)
			if (status < 0 && (rp.flags & DBT_F_QC) && ix.qc && !ix.qc[dst]) { status = DBT_ST_QCFILTERED; if (lane == 0) c_qc += 2 - rm[0] - rm[1]; }
		}

		// =========================== phase 4: assignTRkmc, accumulate, record ===========================
		if (active && !deferred) {
			if (lane == 0) { c_kf += kf[0] + kf[1]; c_h1 += nhash1; }
			if (status < 0 && (rp.flags & DBT_F_THREADING)) {
				if (rp.thr_on) {                                     // restored threading: hand the pair to thread_kernel (cpp:2070-2088)
					status = DBT_ST_THREAD_PENDING;
					if (lane == 0) {
						c_thr0 += 2;
						sv.thr[atomicAdd(sv.ncand + 3, 1u)] = make_uint4(c, p, dst | ((uint32_t)(kf[0] | (kf[1] << 1)) << 28) | ((uint32_t)(rm[0] | (rm[1] << 1)) << 30), 0u);
					}
				} else { status = DBT_ST_NOT_THREADED; if (lane == 0) c_thr0 += 2; }   // cpp:2070-2090 at HEAD
			}
			if (status < 0 && (rp.flags & DBT_F_EXTRACT)) {       // `-e`: the pair is handed out as it is, nothing is counted (cpp:2094-2099)
				if (lane == 0) c_thr += 2;
				status = DBT_ST_COUNTED;
			}
			if (status < 0 && ix.btable.slots) {                 // `-b`: bait filter (cpp:2110-2125): either mate removes the pair
#pragma unroll 1
				for (int m = 0; m < 2; ++m) {
					uint32_t* qm = nullptr;
					if (sv.qual) {                                  // FASTQ: the quality mask of the mate into the (spent) vote table
						qm = lkey;
						__syncwarp();
						if (lane == 0) quality_mask(sv.qual + sv.qoff[2 * (uint64_t)p + m], (int)(nk[m] + K - 1), (int)K, (int)rp.qth, qm, MAXJ);
						__syncwarp();
					}
					if (bait_filtered<B>(ix.btable, pk + m * W64, WB, nk[m], K, dst, lane, qm)) bf[m] = 1;
				}
				if (bf[0] | bf[1]) { if (lane == 0) c_bait += (bf[0] & !rm[0]) + (bf[1] & !rm[1]); rm[0] = rm[1] = 1; }
			}
			if (status < 0) {
				if (lane == 0) c_thr += 2;
				// assignTRkmc (cpp:1450-1556) against the tables of destLocus0
#pragma unroll
				for (int m = 0; m < 2; ++m) {
					if (!okam && rm[m]) continue;                           // cpp:1451
					nas[m] = nk[m];
					uint32_t ntr = 0;
#pragma unroll
					for (int j = 0; j < MAXJ; ++j) {
						const uint32_t i = j * 32 + lane;
						uint32_t cls = 0;
						if (i < nk[m]) cls = classify2(ix, pval[m * NKP + i], paux[m * NKP + i], dst0) >> AUX_CLASS_SHIFT;
						Fm[m].w[j] = __ballot_sync(0xffffffffu, cls == 1);
						Tm[m].w[j] = __ballot_sync(0xffffffffu, cls == 2);
						ntr += __popc(Tm[m].w[j]);
					}
					mo[m].rm = rm[m];
					if (rm[m]) { mo[m].nt = -1; mo[m].bs = -1; mo[m].ti = -1; }   // cpp:1470-1475
					else mo[m] = assign_masks<MAXJ>(Fm[m], Tm[m], (int)nk[m], ntr & 0xFF, rp.nm_tr, mo[m]);   // uint8_t ntr (cpp:1454)
					rm[m] = mo[m].rm;
				}
				// accumulate (cpp:2145-2158)
				if (rm[0] && rm[1]) { dst = (uint32_t)ix.nloci; status = DBT_ST_ASGN_REMOVED; }
				else {
					status = DBT_ST_COUNTED;
					const int nr = 2 - rm[0] - rm[1];
					if (lane == 0) {
						atomicAdd(ix.nmapread + dst, (uint32_t)nr);
						c_asgn += nr;
						const int span = (mo[0].ei - mo[0].si) + (mo[1].ei - mo[1].si);
						atomicAdd(reinterpret_cast<unsigned long long*>(ix.kmc + dst), (unsigned long long)(long long)span);
					}
					uint32_t inc_sum = 0;
					for (uint32_t u = lane; u < U; u += 32) {
						const uint32_t inc = (rm[0] ? 0u : utf[u]) + (rm[1] ? 0u : utr[u]);
						if (!inc) continue;
						const uint32_t ca = classify2(ix, uval[u], uaux[u], dst0);
						if ((ca >> AUX_CLASS_SHIFT) == 2) {
							atomicAdd(reinterpret_cast<unsigned long long*>(ix.counts + (ca & AUX_IDX_MASK)), (unsigned long long)inc);
							inc_sum += inc;
						}
					}
					inc_sum = warp_sum(inc_sum);
					if (lane == 0) c_inc += inc_sum;
				}
			}
			// record
			const bool emit = trace || (okam && (status == DBT_ST_COUNTED || ((rp.flags & DBT_F_SIM) && status == DBT_ST_ASGN_REMOVED)));
			if (!emit && wstat && c < rp.max_records) {
				// the slot of a candidate that prints no kam line must not keep a stale record
				if (lane == 0) reinterpret_cast<dbt_pair_record*>(sv.records + (uint64_t)c * rp.rec_stride)->status = (uint8_t)status;
			}
			if (emit) {
				const uint64_t slot = trace ? (uint64_t)p : (uint64_t)c;
				if (!trace && slot >= rp.max_records) { if (lane == 0) atomicOr(sv.ncand + 1, ERR_RECORD_OVERFLOW); }
				else {
					uint8_t* rec = sv.records + slot * rp.rec_stride;
					if (lane == 0) {
						dbt_pair_record* h = reinterpret_cast<dbt_pair_record*>(rec);
						const uint16_t nh0 = trace ? h->nhash0 : 0;       // left there by subfilter_kernel
						h->pair_index = sv.first_pair_index + p;
						const bool voted = status >= DBT_ST_UNASSIGNED;
						h->dst = voted ? dst : (uint32_t)ix.nloci; h->dst0 = voted ? dst0 : NAN32;
						h->fc = (voted && trace) ? fc : 0; h->rc = (voted && trace) ? rc : 0;   // exact only under DBT_F_TRACE (nm1/nm2 are unused downstream, cpp:2016)
						h->nhash0 = nh0; h->nhash1 = (uint16_t)nhash1; h->status = (uint8_t)status;
						h->pad[0] = h->pad[1] = h->pad[2] = 0;
#pragma unroll
						for (int m = 0; m < 2; ++m) {
							dbt_mate_record& o = h->mate[m];
							o.si = (int16_t)mo[m].si; o.ei = (int16_t)mo[m].ei; o.si_ = (int16_t)mo[m].si_; o.ei_ = (int16_t)mo[m].ei_;
							o.ti = (int16_t)mo[m].ti; o.nt = (int8_t)mo[m].nt; o.bs = (int8_t)mo[m].bs;
							o.flags = (uint8_t)((kf[m] ? DBT_MF_KF : 0) | (hf[m] ? DBT_MF_HF : 0) | (bf[m] ? DBT_MF_BF : 0) | (mo[m].af ? DBT_MF_AF : 0) | (rm[m] ? DBT_MF_RM : 0));
							o.pad = 0; o.nk = (uint16_t)nas[m];
						}
					}
					// 2-bit annotation (cpp:122): byte b = k-mers 4b..4b+3, from the flank/TR masks
#pragma unroll
					for (int m = 0; m < 2; ++m) {
						uint8_t* dsta = rec + sizeof(dbt_pair_record) + (size_t)m * rp.as_stride;
						for (uint32_t b = lane; b < rp.as_stride; b += 32) {
							uint32_t fw = 0, tw = 0;
#pragma unroll
							for (int j = 0; j < MAXJ; ++j) if (j == (int)(b >> 3)) { fw = Fm[m].w[j]; tw = Tm[m].w[j]; }
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
		counter_add(sv.counters, DBT_C_GEN_PAIRS, c_pairs);
		counter_add(sv.counters, DBT_C_GEN_NHASH1, c_h1);
		counter_add(sv.counters, DBT_C_GEN_NTHREADINGREADS, c_thr);
		counter_add(sv.counters, DBT_C_GEN_NINC, c_inc);
	}
}

}  // namespace dbt
