// thread_fast_kernel: the first of the two threading kernels (SURVEY.md §8(a) rows T1-T3, T6 + decision D1 of §8(c)).
//
// Three quarters of the mates handed to isThreadFeasible (cpp:1114-1260) carry no sequencing error: their first k-mer is
// a node of the locus graph and every later k-mer sits on an out-edge of its predecessor.  The reference then anchors at
// k-mer 0 (cpp:1129; find_anchor marks es[0..k) '=', cpp:878-888), walks the whole read through cpp:1167-1180
// (tr[i] = '=' for TR k-mers, '.' otherwise; es[i+k-1] = '=') and returns 1 (cpp:1259: no skip, no correction).  Whether a
// mate is of that kind is a function of per-position data only, so this kernel decides it for BOTH mates of every pair
// with one probe round per mate and nothing else: no walk state, no CIGAR arrays, no correction code -- about 1.5 k warp
// instructions per pair where the all-in-one kernel of round 1 spent 7 k (profiles/r1/s28: instruction-fetch bound, every
// warp of a scheduler in different code).  A clean mate is finished here: its TR k-mers are counted (D1: both edited
// vectors of a pair that threads are counted, cpp:2203-2208 -- a clean mate makes the pair thread) and its half of the
// aln record is written from the ballots.  Mates with an error go to a work list for thread_kernel (thread_kernel.cuh),
// which owns find_anchor / errorCorrection_* / edit_kmers_*; a pair with one clean mate is already known to count, a
// pair with two dirty mates is decided there.
//
// Per mate, ROLLED loops over the 32-position groups (small code; every warp of the SM runs these same loops):
//   A   k-mer of position j*32+lane straight from the packed words (funnel shifts; the little-endian packing makes the
//       reverse complement a NOT + mask, kmer.cuh), canonical form, home bucket of (canonical k-mer, locus) -> 32-byte
//       asynchronous copy into shared memory; all groups of the mate are in flight together.  Also here, from the packed
//       words alone: "window i equals window i+1" (the homopolymer repeat of cpp:1152) -- such mates are not clean;
//   B1  match the staged slots -> node bits + count-array position per position; keys that continue in the stash (7 % at
//       load 0.4) take a second, again batched, round;
//   B2  "on an out-edge of its predecessor" (cpp:1170-1178 with node == kmers[ki-1]) and "TR k-mer" ballots.
#pragma once
#include "thread_kernel.cuh"          // CIGAR codes, kmer_rc, THR2_* work-item bits

namespace dbt {

#ifndef THRF_WARPS_N
#define THRF_WARPS_N 4
#endif
#ifndef THRF_MINBLOCKS
#define THRF_MINBLOCKS 7
#endif
constexpr int THRF_WARPS = THRF_WARPS_N;
constexpr uint64_t THRF_RCFLAG = 1ull << 62;       // W.key: the read meets the k-mer in its reverse-complement orientation (k <= 27: bits 54.. are free)
constexpr uint32_t THRF_F3_SHIFT = 60;             // W.key bits 60-61: last base of the strand-specific k-mer

// continuation of a look-up behind the first stash bucket (rare): out of line
template <int B>
__device__ __noinline__ GNode g_lookup_canon_nl(const Table gt, uint64_t canon, bool is_rc, uint32_t locus, uint64_t b) {
	const uint64_t hk = g_hash_key(canon, locus);
	(void)hk;
	for (;;) {
		const uint4* bp = gt.slots + b * B;
		bool full = true; uint64_t k0 = 0;
#pragma unroll
		for (int q = 0; q < B; ++q) {
			const uint4 sl = ldg128(bp + q);
			const uint64_t kk = slot_key(sl);
			if (q == 0) k0 = kk;
			if (kk == SLOT_EMPTY) full = false;
			else if ((kk & SLOT_KEYMASK) == canon && (sl.z & G_LOCUS_MASK) == locus) return g_decode(sl.z, sl.w, is_rc);
		}
		if (!full || !(k0 & SLOT_OVF)) { GNode n; n.bits = 0; n.tr = 0; return n; }
		b = stash_next(gt, b);
	}
}

template <int B>
__device__ __noinline__ GNode g_lookup_fast_nl(const Table gt, uint64_t fw, uint32_t locus, uint32_t k) { return g_lookup<B>(gt, fw, locus, k); }

// What is left of a mate once every position has its graph entry (W.key / W.nb / W.trx): is it on a path of the graph, can the
// stops be settled here, and if so the counts and its half of the aln record.  Shared by thread_fast_kernel (which probes the graph
// table for the entries) and candidate_fast_kernel (whose index entries carry the graph bits when the graph is merged into the index,
// dbt_load_index).  Returns false when the mate has to go to thread_kernel.
struct ThrfArrays { uint64_t* key; uint32_t* trx; uint16_t* nb; uint32_t* T; uint32_t* ok; uint32_t* P; };
struct ThrfCount { uint32_t inc = 0, cls = 0, probe = 0, fix = 0, ec = 0; };

// HAVE_B2: the caller has already formed the three ballots of the B2 pass (W.T / W.ok / W.P) while it filled the arrays and says
// whether the mate is clean (candidate_fast_kernel: the values are in registers there, a second pass over shared memory is 8 % of
// that kernel's instructions)
template <int B, int MAXJ, bool HAVE_B2 = false>
__device__ __forceinline__ bool thrf_finish_mate(const DevIndex& ix, const RunParams& rp, const SlotView& sv, const ThrfArrays W, int m, uint32_t len, int nkm, uint32_t dst,
                                                 uint8_t* rec, bool slot_ok, bool trace, int lane, ThrfCount& C, bool clean_in = true) {
	const uint32_t K = rp.k;
	const Table gt = ix.gtable;
	const uint64_t kmask = (1ull << (2 * K)) - 1;
	const bool correction1 = rp.correction != 0 && rp.max_ncorrection >= 1;
	const int ng = (nkm + 31) >> 5;
	(void)sv;
		__syncwarp();
		// ---- B2: every k-mer on an out-edge of its predecessor, the first one a node? ----
		bool clean = clean_in;
#pragma unroll 1
		for (int j = 0; j < (HAVE_B2 ? 0 : ng); ++j) {
			const int i = j * 32 + lane;
			bool ok = true, tr = false;
			if (i < nkm) {
				const uint32_t f3 = (uint32_t)(W.key[i] >> THRF_F3_SHIFT) & 3u;
				const uint32_t pb = W.nb[i];                                 // node of position i-1 (i == 0: unused)
				ok = i == 0 ? (W.nb[1] & 1u) != 0 : ((pb & 1u) && ((pb >> (1 + f3)) & 1u));
				tr = W.trx[i] != 0;
			}
			const uint32_t bo = __ballot_sync(0xffffffffu, ok);
			if (bo != 0xffffffffu) clean = false;
			const uint32_t bt = __ballot_sync(0xffffffffu, tr);
			const uint32_t bp = __ballot_sync(0xffffffffu, i < nkm && (W.nb[i + 1] & 1u));
			if (lane == 0) { W.T[j] = bt; W.ok[j] = bo; W.P[j] = bp; }
		}
		__syncwarp();
		int es_beg = 0, tr_beg = 0;                                       // ... and skipped k-mers before the anchor leave the entries before it '*'
		int es_end = (int)len, tr_end = nkm;                              // a walk that ends early (cpp:1182-1185) leaves the entries from here on '*'
		uint64_t xlist = 0; int nfix = 0;                                 // corrected mate: 'X' entries of the CIGAR, 16 bits each: position | graph base (1-4) << 13
		if (!clean && correction1) {
			// ---- substitutions in the interior of the read, graph without a branch there ----
			// The walk (cpp:1167-1180) stops at the first k-mer ki that is not on an out-edge of its predecessor.  If that
			// predecessor has ONE out-edge a, the k-mer with its last base replaced by a is a node whose one out-edge is the
			// next read base, and the k following k-mers with that base replaced (the last of them is the read's own k-mer
			// ki+k) form a path, then errorCorrection_forward (cpp:898-1089) scores the 1-mismatch edit `a` with the full
			// extension k: no other edit can score more (every family's walk is at most k steps long here) and this one comes
			// first in thread_ext_t::get_edit's scan (cpp:628-647).  edit_kmers_forward (cpp:828-862) rewrites exactly those
			// k k-mers and one CIGAR entry, and the walk resumes behind them: the next stop, if any, is handled the same way
			// while corrections (cpp:1186) and skips (cpp:1194) are left.  If every other k-mer is on an out-edge of its
			// predecessor the read threads with nfix corrections (isThreadFeasible returns 2).  Anything else -- an error
			// near an end, an indel, a branch, a missing node -- is thread_kernel's, which starts over from the packed read.
			const uint32_t maxskip = (uint32_t)nkm > rp.thread_cth ? (uint32_t)nkm - rp.thread_cth : 0u;
			int from = 0; bool fail = false;
			int nskip = 0;                                                    // k-mers skipped so far (cpp:1123); every correction adds its edit (cpp:1193)
			// Start of the read.  The walk begins at the first k-mer `an` that is a node (find_anchor, cpp:878-888); the an k-mers
			// before it are skipped, and when there are at least 6 of them errorCorrection_backward (cpp:1091-1106: the same
			// scorer on the reverse complements, from the anchor towards the start) is tried once (cpp:1131-1141).  A wrong base
			// in the first k bases makes exactly the k-mers that hold it non-nodes: an <= k, the mirror image of a stop with
			// an k-mers left -- one 1-mismatch walk of an-1 steps on the reverse strand, the deletion families out of play under
			// the same conditions -- and edit_kmers_backward (cpp:649-826) then rewrites k-mers 0..an-1, marks CIGAR entry an-1
			// 'X' with the complement of the edge taken and the entries before it '=' (its merge rules see a tract of one entry).
			{
				int an = nkm;
#pragma unroll 1
				for (int j = ng - 1; j >= 0; --j) { const uint32_t z = W.P[j]; if (z) an = j * 32 + __ffs(z) - 1; }
				if (an >= nkm) fail = true;                                   // no node at all: the read does not thread (cpp:1129)
				else if (an > 0) {
					if (an >= (int)THR_MSC + 1) {
						const uint32_t nbw = W.nb[an + 1];
						const uint32_t rm = (nbw >> 6) & 0xFu;                     // out-edges of the anchor's reverse complement
						if (an > (int)K || !(nbw & GN_RCPRESENT) || __popc(rm) != 1 || rp.max_ncorrection < 1) fail = true;
						else {
							const uint32_t tb = 3u - ((uint32_t)__ffs(rm) - 1u);      // the base the graph has where the read is wrong (forward strand)
							const int t = lane;                                       // lane t < an: k-mer an-1-t, wrong base = its base number t
							uint64_t c = 0; uint32_t bits = 0, trp = 0, fb = 0;
							if (t < an) {
								const uint64_t kw = W.key[an - 1 - t];
								const uint64_t cn = kw & kmask;
								const uint64_t fw = (kw & THRF_RCFLAG) ? kmer_rc(cn, K) : cn;
								const uint32_t shf = 2 * (K - 1 - (uint32_t)t);
								c = (fw & ~(3ull << shf)) | ((uint64_t)tb << shf);
								const GNode g = g_lookup_fast_nl<B>(gt, c, dst, K);
								bits = g.bits; trp = g.tr;
								++C.probe;
								fb = (uint32_t)(c >> (2 * (K - 1))) & 3u;               // first base of the rewritten k-mer
							}
							const uint32_t fbn = __shfl_down_sync(0xffffffffu, fb, 1);  // first base of k-mer an-2-t: the next step of the reverse-strand walk appends its complement
							bool good = true;
							if (t + 1 < an) {
								const uint32_t mk = gn_rcmask(bits);
								good = (bits & GN_RCPRESENT) && ((mk >> (3u - fbn)) & 1u) && (t != 0 || __popc(mk) == 1);
							}
							{
								const uint64_t kw = W.key[an - 1];
								const uint64_t fw0 = (kw & THRF_RCFLAG) ? kmer_rc(kw & kmask, K) : (kw & kmask);
								const uint32_t oldnt = 3u - ((uint32_t)(fw0 >> (2 * (K - 1))) & 3u);   // the wrong base as the reverse-strand walk reads it
								const uint32_t m0 = __shfl_sync(0xffffffffu, gn_rcmask(bits), 0), m1 = __shfl_sync(0xffffffffu, gn_rcmask(bits), 1);
								if (((m0 >> oldnt) & 1u) || (an >= 2 * (int)THR_MSC + 2 && ((m1 >> oldnt) & 1u))) good = false;
							}
							if (!__all_sync(0xffffffffu, good)) fail = true;
							else {
								__syncwarp();
								if (t < an) {
									const uint64_t rc = kmer_rc(c, K);
									const bool is_rc = rc < c;
									W.key[an - 1 - t] = (is_rc ? rc : c) | ((c & 3) << THRF_F3_SHIFT) | (is_rc ? THRF_RCFLAG : 0ull);
									W.trx[an - 1 - t] = trp;
								}
								__syncwarp();
								xlist |= (uint64_t)((uint32_t)(an - 1) | ((tb + 1) << 13)) << (16 * nfix);
								++nfix; nskip = 1;                                      // cpp:690-694: the skipped k-mers are taken back, the edit stays
								if (lane == 0) C.ec += (uint32_t)an + 1;
							}
						}
					} else { nskip = an; es_beg = an; tr_beg = an; }            // too few k-mers to correct (cpp:1131): they stay skipped
					from = an + 1;
				}
			}
			// Stops near the end of the read.  With nrem = nkm - ki k-mers left: nrem <= 5 ends the walk (cpp:1182-1185: the rest is
			// skipped, no correction); nrem = k+1 is the interior case (every family's walk is still at most k steps); 6 <= nrem <= k:
			// the 1-mismatch walk is nrem-1 steps long and only the two deletion families walk one step further (jb = 0, cpp:961,
			// 1040) -- they are not tried at all unless the wrong base is an out-edge of the corrected k-mer (1 deletion, cpp:957)
			// or of its successor (2 deletions, cpp:1036, needs >= 12 k-mers left), and then the pair is left to thread_kernel.
#pragma unroll 1
			for (; !fail;) {
				int ki = nkm;
#pragma unroll 1
				for (int j = ng - 1; j >= (from >> 5); --j) {
					uint32_t z = ~W.ok[j];
					if (j == (from >> 5)) z &= 0xFFFFFFFFu << (from & 31);
					if (z) ki = j * 32 + __ffs(z) - 1;
				}
				if (ki >= nkm) break;                                         // nothing stops the walk any more
				const int nrem = nkm - ki;
				if (ki >= 1 && nrem <= (int)THR_MSC) {                           // cpp:1182-1185
					nskip += nrem;
					es_end = ki + (int)K - 1; tr_end = ki;
					break;
				}
				if (nfix >= 4 || (uint32_t)nfix >= rp.max_ncorrection || (uint32_t)nskip + 1 > maxskip || ki < 1) { fail = true; break; }
				const int nrw = min(nrem, (int)K);                            // k-mers the edit rewrites: ki .. ki+nrw-1
				const uint32_t pm = (uint32_t)(W.nb[ki] >> 1) & 0xFu;        // out-edges of k-mer ki-1 (never a rewritten one: ki-1 >= from-1)
				if (__popc(pm) != 1) { fail = true; break; }
				const uint32_t a = (uint32_t)__ffs(pm) - 1u;
				const int t = lane;                                           // lane t < k: k-mer ki+t with the wrong base (2t bits up) replaced
				uint64_t c = 0; uint32_t bits = 0, trp = 0, f3n = 0;
				if (t < nrw) {
					const uint64_t kw = W.key[ki + t];
					const uint64_t cn = kw & kmask;
					const uint64_t fw = (kw & THRF_RCFLAG) ? kmer_rc(cn, K) : cn;
					c = (fw & ~(3ull << (2 * t))) | ((uint64_t)a << (2 * t));
					const GNode g = g_lookup_fast_nl<B>(gt, c, dst, K);
					bits = g.bits; trp = g.tr;
					++C.probe;
					if (ki + t + 1 < nkm) f3n = (uint32_t)(W.key[ki + t + 1] >> THRF_F3_SHIFT) & 3u;   // last base of the next k-mer
				}
				// lane t (< k): present, and the next k-mer's last base is an out-edge (lane 0: its ONLY out-edge, the level-2
				// expansion of the scorer then holds just that base, cpp:907-920)
				bool good = true;
				if (t < nrw && ki + t + 1 < nkm) {                             // (the last k-mer of the read is only visited, cpp:553)
					const uint32_t mk = gn_mask(bits);
					good = (bits & GN_PRESENT) && ((mk >> f3n) & 1u) && (t != 0 || __popc(mk) == 1);
				}
				if (nrem <= (int)K) {                                         // deletion families must not be in play (see above)
					const uint32_t oldnt = (uint32_t)(W.key[ki] >> THRF_F3_SHIFT) & 3u;
					const uint32_t m0 = __shfl_sync(0xffffffffu, gn_mask(bits), 0), m1 = __shfl_sync(0xffffffffu, gn_mask(bits), 1);
					if (((m0 >> oldnt) & 1u) || (nrem >= 2 * (int)THR_MSC + 2 && ((m1 >> oldnt) & 1u))) good = false;
				}
				const bool allgood = __all_sync(0xffffffffu, good);
				if (!allgood) { fail = true; break; }
				__syncwarp();
				if (t < nrw) {
					const uint64_t rc = kmer_rc(c, K);
					const bool is_rc = rc < c;
					W.key[ki + t] = (is_rc ? rc : c) | ((c & 3) << THRF_F3_SHIFT) | (is_rc ? THRF_RCFLAG : 0ull);
					W.trx[ki + t] = trp;
				}
				__syncwarp();
				xlist |= (uint64_t)((uint32_t)(ki + (int)K - 1) | ((a + 1) << 13)) << (16 * nfix);
				++nfix; ++nskip;
				from = ki + (int)K + 1;                                       // the walk resumes here (cpp:1197: ki += score - 1, then the loop's ++ki)
				if (lane == 0) C.ec += K + 2;                                 // look-ups of the scorer's winning path: 1 + 1 expansion nodes + k walk steps
			}
			if ((uint32_t)nskip > maxskip) fail = true;                       // cpp:1259: the read does not thread; thread_kernel reports it
			if (!fail && (nfix || tr_end < nkm || tr_beg > 0)) {
#pragma unroll 1
				for (int j = 0; j < ng; ++j) {                                 // TR ballots again: the rewritten k-mers have their own
					const int i = j * 32 + lane;
					const uint32_t bt = __ballot_sync(0xffffffffu, i < nkm && W.trx[i] != 0);
					if (lane == 0) W.T[j] = bt;
				}
				__syncwarp();
				clean = true;
				if (lane == 0) ++C.fix;
			}
		}
		if (!clean) return false;

		// ---- clean mate: alned = 1, es all '=', tr from the TR ballots; count its TR k-mers (D1) ----
		{
			uint32_t inc_sum = 0;
#pragma unroll 1
			for (int j = 0; j < ng; ++j) {
				const int i = j * 32 + lane;
				const uint32_t tx = i < nkm ? W.trx[i] : 0u;
				const uint32_t tm = W.T[j];
				if (!tm) continue;
				inc_sum += __popc(tm);
				if (tx) {
					const uint32_t peers = __match_any_sync(tm, tx);
					if ((uint32_t)lane == (uint32_t)(__ffs(peers) - 1))
						atomicAdd(reinterpret_cast<unsigned long long*>(ix.counts + (tx - 1)), (unsigned long long)__popc(peers));
				}
			}
			if (lane == 0) { C.inc += inc_sum; C.cls += (uint32_t)nkm; }
		}
		if (slot_ok) {                                                   // sam_t.r1 / r2 -> dbt_aln_mate blob
			uint8_t* b = rec + sizeof(dbt_pair_record) + (size_t)m * rp.aln_stride;
			uint8_t* es = b + sizeof(dbt_aln_mate); uint8_t* trb = es + rp.es_cap;
			const int nes = min((int)len, (int)rp.es_cap), ntr = min(nkm, (int)rp.tr_cap);
#pragma unroll 1
			for (int q = lane; q < (int)rp.es_cap / 4; q += 32) {         // four entries per lane and store; entries >= nes read 0
				const int nbb = min(nes, es_end) - 4 * q, nlo = es_beg - 4 * q;
				const uint32_t keep = (nbb >= 4 ? 0xFFFFFFFFu : (nbb <= 0 ? 0u : (1u << 8 * nbb) - 1u)) & (nlo <= 0 ? 0xFFFFFFFFu : (nlo >= 4 ? 0u : 0xFFFFFFFFu << 8 * nlo));
				uint32_t v = 0x01010101u * ET_EQ;
#pragma unroll 1
				for (int f = 0; f < nfix; ++f) {
					const uint32_t e = (uint32_t)(xlist >> (16 * f)) & 0xFFFFu, xp = e & 0x1FFFu, xb = e >> 13;
					if ((int)(xp >> 2) == q) v = (v & ~(0xFFu << (8 * (xp & 3)))) | ((uint32_t)(ET_X | (xb << 3)) << (8 * (xp & 3)));
				}
				reinterpret_cast<uint32_t*>(es)[q] = v & keep;
			}
#pragma unroll 1
			for (int q = lane; q < (int)rp.tr_cap / 4; q += 32) {         // tr[i] = '=' (2) for TR k-mers, '.' (1) otherwise: 4 TR ballot bits -> 4 bytes
				const int nbb = min(ntr, tr_end) - 4 * q, nlo = tr_beg - 4 * q;
				const uint32_t keep = (nbb >= 4 ? 0xFFFFFFFFu : (nbb <= 0 ? 0u : (1u << 8 * nbb) - 1u)) & (nlo <= 0 ? 0xFFFFFFFFu : (nlo >= 4 ? 0u : 0xFFFFFFFFu << 8 * nlo));
				const uint32_t bits = 4 * q < ng * 32 ? (W.T[(4 * q) >> 5] >> ((4 * q) & 31)) & 0xFu : 0u;
				const uint32_t v = 0x01010101u + ((bits & 1u) | ((bits & 2u) << 7) | ((bits & 4u) << 14) | ((bits & 8u) << 21));
				reinterpret_cast<uint32_t*>(trb)[q] = v & keep;
			}
			if (lane == 0) {
				dbt_aln_mate* h = reinterpret_cast<dbt_aln_mate*>(b);
				h->alned = nfix ? 2 : 1; h->err = (uint8_t)((int)len > (int)rp.es_cap || nkm > (int)rp.tr_cap);   // cpp:1259 / 1184
				h->nes = (uint16_t)len; h->ntr = (uint16_t)nkm; h->nkm = (uint16_t)nkm;
				uint64_t hash = 0;
				if (trace) {                                                // FNV-1a over the k-mer vector (parity tests only)
					hash = 0xcbf29ce484222325ull;
#pragma unroll 1
					for (int i = 0; i < nkm; ++i) {
						const uint64_t kw = W.key[i];
						const uint64_t v = (kw & THRF_RCFLAG) ? kmer_rc(kw & kmask, K) : (kw & kmask);
						for (int q = 0; q < 8; ++q) { hash ^= (v >> (8 * q)) & 0xFF; hash *= 0x100000001b3ull; }
					}
				}
				h->kmers_hash = hash;
			}
		}
	
	return true;
}

// pair level: a clean (or corrected) mate makes the pair thread (commented caller, cpp:2082-2087); mates left over go to thread_kernel
__device__ __forceinline__ void thrf_finish_pair(const RunParams& rp, const SlotView& sv, uint32_t c, uint32_t p, uint32_t dw, uint32_t dst, uint32_t dirty,
                                                 uint8_t* rec, bool slot_ok, bool want_rec, bool trace, int lane, uint32_t& c_fea, uint32_t& c_items) {
	if (dirty != 3u) {
		if (lane == 0) c_fea += 2;
		if (want_rec && !slot_ok) { if (lane == 0) atomicOr(sv.ncand + 1, ERR_RECORD_OVERFLOW); }
		else if (slot_ok && lane == 0) {
			dbt_pair_record* h = reinterpret_cast<dbt_pair_record*>(rec);
			if (!trace) {                                                   // under DBT_F_TRACE the generic candidate kernel left the full header
				h->pair_index = sv.first_pair_index + p; h->dst0 = dst; h->fc = 0; h->rc = 0; h->nhash0 = 0; h->nhash1 = 0;
				h->pad[0] = h->pad[1] = h->pad[2] = 0;
#ifdef DBT_DEBUG_DIRTY
				h->pad[0] = (uint8_t)(0x80u | (dirty & 3u));      // development aid (tools/dirty_breakdown.py): which mates go on to thread_kernel
#endif
#pragma unroll 1
				for (int m = 0; m < 2; ++m) {
					dbt_mate_record& o = h->mate[m];
					o.si = o.ei = o.si_ = o.ei_ = o.ti = -1; o.nt = 0; o.bs = 0; o.pad = 0; o.nk = 0;
					o.flags = (uint8_t)((((dw >> (28 + m)) & 1u) ? DBT_MF_KF : 0) | (((dw >> (30 + m)) & 1u) ? DBT_MF_RM : 0));
				}
			}
			h->dst = dst; h->status = (uint8_t)DBT_ST_COUNTED;
		}
	}
	if (dirty && lane == 0) { sv.thr2[atomicAdd(sv.ncand + 4, 1u)] = make_uint4(c, p, dw, dirty | (dirty != 3u ? THR2_ALNED : 0u)); ++c_items; }
}

template <int B, int MAXJ>
struct FastThreadWarp {
	uint4    stage[MAXJ][B][32];               // home (then stash) buckets of every position of the mate being probed
	uint64_t key[MAXJ * 32];                   // canonical k-mer | last base << 60 | THRF_RCFLAG; KMER_INVALID for an invalid window
	uint32_t trx[MAXJ * 32];                   // count-array position + 1 of TR k-mers, else 0
	uint16_t nb[MAXJ * 32 + 32];               // nb[i + 1] = GNode bits 0-9 of position i: present | out-edge mask << 1 | the same of the reverse strand << 5 (nb[0]: unused)
	uint64_t pk[2 * MAXJ + 6];                 // packed record of the mate (W64 <= 1.5 MAXJ + 4 for reads of <= MAXJ*32 k-mers)
	uint32_t T[MAXJ];                          // TR ballots
	uint32_t ok[MAXJ];                         // "on an out-edge of its predecessor" ballots (position 0: is a node)
	uint32_t P[MAXJ];                          // "is a node" ballots
};

template <int B, int MAXJ>
__global__ void __launch_bounds__(THRF_WARPS * 32, THRF_MINBLOCKS) thread_fast_kernel(DevIndex ix, RunParams rp, SlotView sv) {
	extern __shared__ __align__(16) uint8_t thrf_smem[];
	FastThreadWarp<B, MAXJ>& W = reinterpret_cast<FastThreadWarp<B, MAXJ>*>(thrf_smem)[threadIdx.x >> 5];
	const int lane = threadIdx.x & 31;
	const uint32_t gwarp = blockIdx.x * THRF_WARPS + (threadIdx.x >> 5), nwarps = gridDim.x * THRF_WARPS;
	const uint32_t nwork = sv.ncand[3];
	const bool trace = (rp.flags & DBT_F_TRACE) != 0;
	const bool aln_all = (rp.flags & DBT_F_ALN) && !(rp.flags & DBT_F_ALN_MIN), aln_min = (rp.flags & DBT_F_ALN_MIN) != 0;
	const bool want_rec = trace || aln_all || aln_min;
	const uint32_t W64 = rp.lay.W64, WB = rp.lay.WB, K = rp.k;
	const Table gt = ix.gtable;
	const uint64_t kmask = (1ull << (2 * K)) - 1, vmask = (1ull << K) - 1;
	uint32_t c_fea = 0, c_walk = 0, c_items = 0;                         // lane-0 accumulators
	ThrfCount C;                                                         // (C.probe: per lane, reduced at the end)

	uint4 item = gwarp < nwork ? sv.thr[gwarp] : make_uint4(0, 0, 0, 0);
#pragma unroll 1
	for (uint32_t t = gwarp; t < nwork; t += nwarps) {
		const uint32_t c = item.x, p = item.y, dw = item.z;
		if (t + nwarps < nwork) item = sv.thr[t + nwarps];                   // next work item: in flight while this pair is probed
		const uint32_t dst = dw & G_LOCUS_MASK;
		const uint64_t lmix = (uint64_t)dst * 0xC2B2AE3D27D4EB4Full;         // g_hash_key
		const uint64_t slot = trace ? (uint64_t)p : (uint64_t)c;
		const bool slot_ok = want_rec && (trace || slot < rp.max_records);
		uint8_t* rec = sv.records + slot * rp.rec_stride;
		uint32_t dirty = 0;                                                  // bit m: mate m goes to thread_kernel

#pragma unroll 1
		for (int m = 0; m < 2; ++m) {
			const uint32_t len = (m ? sv.meta[1] : sv.meta[0])[p] & META_LENMASK;           // (no dynamic index into the parameter struct: it would be copied to local memory)
			const int nkm = (int)len - (int)K + 1;
			__syncwarp();
			for (uint32_t q = lane; q < W64 && q < 2 * MAXJ + 6; q += 32) W.pk[q] = (m ? sv.packed[1] : sv.packed[0])[(uint64_t)p * W64 + q];
			__syncwarp();
			if (nkm < 1 || nkm > MAXJ * 32 || W64 > 2 * MAXJ + 6) { dirty |= 1u << m; continue; }   // thread_kernel handles (and reports) the odd ones
			const uint32_t* mk32 = reinterpret_cast<const uint32_t*>(W.pk + WB);
			const uint32_t sh = 2 * lane;
			const int ng = (nkm + 31) >> 5;
			// ---- A: k-mers + home-bucket copies ----
			bool plain = true, anyvalid = false;                               // plain: every window valid, none equal to its successor
#pragma unroll 1
			for (int j = 0; j < ng; ++j) {
				const int i = j * 32 + lane;
				const uint64_t wlo = W.pk[j], whi = (j + 1 < (int)WB) ? W.pk[j + 1] : 0ull;
				const uint32_t mlo = mk32[j], mnx = ((uint32_t)(j + 1) < 2 * rp.lay.WM) ? mk32[j + 1] : 0xFFFFFFFFu;
				const uint64_t wv = sh ? ((wlo >> sh) | (whi << (64 - sh))) : wlo;
				const uint64_t mm = (((uint64_t)mnx << 32) | mlo) >> lane;
				uint64_t kk = KMER_INVALID;
				if (i < nkm) {
					if ((mm & vmask) == 0) {
						const uint64_t x = wv & kmask, rcx = (~wv) & kmask, fw = rev2(x) >> (64 - 2 * K);   // read2kmers(canonical = false), h:246-271
						const bool is_rc = rcx < fw;
						const uint64_t canon = is_rc ? rcx : fw;
						kk = canon | ((fw & 3) << THRF_F3_SHIFT) | (is_rc ? THRF_RCFLAG : 0ull);
						const uint4* bp = gt.slots + home_bucket(gt, canon ^ lmix) * B;
#pragma unroll
						for (int q = 0; q < B; ++q) cp_async16(&W.stage[j][q][lane], bp + q);
						++C.probe;
						anyvalid = true;
						if (i + 1 < nkm && ((wv >> 2) & kmask) == x) plain = false;   // kmers[i+1] == kmers[i]: cpp:1152 is thread_kernel's
					} else plain = false;
				}
				W.key[i] = kk;
			}
			cp_async_wait_all();
			__syncwarp();
			anyvalid = __any_sync(0xffffffffu, anyvalid);
			plain = __all_sync(0xffffffffu, plain);
			if (lane == 0 && anyvalid) c_walk += (uint32_t)nkm;               // the reference's walk: one node look-up per k-mer (cpp:1170)
			if (!plain) { dirty |= 1u << m; continue; }
			// ---- B1: match the home buckets ----
			uint32_t need = 0;                                               // bit j: position j*32+lane continues in the stash
#pragma unroll 1
			for (int j = 0; j < ng; ++j) {
				const int i = j * 32 + lane;
				uint32_t bi = 0, trp = 0;
				if (i < nkm) {
					const uint64_t kw = W.key[i];
					const uint64_t canon = kw & kmask;
					const bool is_rc = (kw & THRF_RCFLAG) != 0;
					bool full = true, found = false; uint64_t k0 = 0;
#pragma unroll
					for (int q = 0; q < B; ++q) {
						const uint4 sl = W.stage[j][q][lane];
						const uint64_t kk = slot_key(sl);
						if (q == 0) k0 = kk;
						if (kk == SLOT_EMPTY) full = false;
						else if ((kk & SLOT_KEYMASK) == canon && (sl.z & G_LOCUS_MASK) == dst) { const GNode n = g_decode(sl.z, sl.w, is_rc); bi = n.bits; trp = n.tr; found = true; }
					}
					if (!found && full && (k0 & SLOT_OVF)) need |= 1u << j;
				}
				W.nb[i + 1] = (uint16_t)(bi & 0x3FFu); W.trx[i] = trp;
			}
			if (__any_sync(0xffffffffu, need != 0)) {
				// second round: the first stash bucket of the few positions (7 % of the keys) whose home bucket overflowed.  They
				// are compacted into a list first, so that the round is ONE pass of 32 lanes instead of a pass per 32-position
				// group in which two or three lanes have work (those passes were a sixth of the kernel's instructions).
				__syncwarp();
				uint16_t* list = reinterpret_cast<uint16_t*>(&W.ok[0]);       // (W.ok / W.P are written by B2 below; 2 x MAXJ words hold 4 MAXJ entries)
				int nlist = 0;
#pragma unroll 1
				for (int j = 0; j < ng; ++j) {
					const bool nd = (need >> j) & 1u;
					const uint32_t bm = __ballot_sync(0xffffffffu, nd);
					if (nd) { const int o = nlist + __popc(bm & ((1u << lane) - 1u)); if (o < 4 * MAXJ) list[o] = (uint16_t)(j * 32 + lane); }
					nlist += __popc(bm);
				}
				__syncwarp();
				if (nlist <= 4 * MAXJ && nlist <= 32) {
					const bool act = lane < nlist;
					const int i = act ? (int)list[lane] : 0;
					const uint64_t kw = act ? W.key[i] : 0ull;
					const uint64_t canon = kw & kmask;
					const uint64_t sb = stash_bucket(gt, canon ^ lmix);
					if (act) {
						const uint4* bp = gt.slots + sb * B;
#pragma unroll
						for (int q = 0; q < B; ++q) cp_async16(&W.stage[0][q][lane], bp + q);
						++C.probe;
					}
					cp_async_wait_all();
					if (act) {
						const bool is_rc = (kw & THRF_RCFLAG) != 0;
						uint32_t bi = 0, trp = 0;
						bool full = true, found = false; uint64_t k0 = 0;
#pragma unroll
						for (int q = 0; q < B; ++q) {
							const uint4 sl = W.stage[0][q][lane];
							const uint64_t kk = slot_key(sl);
							if (q == 0) k0 = kk;
							if (kk == SLOT_EMPTY) full = false;
							else if ((kk & SLOT_KEYMASK) == canon && (sl.z & G_LOCUS_MASK) == dst) { const GNode n = g_decode(sl.z, sl.w, is_rc); bi = n.bits; trp = n.tr; found = true; }
						}
						if (!found && full && (k0 & SLOT_OVF)) { const GNode n = g_lookup_canon_nl<B>(gt, canon, is_rc, dst, stash_next(gt, sb)); bi = n.bits; trp = n.tr; }
						W.nb[i + 1] = (uint16_t)(bi & 0x3FFu); W.trx[i] = trp;
					}
				} else {
					// more than 20 of them (never seen at load 0.4): the whole chain by plain look-ups
#pragma unroll 1
					for (int j = 0; j < ng; ++j) {
						if (!((need >> j) & 1u)) continue;
						const int i = j * 32 + lane;
						const uint64_t kw = W.key[i];
						const GNode n = g_lookup_canon_nl<B>(gt, kw & kmask, (kw & THRF_RCFLAG) != 0, dst, stash_bucket(gt, (kw & kmask) ^ lmix));
						W.nb[i + 1] = (uint16_t)(n.bits & 0x3FFu); W.trx[i] = n.tr;
						++C.probe;
					}
				}
			}
			__syncwarp();
			{
				const ThrfArrays A{W.key, W.trx, W.nb, W.T, W.ok, W.P};
				if (!thrf_finish_mate<B, MAXJ>(ix, rp, sv, A, m, len, nkm, dst, rec, slot_ok, trace, lane, C)) dirty |= 1u << m;
			}
		}

		thrf_finish_pair(rp, sv, c, p, dw, dst, dirty, rec, slot_ok, want_rec, trace, lane, c_fea, c_items);
	}
	const uint32_t c_probe = __reduce_add_sync(0xffffffffu, C.probe);
	const uint32_t c_inc = C.inc, c_cls = C.cls, c_fix = C.fix, c_ec = C.ec;
	if (lane == 0) {
		if (c_walk) atomicAdd(sv.counters + DBT_C_THR_WALK, (unsigned long long)c_walk);
		if (c_probe) atomicAdd(sv.counters + DBT_C_THR_REFRESH, (unsigned long long)c_probe);
		if (c_cls) atomicAdd(sv.counters + DBT_C_THR_CLS, (unsigned long long)c_cls);
		if (c_fea) atomicAdd(sv.counters + DBT_C_NFEASIBLEREADS, (unsigned long long)c_fea);
		if (c_inc) atomicAdd(sv.counters + DBT_C_NINC, (unsigned long long)c_inc);
		if (c_items) atomicAdd(sv.counters + DBT_C_THR2_PAIRS, (unsigned long long)c_items);
		if (c_fix) atomicAdd(sv.counters + DBT_C_THR1_FIXED, (unsigned long long)c_fix);
		if (c_ec) atomicAdd(sv.counters + DBT_C_THR_EC, (unsigned long long)c_ec);
	}
}

}  // namespace dbt
