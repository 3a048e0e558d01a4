// Hand-written sm_100a kernels of the `danbing-tk align` hot path (SURVEY.md §8(a), P0 = "asgn" mode).
//
//   pack_kernel        A1/A2   ASCII reads -> 2-bit packed records + validity masks
//   subfilter_kernel   A2+A3   sampled canonical k-mers + index probes, pass bitmap   (cpp:172-188)
//   compact_*          --      ordered stream compaction of the pass bitmap
//   candidate_kernel   A2,A4-A11  one warp per surviving pair: all k-mers, kfilter (cpp:190-245),
//                      countDupRemove/fillstats (cpp:257-329), the early-stopping locus vote
//                      (cpp:364-453), QC gate (cpp:2059), assignTRkmc (cpp:1450-1556), accumulation
//                      with integer atomics (cpp:2145-2158) and the kam record (cpp:2169-2175)
//   index build        A14     bulk insert + per-locus class fill
// "cpp" = src/aQueryFasta_thread.cpp, "h" = src/aQueryFasta_thread.h of the reference.
#include "device_types.cuh"
#include "introsort.cuh"

namespace dbt {

// ============================================================================================
// small helpers
// ============================================================================================
__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
	for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

__device__ __forceinline__ void counter_add(unsigned long long* counters, int which, uint32_t v) {
	if (v) atomicAdd(counters + which, (unsigned long long)v);
}

// zero-filled record with the defaults of km_asgn_read_t (cpp:94-95): si=ei=si_=ei_=ti=-1, nt=bs=0
__device__ void write_plain_record(uint8_t* rec, uint32_t stride, uint64_t pair_index, uint32_t status,
                                   uint32_t nhash0, uint32_t nloci) {
	uint4* q = reinterpret_cast<uint4*>(rec);
	for (uint32_t i = 0; i < stride / 16; ++i) q[i] = make_uint4(0, 0, 0, 0);
	dbt_pair_record* h = reinterpret_cast<dbt_pair_record*>(rec);
	h->pair_index = pair_index; h->dst = nloci; h->dst0 = NAN32; h->nhash0 = (uint16_t)nhash0; h->status = (uint8_t)status;
	for (int m = 0; m < 2; ++m) { h->mate[m].si = h->mate[m].ei = h->mate[m].si_ = h->mate[m].ei_ = h->mate[m].ti = -1; }
}

// ============================================================================================
// pack_kernel: one thread per read.  A1 (device half): 2 bits/base + invalid mask + "no valid window"
// ============================================================================================
__global__ void pack_kernel(const uint8_t* __restrict__ ascii, const uint64_t* __restrict__ offsets, uint64_t nreads,
                            uint32_t k, ReadLayout lay, uint32_t max_read_len, uint64_t npairs,
                            uint64_t* __restrict__ packed0, uint64_t* __restrict__ packed1,
                            uint16_t* __restrict__ meta0, uint16_t* __restrict__ meta1, uint32_t* __restrict__ err) {
	uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (r >= nreads) return;
	uint64_t pair = r >> 1; uint32_t mate = (uint32_t)(r & 1);
	uint64_t off = offsets[r]; uint64_t len64 = offsets[r + 1] - off;
	uint32_t len = (uint32_t)len64;
	if (len64 > max_read_len) { atomicOr(err, ERR_READ_TOO_LONG); len = max_read_len; }
	const uint8_t* s = ascii + off;
	uint64_t* out = (mate ? packed1 : packed0) + pair * lay.W64;
	uint32_t run = 0; bool any = false;
	for (uint32_t w = 0; w < lay.WB; w += 2) {                 // two base words + one mask word per round
		uint64_t b0 = 0, b1 = 0, m = 0;
#pragma unroll 4
		for (uint32_t j = 0; j < 64; ++j) {
			uint32_t i = w * 32 + j;
			uint32_t c = 4;
			if (i < len) c = base_code(__ldg(s + i));
			if (c < 4) { if (j < 32) b0 |= (uint64_t)c << (2 * j); else b1 |= (uint64_t)c << (2 * (j - 32)); ++run; if (run >= k) any = true; }
			else { m |= 1ull << j; run = 0; }
		}
		out[w] = b0;
		if (w + 1 < lay.WB) out[w + 1] = b1;
		out[lay.WB + (w >> 1)] = m;
	}
	for (uint32_t w = lay.WB + lay.WM; w < lay.W64; ++w) out[w] = 0;
	(mate ? meta1 : meta0)[pair] = (uint16_t)(len | (any ? 0 : META_NOVALID));
}

// ============================================================================================
// subfilter_kernel: one thread per pair (cpp:172-188), pass bitmap via ballot.
// ============================================================================================
// sampled-probe loop of one mate; returns pass, adds the reference's nhash increments to nh.
template <int B>
__device__ __forceinline__ bool subfilter_mate(const DevIndex& ix, const RunParams& rp, const uint64_t* rec, uint32_t len, uint32_t& nh) {
	const uint32_t L = len - rp.k + 1, N = rp.n_filter, M = rp.nm_filter;
	const uint32_t S = L / (N - 1);
	uint32_t h = 0;
	for (uint32_t i0 = 0; i0 < N; i0 += 4) {                     // 4 independent probes in flight
		uint64_t key[4]; bool hit[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			uint32_t i = i0 + u;
			key[u] = KMER_INVALID;
			if (i < N) { uint32_t idx = (i != N - 1) ? i * S : L - 1; key[u] = canonical_kmer_at(rec, rp.lay.WB, idx, rp.k); }
		}
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			uint32_t v, a;
			hit[u] = (key[u] != KMER_INVALID) && table_lookup<B>(ix.table, key[u], v, a);
		}
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			if (i0 + u < N) {
				h += hit[u];
				if (h >= M) return true;            // `break` before the loop header's ++nhash (cpp:176)
				++nh;
			}
		}
	}
	return h >= M;
}

template <int B>
__global__ void __launch_bounds__(256) subfilter_kernel(DevIndex ix, RunParams rp, SlotView sv) {
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
	const uint64_t wid = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const uint64_t nwords = (sv.npairs + 31) / 32;
	const bool trace = (rp.flags & DBT_F_TRACE) != 0;
	const bool filter_on = rp.n_filter && rp.nm_filter;
	uint32_t c_short = 0, c_sub = 0, c_h0 = 0;
	for (uint64_t w = wid; w < nwords; w += nwarps) {
		const uint64_t p = w * 32 + lane;
		bool pass = false;
		if (p < sv.npairs) {
			const uint16_t m1 = sv.meta[0][p], m2 = sv.meta[1][p];
			uint32_t nh = 0; int status = -1;
			if ((m1 | m2) & META_NOVALID) { status = DBT_ST_SHORT; ++c_short; }          // cpp:2037
			else if (filter_on) {
				bool ok = subfilter_mate<B>(ix, rp, sv.packed[0] + p * rp.lay.W64, m1 & META_LENMASK, nh);
				if (ok) ok = subfilter_mate<B>(ix, rp, sv.packed[1] + p * rp.lay.W64, m2 & META_LENMASK, nh);
				if (!ok) { status = DBT_ST_SUBFILTERED; ++c_sub; }
			}
			c_h0 += nh;
			pass = status < 0;
			if (trace) {
				uint8_t* rec = sv.records + p * (uint64_t)rp.rec_stride;
				if (!pass) write_plain_record(rec, rp.rec_stride, sv.first_pair_index + p, (uint32_t)status, nh, (uint32_t)ix.nloci);
				else reinterpret_cast<dbt_pair_record*>(rec)->nhash0 = (uint16_t)nh;   // candidate_kernel fills the rest
			}
		}
		uint32_t bits = __ballot_sync(0xffffffffu, pass);
		if (lane == 0) sv.passmask[w] = bits;
	}
	c_short = warp_sum(c_short); c_sub = warp_sum(c_sub); c_h0 = warp_sum(c_h0);
	if (lane == 0) {
		counter_add(sv.counters, DBT_C_NSHORT, c_short);
		counter_add(sv.counters, DBT_C_NSUBFILTERED, 2 * c_sub);
		counter_add(sv.counters, DBT_C_NHASH0, c_h0);
	}
}

// ============================================================================================
// ordered compaction of the pass bitmap: blocksum -> exclusive scan -> scatter
// ============================================================================================
constexpr int COMPACT_WORDS = 1024;      // bitmap words per block (32768 pairs)

__global__ void __launch_bounds__(256) compact_count_kernel(SlotView sv) {
	__shared__ uint32_t ssum[8];
	const uint64_t nwords = (sv.npairs + 31) / 32;
	uint32_t c = 0;
	for (int i = threadIdx.x; i < COMPACT_WORDS; i += 256) {
		uint64_t w = (uint64_t)blockIdx.x * COMPACT_WORDS + i;
		if (w < nwords) c += __popc(sv.passmask[w]);
	}
	c = warp_sum(c);
	if ((threadIdx.x & 31) == 0) ssum[threadIdx.x >> 5] = c;
	__syncthreads();
	if (threadIdx.x == 0) { uint32_t t = 0; for (int i = 0; i < 8; ++i) t += ssum[i]; sv.blocksum[blockIdx.x] = t; }
}

__global__ void __launch_bounds__(1024) compact_scan_kernel(SlotView sv, uint32_t nblocks) {
	// single block; nblocks is small (npairs / 32768)
	__shared__ uint32_t carry;
	__shared__ uint32_t wsum[32];
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	for (uint32_t base = 0; base < nblocks; base += 1024) {
		uint32_t i = base + threadIdx.x;
		uint32_t v = i < nblocks ? sv.blocksum[i] : 0, x = v;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((threadIdx.x & 31) >= o) x += y; }
		if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = x;
		__syncthreads();
		if (threadIdx.x < 32) {
			uint32_t s = wsum[threadIdx.x], t = s;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, t, o); if (threadIdx.x >= o) t += y; }
			wsum[threadIdx.x] = t - s;       // exclusive
		}
		__syncthreads();
		uint32_t excl = carry + wsum[threadIdx.x >> 5] + x - v;
		if (i < nblocks) sv.blocksum[i] = excl;
		__syncthreads();
		if (threadIdx.x == 1023) carry = excl + v;
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		sv.ncand[0] = carry;
		atomicAdd(sv.counters + DBT_C_NCANDIDATES, (unsigned long long)carry);
	}
}

__global__ void __launch_bounds__(256) compact_scatter_kernel(SlotView sv) {
	__shared__ uint32_t wbase[8];
	__shared__ uint32_t run;
	const uint64_t nwords = (sv.npairs + 31) / 32;
	if (threadIdx.x == 0) run = sv.blocksum[blockIdx.x];
	__syncthreads();
	for (int i0 = 0; i0 < COMPACT_WORDS; i0 += 256) {
		uint64_t w = (uint64_t)blockIdx.x * COMPACT_WORDS + i0 + threadIdx.x;
		uint32_t bits = w < nwords ? sv.passmask[w] : 0;
		uint32_t c = __popc(bits), x = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((threadIdx.x & 31) >= o) x += y; }
		if ((threadIdx.x & 31) == 31) wbase[threadIdx.x >> 5] = x;
		__syncthreads();
		uint32_t pre = run;
		for (int j = 0; j < (int)(threadIdx.x >> 5); ++j) pre += wbase[j];
		uint32_t o = pre + x - c;
		while (bits) { int b = __ffs(bits) - 1; bits &= bits - 1; sv.cand[o++] = (uint32_t)(w * 32 + b); }
		__syncthreads();
		if (threadIdx.x == 0) { uint32_t t = 0; for (int j = 0; j < 8; ++j) t += wbase[j]; run += t; }
		__syncthreads();
	}
}

// ============================================================================================
// candidate_kernel: one warp per pair that survived subfilter
// ============================================================================================
struct Asgn { uint32_t idx; uint64_t fc, rc; };               // asgn_t, cpp:146-149

__device__ __forceinline__ void updatetop2(uint64_t cf, uint32_t ind, uint64_t cr, Asgn& top, Asgn& second) {   // cpp:331-347
	if (cf + cr > top.fc + top.rc) {
		if (top.idx != ind) { second = top; top.idx = ind; }
		top.fc = cf; top.rc = cr;
	} else if (cf + cr > second.fc + second.rc) {
		second.idx = ind; second.fc = cf; second.rc = cr;
	}
}

// per-pair hits1/hits2 (cpp:432-434): small shared-memory table, falling back to a per-warp
// epoch-tagged global array when a pair touches more than VOTE_SMEM_SLOTS loci
struct VoteHits {
	uint32_t* lkey; uint32_t* lf; uint32_t* lr;    // shared memory, VOTE_SMEM_SLOTS entries, key = locus+1
	uint64_t* scratch; uint32_t epoch; uint32_t used;
	__device__ __forceinline__ void add(uint32_t locus, uint32_t df, uint32_t dr, uint32_t& f, uint32_t& r) {
		uint32_t h = (locus * 2654435761u) >> 26;            // 6 bits
		for (int probe = 0; probe < VOTE_SMEM_SLOTS; ++probe) {
			uint32_t s = (h + probe) & (VOTE_SMEM_SLOTS - 1);
			uint32_t kk = lkey[s];
			if (kk == locus + 1) { f = (lf[s] += df); r = (lr[s] += dr); return; }
			if (kk == 0) {
				if (used >= VOTE_SMEM_SLOTS - 8) break;          // keep the table sparse; spill to global
				lkey[s] = locus + 1; lf[s] = df; lr[s] = dr; ++used; f = df; r = dr; return;
			}
		}
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

struct MateOut { int si, ei, si_, ei_, ti, nt, bs, af, rm; };

// assignTRkmc state machine (cpp:1477-1555) over the 0/1/2 annotation of one mate; one lane.
__device__ void assign_state_machine(const uint8_t* as, int nk, uint32_t ntr8, uint32_t max_nt, uint32_t nm_tr, MateOut& r) {
	int s = 0, s_ = 0, s__ = 0;
	int ti2 = -1, si1 = -1, ei1 = -1, si2 = -1, ei2 = -1;
	for (int i = 0; i < nk; ++i) {
		s = as[i];
		if (s && s__ && s != s__) {
			++r.nt;
			if ((uint32_t)r.nt > max_nt) { r.af = 1; r.rm = 1; return; }
			if (r.nt == 1) { r.ti = i; if (s_) { si1 = -1; ei1 = -1; } }
			else if (r.nt == 2) {
				if (r.bs == 2) { r.af = 1; r.rm = 1; return; }
				ti2 = i; if (s_) { si2 = -1; ei2 = -1; }
			}
		}
		if (!r.bs && s) r.bs = s;
		if (!s) {
			if (r.nt == 0) { if (!s_) ++ei1; else { si1 = i; ei1 = i + 1; } }
			if (r.nt == 1) { if (!s_) ++ei2; else { si2 = i; ei2 = i + 1; } }
		}
		s_ = s;
		if (s) s__ = s;
	}
	const int ti1 = r.ti;
	if (r.nt == 0) {
		if (r.bs != 2) { r.af = 1; r.rm = 1; return; }
		r.si = 0; r.ei = nk; r.si_ = 0; r.ei_ = nk;
	} else if (r.nt == 1) {
		if (r.bs == 1) { r.si = si1 >= 0 ? (si1 + ei1) / 2 : ti1; r.ei = nk; r.si_ = si1 >= 0 ? ei1 : ti1; r.ei_ = nk; }
		else           { r.si = 0; r.ei = si1 >= 0 ? (si1 + ei1) / 2 : ti1; r.si_ = 0; r.ei_ = si1 >= 0 ? si1 : ti1; }
	} else {
		if (ntr8 < nm_tr) { r.af = 1; r.rm = 1; return; }
		r.si = (si1 >= 0 ? (si1 + ei1) / 2 : ti1);
		r.ei = (si2 >= 0 ? (si2 + ei2) / 2 : ti2);
		r.si_ = ei1 >= 0 ? ei1 : ti1;
		r.ei_ = si2 >= 0 ? si2 : ti2;
	}
}

template <int B, int MAXJ>
__global__ void __launch_bounds__(CAND_WARPS_PER_BLOCK * 32) candidate_kernel(DevIndex ix, RunParams rp, SlotView sv) {
	extern __shared__ __align__(16) uint8_t smem_raw[];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t gwarp = blockIdx.x * CAND_WARPS_PER_BLOCK + wib;
	const uint32_t nwarps = gridDim.x * CAND_WARPS_PER_BLOCK;
	const uint32_t W64 = rp.lay.W64, WB = rp.lay.WB, K = rp.k, NKP = rp.nkp, UMAX = rp.umax, SORTN = rp.sortn;
	const bool okam = (rp.flags & DBT_F_KAM) != 0, trace = (rp.flags & DBT_F_TRACE) != 0;

	// ---- shared-memory carve-up (per warp) ----
	uint8_t* ws = smem_raw + (size_t)wib * rp.warp_smem;
	uint64_t* pk    = reinterpret_cast<uint64_t*>(ws);                  // [2][W64] packed records
	uint64_t* sortk = pk + 2 * W64;                                     // [SORTN]  key<<10 | mate<<9 | pos
	uint32_t* pval  = reinterpret_cast<uint32_t*>(sortk + SORTN);       // [2][NKP] index value per position
	uint32_t* paux  = pval + 2 * NKP;                                   // [2][NKP] aux word or AUX_MISS
	uint32_t* uval  = paux + 2 * NKP;                                   // [UMAX]   unique k-mers: value
	uint32_t* uaux  = uval + UMAX;
	uint32_t* unml  = uaux + UMAX;                                      // number of mapped loci (cpp:314-317)
	uint32_t* lkey  = unml + UMAX;                                      // vote table
	uint32_t* lf    = lkey + VOTE_SMEM_SLOTS;
	uint32_t* lr    = lf + VOTE_SMEM_SLOTS;
	uint16_t* utf   = reinterpret_cast<uint16_t*>(lr + VOTE_SMEM_SLOTS);// true multiplicity in mate 0 / mate 1
	uint16_t* utr   = utf + UMAX;
	uint16_t* ustart = utr + UMAX;                                      // [UMAX+2] group start in the sorted array
	uint16_t* umid  = ustart + UMAX + 2;                                // first mate-1 element of the group
	uint16_t* perm  = umid + UMAX;                                      // visiting order (cpp:320-327)
	uint8_t*  as    = reinterpret_cast<uint8_t*>(perm + UMAX);          // [2][NKP] 0 unknown / 1 flank / 2 TR

	uint32_t c_kf = 0, c_laf = 0, c_qc = 0, c_thr = 0, c_asgn = 0, c_h1 = 0, c_inc = 0;   // lane-0 accumulators
	uint32_t epoch = ix.vote_epoch[gwarp];
	uint64_t* scratch = ix.vote_scratch + (uint64_t)gwarp * (ix.nloci + 1);
	const uint32_t ncand = sv.ncand[0];

	for (uint32_t c = gwarp; c < ncand; c += nwarps) {
		const uint32_t p = sv.cand[c];
		__syncwarp();
		// ---- 1. packed records -> shared ----
		for (uint32_t i = lane; i < 2 * W64; i += 32) {
			uint32_t m = i >= W64;
			pk[i] = sv.packed[m][(uint64_t)p * W64 + (i - m * W64)];
		}
		const uint32_t len0 = sv.meta[0][p] & META_LENMASK, len1 = sv.meta[1][p] & META_LENMASK;
		const uint32_t nk[2] = { len0 - K + 1, len1 - K + 1 };
		for (uint32_t i = lane; i < VOTE_SMEM_SLOTS; i += 32) lkey[i] = 0;
		__syncwarp();

		// ---- 2. all canonical k-mers + index probes (A2 + the lookups of kfilter, cpp:207/218) ----
		uint64_t key[2][MAXJ];
#pragma unroll
		for (int m = 0; m < 2; ++m) {
#pragma unroll
			for (int j = 0; j < MAXJ; ++j) {
				uint32_t i = j * 32 + lane;
				key[m][j] = (i < nk[m]) ? canonical_kmer_at(pk + m * W64, WB, i, K) : KMER_INVALID;
			}
#pragma unroll
			for (int j = 0; j < MAXJ; ++j) {
				uint32_t i = j * 32 + lane;
				uint32_t v = 0, a = AUX_MISS;
				if (key[m][j] != KMER_INVALID) { if (!table_lookup<B>(ix.table, key[m][j], v, a)) a = AUX_MISS; }
				if (i < NKP) { pval[m * NKP + i] = v; paux[m * NKP + i] = a; }
			}
		}
		__syncwarp();

		// ---- 3. kfilter (cpp:190-224): abort a mate at its (nk-cth+1)-th miss ----
		int kf[2], rm[2], hf[2] = {0, 0};
		kf[0] = nk[0] < rp.cth; kf[1] = nk[1] < rp.cth;
		rm[0] = kf[0]; rm[1] = kf[1];
		uint32_t nhash1 = 0;
		if (!(rm[0] && rm[1])) {
#pragma unroll
			for (int m = 0; m < 2; ++m) {
				if (rm[m]) continue;
				const uint32_t max_ns = nk[m] - rp.cth;
				uint32_t cum = 0, probes = nk[m]; bool aborted = false;
#pragma unroll
				for (int j = 0; j < MAXJ; ++j) {
					uint32_t i = j * 32 + lane;
					bool miss = (i < nk[m]) && (paux[m * NKP + i] == AUX_MISS);
					uint32_t mm = __ballot_sync(0xffffffffu, miss);
					if (!aborted && cum + __popc(mm) > max_ns) {
						aborted = true;
						probes = j * 32 + __fns(mm, 0, max_ns + 1 - cum) + 1;
					}
					cum += __popc(mm);
				}
				nhash1 += probes;
				kf[m] = aborted; rm[m] |= kf[m];
			}
		}
		if (lane == 0) { c_kf += kf[0] + kf[1]; c_h1 += nhash1; }

		int status = -1;
		uint32_t dst = (uint32_t)ix.nloci, dst0 = NAN32, fc = 0, rc = 0;
		MateOut mo[2];
#pragma unroll
		for (int m = 0; m < 2; ++m) { mo[m].si = mo[m].ei = mo[m].si_ = mo[m].ei_ = mo[m].ti = -1; mo[m].nt = 0; mo[m].bs = 0; mo[m].af = 0; mo[m].rm = 0; }
		uint32_t nas[2] = {0, 0};
		if (rm[0] && rm[1]) status = DBT_ST_KFILTERED;

		if (status < 0) {
			// ---- 4. hits of the surviving mates -> sort array (countDupRemove, cpp:263-266) ----
			uint32_t n = 0;
#pragma unroll
			for (int m = 0; m < 2; ++m) {
				if (rm[m]) continue;
#pragma unroll
				for (int j = 0; j < MAXJ; ++j) {
					uint32_t i = j * 32 + lane;
					bool hit = (i < nk[m]) && (paux[m * NKP + i] != AUX_MISS);
					uint32_t hm = __ballot_sync(0xffffffffu, hit);
					if (hit) sortk[n + __popc(hm & ((1u << lane) - 1))] = (key[m][j] << 10) | ((uint64_t)m << 9) | i;
					n += __popc(hm);
				}
			}
			__syncwarp();
			// ---- 5. bitonic sort, ascending, virtual +inf padding above n (flip formulation) ----
			uint32_t npow = 2; while (npow < n) npow <<= 1;
			for (uint32_t size = 2; size <= npow; size <<= 1) {
				for (uint32_t t = lane; t < npow / 2; t += 32) {            // flip step
					uint32_t blk = t / (size / 2), o = t % (size / 2);
					uint32_t a = blk * size + o, b = blk * size + size - 1 - o;
					if (b < n) { uint64_t x = sortk[a], y = sortk[b]; if (x > y) { sortk[a] = y; sortk[b] = x; } }
				}
				__syncwarp();
				for (uint32_t stride = size / 4; stride >= 1; stride >>= 1) {
					for (uint32_t t = lane; t < npow / 2; t += 32) {
						uint32_t a = 2 * stride * (t / stride) + (t % stride), b = a + stride;
						if (b < n) { uint64_t x = sortk[a], y = sortk[b]; if (x > y) { sortk[a] = y; sortk[b] = x; } }
					}
					__syncwarp();
				}
			}
			// ---- 6. collapse equal keys (cpp:278-295): U unique entries with per-mate multiplicities ----
			uint32_t U = 0;
			for (uint32_t base = 0; base < n; base += 32) {
				uint32_t i = base + lane;
				bool head = false, mid = false; uint64_t e = 0;
				if (i < n) {
					e = sortk[i];
					uint64_t prev = i ? sortk[i - 1] : 0;
					head = (i == 0) || ((prev >> 10) != (e >> 10));
					mid = ((e >> 9) & 1) && (head || !((prev >> 9) & 1));
				}
				uint32_t hm = __ballot_sync(0xffffffffu, head);
				uint32_t u = U + __popc(hm & ((2u << lane) - 1)) - 1;       // group of element i
				if (head) {
					ustart[u] = (uint16_t)i; umid[u] = 0xFFFF;
					uint32_t m = (uint32_t)(e >> 9) & 1, pos = (uint32_t)e & 511;
					uval[u] = pval[m * NKP + pos]; uaux[u] = paux[m * NKP + pos];
				}
				__syncwarp();
				if (mid) umid[u] = (uint16_t)i;
				U += __popc(hm);
			}
			if (lane == 0) ustart[U] = (uint16_t)n;
			__syncwarp();
			bool odd = false; uint32_t wsum8 = 0;
			for (uint32_t u = lane; u < U; u += 32) {
				uint32_t s = ustart[u], e = ustart[u + 1], md = umid[u] == 0xFFFF ? e : umid[u];
				utf[u] = (uint16_t)(md - s); utr[u] = (uint16_t)(e - md);
				wsum8 += ((md - s) & 0xFF) + ((e - md) & 0xFF);              // PE_KMC is uint8 (cpp:42)
				uint32_t v = uval[u];
				uint32_t nm = 1;
				if (v & 1) { nm = __ldg(ix.vv + (v >> 1)); odd = true; }
				unml[u] = nm;
			}
			const bool any_odd = __any_sync(0xffffffffu, odd);
			const uint32_t total = warp_sum(wsum8);                           // dupsum of countRemain (cpp:300)
			// ---- 7. visiting order = std::sort by number of mapped loci (cpp:320-327) ----
			if (!any_odd && U <= ix.permtab_maxU) {
				const uint16_t* pt = ix.permtab + (size_t)U * (U - 1) / 2;
				for (uint32_t u = lane; u < U; u += 32) perm[u] = pt[u];
			} else {
				for (uint32_t u = lane; u < U; u += 32) perm[u] = (uint16_t)u;
				__syncwarp();
				if (lane == 0) { IntroSort<uint16_t, uint32_t> s{perm, unml}; s.sort((int)U); }
			}
			__syncwarp();
			// ---- 8. find_matching_locus (cpp:364-422), one lane ----
			Asgn top = {NAN32, 0, 0}, second = {NAN32, 0, 0};
			if (lane == 0) {
				VoteHits vh{lkey, lf, lr, scratch, epoch + 1, 0};
				uint64_t rem = total;
				const uint64_t cth = rp.cth;
				for (uint32_t i = 0; i < U; ++i) {
					const uint32_t u = perm[i];
					const uint32_t vi = uval[u], df = utf[u] & 0xFF, dr = utr[u] & 0xFF;
					rem = (i + 1 < U) ? rem - df - dr : 0;                    // remain[i] (cpp:302-305; last entry stays 0)
					if (vi & 1) {
						uint32_t j0 = (vi >> 1) + 1, j1 = j0 + __ldg(ix.vv + (vi >> 1));
						for (; j0 < j1; ++j0) {
							uint32_t locus = __ldg(ix.vv + j0), f, r;
							vh.add(locus, df, dr, f, r);
							updatetop2(f, locus, r, top, second);
						}
					} else {
						uint32_t locus = vi >> 1, f, r;
						vh.add(locus, df, dr, f, r);
						updatetop2(f, locus, r, top, second);
					}
					if (!((top.fc + top.rc - second.fc - second.rc) < rem)) {        // not get_acm2 (cpp:383)
						uint32_t j = i; uint64_t remj = rem;
						while ((top.fc < cth && cth - top.fc <= remj) || (top.rc < cth && cth - top.rc <= remj)) {   // get_acm1
							if (++j >= U) break;
							const uint32_t uj = perm[j];
							const uint32_t vj = uval[uj], dfj = utf[uj] & 0xFF, drj = utr[uj] & 0xFF;
							remj = (j + 1 < U) ? remj - dfj - drj : 0;
							if (vj & 1) {
								uint32_t j0 = (vj >> 1) + 1, j1 = j0 + __ldg(ix.vv + (vj >> 1));
								for (; j0 < j1; ++j0) if (__ldg(ix.vv + j0) == top.idx) { top.fc += dfj; top.rc += drj; break; }
							} else if ((vj >> 1) == top.idx) { top.fc += dfj; top.rc += drj; }
						}
						break;
					}
				}
			}
			// the global fallback may have been written under epoch+1: retire that epoch
			epoch += 1;
			if (epoch >= 0xFFFFF0u) {                       // 24-bit tag about to wrap: clear this warp's scratch
				for (uint64_t i = lane; i < ix.nloci + 1; i += 32) scratch[i] = 0;
				epoch = 0;
			}
			dst0 = __shfl_sync(0xffffffffu, top.idx, 0);
			fc = (uint32_t)__shfl_sync(0xffffffffu, (uint32_t)top.fc, 0);
			rc = (uint32_t)__shfl_sync(0xffffffffu, (uint32_t)top.rc, 0);
			// ---- 9. countHit tail (cpp:436-452) ----
			const bool test1 = fc >= rp.cth && rc >= rp.cth, test2 = (fc + rc) >= 2 * rp.cth;
			if ((test1 || test2) && dst0 != NAN32) dst = dst0;
			else { hf[0] = 1 & !rm[0]; hf[1] = 1 & !rm[1]; rm[0] = rm[1] = 1; status = DBT_ST_UNASSIGNED; if (lane == 0) c_laf += hf[0] + hf[1]; }

			// ---- 10. QC gate (cpp:2059-2062) ----
			if (status < 0 && (rp.flags & DBT_F_QC) && ix.qc && !ix.qc[dst]) { status = DBT_ST_QCFILTERED; if (lane == 0) c_qc += 2 - rm[0] - rm[1]; }

			if (status < 0) {
				if (lane == 0) c_thr += 2;
				// ---- 11. assignTRkmc (cpp:1450-1556) against the tables of destLocus0 ----
				uint32_t ntr8[2] = {0, 0};
#pragma unroll
				for (int m = 0; m < 2; ++m) {
					if (!okam && rm[m]) continue;                           // cpp:1451
					nas[m] = nk[m];
					uint32_t cnt = 0;
					for (uint32_t i = lane; i < NKP; i += 32) {
						uint32_t cls = 0;
						if (i < nk[m]) cls = classify(ix, pval[m * NKP + i], paux[m * NKP + i], dst0) >> AUX_CLASS_SHIFT;
						as[m * NKP + i] = (uint8_t)cls;
						cnt += (cls == 2);
					}
					ntr8[m] = warp_sum(cnt) & 0xFF;                         // uint8_t ntr (cpp:1454)
				}
				__syncwarp();
				if (lane < 2) {
					const int m = lane;
					MateOut r = mo[m];
					r.rm = rm[m];
					if (!(!okam && rm[m])) {
						if (rm[m]) { r.nt = -1; r.bs = -1; r.ti = -1; }       // cpp:1470-1475
						else assign_state_machine(as + m * NKP, (int)nk[m], ntr8[m], rp.max_nt, rp.nm_tr, r);
					}
					mo[m] = r;
				}
#pragma unroll
				for (int m = 0; m < 2; ++m) {
					mo[m].si = __shfl_sync(0xffffffffu, mo[m].si, m); mo[m].ei = __shfl_sync(0xffffffffu, mo[m].ei, m);
					mo[m].si_ = __shfl_sync(0xffffffffu, mo[m].si_, m); mo[m].ei_ = __shfl_sync(0xffffffffu, mo[m].ei_, m);
					mo[m].ti = __shfl_sync(0xffffffffu, mo[m].ti, m); mo[m].nt = __shfl_sync(0xffffffffu, mo[m].nt, m);
					mo[m].bs = __shfl_sync(0xffffffffu, mo[m].bs, m); mo[m].af = __shfl_sync(0xffffffffu, mo[m].af, m);
					mo[m].rm = __shfl_sync(0xffffffffu, mo[m].rm, m);
					rm[m] = mo[m].rm;
				}
				// ---- 12. accumulate (cpp:2145-2158) ----
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
						uint32_t inc = (rm[0] ? 0u : utf[u]) + (rm[1] ? 0u : utr[u]);
						if (!inc) continue;
						uint32_t ca = classify(ix, uval[u], uaux[u], dst0);
						if ((ca >> AUX_CLASS_SHIFT) == 2) {
							atomicAdd(reinterpret_cast<unsigned long long*>(ix.counts + (ca & AUX_IDX_MASK)), (unsigned long long)inc);
							inc_sum += inc;
						}
					}
					inc_sum = warp_sum(inc_sum);
					if (lane == 0) c_inc += inc_sum;
				}
			}
		}

		// ---- 13. record ----
		const bool emit = trace || (okam && status == DBT_ST_COUNTED);
		if (!emit && okam && c < rp.max_records) {
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
					h->fc = voted ? fc : 0; h->rc = voted ? rc : 0;
					h->nhash0 = nh0; h->nhash1 = (uint16_t)nhash1; h->status = (uint8_t)status;
					h->pad[0] = h->pad[1] = h->pad[2] = 0;
#pragma unroll
					for (int m = 0; m < 2; ++m) {
						dbt_mate_record& o = h->mate[m];
						o.si = (int16_t)mo[m].si; o.ei = (int16_t)mo[m].ei; o.si_ = (int16_t)mo[m].si_; o.ei_ = (int16_t)mo[m].ei_;
						o.ti = (int16_t)mo[m].ti; o.nt = (int8_t)mo[m].nt; o.bs = (int8_t)mo[m].bs;
						o.flags = (uint8_t)((kf[m] ? DBT_MF_KF : 0) | (hf[m] ? DBT_MF_HF : 0) | (mo[m].af ? DBT_MF_AF : 0) | (rm[m] ? DBT_MF_RM : 0));
						o.pad = 0; o.nk = (uint16_t)nas[m];
					}
				}
#pragma unroll
				for (int m = 0; m < 2; ++m) {
					uint8_t* dsta = rec + sizeof(dbt_pair_record) + (size_t)m * rp.as_stride;
					for (uint32_t b = lane; b < rp.as_stride; b += 32) {
						uint32_t v = 0;
#pragma unroll
						for (int q = 0; q < 4; ++q) { uint32_t i = 4 * b + q; if (i < nas[m]) v |= (uint32_t)as[m * NKP + i] << (2 * q); }
						dsta[b] = (uint8_t)v;
					}
				}
			}
		}
	}
	ix.vote_epoch[gwarp] = epoch;
	if (lane == 0) {
		counter_add(sv.counters, DBT_C_NKMERFILTERED, c_kf);
		counter_add(sv.counters, DBT_C_NLOCUSASSIGNFILTERED, c_laf);
		counter_add(sv.counters, DBT_C_NQCFILTERED, c_qc);
		counter_add(sv.counters, DBT_C_NTHREADINGREADS, c_thr);
		counter_add(sv.counters, DBT_C_NFEASIBLEREADS, c_thr);
		counter_add(sv.counters, DBT_C_NASGNREADS, c_asgn);
		counter_add(sv.counters, DBT_C_NHASH1, c_h1);
		counter_add(sv.counters, DBT_C_NINC, c_inc);
	}
}

// ============================================================================================
// index build (A14 device half)
// ============================================================================================
template <int B>
__global__ void insert_kernel(Table t, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, uint64_t n, uint32_t* err) {
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (table_insert<B>(t, keys[i], vals[i])) atomicOr(err, ERR_INDEX_DUPKEY);
}

// Attaches class cls (1 flank, 2 TR) and, for TR, the count-array position to the index entry of
// every (locus, k-mer).  TR pass runs first, flank pass second: flank wins (cpp:1467-1468).
template <int B>
__global__ void fill_aux_kernel(Table t, const uint32_t* __restrict__ vv, uint32_t* __restrict__ vv_aux,
                                const uint64_t* __restrict__ kmers, const uint32_t* __restrict__ locus_of, uint64_t n,
                                uint32_t cls, uint32_t* err) {
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint64_t key = kmers[i]; const uint32_t loc = locus_of[i];
	const uint32_t word = (cls << AUX_CLASS_SHIFT) | (cls == 2 ? (uint32_t)i : 0u);
	uint4* s = table_find_slot<B>(t, key);
	if (!s) { atomicOr(err, ERR_INDEX_MISSING); return; }
	uint32_t* w = reinterpret_cast<uint32_t*>(s);
	const uint32_t val = w[2];
	if (!(val & 1)) {
		if ((val >> 1) != loc) { atomicOr(err, ERR_INDEX_MISSING); return; }
		if (cls == 2 && (w[3] >> AUX_CLASS_SHIFT) == 2) atomicOr(err, ERR_INDEX_DUPTR);
		w[3] = word;
	} else {
		uint32_t off = val >> 1, cnt = vv[off];
		for (uint32_t j = 1; j <= cnt; ++j) if (vv[off + j] == loc) {
			if (cls == 2 && (vv_aux[off + j] >> AUX_CLASS_SHIFT) == 2) atomicOr(err, ERR_INDEX_DUPTR);
			vv_aux[off + j] = word; return;
		}
		atomicOr(err, ERR_INDEX_MISSING);
	}
}

// table occupancy statistics: [0] used slots, [1] buckets with the overflow flag, [2] keys not in their home bucket
template <int B>
__global__ void table_stats_kernel(Table t, unsigned long long* out) {
	uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (b >= t.nbuckets) return;
	uint32_t used = 0, displaced = 0; bool ovf = false;
	for (int i = 0; i < B; ++i) {
		uint64_t k = slot_key(t.slots[b * B + i]);
		if (k == SLOT_EMPTY) continue;
		++used;
		if (i == 0 && (k & SLOT_OVF)) ovf = true;
		if (home_bucket(t, k & SLOT_KEYMASK) != b) ++displaced;
	}
	if (used) atomicAdd(out + 0, (unsigned long long)used);
	if (ovf) atomicAdd(out + 1, 1ull);
	if (displaced) atomicAdd(out + 2, (unsigned long long)displaced);
}

// R_rand microbenchmark: random aligned gathers of sector_bytes from a big table
template <int WORDS16>
__global__ void random_gather_kernel(const uint4* __restrict__ table, uint64_t nchunks, uint64_t ngathers, uint64_t seed, uint32_t* sink) {
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	uint32_t acc = 0;
	for (; i < ngathers; i += stride * 4) {
		uint4 v[4][WORDS16];
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			uint64_t g = i + u * stride;
			uint64_t c = __umul64hi(mix64(g ^ seed), nchunks);
#pragma unroll
			for (int q = 0; q < WORDS16; ++q) v[u][q] = (g < ngathers) ? __ldg(table + c * WORDS16 + q) : make_uint4(0, 0, 0, 0);
		}
#pragma unroll
		for (int u = 0; u < 4; ++u)
#pragma unroll
			for (int q = 0; q < WORDS16; ++q) acc += v[u][q].x ^ v[u][q].w;
	}
	if (acc == 0x12345678u) *sink = acc;
}

// ============================================================================================
// host-callable launchers (used by dbt_abi.cu)
// ============================================================================================
#define DISPATCH_B(B, ...) switch (B) { case 2: { constexpr int BB = 2; __VA_ARGS__; } break; case 4: { constexpr int BB = 4; __VA_ARGS__; } break; case 8: { constexpr int BB = 8; __VA_ARGS__; } break; default: return cudaErrorInvalidValue; }

cudaError_t launch_pack(cudaStream_t st, const uint8_t* ascii, const uint64_t* offsets, uint64_t npairs, const RunParams& rp,
                        uint64_t* p0, uint64_t* p1, uint16_t* m0, uint16_t* m1, uint32_t* err) {
	if (!npairs) return cudaSuccess;
	uint64_t nreads = 2 * npairs;
	pack_kernel<<<(unsigned)((nreads + 127) / 128), 128, 0, st>>>(ascii, offsets, nreads, rp.k, rp.lay, rp.max_read_len, npairs, p0, p1, m0, m1, err);
	return cudaGetLastError();
}

cudaError_t launch_subfilter(cudaStream_t st, int B, int nsm, const DevIndex& ix, const RunParams& rp, const SlotView& sv) {
	if (!sv.npairs) return cudaSuccess;
	uint64_t nwords = (sv.npairs + 31) / 32;
	unsigned grid = (unsigned)((nwords + 7) / 8);
	unsigned cap = (unsigned)nsm * 8 * 4;
	if (grid > cap) grid = cap;
	DISPATCH_B(B, subfilter_kernel<BB><<<grid, 256, 0, st>>>(ix, rp, sv));
	return cudaGetLastError();
}

cudaError_t launch_compact(cudaStream_t st, const SlotView& sv) {
	uint64_t nwords = (sv.npairs + 31) / 32;
	unsigned nblocks = (unsigned)((nwords + COMPACT_WORDS - 1) / COMPACT_WORDS);
	if (!nblocks) nblocks = 1;
	compact_count_kernel<<<nblocks, 256, 0, st>>>(sv);
	compact_scan_kernel<<<1, 1024, 0, st>>>(sv, nblocks);
	compact_scatter_kernel<<<nblocks, 256, 0, st>>>(sv);
	return cudaGetLastError();
}

template <int B, int MAXJ>
static cudaError_t launch_candidate_t(cudaStream_t st, unsigned grid, size_t smem, const DevIndex& ix, const RunParams& rp, const SlotView& sv) {
	cudaError_t e = cudaFuncSetAttribute(candidate_kernel<B, MAXJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	if (e != cudaSuccess) return e;
	candidate_kernel<B, MAXJ><<<grid, CAND_WARPS_PER_BLOCK * 32, smem, st>>>(ix, rp, sv);
	return cudaGetLastError();
}

cudaError_t launch_candidate(cudaStream_t st, int B, unsigned grid, const DevIndex& ix, const RunParams& rp, const SlotView& sv) {
	if (!sv.npairs) return cudaSuccess;
	size_t smem = (size_t)rp.warp_smem * CAND_WARPS_PER_BLOCK;
	int maxj = (int)(rp.nkp / 32);
	if (maxj <= 5)      { DISPATCH_B(B, return launch_candidate_t<BB, 5>(st, grid, smem, ix, rp, sv)); }
	else if (maxj <= 8) { DISPATCH_B(B, return launch_candidate_t<BB, 8>(st, grid, smem, ix, rp, sv)); }
	else if (maxj <= 16){ DISPATCH_B(B, return launch_candidate_t<BB, 16>(st, grid, smem, ix, rp, sv)); }
	return cudaErrorInvalidValue;
}

cudaError_t launch_insert(cudaStream_t st, int B, Table t, const uint64_t* keys, const uint32_t* vals, uint64_t n, uint32_t* err) {
	if (!n) return cudaSuccess;
	DISPATCH_B(B, insert_kernel<BB><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, keys, vals, n, err));
	return cudaGetLastError();
}

cudaError_t launch_fill_aux(cudaStream_t st, int B, Table t, const uint32_t* vv, uint32_t* vv_aux, const uint64_t* kmers,
                            const uint32_t* locus_of, uint64_t n, uint32_t cls, uint32_t* err) {
	if (!n) return cudaSuccess;
	DISPATCH_B(B, fill_aux_kernel<BB><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, vv, vv_aux, kmers, locus_of, n, cls, err));
	return cudaGetLastError();
}

cudaError_t launch_table_stats(cudaStream_t st, int B, Table t, unsigned long long* out) {
	DISPATCH_B(B, table_stats_kernel<BB><<<(unsigned)((t.nbuckets + 255) / 256), 256, 0, st>>>(t, out));
	return cudaGetLastError();
}

cudaError_t launch_random_gather(cudaStream_t st, int nsm, const uint4* table, uint64_t table_bytes, uint32_t sector_bytes,
                                 uint64_t ngathers, uint64_t seed, uint32_t* sink) {
	unsigned grid = (unsigned)nsm * 8;
	uint64_t nchunks = table_bytes / sector_bytes;
	switch (sector_bytes) {
	case 16:  random_gather_kernel<1><<<grid, 256, 0, st>>>(table, nchunks, ngathers, seed, sink); break;
	case 32:  random_gather_kernel<2><<<grid, 256, 0, st>>>(table, nchunks, ngathers, seed, sink); break;
	case 64:  random_gather_kernel<4><<<grid, 256, 0, st>>>(table, nchunks, ngathers, seed, sink); break;
	case 128: random_gather_kernel<8><<<grid, 256, 0, st>>>(table, nchunks, ngathers, seed, sink); break;
	default: return cudaErrorInvalidValue;
	}
	return cudaGetLastError();
}

}  // namespace dbt
