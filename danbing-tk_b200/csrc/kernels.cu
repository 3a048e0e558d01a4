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
#include <algorithm>
#include "device_types.cuh"
#include "pack_read.cuh"

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
// pack_kernel (A1, device half; north_star item 1): ASCII reads -> 2 bits/base + invalid mask + "no valid window".
// One warp per TILE of 32 consecutive reads (16 pairs).  The tile's bytes -- one contiguous range of the ASCII stream --
// are staged into shared memory with coalesced 128-bit asynchronous copies (cp.async, 16 B per lane); each lane then
// packs one read out of shared memory four bases at a time with 32-bit SWAR arithmetic (alphabet test = exactly the
// upper-case "ACGT" of h:52,297; code = ((c >> 1) ^ (c >> 2)) & 3 = encodeSeq's A0 C1 G2 T3, h:126-132) into a second
// shared-memory tile, which the warp writes out with coalesced 128-bit stores: the 16 records of mate 0 and the 16 of
// mate 1 are each one contiguous kilobyte of the SoA output.  HBM traffic = the algorithmic 150 B in + 64 B out per read.
// ============================================================================================
constexpr int PACK_WARPS = 4;

// byte-wise fallback for a tile that holds a read longer than max_read_len (the batch is rejected: ERR_READ_TOO_LONG)
__device__ __noinline__ void pack_read_slow(const uint8_t* __restrict__ s, uint32_t len, uint32_t k, ReadLayout lay, uint64_t* __restrict__ out, uint16_t* meta) {
	uint32_t run = 0; bool any = false;
	for (uint32_t w = 0; w < lay.WB; w += 2) {
		uint64_t b0 = 0, b1 = 0, m = 0;
		for (uint32_t j = 0; j < 64; ++j) {
			const uint32_t i = w * 32 + j;
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
	*meta = (uint16_t)(len | (any ? 0 : META_NOVALID));
}

// in_cap: bytes of one warp's input tile (multiple of 16, >= 32 * max_read_len + 15 alignment + 8 over-read)
// read r = ascii[starts[r], ends[r]): one u64 offsets array (starts = offsets, ends = offsets + 1) for concatenated reads, or
// the two arrays the FASTA/FASTQ record splitter fills (fastx.cu) when the reads sit between title lines of the raw text
__global__ void __launch_bounds__(PACK_WARPS * 32) pack_kernel(const uint8_t* __restrict__ ascii, const uint64_t* __restrict__ starts, const uint64_t* __restrict__ ends, uint64_t nreads,
                                                             uint32_t k, ReadLayout lay, uint32_t max_read_len, uint32_t in_cap,
                                                             uint64_t* __restrict__ packed0, uint64_t* __restrict__ packed1,
                                                             uint16_t* __restrict__ meta0, uint16_t* __restrict__ meta1, uint32_t* __restrict__ err) {
	extern __shared__ __align__(16) uint8_t pack_smem[];
	const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const uint32_t W64 = lay.W64;
	uint8_t* tin = pack_smem + (size_t)wib * (in_cap + 32 * W64 * 8);
	uint64_t* tout = reinterpret_cast<uint64_t*>(tin + in_cap);                       // [mate][16 pairs][W64]
	const uint64_t ntiles = (nreads + 31) / 32;
#pragma unroll 1
	for (uint64_t tile = (uint64_t)blockIdx.x * PACK_WARPS + wib; tile < ntiles; tile += (uint64_t)gridDim.x * PACK_WARPS) {
		const uint64_t r = tile * 32 + lane;
		const bool active = r < nreads;
		const uint64_t off = active ? starts[r] : ~0ull, end = active ? ends[r] : 0;
		const uint32_t nact = (uint32_t)min((uint64_t)32, nreads - tile * 32);
		uint64_t off0 = off, endN = end;                                              // byte range of the tile: min start, max end
#pragma unroll
		for (int o = 16; o; o >>= 1) { off0 = min(off0, __shfl_xor_sync(0xffffffffu, off0, o)); endN = max(endN, __shfl_xor_sync(0xffffffffu, endN, o)); }
		const uint64_t len64 = end - off;
		const uint8_t* a0 = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(ascii + off0) & ~(uintptr_t)15);
		const uint64_t nbytes = ((uint64_t)((ascii + endN) - a0) + 15) & ~15ull;
		const uint64_t pair = r >> 1; const uint32_t mate = (uint32_t)(r & 1);
		if (__any_sync(0xffffffffu, active && (len64 > max_read_len || end < off)) || nbytes + 8 > in_cap) {
			if (__any_sync(0xffffffffu, active && (len64 > max_read_len || end < off)) && lane == 0) atomicOr(err, ERR_READ_TOO_LONG);
			if (active) pack_read_slow(ascii + off, end < off ? 0u : (uint32_t)min(len64, (uint64_t)max_read_len), k, lay, (mate ? packed1 : packed0) + pair * W64, (mate ? meta1 : meta0) + pair);
			continue;
		}
		// ---- stage the tile: coalesced 16-byte asynchronous copies ----
#pragma unroll 1
		for (uint32_t c = lane; c < (uint32_t)(nbytes >> 4); c += 32) cp_async16(tin + 16 * c, a0 + 16 * (uint64_t)c);
		cp_async_wait_all();
		__syncwarp();
		// ---- pack one read per lane ----
		uint64_t* orec = tout + ((size_t)mate * 16 + (lane >> 1)) * W64;
		if (active) {
			const uint32_t so = (uint32_t)((ascii + off) - a0);                         // byte offset of the read inside the tile
			const uint32_t* sw = reinterpret_cast<const uint32_t*>(tin + (so & ~3u));
			uint16_t mt;
			if (!pack_read_fast(sw, (so & 3u) * 8, (uint32_t)len64, k, lay, orec, &mt))      // a byte outside ACGT: the exact path
				mt = pack_read_swar(sw, (so & 3u) * 8, (uint32_t)len64, k, lay, orec);
			(mate ? meta1 : meta0)[pair] = mt;
		}
		__syncwarp();
		// ---- write the two record blocks out: coalesced 128-bit stores ----
		{
			const uint64_t pair0 = tile * 16;
			const uint32_t npair_m[2] = {(nact + 1) / 2, nact / 2};
#pragma unroll
			for (int m = 0; m < 2; ++m) {
				const uint4* src = reinterpret_cast<const uint4*>(tout + (size_t)m * 16 * W64);
				uint4* dst = reinterpret_cast<uint4*>((m ? packed1 : packed0) + pair0 * W64);
				const uint32_t n16 = npair_m[m] * W64 / 2;
#pragma unroll 1
				for (uint32_t q = lane; q < n16; q += 32) dst[q] = src[q];
			}
		}
		__syncwarp();
	}
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
		// a clear bit of the presence bitmap (L2-resident: 4 bytes per key against the table's 48) answers the probe of a background
		// k-mer without a fetch from the table
		bool maybe[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			maybe[u] = key[u] != KMER_INVALID;
			if (maybe[u] && ix.pre_bits) { const uint32_t b = pre_bit(key[u], ix.pre_bits); maybe[u] = (__ldg(ix.pre + (b >> 5)) >> (b & 31u)) & 1u; }
		}
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			uint32_t v, a;
			hit[u] = maybe[u] && table_lookup<B>(ix.table, key[u], v, a);
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

#ifndef SUBF_MINBLOCKS
#define SUBF_MINBLOCKS 6       // <= 42 registers, no spills: the kernel waits on dependent loads (meta -> record -> bitmap / bucket), so resident warps count
#endif
template <int B>
__global__ void __launch_bounds__(256, SUBF_MINBLOCKS) subfilter_kernel(DevIndex ix, RunParams rp, SlotView sv) {
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t nwarps = (uint64_t)gridDim.x * (blockDim.x >> 5);
	const uint64_t wid = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const uint64_t nwords = (sv.npairs + 31) / 32;
	const bool trace = (rp.flags & DBT_F_TRACE) != 0;
	const bool filter_on = rp.n_filter && rp.nm_filter;
	const bool coop2 = rp.n_filter <= 4;                                   // mate 2 by four lanes per pair (the default -kf 4 1 and smaller)
	uint32_t c_short = 0, c_sub = 0, c_h0 = 0;
	// a batch the packer flagged (a read longer than max_read_len: dbt_wait reports DBT_E_CAPACITY) must not be counted at all --
	// the caller rebuilds the context and submits it again: no pair passes, nothing downstream runs on it
	const bool dead = (sv.ncand[1] & ERR_READ_TOO_LONG) != 0;
	for (uint64_t w = wid; w < nwords; w += nwarps) {
		const uint64_t p = w * 32 + lane;
		bool pass = false, need2 = false;
		uint32_t nh = 0; int status = -1; uint16_t m2v = 0;
		if (p < sv.npairs && !dead) {
			const uint16_t m1 = sv.meta[0][p], m2 = sv.meta[1][p];
			if ((m1 | m2) & META_NOVALID) { status = DBT_ST_SHORT; ++c_short; }          // cpp:2037
			else if (filter_on) {
				bool ok = subfilter_mate<B>(ix, rp, sv.packed[0] + p * rp.lay.W64, m1 & META_LENMASK, nh);
				if (ok && coop2) need2 = true;                                   // mate 2: by the whole warp, below
				else {
					if (ok) ok = subfilter_mate<B>(ix, rp, sv.packed[1] + p * rp.lay.W64, m2 & META_LENMASK, nh);
					if (!ok) { status = DBT_ST_SUBFILTERED; ++c_sub; }
				}
			}
			m2v = m2;
		}
		// ---- mate 2 of the pairs whose mate 1 passed (a few lanes per warp): four lanes per pair, one sampled k-mer each, up to
		// eight pairs per round -- instead of those lanes running the whole per-thread loop while the rest of the warp waits ----
		uint32_t pend = __ballot_sync(0xffffffffu, need2);
		while (pend) {
			const uint32_t slot = lane >> 2, u = lane & 3u;
			const uint32_t src = __fns(pend, 0, slot + 1);                       // lane that owns the slot-th pending pair, or 0xffffffff
			const bool act = src != 0xffffffffu;
			const uint32_t mm = __shfl_sync(0xffffffffu, (uint32_t)m2v, src & 31u);
			bool hit = false;
			if (act && u < rp.n_filter) {
				const uint32_t L = (mm & META_LENMASK) - rp.k + 1, N = rp.n_filter;
				const uint32_t idx = (u != N - 1) ? u * (L / (N - 1)) : L - 1;
				const uint64_t key = canonical_kmer_at(sv.packed[1] + (w * 32 + src) * rp.lay.W64, rp.lay.WB, idx, rp.k);
				uint32_t v, a2;
				hit = key != KMER_INVALID;
				if (hit && ix.pre_bits) { const uint32_t b = pre_bit(key, ix.pre_bits); hit = (__ldg(ix.pre + (b >> 5)) >> (b & 31u)) & 1u; }
				hit = hit && table_lookup<B>(ix.table, key, v, a2);
			}
			const uint32_t hm = (__ballot_sync(0xffffffffu, hit) >> (4 * slot)) & 0xFu;
			// the reference's loop over the N samples of this mate (cpp:174-177): break at the M-th hit, before the header's ++nhash
			uint32_t h = 0, dnh = 0; bool ok2 = false;
			for (uint32_t i = 0; i < rp.n_filter; ++i) { h += (hm >> i) & 1u; if (h >= rp.nm_filter) { ok2 = true; break; } ++dnh; }
			// back to the owner: the r-th pending lane reads slot r
			const uint32_t r = __popc(pend & ((1u << lane) - 1u));
			const uint32_t res = __shfl_sync(0xffffffffu, (ok2 ? 0x80000000u : 0u) | dnh, (4 * r) & 31u);
			if (need2 && r < 8) {
				need2 = false;
				nh += res & 0x7FFFFFFFu;
				if (!(res >> 31)) { status = DBT_ST_SUBFILTERED; ++c_sub; }
			}
			for (int q = 0; q < 8 && pend; ++q) pend &= pend - 1;                // the eight lowest pending pairs are done
		}
		if (p < sv.npairs && !dead) {
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

// generic over (bitmap, nbits): used for the pass bitmap -> candidate list and for the record emit bitmap -> emit list
__global__ void __launch_bounds__(256) compact_count_kernel(const uint32_t* __restrict__ mask, uint64_t nbits, uint32_t* __restrict__ blocksum) {
	__shared__ uint32_t ssum[8];
	const uint64_t nwords = (nbits + 31) / 32;
	uint32_t c = 0;
	for (int i = threadIdx.x; i < COMPACT_WORDS; i += 256) {
		uint64_t w = (uint64_t)blockIdx.x * COMPACT_WORDS + i;
		if (w < nwords) c += __popc(mask[w]);
	}
	c = warp_sum(c);
	if ((threadIdx.x & 31) == 0) ssum[threadIdx.x >> 5] = c;
	__syncthreads();
	if (threadIdx.x == 0) { uint32_t t = 0; for (int i = 0; i < 8; ++i) t += ssum[i]; blocksum[blockIdx.x] = t; }
}

__global__ void __launch_bounds__(1024) compact_scan_kernel(uint32_t* __restrict__ blocksum, uint32_t nblocks, uint32_t* __restrict__ out_count,
                                                           unsigned long long* __restrict__ counter) {
	// single block; nblocks is small (nbits / 32768)
	__shared__ uint32_t carry;
	__shared__ uint32_t wsum[32];
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	for (uint32_t base = 0; base < nblocks; base += 1024) {
		uint32_t i = base + threadIdx.x;
		uint32_t v = i < nblocks ? blocksum[i] : 0, x = v;
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
		if (i < nblocks) blocksum[i] = excl;
		__syncthreads();
		if (threadIdx.x == 1023) carry = excl + v;
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		*out_count = carry;
		if (counter) atomicAdd(counter, (unsigned long long)carry);
	}
}

__global__ void __launch_bounds__(256) compact_scatter_kernel(const uint32_t* __restrict__ mask, uint64_t nbits, const uint32_t* __restrict__ blocksum,
                                                             uint32_t* __restrict__ out, uint32_t cap = 0xFFFFFFFFu) {
	__shared__ uint32_t wbase[8];
	__shared__ uint32_t run;
	const uint64_t nwords = (nbits + 31) / 32;
	if (threadIdx.x == 0) run = blocksum[blockIdx.x];
	__syncthreads();
	for (int i0 = 0; i0 < COMPACT_WORDS; i0 += 256) {
		uint64_t w = (uint64_t)blockIdx.x * COMPACT_WORDS + i0 + threadIdx.x;
		uint32_t bits = w < nwords ? mask[w] : 0;
		uint32_t c = __popc(bits), x = c;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if ((threadIdx.x & 31) >= o) x += y; }
		if ((threadIdx.x & 31) == 31) wbase[threadIdx.x >> 5] = x;
		__syncthreads();
		uint32_t pre = run;
		for (int j = 0; j < (int)(threadIdx.x >> 5); ++j) pre += wbase[j];
		uint32_t o = pre + x - c;
		while (bits) { int b = __ffs(bits) - 1; bits &= bits - 1; if (o < cap) out[o] = (uint32_t)(w * 32 + b); ++o; }
		__syncthreads();
		if (threadIdx.x == 0) { uint32_t t = 0; for (int j = 0; j < 8; ++j) t += wbase[j]; run += t; }
		__syncthreads();
	}
}

// ============================================================================================
// record emission (dbt_fetch): which record slots become kam / aln records -> ordered emit list -> dense copy.
// Replaces a per-record host memcpy loop in dbt_wait (360 MB per 10 M-pair step in threading mode).
// ============================================================================================
// slot i < min(*nslots, cap): emit iff COUNTED, or THREAD_FAILED when `emit_failed` & 1 (`-a`: every pair that entered threading,
// cpp:2232-2239), or ASGN_REMOVED when `emit_failed` & 2 (`-s`: kam lines of removed pairs with a known source locus, cpp:2169)
__global__ void __launch_bounds__(256) record_flag_kernel(const uint8_t* __restrict__ records, uint32_t stride, const uint32_t* __restrict__ nslots,
                                                         uint32_t cap, int emit_failed, uint32_t* __restrict__ mask) {
	const uint32_t n = min(*nslots, cap);
	const uint32_t i = blockIdx.x * 256 + threadIdx.x;
	bool e = false;
	if (i < n) {
		const uint8_t st = reinterpret_cast<const dbt_pair_record*>(records + (size_t)i * stride)->status;
		e = st == DBT_ST_COUNTED || ((emit_failed & 1) && st == DBT_ST_THREAD_FAILED) || ((emit_failed & 2) && st == DBT_ST_ASGN_REMOVED);
	}
	const uint32_t b = __ballot_sync(0xffffffffu, e);
	if ((threadIdx.x & 31) == 0 && i < ((cap + 31) / 32) * 32) mask[i >> 5] = b;
}

// one warp per emitted record: out[j] = records[list[j]]; strides are multiples of 8 bytes
__global__ void __launch_bounds__(256) record_gather_kernel(const uint8_t* __restrict__ records, uint32_t stride, const uint32_t* __restrict__ list,
                                                           const uint32_t* __restrict__ nemit, uint8_t* __restrict__ out) {
	const uint32_t n = *nemit;
	const uint32_t lane = threadIdx.x & 31, nw = stride / 8;
	for (uint32_t j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
		const uint64_t* src = reinterpret_cast<const uint64_t*>(records + (size_t)list[j] * stride);
		uint64_t* dst = reinterpret_cast<uint64_t*>(out + (size_t)j * stride);
		for (uint32_t q = lane; q < nw; q += 32) dst[q] = src[q];
	}
}

}  // namespace dbt

#include "candidate_fast.cuh"

namespace dbt {

// ============================================================================================
// index build (A14 device half)
// ============================================================================================
template <int B>
__global__ void insert_kernel(Table t, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, uint64_t n, uint32_t* err) {
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int r = table_insert<B>(t, keys[i], vals[i]);
	if (r) atomicOr(err, r == 1 ? ERR_INDEX_DUPKEY : ERR_INDEX_FULL);
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
		w[3] = (cls == 1 && (w[3] >> AUX_CLASS_SHIFT) == 2 && (w[3] & AUX_IDX_MASK) < AUX_FLANK_TR) ? (word | AUX_FLANK_TR | (w[3] & AUX_IDX_MASK)) : word;
	} else {
		uint32_t off = val >> 1, cnt = vv[off];
		for (uint32_t j = 1; j <= cnt; ++j) if (vv[off + j] == loc) {
			if (cls == 2 && (vv_aux[off + j] >> AUX_CLASS_SHIFT) == 2) atomicOr(err, ERR_INDEX_DUPTR);
			const uint32_t old = vv_aux[off + j];
			vv_aux[off + j] = (cls == 1 && (old >> AUX_CLASS_SHIFT) == 2 && (old & AUX_IDX_MASK) < AUX_FLANK_TR) ? (word | AUX_FLANK_TR | (old & AUX_IDX_MASK)) : word; return;
		}
		atomicOr(err, ERR_INDEX_MISSING);
	}
}

// Merge of the de Bruijn graphs into the index (load time).  One thread per graph node (strand-specific k-mer, locus, mask):
// its five bits (present | mask << 1) go into the index entry of (canonical k-mer, locus) -- the top bits of the value word
// for a k-mer unique to the locus, vv_g for a k-mer listed in several loci -- in the half that belongs to its orientation.
// A node that is not an index key of its locus raises ERR_INDEX_MISSING: the caller then leaves the merge switched off.
template <int B>
__global__ void graph_merge_kernel(Table t, const uint32_t* __restrict__ vv, uint32_t* __restrict__ vv_g, const uint64_t* __restrict__ kmers,
                                   const uint8_t* __restrict__ masks, const uint32_t* __restrict__ locus_of, uint64_t n, uint32_t k, uint32_t* err) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint64_t fw = kmers[i]; const uint32_t loc = locus_of[i];
	const uint64_t rc = rev2(~fw) >> (64 - 2 * k);
	const bool is_rc = rc < fw;
	const uint64_t canon = is_rc ? rc : fw;
	uint32_t g = 1u | ((uint32_t)(masks[i] & 0xFu) << 1);
	g = (rc == fw) ? (g | (g << 5)) : (is_rc ? g << 5 : g);                   // palindrome (even k): both orientations are this node
	uint4* s = table_find_slot<B>(t, canon);
	if (!s) { atomicOr(err, 1u); return; }                                 // node is no index key at all
	uint32_t* w = reinterpret_cast<uint32_t*>(s);
	const uint32_t z = w[2];
	if (!(z & 1u)) {
		if (((z & VAL_LOCUS_MASK) >> 1) != loc) { atomicOr(err, 2u); return; }  // ... is a key of another locus only
		atomicOr(w + 2, g << VAL_LOCUS_BITS);
	} else {
		const uint32_t off = z >> 1, cnt = vv[off];
		for (uint32_t j = 1; j <= cnt; ++j) if (vv[off + j] == loc) { atomicOr(vv_g + off + j, g); return; }
		atomicOr(err, 4u);                                                    // ... is shared by other loci, not this one
	}
}

// ... and every TR k-mer must still carry its count-array position in the index entry of its locus (a k-mer that is also a flank
// k-mer of the locus is classed flank there, cpp:1467-1468, and has lost it)
template <int B>
__global__ void tr_class_check_kernel(Table t, const uint32_t* __restrict__ vv, const uint32_t* __restrict__ vv_aux, const uint64_t* __restrict__ kmers,
                                      const uint32_t* __restrict__ locus_of, uint64_t n, uint32_t* err) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t loc = locus_of[i];
	const uint4* s = table_find_slot<B>(t, kmers[i]);
	if (!s) { atomicOr(err, 8u); return; }
	const uint32_t* w = reinterpret_cast<const uint32_t*>(s);
	const uint32_t z = w[2];
	uint32_t word = 0;
	if (!(z & 1u)) word = (((z & VAL_LOCUS_MASK) >> 1) == loc) ? w[3] : 0u;
	else { const uint32_t off = z >> 1, cnt = vv[off]; for (uint32_t j = 1; j <= cnt; ++j) if (vv[off + j] == loc) word = vv_aux[off + j]; }
	if (word != ((2u << AUX_CLASS_SHIFT) | (uint32_t)i) && word != ((1u << AUX_CLASS_SHIFT) | AUX_FLANK_TR | (uint32_t)i)) atomicOr(err, 16u);   // a TR k-mer classed flank (or missing): its count position is not in the entry
}

// table occupancy statistics: [0] used slots, [1] main buckets with the overflow flag, [2] keys in the stash,
// [3] stash buckets with the overflow flag (look-ups that need a third bucket)
template <int B>
__global__ void table_stats_kernel(Table t, unsigned long long* out) {
	uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (b >= t.nbuckets + t.nstash) return;
	uint32_t used = 0; bool ovf = false;
	for (int i = 0; i < B; ++i) {
		uint64_t k = slot_key(t.slots[b * B + i]);
		if (k == SLOT_EMPTY) continue;
		++used;
		if (i == 0 && (k & SLOT_OVF)) ovf = true;
	}
	if (used) atomicAdd(out + 0, (unsigned long long)used);
	if (ovf) atomicAdd(out + (b < t.nbuckets ? 1 : 3), 1ull);
	if (used && b >= t.nbuckets) atomicAdd(out + 2, (unsigned long long)used);
}

// presence bitmap over the index keys (load time)
__global__ void prefilter_build_kernel(const uint64_t* __restrict__ keys, uint64_t n, uint32_t* __restrict__ bitmap, uint32_t bits) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t b = pre_bit(keys[i], bits);
	atomicOr(bitmap + (b >> 5), 1u << (b & 31u));
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

// the same gathers done the way the probe kernels stage a bucket: LANES adjacent lanes take 16 bytes each of one chunk, so a chunk
// is ONE coalesced request of LANES * 16 bytes (32 / LANES gathers per load instruction) instead of LANES requests of one thread
template <int LANES>
__global__ void random_gather_coop_kernel(const uint4* __restrict__ table, uint64_t nchunks, uint64_t ngathers, uint64_t seed, uint32_t* sink) {
	const uint32_t sub = threadIdx.x % LANES;
	uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x / LANES;
	uint32_t acc = 0;
	for (; i < ngathers; i += stride * 8) {
		uint4 v[8];
#pragma unroll
		for (int u = 0; u < 8; ++u) {
			const uint64_t g = i + u * stride;
			const uint64_t c = __umul64hi(mix64(g ^ seed), nchunks);
			v[u] = (g < ngathers) ? __ldg(table + c * LANES + sub) : make_uint4(0, 0, 0, 0);
		}
#pragma unroll
		for (int u = 0; u < 8; ++u) acc += v[u].x ^ v[u].w;
	}
	if (acc == 0x12345678u) *sink = acc;
}

// ... and 32-byte chunks by ONE 256-bit load per thread (what table_lookup / g_lookup issue)
__global__ void random_gather_ldg256_kernel(const uint4* __restrict__ table, uint64_t nchunks, uint64_t ngathers, uint64_t seed, uint32_t* sink) {
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	uint32_t acc = 0;
	for (; i < ngathers; i += stride * 4) {
		uint4 a[4], b[4];
#pragma unroll
		for (int u = 0; u < 4; ++u) {
			const uint64_t g = i + u * stride;
			const uint64_t c = __umul64hi(mix64(g ^ seed), nchunks);
			if (g < ngathers) ldg256(table + c * 2, a[u], b[u]); else a[u] = b[u] = make_uint4(0, 0, 0, 0);
		}
#pragma unroll
		for (int u = 0; u < 4; ++u) acc += a[u].x ^ b[u].w;
	}
	if (acc == 0x12345678u) *sink = acc;
}

// ============================================================================================
// host-callable launchers (used by dbt_abi.cu)
// ============================================================================================
#define DISPATCH_B(B, ...) switch (B) { case 2: { constexpr int BB = 2; __VA_ARGS__; } break; case 4: { constexpr int BB = 4; __VA_ARGS__; } break; default: return cudaErrorInvalidValue; }

cudaError_t launch_pack(cudaStream_t st, const uint8_t* ascii, const uint64_t* starts, const uint64_t* ends, uint64_t npairs, const RunParams& rp,
                        uint32_t tile_bytes, uint64_t* p0, uint64_t* p1, uint16_t* m0, uint16_t* m1, uint32_t* err) {
	if (!npairs) return cudaSuccess;
	const uint64_t nreads = 2 * npairs;
	// tile_bytes: the byte range 32 consecutive reads span (0: concatenated reads); capped so that four warps fit a block --
	// a tile that does not fit its stage takes the byte-wise per-lane path
	const uint32_t want = std::max(32u * rp.max_read_len, std::min(tile_bytes, 40u * 1024u));
	const uint32_t in_cap = (want + 15u + 8u + 15u) & ~15u;
	const size_t smem = (size_t)PACK_WARPS * (in_cap + 32 * rp.lay.W64 * 8);
	static bool attr_set = false;
	if (!attr_set) {
		cudaError_t e = cudaFuncSetAttribute(pack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
		if (e != cudaSuccess) return e;
		attr_set = true;
	}
	const uint64_t ntiles = (nreads + 31) / 32;
	const unsigned grid = (unsigned)std::min<uint64_t>((ntiles + PACK_WARPS - 1) / PACK_WARPS, 148ull * 64);
	pack_kernel<<<grid, PACK_WARPS * 32, smem, st>>>(ascii, starts, ends, nreads, rp.k, rp.lay, rp.max_read_len, in_cap, p0, p1, m0, m1, err);
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
	compact_count_kernel<<<nblocks, 256, 0, st>>>(sv.passmask, sv.npairs, sv.blocksum);
	compact_scan_kernel<<<1, 1024, 0, st>>>(sv.blocksum, nblocks, sv.ncand, sv.counters + DBT_C_NCANDIDATES);
	compact_scatter_kernel<<<nblocks, 256, 0, st>>>(sv.passmask, sv.npairs, sv.blocksum, sv.cand);
	return cudaGetLastError();
}

// records[0 .. min(*nslots, cap)) -> out[0 .. *nemit) in slot order; 5 launches
// generic: set bits of mask[0, nbits) -> ascending positions in list (at most cap are written), their number in *count
cudaError_t launch_compact_bits(cudaStream_t st, const uint32_t* mask, uint64_t nbits, uint32_t* blocksum, uint32_t* list, uint32_t* count, uint32_t cap) {
	const uint64_t nwords = (nbits + 31) / 32;
	unsigned nblocks = (unsigned)((nwords + COMPACT_WORDS - 1) / COMPACT_WORDS);
	if (!nblocks) nblocks = 1;
	compact_count_kernel<<<nblocks, 256, 0, st>>>(mask, nbits, blocksum);
	compact_scan_kernel<<<1, 1024, 0, st>>>(blocksum, nblocks, count, nullptr);
	compact_scatter_kernel<<<nblocks, 256, 0, st>>>(mask, nbits, blocksum, list, cap);
	return cudaGetLastError();
}

cudaError_t launch_record_emit(cudaStream_t st, int nsm, const uint8_t* records, uint32_t stride, const uint32_t* nslots, uint32_t cap, int emit_failed,
                               uint32_t* mask, uint32_t* blocksum, uint32_t* list, uint32_t* nemit, uint8_t* out) {
	if (!cap) return cudaSuccess;
	const unsigned fb = (cap + 255) / 256;
	record_flag_kernel<<<fb, 256, 0, st>>>(records, stride, nslots, cap, emit_failed, mask);
	const uint64_t nwords = ((uint64_t)cap + 31) / 32;
	unsigned nblocks = (unsigned)((nwords + COMPACT_WORDS - 1) / COMPACT_WORDS);
	if (!nblocks) nblocks = 1;
	compact_count_kernel<<<nblocks, 256, 0, st>>>(mask, cap, blocksum);
	compact_scan_kernel<<<1, 1024, 0, st>>>(blocksum, nblocks, nemit, nullptr);
	compact_scatter_kernel<<<nblocks, 256, 0, st>>>(mask, cap, blocksum, list);
	record_gather_kernel<<<(unsigned)nsm * 8, 256, 0, st>>>(records, stride, list, nemit, out);
	return cudaGetLastError();
}

template <int B, int MAXJ>
static cudaError_t launch_candidate_t(cudaStream_t st, unsigned nsm, unsigned grid_slow, size_t warp_smem, const DevIndex& ix, const RunParams& rp, const SlotView& sv, cudaEvent_t mid) {
	const size_t smem_slow = warp_smem * CAND_WARPS_SLOW;
	cudaError_t e = cudaFuncSetAttribute(candidate_kernel<B, MAXJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_slow);
	if (e != cudaSuccess) return e;
	// DBT_F_TRACE wants the exact (top.fc, top.rc) of every pair; reads with > 255 k-mers overflow the
	// reference's uint8 multiplicities (cpp:42), which only the generic kernel reproduces
	const bool all = (rp.flags & DBT_F_TRACE) != 0 || MAXJ > 8 || rp.nk_max > 255;
	if constexpr (MAXJ <= 8) {
		if (!all) {
			constexpr size_t smem_fast = sizeof(FastWarp<B, MAXJ>) * CAND_FAST_WARPS;
			auto launch = [&](auto kern) -> cudaError_t {
				cudaError_t e2 = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_fast);
				if (e2 != cudaSuccess) return e2;
				int per_sm = 0;
				e2 = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, CAND_FAST_WARPS * 32, smem_fast);
				if (e2 != cudaSuccess) return e2;
				if (per_sm < 1) return cudaErrorInvalidConfiguration;
				kern<<<nsm * (unsigned)per_sm, CAND_FAST_WARPS * 32, smem_fast, st>>>(ix, rp, sv);
				return cudaSuccess;
			};
			e = rp.coop_stage ? launch(candidate_fast_kernel<B, MAXJ, true>) : launch(candidate_fast_kernel<B, MAXJ, false>);
			if (e != cudaSuccess) return e;
		}
	}
	if (mid) { e = cudaEventRecord(mid, st); if (e != cudaSuccess) return e; }
	candidate_kernel<B, MAXJ><<<grid_slow, CAND_WARPS_SLOW * 32, smem_slow, st>>>(ix, rp, sv, all ? 1 : 0);
	return cudaGetLastError();
}

cudaError_t launch_candidate(cudaStream_t st, int B, unsigned grid_fast, unsigned grid_slow, const DevIndex& ix, const RunParams& rp, const SlotView& sv, cudaEvent_t mid) {
	if (!sv.npairs) return mid ? cudaEventRecord(mid, st) : cudaSuccess;
	int maxj = (int)(rp.nkp / 32);
	if (maxj <= 5)      { DISPATCH_B(B, return launch_candidate_t<BB, 5>(st, grid_fast, grid_slow, rp.warp_smem, ix, rp, sv, mid)); }
	else if (maxj <= 8) { DISPATCH_B(B, return launch_candidate_t<BB, 8>(st, grid_fast, grid_slow, rp.warp_smem, ix, rp, sv, mid)); }
	else if (maxj <= 16){ DISPATCH_B(B, return launch_candidate_t<BB, 16>(st, grid_fast, grid_slow, rp.warp_smem, ix, rp, sv, mid)); }
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

cudaError_t launch_graph_merge(cudaStream_t st, int B, Table t, const uint32_t* vv, uint32_t* vv_g, const uint64_t* kmers, const uint8_t* masks,
                               const uint32_t* locus_of, uint64_t n, uint32_t k, uint32_t* err) {
	if (!n) return cudaSuccess;
	DISPATCH_B(B, graph_merge_kernel<BB><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, vv, vv_g, kmers, masks, locus_of, n, k, err));
	return cudaGetLastError();
}

cudaError_t launch_tr_class_check(cudaStream_t st, int B, Table t, const uint32_t* vv, const uint32_t* vv_aux, const uint64_t* kmers, const uint32_t* locus_of, uint64_t n, uint32_t* err) {
	if (!n) return cudaSuccess;
	DISPATCH_B(B, tr_class_check_kernel<BB><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, vv, vv_aux, kmers, locus_of, n, err));
	return cudaGetLastError();
}

cudaError_t launch_bait_insert(cudaStream_t st, int B, Table t, const uint64_t* kmers, const uint16_t* vals, const uint32_t* locus_of, uint64_t n, uint32_t* err) {
	if (!n) return cudaSuccess;
	DISPATCH_B(B, bait_insert_kernel<BB><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(t, kmers, vals, locus_of, n, err));
	return cudaGetLastError();
}

cudaError_t launch_table_stats(cudaStream_t st, int B, Table t, unsigned long long* out) {
	DISPATCH_B(B, table_stats_kernel<BB><<<(unsigned)((t.nbuckets + t.nstash + 255) / 256), 256, 0, st>>>(t, out));
	return cudaGetLastError();
}

cudaError_t launch_prefilter_build(cudaStream_t st, const uint64_t* keys, uint64_t n, uint32_t* bitmap, uint32_t bits) {
	if (!n) return cudaSuccess;
	prefilter_build_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(keys, n, bitmap, bits);
	return cudaGetLastError();
}

cudaError_t launch_random_gather(cudaStream_t st, int nsm, const uint4* table, uint64_t table_bytes, uint32_t sector_bytes,
                                 uint64_t ngathers, uint64_t seed, uint32_t* sink) {
	unsigned grid = (unsigned)nsm * 8;
	if (sector_bytes == (DBT_GATHER_WIDE_LOAD | 32u)) {
		random_gather_ldg256_kernel<<<grid, 256, 0, st>>>(table, table_bytes / 32, ngathers, seed, sink);
		return cudaGetLastError();
	}
	if (sector_bytes & DBT_GATHER_COOPERATIVE) {
		sector_bytes &= ~DBT_GATHER_COOPERATIVE;
		const uint64_t nch = table_bytes / sector_bytes;
		switch (sector_bytes) {
		case 32:  random_gather_coop_kernel<2><<<grid, 256, 0, st>>>(table, nch, ngathers, seed, sink); break;
		case 64:  random_gather_coop_kernel<4><<<grid, 256, 0, st>>>(table, nch, ngathers, seed, sink); break;
		case 128: random_gather_coop_kernel<8><<<grid, 256, 0, st>>>(table, nch, ngathers, seed, sink); break;
		default: return cudaErrorInvalidValue;
		}
		return cudaGetLastError();
	}
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
