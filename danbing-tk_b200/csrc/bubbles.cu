// `-bu`: novel (k+1)-mer edges of the reads inside the observed TR segment.
//
// Reference: for every mate of a counted pair that assignTRkmc kept, countNovelEdges (src/aQueryFasta_thread.cpp:1559-1567)
// walks the canonical (k+1)-mer edges i = si_ .. ei_-2 of the read (read2kmers_edges, src/aQueryFasta_thread.h:287-296: the
// edge between k-mers i and i+1 is the canonical form of the k+1 bases from i; absent when one of them is not ACGT) and counts
// the ones that are not in trEdgeDB[destLocus] (PREF.tre.kdb) in a per-batch unordered_map, merged into bubbleDB at the end of
// the batch (accumBubbles, cpp:1599-1606) and written as OUT.bub.kmdb with the entries counted at least five times
// (dumpBubbles, h:1006-1008).  The file lists every locus' unordered_map in ITERATION order, so what the device hands back
// is not a count table but the EVENTS -- (pair, mate, position, locus, edge) of every novel edge -- which the host replays
// in pair order into the same containers (host/bubbles.hpp).  Events are rare: an edge is novel where a read differs from
// every haplotype of the graph.
//
//   edge_insert_kernel   PREF.tre.kdb -> one open-addressing table keyed (canonical (k+1)-mer, locus), the bucket / stash
//                        scheme of graph_table.cuh
//   bubble_kernel        one warp per candidate record the candidate kernels left COUNTED: lanes take the edge positions of
//                        the segment of each kept mate, one probe each, novel edges are appended to the slot's event list
#include <algorithm>
#include "device_types.cuh"
#include "graph_table.cuh"
#include "launchers.h"

namespace dbt {

template <int B>
__global__ void edge_insert_kernel(Table t, const uint64_t* __restrict__ edges, const uint32_t* __restrict__ locus_of, uint64_t n, uint32_t nbits, uint32_t* err) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint64_t e = edges[i]; const uint32_t loc = locus_of[i];
	if (nbits < 64 && (e >> nbits)) { atomicOr(err, ERR_INDEX_MISSING); return; }
	if (!g_claim<B>(t, e, loc, loc | G_PRESENT_CA, 0)) atomicOr(err, ERR_INDEX_FULL);
}

template <int B>
__global__ void __launch_bounds__(256) bubble_kernel(Table tt, RunParams rp, SlotView sv, uint32_t nloci, uint64_t* __restrict__ ev_key, uint64_t* __restrict__ ev_edge,
                                                     uint32_t* __restrict__ ev_locus, uint32_t* __restrict__ nev, uint32_t cap) {
	const int lane = threadIdx.x & 31;
	const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
	const uint32_t ncand = sv.ncand[0];
	const bool trace = (rp.flags & DBT_F_TRACE) != 0;
	const uint32_t W64 = rp.lay.W64, WB = rp.lay.WB, K1 = rp.k + 1;
	for (uint32_t c = gwarp; c < ncand; c += nwarps) {
		const uint32_t p = sv.cand[c];
		const uint64_t slot = trace ? (uint64_t)p : (uint64_t)c;
		if (!trace && slot >= rp.max_records) continue;                     // (the batch fails with ERR_RECORD_OVERFLOW anyway)
		const dbt_pair_record* h = reinterpret_cast<const dbt_pair_record*>(sv.records + slot * rp.rec_stride);
		if (h->status != DBT_ST_COUNTED) continue;
		const uint32_t dst = h->dst;
		if (dst >= nloci) continue;
#pragma unroll 1
		for (int m = 0; m < 2; ++m) {
			const dbt_mate_record& r = h->mate[m];
			if (r.flags & DBT_MF_RM) continue;                                  // cpp:2163-2164
			const int si = r.si_, ei = (int)r.ei_ - 1;                          // cpp:1560-1561
			const uint64_t* rec = (m ? sv.packed[1] : sv.packed[0]) + (uint64_t)p * W64;
			const int len = (int)((m ? sv.meta[1] : sv.meta[0])[p] & META_LENMASK);
#pragma unroll 1
			for (int i0 = si; i0 < ei; i0 += 32) {
				const int i = i0 + lane;
				bool novel = false; uint64_t e = 0;
				if (i >= 0 && i < ei && i + (int)K1 <= len) {
					e = canonical_kmer_at(rec, WB, (uint32_t)i, K1);
					novel = e != KMER_INVALID && g_find_slot<B>(tt, e, dst) == nullptr;
				}
				const uint32_t bm = __ballot_sync(0xffffffffu, novel);
				if (!bm) continue;
				uint32_t base = 0;
				if (lane == 0) base = atomicAdd(nev, (uint32_t)__popc(bm));
				base = __shfl_sync(0xffffffffu, base, 0);
				if (novel) {
					const uint32_t o = base + __popc(bm & ((1u << lane) - 1u));
					if (o < cap) { ev_key[o] = ((uint64_t)p << 10) | ((uint64_t)m << 9) | (uint64_t)i; ev_edge[o] = e; ev_locus[o] = dst; }
				}
			}
		}
	}
}

cudaError_t launch_edge_insert(cudaStream_t st, int B, Table t, const uint64_t* edges, const uint32_t* locus_of, uint64_t n, uint32_t k1, uint32_t* err) {
	if (!n) return cudaSuccess;
	const unsigned g = (unsigned)((n + 255) / 256);
	if (B == 2) edge_insert_kernel<2><<<g, 256, 0, st>>>(t, edges, locus_of, n, 2 * k1, err);
	else edge_insert_kernel<4><<<g, 256, 0, st>>>(t, edges, locus_of, n, 2 * k1, err);
	return cudaGetLastError();
}

cudaError_t launch_bubbles(cudaStream_t st, int B, int nsm, Table tt, const RunParams& rp, const SlotView& sv, uint32_t nloci,
                           uint64_t* ev_key, uint64_t* ev_edge, uint32_t* ev_locus, uint32_t* nev, uint32_t cap) {
	if (!sv.npairs) return cudaSuccess;
	const unsigned g = (unsigned)nsm * 8;
	if (B == 2) bubble_kernel<2><<<g, 256, 0, st>>>(tt, rp, sv, nloci, ev_key, ev_edge, ev_locus, nev, cap);
	else bubble_kernel<4><<<g, 256, 0, st>>>(tt, rp, sv, nloci, ev_key, ev_edge, ev_locus, nev, cap);
	return cudaGetLastError();
}

}  // namespace dbt
