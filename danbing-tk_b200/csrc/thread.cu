// Restored graph threading (P1): thread_kernel + the graph-table build kernels, in their own translation unit.
#include "thread_kernel.cuh"
#include "thread_fast.cuh"
#include "launchers.h"

namespace dbt {

#define DISPATCH_B(B, ...) switch (B) { case 2: { constexpr int BB = 2; __VA_ARGS__; } break; case 4: { constexpr int BB = 4; __VA_ARGS__; } break; default: return cudaErrorInvalidValue; }

template <int B, int MAXJ>
static cudaError_t launch_thread_t(cudaStream_t st, int nsm, const DevIndex& ix, const RunParams& rp, const SlotView& sv, cudaEvent_t mid) {
	// 1. thread_fast_kernel: every pair; finishes the error-free mates, lists the others in sv.thr2
	{
		constexpr size_t smem = sizeof(FastThreadWarp<B, MAXJ>) * THRF_WARPS;
		cudaError_t e = cudaFuncSetAttribute(thread_fast_kernel<B, MAXJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		if (e != cudaSuccess) return e;
		int per_sm = 0;
		e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, thread_fast_kernel<B, MAXJ>, THRF_WARPS * 32, smem);
		if (e != cudaSuccess) return e;
		if (per_sm < 1) return cudaErrorInvalidConfiguration;
		thread_fast_kernel<B, MAXJ><<<(unsigned)nsm * (unsigned)per_sm, THRF_WARPS * 32, smem, st>>>(ix, rp, sv);
		e = cudaGetLastError();
		if (e != cudaSuccess) return e;
	}
	if (mid) { cudaError_t e = cudaEventRecord(mid, st); if (e != cudaSuccess) return e; }
	// 2. thread_kernel: the mates with errors
	constexpr size_t smem = sizeof(ThreadWarp<MAXJ, B>) * THR_WARPS;
	cudaError_t e = cudaFuncSetAttribute(thread_kernel<B, MAXJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	if (e != cudaSuccess) return e;
	int per_sm = 0;
	e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, thread_kernel<B, MAXJ>, THR_WARPS * 32, smem);
	if (e != cudaSuccess) return e;
	if (per_sm < 1) return cudaErrorInvalidConfiguration;
	thread_kernel<B, MAXJ><<<(unsigned)nsm * (unsigned)per_sm, THR_WARPS * 32, smem, st>>>(ix, rp, sv);
	return cudaGetLastError();
}

// MAXJ*32 must hold the k-mer vector plus the k-mers deletions insert (RunParams::tr_cap)
cudaError_t launch_thread(cudaStream_t st, int B, int nsm, const DevIndex& ix, const RunParams& rp, const SlotView& sv, cudaEvent_t mid) {
	if (!sv.npairs) return mid ? cudaEventRecord(mid, st) : cudaSuccess;
	const uint32_t need = rp.tr_cap;
	if (need <= 5 * 32)       { DISPATCH_B(B, return launch_thread_t<BB, 5>(st, nsm, ix, rp, sv, mid)); }
	else if (need <= 8 * 32)  { DISPATCH_B(B, return launch_thread_t<BB, 8>(st, nsm, ix, rp, sv, mid)); }
	else if (need <= 16 * 32) { DISPATCH_B(B, return launch_thread_t<BB, 16>(st, nsm, ix, rp, sv, mid)); }
	return cudaErrorInvalidValue;
}

cudaError_t launch_graph_build(cudaStream_t st, int B, Table t, const uint64_t* kmers, const uint8_t* masks, const uint32_t* locus_of,
                               uint64_t n, uint32_t k, int pass, uint32_t* err) {
	if (!n) return cudaSuccess;
	const unsigned grid = (unsigned)((n + 255) / 256);
	if (pass == 3) { DISPATCH_B(B, graph_verify_kernel<BB><<<grid, 256, 0, st>>>(t, kmers, masks, locus_of, n, k, err)); }
	else           { DISPATCH_B(B, graph_insert_kernel<BB><<<grid, 256, 0, st>>>(t, kmers, masks, locus_of, n, k, pass, err)); }
	return cudaGetLastError();
}

}  // namespace dbt
