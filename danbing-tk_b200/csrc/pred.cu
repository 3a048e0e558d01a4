// dbt_pred: `danbing-tk-pred` bias correction on the GPU (reference src/pred.cpp:58-80, src/pred.h:141-195).
//
// Input = what the align path writes: one OUT.trkmc.ar count array per sample (u64[nk]) + read depths + the invariant k-mer
// table of the RPGG.  Dense float32 streaming work, HBM-bound:
//   pred_norm_kernel    load_eachBinGT + norm_rd (pred.h:141-171): raw[s + k*ns] = float(kmc[s][k]) / rd[s].  A transpose of a
//                       group of <= 32 staged sample arrays (coalesced 8-byte reads along k) into the sample-major matrix
//                       (coalesced 4-byte writes along s) through a padded shared-memory tile.
//   pred_bias_kernel    bias_correction (pred.h:173-195): one block per locus.  bias[s] = mean_j raw[s, iki[j]] / ikmc[j]; then
//                       bias /= mean_s bias; corrected[s, si:ei] = raw[s, si:ei] / bias[s].
// Bit-exact float32 against the reference: every operation is a single IEEE add or divide (__fadd_rn / __fdiv_rn: no
// contraction, no fast division) issued in the order the reference's Eigen 3.4 expressions evaluate them --
//   * rowwise().mean() over the indexed view gt(all, ikis): a scalar left-to-right sum per sample, divided by float(n);
//   * bias.mean(): Eigen's vectorised full reduction with SSE packets of 4: two packet accumulators over strides of 8, their
//     sum, one more packet if the length allows, lanes combined as (a0+a2)+(a1+a3), then the scalar tail; divided by float(ns).
// tests/pred_common.py restates the same order in numpy and is pinned against the compiled reference.
#include <algorithm>
#include <cstdio>
#include <vector>
#include "../../include/danbing_b200.h"

namespace dbt { int abi_fail(int code, const char* fmt, ...); }

namespace {

constexpr int PG = 32;                                     // samples staged and transposed together

// stage: [g][nk] counts of samples s0 .. s0+ng-1; raw: [k][ns] floats (column-major ns x nk)
__global__ void __launch_bounds__(256) pred_norm_kernel(const uint64_t* __restrict__ stage, uint64_t nk, const float* __restrict__ rd,
                                                        uint32_t s0, uint32_t ng, uint64_t ns, float* __restrict__ raw) {
	__shared__ float tile[PG][33];
	const uint32_t tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
	const uint64_t ntiles = (nk + 31) / 32;
	for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
		const uint64_t k0 = t * 32;
		for (uint32_t g = ty; g < ng; g += 8) {                        // row g of the tile: 32 consecutive k-mers of sample s0+g
			const uint64_t k = k0 + tx;
			if (k < nk) tile[g][tx] = __fdiv_rn((float)stage[(uint64_t)g * nk + k], rd[s0 + g]);
		}
		__syncthreads();
		for (uint32_t kk = ty; kk < 32; kk += 8) {                     // column kk of the tile: ng consecutive samples of k-mer k0+kk
			const uint64_t k = k0 + kk;
			if (k < nk && tx < ng) raw[k * ns + s0 + tx] = tile[tx][kk];
		}
		__syncthreads();
	}
}

// Eigen 3.4 redux<sum> of a contiguous, 16-byte aligned float array with 4-float packets (Redux.h, LinearVectorizedTraversal),
// by one warp: the eight lanes of the two packet accumulators are eight independent chains, one per lane 0..7; the rest of the
// order (accumulators added, one more packet, (a0+a2)+(a1+a3), scalar tail) is replayed with shuffles.  Result in every lane.
__device__ float eigen_sum4_warp(const float* a, uint32_t n, uint32_t lane) {
	const uint32_t n2 = (n / 8) * 8, n1 = (n / 4) * 4;
	float r;
	if (!n1) {
		r = a[0];
		for (uint32_t i = 1; i < n; ++i) r = __fadd_rn(r, a[i]);
		return r;
	}
	float acc = 0.f;
	if (n1 > 4) {                                                      // two accumulators: lane l < 4 is p0[l], lane 4 + l is p1[l]
		if (lane < 8) { acc = a[lane]; for (uint32_t i = 8; i < n2; i += 8) acc = __fadd_rn(acc, a[i + lane]); }
		const float other = __shfl_down_sync(0xffffffffu, acc, 4);
		acc = __fadd_rn(acc, other);                                   // p0 += p1 (lanes 0..3)
		if (n1 > n2 && lane < 4) acc = __fadd_rn(acc, a[n2 + lane]);
	} else if (lane < 4) acc = a[lane];
	const float a2 = __shfl_down_sync(0xffffffffu, acc, 2);            // lane 0: p0[2], lane 1: p0[3]
	const float h = __fadd_rn(acc, a2);                                // lane 0: a0+a2, lane 1: a1+a3
	const float h1 = __shfl_down_sync(0xffffffffu, h, 1);
	r = __fadd_rn(h, h1);
	for (uint32_t i = n1; i < n; ++i) r = __fadd_rn(r, a[i]);
	return __shfl_sync(0xffffffffu, r, 0);
}

// one block per locus.  bias: [locus][ns] (column-major ns x n_tr); skipped loci (no k-mers or no invariant k-mers) leave their
// column untouched and their block of `cor` equal to `raw`, pred.h:181.
// The per-sample sum over the invariant k-mers must be added up left to right, but the gathers and divisions feeding it are
// independent: with ns <= 256 the block first fills a shared-memory tile B[c][s] = raw[s, iki[j0+c]] / ikmc[j0+c] for 256/ns
// invariant k-mers at a time (all threads, coalesced along s), then thread s adds the tile's column in order.
__global__ void __launch_bounds__(256) pred_bias_kernel(const float* __restrict__ raw, float* __restrict__ cor, float* __restrict__ bias, uint64_t ns, uint32_t n_tr,
                                                        const uint32_t* __restrict__ nk_cum, const uint32_t* __restrict__ nik_cum,
                                                        const uint32_t* __restrict__ iki, const uint8_t* __restrict__ ikmc) {
	__shared__ float s_mean;
	__shared__ float s_b[256];
	const uint32_t tid = threadIdx.x;
	for (uint32_t t = blockIdx.x; t < n_tr; t += gridDim.x) {
		const uint32_t si = t ? nk_cum[t - 1] : 0, ei = nk_cum[t], isi = t ? nik_cum[t - 1] : 0, iei = nik_cum[t];
		if (si == ei) continue;
		const uint64_t n = (uint64_t)(ei - si) * ns, o = (uint64_t)si * ns;    // the locus' block is contiguous in the column-major matrix
		if (isi == iei) {                                                // skipped locus: its block stays uncorrected
			for (uint64_t i = tid; i < n; i += blockDim.x) cor[o + i] = raw[o + i];
			continue;
		}
		float* b = bias + (uint64_t)t * ns;
		const float fn = (float)(iei - isi);
		if (ns <= blockDim.x) {
			const uint32_t nsv = (uint32_t)ns, cpb = blockDim.x / nsv, s = tid % nsv, c = tid / nsv;
			float acc = 0.f;
			for (uint32_t j0 = isi; j0 < iei; j0 += cpb) {
				const uint32_t cnt = min(cpb, iei - j0);
				if (c < cnt) s_b[c * nsv + s] = __fdiv_rn(raw[(uint64_t)iki[j0 + c] * ns + s], (float)ikmc[j0 + c]);
				__syncthreads();
				if (c == 0) {
					uint32_t q = 0;
					if (j0 == isi) { acc = s_b[s]; q = 1; }
					for (; q < cnt; ++q) acc = __fadd_rn(acc, s_b[q * nsv + s]);
				}
				__syncthreads();
			}
			if (c == 0) b[s] = __fdiv_rn(acc, fn);
		} else {
			for (uint64_t s = tid; s < ns; s += blockDim.x) {
				float acc = __fdiv_rn(raw[(uint64_t)iki[isi] * ns + s], (float)ikmc[isi]);
				for (uint32_t j = isi + 1; j < iei; ++j) acc = __fadd_rn(acc, __fdiv_rn(raw[(uint64_t)iki[j] * ns + s], (float)ikmc[j]));
				b[s] = __fdiv_rn(acc, fn);
			}
		}
		__syncthreads();
		if (tid < 32) { const float m = __fdiv_rn(eigen_sum4_warp(b, (uint32_t)ns, tid), (float)ns); if (tid == 0) s_mean = m; }
		__syncthreads();
		const float m = s_mean;
		for (uint64_t s = tid; s < ns; s += blockDim.x) b[s] = __fdiv_rn(b[s], m);
		__syncthreads();
		const float* rp = raw + o;
		float* cp = cor + o;
		if (ns <= blockDim.x) {                                          // several columns per pass, each thread keeps its sample
			const uint32_t cpb = blockDim.x / (uint32_t)ns, s = tid % (uint32_t)ns, c = tid / (uint32_t)ns;
			if (c < cpb) {
				const float bs = b[s];
				for (uint64_t k = c; k < ei - si; k += cpb) cp[k * ns + s] = __fdiv_rn(rp[k * ns + s], bs);
			}
		} else {
			for (uint64_t s = tid; s < ns; s += blockDim.x) {
				const float bs = b[s];
				for (uint64_t k = 0; k < ei - si; ++k) cp[k * ns + s] = __fdiv_rn(rp[k * ns + s], bs);
			}
		}
		__syncthreads();
	}
}

#define PR_CU(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { cudaGetLastError(); rc = dbt::abi_fail(DBT_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); goto done; } } while (0)

}  // namespace

extern "C" int dbt_pred(int device, const dbt_pred_input* in, float* gt_raw, float* gt_corrected, float* bias, float kernel_ms[2]) {
	if (!in || !gt_raw || !gt_corrected || !bias) return dbt::abi_fail(DBT_E_BADARG, "null argument");
	const uint64_t ns = in->ns, nk = in->nk, n_tr = in->n_tr, n_ik = in->n_ik;
	if (!ns || !nk || !n_tr) return dbt::abi_fail(DBT_E_BADARG, "empty input (ns=%llu nk=%llu n_tr=%llu)", (unsigned long long)ns, (unsigned long long)nk, (unsigned long long)n_tr);
	if (!in->kmc || !in->rd || !in->nk_cum || !in->nik_cum || (n_ik && (!in->iki || !in->ikmc))) return dbt::abi_fail(DBT_E_BADARG, "null array in dbt_pred_input");
	if (ns > 0xFFFFFFFFull || nk > 0xFFFFFFFFull || n_tr > 0xFFFFFFFFull) return dbt::abi_fail(DBT_E_BADARG, "ns / nk / n_tr beyond 32 bits (the ikmer table holds u32 indices)");
	// the table must be consistent: cumulative counts ascending and ending at nk / n_ik; every invariant k-mer inside its own
	// locus (the loci are corrected independently; the reference, which corrects them one after the other in place, would read
	// already-corrected counts through an index into an earlier locus)
	if (in->nk_cum[n_tr - 1] != nk || in->nik_cum[n_tr - 1] != n_ik) return dbt::abi_fail(DBT_E_INDEX, "ikmer table: cumulative counts end at %u / %u, expected nk=%llu / n_ik=%llu",
	                                                                                      in->nk_cum[n_tr - 1], in->nik_cum[n_tr - 1], (unsigned long long)nk, (unsigned long long)n_ik);
	for (uint64_t t = 0; t < n_tr; ++t) {
		const uint32_t si = t ? in->nk_cum[t - 1] : 0, ei = in->nk_cum[t], isi = t ? in->nik_cum[t - 1] : 0, iei = in->nik_cum[t];
		if (ei < si || iei < isi) return dbt::abi_fail(DBT_E_INDEX, "ikmer table: cumulative counts decrease at locus %llu", (unsigned long long)t);
		for (uint32_t j = isi; j < iei; ++j) {
			if (in->iki[j] < si || in->iki[j] >= ei) return dbt::abi_fail(DBT_E_INDEX, "ikmer table: invariant k-mer %u of locus %llu lies outside the locus' k-mers [%u, %u)", in->iki[j], (unsigned long long)t, si, ei);
			if (!in->ikmc[j]) return dbt::abi_fail(DBT_E_INDEX, "ikmer table: expected count 0 for invariant k-mer %u", in->iki[j]);
		}
	}
	for (uint64_t s = 0; s < ns; ++s) if (!in->kmc[s]) return dbt::abi_fail(DBT_E_BADARG, "null count array for sample %llu", (unsigned long long)s);

	int rc = DBT_OK;
	uint64_t* d_stage = nullptr; float *d_raw = nullptr, *d_cor = nullptr, *d_bias = nullptr, *d_rd = nullptr;
	uint32_t *d_nk = nullptr, *d_nik = nullptr, *d_iki = nullptr; uint8_t* d_ikmc = nullptr;
	cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
	cudaStream_t st = nullptr;
	cudaDeviceProp prop;
	float t_norm = 0.f;
	const uint32_t G = (uint32_t)std::min<uint64_t>(PG, ns);
	{
		cudaError_t e = cudaSetDevice(device);
		if (e != cudaSuccess) return dbt::abi_fail(DBT_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
	}
	PR_CU(cudaGetDeviceProperties(&prop, device));
	{
		size_t fr = 0, tot = 0;
		PR_CU(cudaMemGetInfo(&fr, &tot));
		const size_t need = (size_t)ns * nk * 8 + (size_t)G * nk * 8 + (size_t)ns * n_tr * 4 + n_ik * 5 + n_tr * 8 + ns * 4 + (64u << 20);
		if (need > fr) { rc = dbt::abi_fail(DBT_E_NOMEM, "dbt_pred: %llu samples x %llu k-mers need %zu bytes of device memory, %zu are free", (unsigned long long)ns, (unsigned long long)nk, need, fr); goto done; }
	}
	PR_CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
	for (auto& e : ev) PR_CU(cudaEventCreate(&e));
	PR_CU(cudaMalloc(&d_stage, (size_t)G * nk * 8));
	PR_CU(cudaMalloc(&d_raw, (size_t)ns * nk * 4));
	PR_CU(cudaMalloc(&d_cor, (size_t)ns * nk * 4));
	PR_CU(cudaMalloc(&d_bias, (size_t)ns * n_tr * 4));
	PR_CU(cudaMalloc(&d_rd, ns * 4));
	PR_CU(cudaMalloc(&d_nk, n_tr * 4)); PR_CU(cudaMalloc(&d_nik, n_tr * 4));
	PR_CU(cudaMalloc(&d_iki, std::max<uint64_t>(n_ik, 1) * 4)); PR_CU(cudaMalloc(&d_ikmc, std::max<uint64_t>(n_ik, 1)));
	PR_CU(cudaMemcpyAsync(d_rd, in->rd, ns * 4, cudaMemcpyHostToDevice, st));
	PR_CU(cudaMemcpyAsync(d_nk, in->nk_cum, n_tr * 4, cudaMemcpyHostToDevice, st));
	PR_CU(cudaMemcpyAsync(d_nik, in->nik_cum, n_tr * 4, cudaMemcpyHostToDevice, st));
	if (n_ik) { PR_CU(cudaMemcpyAsync(d_iki, in->iki, n_ik * 4, cudaMemcpyHostToDevice, st)); PR_CU(cudaMemcpyAsync(d_ikmc, in->ikmc, n_ik, cudaMemcpyHostToDevice, st)); }
	// the caller's bias matrix comes back unchanged where a locus is skipped (the reference leaves those columns uninitialised)
	PR_CU(cudaMemcpyAsync(d_bias, bias, (size_t)ns * n_tr * 4, cudaMemcpyHostToDevice, st));
	for (uint64_t s0 = 0; s0 < ns; s0 += G) {
		const uint32_t ng = (uint32_t)std::min<uint64_t>(G, ns - s0);
		for (uint32_t g = 0; g < ng; ++g) PR_CU(cudaMemcpyAsync(d_stage + (size_t)g * nk, in->kmc[s0 + g], nk * 8, cudaMemcpyHostToDevice, st));
		PR_CU(cudaEventRecord(ev[0], st));
		pred_norm_kernel<<<(unsigned)std::min<uint64_t>((nk + 31) / 32, (uint64_t)prop.multiProcessorCount * 16), 256, 0, st>>>(d_stage, nk, d_rd, (uint32_t)s0, ng, ns, d_raw);
		PR_CU(cudaGetLastError());
		PR_CU(cudaEventRecord(ev[1], st));
		PR_CU(cudaEventSynchronize(ev[1]));                         // the stage is reused by the next group
		float ms = 0.f; PR_CU(cudaEventElapsedTime(&ms, ev[0], ev[1])); t_norm += ms;
	}
	PR_CU(cudaEventRecord(ev[2], st));
	pred_bias_kernel<<<(unsigned)std::min<uint64_t>(n_tr, (uint64_t)prop.multiProcessorCount * 8), 256, 0, st>>>(d_raw, d_cor, d_bias, ns, (uint32_t)n_tr, d_nk, d_nik, d_iki, d_ikmc);
	PR_CU(cudaGetLastError());
	PR_CU(cudaEventRecord(ev[3], st));
	PR_CU(cudaMemcpyAsync(gt_raw, d_raw, (size_t)ns * nk * 4, cudaMemcpyDeviceToHost, st));
	PR_CU(cudaMemcpyAsync(gt_corrected, d_cor, (size_t)ns * nk * 4, cudaMemcpyDeviceToHost, st));
	PR_CU(cudaMemcpyAsync(bias, d_bias, (size_t)ns * n_tr * 4, cudaMemcpyDeviceToHost, st));
	PR_CU(cudaStreamSynchronize(st));
	if (kernel_ms) { kernel_ms[0] = t_norm; PR_CU(cudaEventElapsedTime(&kernel_ms[1], ev[2], ev[3])); }
done:
	cudaFree(d_stage); cudaFree(d_raw); cudaFree(d_cor); cudaFree(d_bias); cudaFree(d_rd); cudaFree(d_nk); cudaFree(d_nik); cudaFree(d_iki); cudaFree(d_ikmc);
	for (auto& e : ev) if (e) cudaEventDestroy(e);
	if (st) cudaStreamDestroy(st);
	return rc;
}
