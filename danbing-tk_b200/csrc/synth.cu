// libdbt_synth.so -- WORKLOAD GENERATOR for bench.py / tests (not part of the product library).
// Counter-based simulated read pairs: pair i is a pure function of
// (seed, i), so any shard can be regenerated on any rank.  A fraction `frac_locus` of the pairs
// is an error-free 150 bp paired-end fragment of a random haplotype (forward read + reverse-
// complemented far end, like src/sim_reads.cpp:74-104 of the reference), the rest is iid-uniform
// ACGT background.  Output = fixed-length ASCII reads in the ABI's batch order
// (read 2i = late mate = "/2", read 2i+1 = "/1").
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t smix(uint64_t x) {
	x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}

__global__ void synth_reads_kernel(const uint8_t* __restrict__ pool, const uint64_t* __restrict__ hap_off, uint32_t nhap,
                                   uint64_t first_pair, uint64_t npairs, uint64_t seed, uint32_t locus_per_million,
                                   uint32_t rlen, uint32_t flen, uint8_t* __restrict__ out) {
	const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= npairs) return;
	const uint64_t g = first_pair + p;
	uint64_t r0 = smix(seed ^ smix(g));
	uint8_t* o2 = out + p * 2 * rlen;        // read 2i   = "/2" mate (reverse strand end)
	uint8_t* o1 = o2 + rlen;                 // read 2i+1 = "/1" mate (forward)
	const char* ACGT = "ACGT";
	if ((uint32_t)(r0 % 1000000u) < locus_per_million) {
		uint64_t r1 = smix(r0);
		uint32_t h = (uint32_t)(r1 % nhap);
		uint64_t b = hap_off[h], len = hap_off[h + 1] - b;
		uint64_t beg = (len > flen) ? smix(r1) % (len - flen + 1) : 0;
		const uint8_t* s = pool + b + beg;
		for (uint32_t i = 0; i < rlen; ++i) o1[i] = s[i];
		for (uint32_t i = 0; i < rlen; ++i) {
			uint8_t c = s[flen - 1 - i];
			o2[i] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N';
		}
	} else {
		uint64_t x = r0;
		for (uint32_t i = 0; i < 2 * rlen; i += 32) {
			x = smix(x);
			for (uint32_t j = 0; j < 32 && i + j < 2 * rlen; ++j) o2[i + j] = ACGT[(x >> (2 * j)) & 3];
		}
	}
}

// pool/hap_off: host pointers; out: host pointer (pinned or pageable) OR, when out_is_device != 0, a device pointer.
extern "C" int dbt_synth_reads(int device, const uint8_t* pool, const uint64_t* hap_off, uint32_t nhap, uint64_t first_pair,
                               uint64_t npairs, uint64_t seed, double frac_locus, uint32_t rlen, uint32_t flen, uint8_t* out) {
	if (cudaSetDevice(device) != cudaSuccess) return -1;
	uint8_t* dpool = nullptr; uint64_t* doff = nullptr; uint8_t* dout = nullptr;
	const uint64_t pool_bytes = hap_off[nhap];
	if (cudaMalloc(&dpool, pool_bytes) != cudaSuccess || cudaMalloc(&doff, (nhap + 1) * 8) != cudaSuccess) return -2;
	cudaMemcpy(dpool, pool, pool_bytes, cudaMemcpyHostToDevice);
	cudaMemcpy(doff, hap_off, (nhap + 1) * 8, cudaMemcpyHostToDevice);
	const uint64_t chunk = 1ull << 22;
	if (cudaMalloc(&dout, chunk * 2 * rlen) != cudaSuccess) return -2;
	int rc = 0;
	for (uint64_t p0 = 0; p0 < npairs && !rc; p0 += chunk) {
		uint64_t n = npairs - p0 < chunk ? npairs - p0 : chunk;
		synth_reads_kernel<<<(unsigned)((n + 127) / 128), 128>>>(dpool, doff, nhap, first_pair + p0, n, seed,
		                                                          (uint32_t)(frac_locus * 1e6 + 0.5), rlen, flen, dout);
		if (cudaMemcpy(out + p0 * 2 * rlen, dout, n * 2 * rlen, cudaMemcpyDeviceToHost) != cudaSuccess) rc = -3;
	}
	cudaFree(dpool); cudaFree(doff); cudaFree(dout);
	return rc;
}
