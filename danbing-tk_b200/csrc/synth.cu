// libdbt_synth.so -- WORKLOAD GENERATOR for bench.py / tests (not part of the product library).
// Counter-based simulated read pairs: pair i is a pure function of
// (seed, i), so any shard can be regenerated on any rank.  A fraction `frac_locus` of the pairs
// is an error-free 150 bp paired-end fragment of a random haplotype (forward read + reverse-
// complemented far end, like src/sim_reads.cpp:74-104 of the reference), the rest is iid-uniform
// ACGT background.  Output = fixed-length ASCII reads in the ABI's batch order
// (read 2i = late mate = "/2", read 2i+1 = "/1").
// dbt_synth_reads2 adds sequencing errors to the locus reads for the threading (P1) runs of S2/S5:
// per emitted base a substitution with probability sub_ppm/1e6 and an insertion or deletion with
// probability indel_ppm/1e6 each half (counter-based as well: read = f(seed, pair, mate)).
#include <cstdint>
#include <cstdio>
#include <thread>
#include <vector>
#include <cuda_runtime.h>

__host__ __device__ __forceinline__ uint64_t smix(uint64_t x) {
	x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}

// pair p of the batch (global index first_pair + p): the same function on the device and on the host
__host__ __device__ __forceinline__ void synth_pair(const uint8_t* __restrict__ pool, const uint64_t* __restrict__ hap_off, uint32_t nhap,
                                                    uint64_t first_pair, uint64_t p, uint64_t seed, uint32_t locus_per_million,
                                                    uint32_t rlen, uint32_t flen, uint8_t* __restrict__ out,
                                                    uint32_t sub_ppm, uint32_t indel_ppm) {
	const uint64_t g = first_pair + p;
	uint64_t r0 = smix(seed ^ smix(g));
	uint8_t* o2 = out + p * 2 * rlen;        // read 2i   = "/2" mate (reverse strand end)
	uint8_t* o1 = o2 + rlen;                 // read 2i+1 = "/1" mate (forward)
	const char ACGT[5] = {'A', 'C', 'G', 'T', 0};
	if ((uint32_t)(r0 % 1000000u) < locus_per_million) {
		uint64_t r1 = smix(r0);
		uint32_t h = (uint32_t)(r1 % nhap);
		uint64_t b = hap_off[h], len = hap_off[h + 1] - b;
		uint64_t beg = (len > flen) ? smix(r1) % (len - flen + 1) : 0;
		const uint8_t* s = pool + b + beg;
		if (sub_ppm == 0 && indel_ppm == 0) {
			for (uint32_t i = 0; i < rlen; ++i) o1[i] = s[i];
			for (uint32_t i = 0; i < rlen; ++i) {
				uint8_t c = s[flen - 1 - i];
				o2[i] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N';
			}
		} else {
			const uint32_t span = len < flen ? (uint32_t)len : flen;       // source bases available to either mate
			for (int m = 0; m < 2; ++m) {
				uint8_t* o = m ? o2 : o1;
				uint64_t x = smix(r1 ^ (0xA5A5ull + m));
				uint32_t src = 0;
				for (uint32_t i = 0; i < rlen; ) {
					x = smix(x);
					const uint32_t u = (uint32_t)(x % 1000000u), v = (uint32_t)(x >> 40) & 3u;
					if (u < indel_ppm / 2 && src + 1 < span) { ++src; continue; }                 // deletion: skip a source base
					if (u < indel_ppm) { o[i++] = ACGT[v]; continue; }                            // insertion
					uint8_t c = src < span ? (m ? s[span - 1 - src] : s[src]) : (uint8_t)'N';
					if (m) c = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N';
					++src;
					if (u - indel_ppm < sub_ppm && c != 'N') { uint8_t d = ACGT[v]; if (d == c) d = ACGT[(v + 1) & 3]; c = d; }
					o[i++] = c;
				}
			}
		}
	} else {
		uint64_t x = r0;
		for (uint32_t i = 0; i < 2 * rlen; i += 32) {
			x = smix(x);
			for (uint32_t j = 0; j < 32 && i + j < 2 * rlen; ++j) o2[i + j] = ACGT[(x >> (2 * j)) & 3];
		}
	}
}

__global__ void synth_reads_kernel(const uint8_t* __restrict__ pool, const uint64_t* __restrict__ hap_off, uint32_t nhap,
                                   uint64_t first_pair, uint64_t npairs, uint64_t seed, uint32_t locus_per_million,
                                   uint32_t rlen, uint32_t flen, uint8_t* __restrict__ out,
                                   uint32_t sub_ppm, uint32_t indel_ppm) {
	const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (p >= npairs) return;
	synth_pair(pool, hap_off, nhap, first_pair, p, seed, locus_per_million, rlen, flen, out, sub_ppm, indel_ppm);
}

// The same generator on the host cores, without touching CUDA: bench.py's reference arm (`--impl reference`) builds its
// reads with this so that no GPU code runs anywhere in that arm.  Identical bytes to dbt_synth_reads2.
extern "C" int dbt_synth_reads_host(const uint8_t* pool, const uint64_t* hap_off, uint32_t nhap, uint64_t first_pair,
                                    uint64_t npairs, uint64_t seed, double frac_locus, uint32_t rlen, uint32_t flen, uint8_t* out,
                                    uint32_t sub_ppm, uint32_t indel_ppm, int nthreads) {
	if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
	if (nthreads < 1) nthreads = 1;
	const uint32_t lpm = (uint32_t)(frac_locus * 1e6 + 0.5);
	auto work = [&](uint64_t a, uint64_t b) {
		for (uint64_t p = a; p < b; ++p) synth_pair(pool, hap_off, nhap, first_pair, p, seed, lpm, rlen, flen, out, sub_ppm, indel_ppm);
	};
	std::vector<std::thread> th;
	const uint64_t chunk = (npairs + nthreads - 1) / nthreads;
	for (int t = 1; t < nthreads; ++t) { const uint64_t a = std::min<uint64_t>(npairs, t * chunk), b = std::min<uint64_t>(npairs, (t + 1) * chunk); if (a < b) th.emplace_back(work, a, b); }
	work(0, std::min<uint64_t>(npairs, chunk));
	for (auto& x : th) x.join();
	return 0;
}

// pool/hap_off: host pointers; out: host pointer (pinned or pageable) OR, when out_is_device != 0, a device pointer.
extern "C" int dbt_synth_reads2(int device, const uint8_t* pool, const uint64_t* hap_off, uint32_t nhap, uint64_t first_pair,
                                uint64_t npairs, uint64_t seed, double frac_locus, uint32_t rlen, uint32_t flen, uint8_t* out,
                                uint32_t sub_ppm, uint32_t indel_ppm) {
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
		                                                          (uint32_t)(frac_locus * 1e6 + 0.5), rlen, flen, dout, sub_ppm, indel_ppm);
		if (cudaMemcpy(out + p0 * 2 * rlen, dout, n * 2 * rlen, cudaMemcpyDeviceToHost) != cudaSuccess) rc = -3;
	}
	cudaFree(dpool); cudaFree(doff); cudaFree(dout);
	return rc;
}

extern "C" int dbt_synth_reads(int device, const uint8_t* pool, const uint64_t* hap_off, uint32_t nhap, uint64_t first_pair,
                               uint64_t npairs, uint64_t seed, double frac_locus, uint32_t rlen, uint32_t flen, uint8_t* out) {
	return dbt_synth_reads2(device, pool, hap_off, nhap, first_pair, npairs, seed, frac_locus, rlen, flen, out, 0, 0);
}

// PREF.*.kmers text writer (`>locus` / `kmer<TAB>value`, h:926-937 format) for the workload generator: the Python loop
// takes minutes at 20 000 loci.  vals may be NULL (value 0).
extern "C" int dbt_write_kmers_text(const char* fn, uint64_t nloci, const uint64_t* n, const uint64_t* ks, const uint8_t* vals) {
	FILE* f = fopen(fn, "wb");
	if (!f) return -1;
	std::vector<char> buf(1 << 22);
	size_t o = 0; uint64_t j = 0;
	auto put_u64 = [&](uint64_t v) { char t[24]; int q = 0; do { t[q++] = (char)('0' + v % 10); v /= 10; } while (v); while (q) buf[o++] = t[--q]; };
	for (uint64_t l = 0; l < nloci; ++l) {
		buf[o++] = '>'; put_u64(l); buf[o++] = '\n';
		for (uint64_t i = 0; i < n[l]; ++i, ++j) {
			put_u64(ks[j]); buf[o++] = '\t'; put_u64(vals ? vals[j] : 0); buf[o++] = '\n';
			if (o + 64 > buf.size()) { if (fwrite(buf.data(), 1, o, f) != o) { fclose(f); return -2; } o = 0; }
		}
		if (o + 64 > buf.size()) { if (fwrite(buf.data(), 1, o, f) != o) { fclose(f); return -2; } o = 0; }
	}
	if (o && fwrite(buf.data(), 1, o, f) != o) { fclose(f); return -2; }
	return fclose(f) ? -2 : 0;
}
