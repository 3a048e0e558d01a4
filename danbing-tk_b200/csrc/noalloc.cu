// dbt_pack_reads_device_noalloc: ASCII reads -> 2-bit packed records on the GPU without allocating anything.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is pack_kernel (kernels.cu) run on a FIXED
// device scratch area: the staged ASCII bytes, the read offsets, the packed records, the meta words and the error word all live
// in one file-scope `static __device__` array of DBT_NOALLOC_SCRATCH_BYTES, which every GPU that loads the library holds until
// the process exits.  The caller passes host inputs and a host output buffer, laid out exactly like dbt_pack_reads' (the host
// packer), so the two entry points are interchangeable and are compared bit for bit in tests/test_zz_no_alloc.py.
//
// The entry point calls no cudaMalloc / cudaFree / cudaMallocHost / stream or event creation: copies and the launch go to the
// legacy default stream of the current device.  All calls share the array, which keeps the previous call's contents -- the error
// word is cleared by every call, everything else is overwritten before it is read.  Calls that overlap on one GPU race: a mutex
// serialises callers inside this process; work another process (or another library) has in flight on the same GPU is the
// caller's business.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include "../../include/danbing_b200.h"
#include "device_types.cuh"
#include "launchers.h"

namespace dbt { int abi_fail(int code, const char* fmt, ...); }   // dbt_abi.cu: sets dbt_last_error(), returns code

// the scratch itself.  In HOST code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address a
// kernel may use comes from cudaGetSymbolAddress (scratch_of below).
static __device__ __align__(256) uint8_t g_noalloc_scratch[DBT_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
uint8_t* g_scratch_ptr[MAX_DEVICES] = {};      // device address of the array on each device, resolved once per device

// device address of the scratch array on `device` (which must be current)
int scratch_of(int device, uint8_t** out) {
	if (device < 0 || device >= MAX_DEVICES) return dbt::abi_fail(DBT_E_BADARG, "device %d out of range", device);
	if (!g_scratch_ptr[device]) {
		void* p = nullptr;
		cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if (e != cudaSuccess || !p) {
			cudaGetLastError();
			return dbt::abi_fail(DBT_E_CUDA, "cudaGetSymbolAddress(no-alloc scratch) on device %d: %s", device, cudaGetErrorString(e));
		}
		g_scratch_ptr[device] = static_cast<uint8_t*>(p);
	}
	*out = g_scratch_ptr[device];
	return DBT_OK;
}

size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

#define NA_CU(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { cudaGetLastError(); return dbt::abi_fail(DBT_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); } } while (0)

extern "C" size_t dbt_noalloc_scratch_needed(uint32_t max_read_len, uint64_t npairs, uint64_t nbases) {
	const dbt::ReadLayout lay = dbt::make_layout(max_read_len);
	return up256(nbases + 64) + up256((2 * npairs + 1) * 8) + 2 * up256(npairs * lay.W64 * 8) + 2 * up256(npairs * 2) + 256;
}

extern "C" int dbt_pack_reads_device_noalloc(int device, uint32_t k, uint32_t max_read_len, uint64_t npairs, const uint8_t* ascii,
                                             const uint64_t* offsets, uint8_t* packed_out) {
	if (!ascii || !offsets || !packed_out) return dbt::abi_fail(DBT_E_BADARG, "null argument");
	if (k < 1 || k > 32 || max_read_len < k || max_read_len > 0x7FFF) return dbt::abi_fail(DBT_E_BADARG, "k=%u / max_read_len=%u out of range", k, max_read_len);
	if (!npairs) return DBT_OK;
	if (offsets[2 * npairs] < offsets[0]) return dbt::abi_fail(DBT_E_BADARG, "offsets are not ascending");
	const uint64_t nbases = offsets[2 * npairs] - offsets[0];
	// the whole job must fit the fixed scratch: refuse anything larger, never fall back to an allocation
	if (npairs > DBT_NOALLOC_MAX_PAIRS) return dbt::abi_fail(DBT_E_CAPACITY, "dbt_pack_reads_device_noalloc: %llu pairs, the fixed scratch takes at most %u per call",
	                                                         (unsigned long long)npairs, (unsigned)DBT_NOALLOC_MAX_PAIRS);
	const size_t need = dbt_noalloc_scratch_needed(max_read_len, npairs, nbases);
	if (need > (size_t)DBT_NOALLOC_SCRATCH_BYTES)
		return dbt::abi_fail(DBT_E_CAPACITY, "dbt_pack_reads_device_noalloc: %llu pairs / %llu bases at max_read_len %u need %zu bytes of scratch, the fixed array holds %zu",
		                     (unsigned long long)npairs, (unsigned long long)nbases, max_read_len, need, (size_t)DBT_NOALLOC_SCRATCH_BYTES);
	for (uint64_t r = 0; r < 2 * npairs; ++r)
		if (offsets[r + 1] < offsets[r]) return dbt::abi_fail(DBT_E_BADARG, "offsets are not ascending at read %llu", (unsigned long long)r);

	std::lock_guard<std::mutex> lock(g_mu);
	NA_CU(cudaSetDevice(device));
	uint8_t* base = nullptr;
	if (int rc = scratch_of(device, &base)) return rc;

	const dbt::ReadLayout lay = dbt::make_layout(max_read_len);
	// carve-up (every part 256-byte aligned): error word | ASCII (+64 over-read) | offsets | packed0 | packed1 | meta0 | meta1
	uint8_t* q = base;
	uint32_t* d_err = reinterpret_cast<uint32_t*>(q);            q += 256;
	uint8_t* d_ascii = q;                                        q += up256(nbases + 64);
	uint64_t* d_off = reinterpret_cast<uint64_t*>(q);            q += up256((2 * npairs + 1) * 8);
	uint64_t* d_p0 = reinterpret_cast<uint64_t*>(q);             q += up256(npairs * lay.W64 * 8);
	uint64_t* d_p1 = reinterpret_cast<uint64_t*>(q);             q += up256(npairs * lay.W64 * 8);
	uint16_t* d_m0 = reinterpret_cast<uint16_t*>(q);             q += up256(npairs * 2);
	uint16_t* d_m1 = reinterpret_cast<uint16_t*>(q);             q += up256(npairs * 2);
	if ((size_t)(q - base) > (size_t)DBT_NOALLOC_SCRATCH_BYTES) return dbt::abi_fail(DBT_E_CAPACITY, "no-alloc scratch carve-up overflow");

	cudaStream_t st = 0;                                          // legacy default stream: nothing to create
	NA_CU(cudaMemsetAsync(d_err, 0, 256, st));                    // the one thing that must start at zero (the array keeps old contents)
	NA_CU(cudaMemcpyAsync(d_ascii, ascii + offsets[0], nbases, cudaMemcpyHostToDevice, st));
	NA_CU(cudaMemcpyAsync(d_off, offsets, (2 * npairs + 1) * 8, cudaMemcpyHostToDevice, st));
	dbt::RunParams rp{};
	rp.k = k; rp.lay = lay; rp.max_read_len = max_read_len;
	// offsets stay relative to the caller's `ascii`: shift the base pointer instead (as dbt_submit does)
	NA_CU(dbt::launch_pack(st, d_ascii - offsets[0], d_off, d_off + 1, npairs, rp, 0, d_p0, d_p1, d_m0, d_m1, d_err));
	uint64_t* h_p0 = reinterpret_cast<uint64_t*>(packed_out);
	uint64_t* h_p1 = h_p0 + npairs * lay.W64;
	uint16_t* h_m0 = reinterpret_cast<uint16_t*>(h_p1 + npairs * lay.W64);
	uint16_t* h_m1 = h_m0 + npairs;
	uint32_t h_err = 0;
	NA_CU(cudaMemcpyAsync(h_p0, d_p0, npairs * lay.W64 * 8, cudaMemcpyDeviceToHost, st));
	NA_CU(cudaMemcpyAsync(h_p1, d_p1, npairs * lay.W64 * 8, cudaMemcpyDeviceToHost, st));
	NA_CU(cudaMemcpyAsync(h_m0, d_m0, npairs * 2, cudaMemcpyDeviceToHost, st));
	NA_CU(cudaMemcpyAsync(h_m1, d_m1, npairs * 2, cudaMemcpyDeviceToHost, st));
	NA_CU(cudaMemcpyAsync(&h_err, d_err, 4, cudaMemcpyDeviceToHost, st));
	NA_CU(cudaStreamSynchronize(st));
	if (h_err & dbt::ERR_READ_TOO_LONG) return dbt::abi_fail(DBT_E_CAPACITY, "a read is longer than max_read_len=%u", max_read_len);
	if (h_err) return dbt::abi_fail(DBT_E_CUDA, "pack_kernel reported error bits 0x%x", h_err);
	return DBT_OK;
}
