// Host read packer (host_pack.cpp): one read -> one packed record + meta word, the same bytes pack_kernel writes.
#pragma once
#include <stdint.h>
#include "kmer.cuh"

namespace dbt {
enum { HOST_PACK_AUTO = 0, HOST_PACK_SWAR = 1 };   // AUTO: AVX2 when the CPU has it
// avail >= len: bytes readable from s (lets the SIMD path load whole vectors across the end of the read)
void host_pack_read(const uint8_t* s, uint32_t len, uint64_t avail, uint32_t k, const ReadLayout& lay, uint64_t* out, uint16_t* meta, int impl);
}
