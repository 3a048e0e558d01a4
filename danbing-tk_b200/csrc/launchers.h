// host-callable launchers implemented in kernels.cu
#pragma once
#include "device_types.cuh"
namespace dbt {
// read r = ascii[starts[r], ends[r]); tile_bytes: span of 32 consecutive reads in the text (0 = concatenated reads)
cudaError_t launch_pack(cudaStream_t st, const uint8_t* ascii, const uint64_t* starts, const uint64_t* ends, uint64_t npairs, const RunParams& rp,
                        uint32_t tile_bytes, uint64_t* p0, uint64_t* p1, uint16_t* m0, uint16_t* m1, uint32_t* err);
cudaError_t launch_compact_bits(cudaStream_t st, const uint32_t* mask, uint64_t nbits, uint32_t* blocksum, uint32_t* list, uint32_t* count, uint32_t cap);
// raw FASTA/FASTQ text -> newline list, records, adjacent-mate pairs (fastx.cu)
cudaError_t launch_fastx_split(cudaStream_t st, int nsm, const RawView& v, uint64_t nbytes, uint32_t lines_per_record, uint32_t min_len);
cudaError_t launch_subfilter(cudaStream_t st, int B, int nsm, const DevIndex& ix, const RunParams& rp, const SlotView& sv);
cudaError_t launch_compact(cudaStream_t st, const SlotView& sv);
cudaError_t launch_record_emit(cudaStream_t st, int nsm, const uint8_t* records, uint32_t stride, const uint32_t* nslots, uint32_t cap, int emit_failed,
                               uint32_t* mask, uint32_t* blocksum, uint32_t* list, uint32_t* nemit, uint8_t* out);
cudaError_t launch_candidate(cudaStream_t st, int B, unsigned grid_fast, unsigned grid_slow, const DevIndex& ix, const RunParams& rp, const SlotView& sv, cudaEvent_t mid);
cudaError_t launch_prefilter_build(cudaStream_t st, const uint64_t* keys, uint64_t n, uint32_t* bitmap, uint32_t bits);   // presence bitmap of the index keys
cudaError_t launch_insert(cudaStream_t st, int B, Table t, const uint64_t* keys, const uint32_t* vals, uint64_t n, uint32_t* err);
cudaError_t launch_fill_aux(cudaStream_t st, int B, Table t, const uint32_t* vv, uint32_t* vv_aux, const uint64_t* kmers,
                            const uint32_t* locus_of, uint64_t n, uint32_t cls, uint32_t* err);
cudaError_t launch_thread(cudaStream_t st, int B, int nsm, const DevIndex& ix, const RunParams& rp, const SlotView& sv, cudaEvent_t mid);   // thread_fast_kernel, event `mid`, thread_kernel
// pass 0/1: graph nodes (canonical / reverse-complement orientation), pass 2: TR k-mers (masks == nullptr), pass 3: verify
cudaError_t launch_graph_build(cudaStream_t st, int B, Table t, const uint64_t* kmers, const uint8_t* masks, const uint32_t* locus_of,
                               uint64_t n, uint32_t k, int pass, uint32_t* err);
// graphs merged into the index entries (graph_merge_kernel) + the check that every TR k-mer keeps its count position there
cudaError_t launch_graph_merge(cudaStream_t st, int B, Table t, const uint32_t* vv, uint32_t* vv_g, const uint64_t* kmers, const uint8_t* masks,
                               const uint32_t* locus_of, uint64_t n, uint32_t k, uint32_t* err);
cudaError_t launch_tr_class_check(cudaStream_t st, int B, Table t, const uint32_t* vv, const uint32_t* vv_aux, const uint64_t* kmers, const uint32_t* locus_of, uint64_t n, uint32_t* err);
cudaError_t launch_bait_insert(cudaStream_t st, int B, Table t, const uint64_t* kmers, const uint16_t* vals, const uint32_t* locus_of, uint64_t n, uint32_t* err);   // `-b` (bait.cuh)
cudaError_t launch_table_stats(cudaStream_t st, int B, Table t, unsigned long long* out);
cudaError_t launch_random_gather(cudaStream_t st, int nsm, const uint4* table, uint64_t table_bytes, uint32_t sector_bytes,
                                 uint64_t ngathers, uint64_t seed, uint32_t* sink);
// `-bu` (bubbles.cu): PREF.tre.kdb -> edge table; novel-edge events of the COUNTED records of a batch
cudaError_t launch_edge_insert(cudaStream_t st, int B, Table t, const uint64_t* edges, const uint32_t* locus_of, uint64_t n, uint32_t k1, uint32_t* err);
cudaError_t launch_bubbles(cudaStream_t st, int B, int nsm, Table tt, const RunParams& rp, const SlotView& sv, uint32_t nloci,
                           uint64_t* ev_key, uint64_t* ev_edge, uint32_t* ev_locus, uint32_t* nev, uint32_t cap);
constexpr int COMPACT_WORDS_H = 1024;
}
