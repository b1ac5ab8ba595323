// coral_b200/csrc/relops.cuh — small relational helpers (stable radix sort, scan, ordered compaction)
// for the candidate/cluster tables.  CUB device primitives are library code; the kernels that carry
// the reference's semantics are ours and live in the .cu files.
#pragma once
#include <cub/cub.cuh>

#include "ctx.h"

namespace coral {

// Stable sort of (key u64, value u32) pairs on bits [begin_bit, end_bit).
inline int sort_pairs_u64(coral_ctx* c, const uint64_t* k_in, uint64_t* k_out, const uint32_t* v_in, uint32_t* v_out,
                          int64_t n, int begin_bit = 0, int end_bit = 64) {
    if (n <= 0) return CORAL_OK;
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, bytes, k_in, k_out, v_in, v_out, (int)n, begin_bit, end_bit, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, bytes, k_in, k_out, v_in, v_out, (int)n, begin_bit, end_bit, c->stream));
    c->n_launches += 3;
    return CORAL_OK;
}

inline int exclusive_sum_u32(coral_ctx* c, const uint32_t* in, uint32_t* out, int64_t n) {
    if (n <= 0) return CORAL_OK;
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int)n, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceScan::ExclusiveSum(c->cub_tmp.p, bytes, in, out, (int)n, c->stream));
    c->n_launches += 1;
    return CORAL_OK;
}

inline int inclusive_sum_u32(coral_ctx* c, const uint32_t* in, uint32_t* out, int64_t n) {
    if (n <= 0) return CORAL_OK;
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(nullptr, bytes, in, out, (int)n, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(c->cub_tmp.p, bytes, in, out, (int)n, c->stream));
    c->n_launches += 1;
    return CORAL_OK;
}

// out[j] = i for every i in [0, n) with flags[i] != 0, in increasing i; *d_count = number selected.
inline int select_flagged_iota(coral_ctx* c, const uint8_t* flags, uint32_t* out, uint32_t* d_count, int64_t n) {
    if (n <= 0) {
        CORAL_CUDA(c, cudaMemsetAsync(d_count, 0, 4, c->stream));
        return CORAL_OK;
    }
    cub::CountingInputIterator<uint32_t> iota(0);
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceSelect::Flagged(nullptr, bytes, iota, flags, out, d_count, (int)n, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceSelect::Flagged(c->cub_tmp.p, bytes, iota, flags, out, d_count, (int)n, c->stream));
    c->n_launches += 1;
    return CORAL_OK;
}

template <class T>
inline int d2h(coral_ctx* c, T* dst, const void* src, size_t n) {
    if (n == 0) return CORAL_OK;
    CORAL_CUDA(c, cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    return CORAL_OK;
}
template <class T>
inline int h2d(coral_ctx* c, void* dst, const T* src, size_t n) {
    if (n == 0) return CORAL_OK;
    CORAL_CUDA(c, cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return CORAL_OK;
}

}  // namespace coral
