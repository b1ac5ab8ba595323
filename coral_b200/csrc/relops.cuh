// coral_b200/csrc/relops.cuh — small relational helpers (stable radix sort, scan, ordered compaction)
// for the candidate/cluster tables.  CUB device primitives are library code; the kernels that carry
// the reference's semantics are ours and live in the .cu files.
#pragma once
#include <cub/cub.cuh>

#include <cstring>

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

// ---- pipelined pass helpers (PassState, ctx.h) ---------------------------------------------------------------------
static __global__ void k_pass_check(const void* d, unsigned long long expect, int nbytes, uint32_t* flag) {
    const unsigned long long v = nbytes == 8 ? *static_cast<const unsigned long long*>(d)
                                             : (unsigned long long)*static_cast<const uint32_t*>(d);
    if (v != expect) atomicAdd(flag, 1u);
}

inline uint64_t pass_hash(uint64_t h, const void* p, size_t bytes) {      // FNV-1a over host input arrays
    const uint8_t* b = static_cast<const uint8_t*>(p);
    for (size_t i = 0; i < bytes; i++) { h ^= b[i]; h *= 0x100000001B3ull; }
    return h;
}
constexpr uint64_t PASS_H0 = 0xCBF29CE484222325ull;

// Every entry point that can run inside a pass announces itself with a key over its host inputs.  A replayed pass
// whose call sequence departs from the recorded one continues as a recording pass from that call on.
inline void pass_call(coral_ctx* c, uint64_t key) {
    PassState& P = c->pass;
    if (!P.active) return;
    if (P.replay) {
        if (P.kcur < P.keys.size() && P.keys[P.kcur] == key) { P.kcur++; return; }
        P.replay = false;
        P.vals.resize(P.cur);
        P.keys.resize(P.kcur);
    }
    P.keys.push_back(key);
    P.kcur = P.keys.size();
}

// A device scalar (4 or 8 bytes) that sizes later launches.  Outside a pass and while recording: read back now.
// While replaying: the recorded value, checked on the device.
template <class T>
inline int pass_resolve(coral_ctx* c, T* out, const void* d_src) {
    static_assert(sizeof(T) == 4 || sizeof(T) == 8, "4- or 8-byte scalars");
    PassState& P = c->pass;
    if (P.active && P.replay && P.cur < P.vals.size()) {
        const uint64_t v = P.vals[P.cur++];
        T t;
        memcpy(&t, &v, sizeof(T));
        *out = t;
        k_pass_check<<<1, 1, 0, c->stream>>>(d_src, (unsigned long long)v, (int)sizeof(T), P.d_flag);
        c->n_launches++;
        return CORAL_OK;
    }
    if (P.active && P.replay) {       // the recorded sequence is shorter than this pass: record from here on
        P.replay = false;
        P.keys.resize(P.kcur);
    }
    CORAL_CUDA(c, cudaMemcpyAsync(out, d_src, sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    if (P.active) {
        uint64_t v = 0;
        memcpy(&v, out, sizeof(T));
        P.vals.push_back(v);
        P.cur = P.vals.size();
    }
    return CORAL_OK;
}

inline uint8_t* pass_stage(coral_ctx* c, size_t bytes) {          // page-locked staging for one deferred copy
    PassState& P = c->pass;
    bytes = (bytes + 63) & ~(size_t)63;
    for (auto& ch : P.pinned)
        if (ch.cap - ch.used >= bytes) { uint8_t* r = ch.p + ch.used; ch.used += bytes; return r; }
    const size_t cap = std::max<size_t>(bytes, (size_t)8 << 20);
    uint8_t* p = nullptr;
    if (cudaHostAlloc(reinterpret_cast<void**>(&p), cap, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    P.pinned.push_back(PassState::Chunk{p, cap, bytes});
    return p;
}

// Copy of a RESULT to the caller's array: now (the caller synchronises) or, inside a pass, at coral_pass_end.
template <class T>
inline int d2h_result(coral_ctx* c, T* dst, const void* src, size_t n) {
    if (n == 0) return CORAL_OK;
    PassState& P = c->pass;
    if (P.active) {
        uint8_t* st = pass_stage(c, n * sizeof(T));
        if (!st) { c->err = "cudaHostAlloc failed for the pass staging buffer"; return CORAL_E_NOMEM; }
        CORAL_CUDA(c, cudaMemcpyAsync(st, src, n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
        P.copies.push_back(PassState::Copy{dst, st, n * sizeof(T)});
        return CORAL_OK;
    }
    CORAL_CUDA(c, cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    return CORAL_OK;
}
// The synchronisation that ends an entry point: skipped inside a pass (coral_pass_end synchronises once).
inline int pass_sync(coral_ctx* c) {
    if (c->pass.active) return CORAL_OK;
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}
// Host post-processing of results that were just copied: now, or after the pass's synchronisation.
inline int pass_after(coral_ctx* c, std::function<int()> fn) {
    if (c->pass.active) { c->pass.after.push_back(std::move(fn)); return CORAL_OK; }
    return fn();
}

// Immediate copy (the caller synchronises right after): entry points that do not take part in pipelined passes.
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
