// coral_b200/csrc/dev_util.cuh — small device helpers shared by the kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#define FULL_MASK 0xffffffffu

namespace coral {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// Position-keyed additive 128-bit string hash: h = sum_i mix(K ^ (i << 8 | byte_i)) + mix(len).
// Additive => lanes can hash disjoint byte subsets and warp-reduce.
#define CORAL_HK1 0x243F6A8885A308D3ull
#define CORAL_HK2 0x13198A2E03707344ull
__host__ __device__ __forceinline__ uint64_t hterm1(uint32_t i, uint32_t b) { return mix64(CORAL_HK1 ^ (((uint64_t)i << 8) | b)); }
__host__ __device__ __forceinline__ uint64_t hterm2(uint32_t i, uint32_t b) { return mix64(CORAL_HK2 ^ (((uint64_t)i << 8) | b)); }

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }

// 32-bit little-endian load from an arbitrarily aligned address (two aligned word loads).
// The record buffer carries >= 64 bytes of tail padding, so the second word is always readable.
__device__ __forceinline__ uint32_t ld_u32_any(const uint8_t* p) {
    uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
    uint32_t sh = (uint32_t)(a & 3) * 8;
    uint32_t lo = __ldg(w);
    if (sh == 0) return lo;
    uint32_t hi = __ldg(w + 1);
    return __funnelshift_r(lo, hi, sh);
}

__device__ __forceinline__ int64_t warp_sum_i64(int64_t v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL_MASK, v, d);
    return v;
}
__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL_MASK, v, d);
    return v;
}
__device__ __forceinline__ int32_t warp_max_i32(int32_t v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = max(v, __shfl_xor_sync(FULL_MASK, v, d));
    return v;
}
// inclusive prefix sum across the warp
__device__ __forceinline__ int32_t warp_incl_sum_i32(int32_t v) {
    const uint32_t lane = lane_id();
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int32_t t = __shfl_up_sync(FULL_MASK, v, d);
        if (lane >= (uint32_t)d) v += t;
    }
    return v;
}
// inclusive prefix max across the warp
__device__ __forceinline__ int32_t warp_incl_max_i32(int32_t v) {
    const uint32_t lane = lane_id();
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int32_t t = __shfl_up_sync(FULL_MASK, v, d);
        if (lane >= (uint32_t)d) v = max(v, t);
    }
    return v;
}

// number of nibbles in `w` (16 nibbles) whose value is NOT one of {1,2,4,8} (BAM codes A,C,G,T)
__device__ __forceinline__ uint32_t non_acgt_nibbles(uint64_t w) {
    uint64_t c = w - ((w >> 1) & 0x7777777777777777ull) - ((w >> 2) & 0x3333333333333333ull) -
                 ((w >> 3) & 0x1111111111111111ull);          // per-nibble popcount
    uint64_t t = c ^ 0x1111111111111111ull;                    // zero nibble <=> popcount == 1
    uint64_t z = (t | (t >> 1) | (t >> 2) | (t >> 3)) & 0x1111111111111111ull;
    return (uint32_t)__popcll(z);
}
#endif  // __CUDACC__

}  // namespace coral
