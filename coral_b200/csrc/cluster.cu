// coral_b200/csrc/cluster.cu — breakpoint clustering and exact-breakpoint calling.
//
//  * cluster_bp_list (reference breakpoint_utilities.py:252-286): group by (chr1, chr2, o1, o2) in
//    first-appearance order, then ORDER-DEPENDENT greedy first-fit with a Chebyshev window.  Exact GPU
//    form: candidates in different connected components of the "both |d| < cutoff"
//    graph never interact, so sort by pos1 / cut, sort by pos2 / cut => (super-)components; a component
//    whose extent is < cutoff in both coordinates is a clique and yields one cluster whatever the order;
//    any other component runs the literal greedy, in input order, inside one thread block.
//  * bpc2bp + bp_match + the peel loop (breakpoint_utilities.py:299-416, infer_breakpoint_graph.py:
//    698-716): one block per cluster; 128-bit integer sums, correctly rounded int->double, FMA-free
//    double arithmetic, unique-mode / median position calling on a block-wide bitonic sort.
#include <vector>

#include "ctx.h"
#include "dev_util.cuh"
#include "relops.cuh"

namespace coral {

struct CandIn {
    const int32_t *chr1, *pos1, *chr2, *pos2, *rgap, *mq1, *mq2;
    const uint8_t *o1, *o2, *flag;
};

__global__ void k_group_key(int64_t n, CandIn C, uint64_t* __restrict__ key, uint32_t* __restrict__ idx) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    key[i] = ((uint64_t)(uint32_t)(C.chr1[i] & 0xffff) << 18) | ((uint64_t)(uint32_t)(C.chr2[i] & 0xffff) << 2) |
             ((uint64_t)(C.o1[i] & 1) << 1) | (uint64_t)(C.o2[i] & 1);
    idx[i] = (uint32_t)i;
}

// generic: head[i] = (i == 0) || key[i] != key[i-1]
__global__ void k_key_heads(int64_t n, const uint64_t* __restrict__ key, uint32_t* __restrict__ head) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    head[i] = (i == 0 || key[i] != key[i - 1]) ? 1u : 0u;
}

// from sorted-by-group order: dense gid per candidate, per-group first index / start offset
__global__ void k_group_assign(int64_t n, const uint32_t* __restrict__ sidx, const uint32_t* __restrict__ head_incl,
                               uint32_t* __restrict__ gid, uint32_t* __restrict__ gstart, uint32_t* __restrict__ gfirst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = head_incl[i] - 1;
    gid[sidx[i]] = g;
    if (i == 0 || head_incl[i - 1] != head_incl[i]) { gstart[g] = (uint32_t)i; gfirst[g] = sidx[i]; }
}

// key = (seg id << 32) | biased coordinate, for candidates in arbitrary order `idx`
__global__ void k_seg_pos_key(int64_t n, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ seg,
                              const int32_t* __restrict__ pos, uint64_t* __restrict__ key) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t c = idx[i];
    key[i] = ((uint64_t)seg[c] << 32) | (uint64_t)((uint32_t)pos[c] ^ 0x80000000u);
}

// cut where the segment changes, or (segment belongs to a big group and the coordinate gap >= cutoff)
__global__ void k_gap_heads(int64_t n, const uint64_t* __restrict__ key, const uint32_t* __restrict__ idx,
                            const uint32_t* __restrict__ gid, const uint8_t* __restrict__ gbig, int32_t cutoff,
                            uint32_t* __restrict__ head) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == 0) { head[0] = 1; return; }
    const uint64_t a = key[i - 1], b = key[i];
    bool cut = (a >> 32) != (b >> 32);
    if (!cut && gbig[gid[idx[i]]]) {
        const int64_t pa = (int64_t)(int32_t)((uint32_t)a ^ 0x80000000u), pb = (int64_t)(int32_t)((uint32_t)b ^ 0x80000000u);
        cut = (pb - pa) >= (int64_t)cutoff;
    }
    head[i] = cut ? 1u : 0u;
}

__global__ void k_scatter_seg(int64_t n, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ head_incl,
                              uint32_t* __restrict__ seg) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) seg[idx[i]] = head_incl[i] - 1;
}

__global__ void k_seg_idx_key(int64_t n, const uint32_t* __restrict__ seg, uint64_t* __restrict__ key, uint32_t* __restrict__ idx) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    key[i] = ((uint64_t)seg[i] << 32) | (uint64_t)i;
    idx[i] = (uint32_t)i;
}

__global__ void k_seg_starts(int64_t n, const uint64_t* __restrict__ key, uint32_t* __restrict__ start, uint32_t n_seg) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == 0 || (key[i] >> 32) != (key[i - 1] >> 32)) start[key[i] >> 32] = (uint32_t)i;
    if (i == n - 1) start[n_seg] = (uint32_t)n;
}

__global__ void k_group_big(int64_t n_groups, const uint32_t* __restrict__ gstart, int64_t n, double min_cluster_size,
                            uint8_t* __restrict__ gbig) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_groups) return;
    const uint32_t e = (g + 1 < n_groups) ? gstart[g + 1] : (uint32_t)n;
    gbig[g] = ((double)(e - gstart[g]) >= min_cluster_size) ? 1 : 0;  // len(group) >= min_cluster_size, bu:266
}

// One block per component (members in input order).  Small groups and cliques: one cluster.
// Otherwise the literal greedy first-fit of bu:267-283; label = input index of the cluster's first member.
constexpr int CL_THREADS = 256;
__global__ void __launch_bounds__(CL_THREADS)
k_component_cluster(const uint32_t* __restrict__ comp_start, const uint32_t* __restrict__ members, CandIn C,
                    const uint32_t* __restrict__ gid, const uint8_t* __restrict__ gbig, int32_t cutoff,
                    uint32_t* __restrict__ label, uint32_t* __restrict__ scratch) {
    __shared__ int32_t smin1[CL_THREADS], smax1[CL_THREADS], smin2[CL_THREADS], smax2[CL_THREADS];
    __shared__ uint32_t sbest[CL_THREADS];
    const uint32_t b = comp_start[blockIdx.x], e = comp_start[blockIdx.x + 1];
    const uint32_t k = e - b;
    const uint32_t first = members[b];
    bool single = !gbig[gid[first]];
    if (!single) {
        int32_t mn1 = INT32_MAX, mx1 = INT32_MIN, mn2 = INT32_MAX, mx2 = INT32_MIN;
        for (uint32_t i = threadIdx.x; i < k; i += CL_THREADS) {
            const uint32_t m = members[b + i];
            mn1 = min(mn1, C.pos1[m]); mx1 = max(mx1, C.pos1[m]);
            mn2 = min(mn2, C.pos2[m]); mx2 = max(mx2, C.pos2[m]);
        }
        smin1[threadIdx.x] = mn1; smax1[threadIdx.x] = mx1; smin2[threadIdx.x] = mn2; smax2[threadIdx.x] = mx2;
        __syncthreads();
        for (int d = CL_THREADS / 2; d > 0; d >>= 1) {
            if (threadIdx.x < d) {
                smin1[threadIdx.x] = min(smin1[threadIdx.x], smin1[threadIdx.x + d]);
                smax1[threadIdx.x] = max(smax1[threadIdx.x], smax1[threadIdx.x + d]);
                smin2[threadIdx.x] = min(smin2[threadIdx.x], smin2[threadIdx.x + d]);
                smax2[threadIdx.x] = max(smax2[threadIdx.x], smax2[threadIdx.x + d]);
            }
            __syncthreads();
        }
        single = ((int64_t)smax1[0] - smin1[0] < cutoff) && ((int64_t)smax2[0] - smin2[0] < cutoff);
    }
    if (single) {
        for (uint32_t i = threadIdx.x; i < k; i += CL_THREADS) label[members[b + i]] = first;
        return;
    }
    // greedy: scratch[b + i] = label of member i (visible to the whole block after each barrier)
    for (uint32_t i = 0; i < k; i++) {
        const uint32_t m = members[b + i];
        const int32_t p1 = C.pos1[m], p2 = C.pos2[m];
        uint32_t best = 0xFFFFFFFFu;
        for (uint32_t j = threadIdx.x; j < i; j += CL_THREADS) {
            const uint32_t o = members[b + j];
            int64_t d1 = (int64_t)p1 - C.pos1[o], d2 = (int64_t)p2 - C.pos2[o];
            if (d1 < 0) d1 = -d1;
            if (d2 < 0) d2 = -d2;
            if (d1 < cutoff && d2 < cutoff) best = min(best, scratch[b + j]);
        }
        sbest[threadIdx.x] = best;
        __syncthreads();
        for (int d = CL_THREADS / 2; d > 0; d >>= 1) {
            if (threadIdx.x < d) sbest[threadIdx.x] = min(sbest[threadIdx.x], sbest[threadIdx.x + d]);
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            const uint32_t l = sbest[0] == 0xFFFFFFFFu ? m : sbest[0];
            scratch[b + i] = l;
            label[m] = l;
        }
        __syncthreads();
    }
}

// cluster key = (first index of the group << 32) | label  => reference output order (bu:283-285)
__global__ void k_cluster_key(int64_t n, const uint32_t* __restrict__ gid, const uint32_t* __restrict__ gfirst,
                              const uint32_t* __restrict__ label, uint64_t* __restrict__ key, uint32_t* __restrict__ idx) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    key[i] = ((uint64_t)gfirst[gid[i]] << 32) | (uint64_t)label[i];
    idx[i] = (uint32_t)i;
}

__global__ void k_cluster_ptr(int64_t n, const uint32_t* __restrict__ head_incl, uint32_t* __restrict__ cptr, uint32_t n_clusters) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == 0 || head_incl[i] != head_incl[i - 1]) cptr[head_incl[i] - 1] = (uint32_t)i;
    if (i == n - 1) cptr[n_clusters] = (uint32_t)n;
}

}  // namespace coral

namespace coral {

// ---- bpc2bp + peel ------------------------------------------------------------------------------------
typedef unsigned __int128 u128;

// Python int -> float: correctly rounded (round-half-even) conversion of an exact integer sum.
__device__ __forceinline__ double u128_to_double(u128 v) {
    const uint64_t hi = (uint64_t)(v >> 64), lo = (uint64_t)v;
    if (hi == 0) return __ull2double_rn(lo);
    const int shift = 64 - __clzll((long long)hi);
    uint64_t top = (uint64_t)(v >> shift);
    if ((v & (((u128)1 << shift) - 1)) != 0) top |= 1ull;  // sticky
    return ldexp(__ull2double_rn(top), shift);
}

// bp_match(bp_, bp, rgap = bp_[7] * 1.2, [c, c]) for members of one cluster (chr/orientation equal), bu:391-416
// general form: any rgap (double), one cutoff per coordinate
__device__ __forceinline__ bool bp_match_gen(int64_t p1, int64_t p4, bool plus1, bool minus1, bool plus2, bool minus2,
                                             double rgap, int64_t b1, int64_t b4, int64_t c1, int64_t c2) {
    const int64_t d1 = p1 - b1, d2 = p4 - b4;
    const bool near1 = (d1 < 0 ? -d1 : d1) < c1, near2 = (d2 < 0 ? -d2 : d2) < c2;
    if (rgap <= 0) return near1 && near2;
    double slack = rgap;
    bool used1 = false, used2 = false;
    if (plus1 && d1 <= -c1) { slack = __dsub_rn(slack, (double)(-d1 - c1 + 1)); used1 = true; }
    if (minus1 && d1 >= c1) { slack = __dsub_rn(slack, (double)(d1 - c1 + 1)); used1 = true; }
    if (plus2 && d2 <= -c2) { slack = __dsub_rn(slack, (double)(-d2 - c2 + 1)); used2 = true; }
    if (minus2 && d2 >= c2) { slack = __dsub_rn(slack, (double)(d2 - c2 + 1)); used2 = true; }
    return ((used1 && slack >= 0) || near1) && ((used2 && slack >= 0) || near2);
}
__device__ __forceinline__ bool bp_match_dev(int32_t p1, int32_t p4, bool minus1, bool minus2, int32_t rgap_i,
                                             int32_t b1, int32_t b4, int32_t c) {
    return bp_match_gen(p1, p4, !minus1, minus1, !minus2, minus2, __dmul_rn((double)rgap_i, 1.2), b1, b4, c, c);
}

// operator form of bp_match (bu:391-416) over n pairs
__global__ void k_bp_match(int64_t n, const int64_t* __restrict__ a1, const int64_t* __restrict__ a4,
                           const uint8_t* __restrict__ m1, const uint8_t* __restrict__ m2, const double* __restrict__ rgap,
                           const int64_t* __restrict__ b1, const int64_t* __restrict__ b4, int64_t c1, int64_t c2,
                           uint8_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)       // orientation codes: 0 '+', 1 '-', anything else = neither (the reference compares strings)
        out[i] = bp_match_gen(a1[i], a4[i], m1[i] == 0, m1[i] == 1, m2[i] == 0, m2[i] == 1, rgap[i], b1[i], b4[i], c1, c2) ? 1 : 0;
}

// signed variant (positions may be negative in synthetic operator tests; Python sums are exact either way)
__device__ __forceinline__ double i128_to_double(u128 two_c) {
    const __int128 v = (__int128)two_c;
    return v < 0 ? -u128_to_double((u128)(-v)) : u128_to_double((u128)v);
}

constexpr int PEEL_THREADS = 256;

__device__ __forceinline__ void block_bitonic_sort(int32_t* a, uint32_t P) {
    for (uint32_t size = 2; size <= P; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            for (uint32_t t = threadIdx.x; t < (P >> 1); t += PEEL_THREADS) {
                const uint32_t lo = 2 * t - (t & (stride - 1));
                const uint32_t hi = lo + stride;
                const bool up = (lo & size) == 0;
                const int32_t x = a[lo], y = a[hi];
                if ((x > y) == up) { a[lo] = y; a[hi] = x; }
            }
            __syncthreads();
        }
    }
}

// Position call on a sorted array (bu:336-357): unique mode, else median (ceil for '+', floor for '-').
// All threads return the same value.
__device__ int32_t block_call_position(const int32_t* a, uint32_t k, bool minus, int32_t* s_i, uint32_t* s_u) {
    // pass 1: maximum run length
    uint32_t best = 0;
    for (uint32_t i = threadIdx.x; i < k; i += PEEL_THREADS) {
        if (i == 0 || a[i] != a[i - 1]) {
            uint32_t lo = i, hi = k;  // upper bound of a[i]
            const int32_t v = a[i];
            while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (a[mid] <= v) lo = mid + 1; else hi = mid; }
            best = max(best, lo - i);
        }
    }
    s_u[threadIdx.x] = best;
    __syncthreads();
    for (int d = PEEL_THREADS / 2; d > 0; d >>= 1) {
        if (threadIdx.x < d) s_u[threadIdx.x] = max(s_u[threadIdx.x], s_u[threadIdx.x + d]);
        __syncthreads();
    }
    const uint32_t M = s_u[0];
    __syncthreads();
    // pass 2: how many runs reach M, and the (smallest) value of such a run
    uint32_t cntM = 0;
    int32_t valM = INT32_MAX;
    for (uint32_t i = threadIdx.x; i < k; i += PEEL_THREADS) {
        if (i == 0 || a[i] != a[i - 1]) {
            if (i + M <= k && a[i + M - 1] == a[i] && (i + M == k || a[i + M] != a[i])) { cntM++; valM = min(valM, a[i]); }
        }
    }
    s_u[threadIdx.x] = cntM;
    s_i[threadIdx.x] = valM;
    __syncthreads();
    for (int d = PEEL_THREADS / 2; d > 0; d >>= 1) {
        if (threadIdx.x < d) {
            s_u[threadIdx.x] += s_u[threadIdx.x + d];
            s_i[threadIdx.x] = min(s_i[threadIdx.x], s_i[threadIdx.x + d]);
        }
        __syncthreads();
    }
    const uint32_t nM = s_u[0];
    const int32_t vM = s_i[0];
    __syncthreads();
    if (nM == 1) return vM;
    if (k & 1) return a[k >> 1];
    const int64_t s = (int64_t)a[(k >> 1) - 1] + (int64_t)a[k >> 1];
    // np.median = s / 2.0; ceil for '+', floor for '-' (floor division handles negative sums too)
    const int64_t fl = (s >= 0) ? (s >> 1) : -((-s + 1) >> 1);
    return (int32_t)(minus ? fl : fl + (s & 1));
}

struct PeelOut {
    int32_t *pos1, *pos2, *nb, *n_sub, *sub_of;
    uint32_t* first_alive;
    double* stats;  // 6 per row
};

__global__ void __launch_bounds__(PEEL_THREADS)
k_peel(const uint32_t* __restrict__ cptr, const uint32_t* __restrict__ order, CandIn C, double min_cluster_cutoff,
       int32_t match_cutoff, int32_t max_peels, uint8_t* __restrict__ alive, int32_t* __restrict__ sort1,
       int32_t* __restrict__ sort2, PeelOut o) {
    __shared__ u128 r_s1[PEEL_THREADS], r_q1[PEEL_THREADS], r_s2[PEEL_THREADS], r_q2[PEEL_THREADS];
    __shared__ long long r_m1[PEEL_THREADS], r_m2[PEEL_THREADS];
    __shared__ uint32_t r_u[PEEL_THREADS];
    __shared__ int32_t r_i[PEEL_THREADS];
    __shared__ double sh_lo1, sh_hi1, sh_lo2, sh_hi2;
    __shared__ uint32_t sh_kept;
    const uint32_t c = blockIdx.x;
    const uint32_t b = cptr[c], e = cptr[c + 1];
    const uint32_t k = e - b;
    if (threadIdx.x == 0) o.n_sub[c] = 0;
    for (uint32_t i = threadIdx.x; i < k; i += PEEL_THREADS) { alive[b + i] = 1; o.sub_of[order[b + i]] = -1; }
    if ((double)k < min_cluster_cutoff) return;
    __syncthreads();
    int32_t* a1 = sort1 + 2 * (size_t)b;
    int32_t* a2 = sort2 + 2 * (size_t)b;
    uint32_t m = k;
    int32_t iter = 0;
    const double floor_sd = __ddiv_rn((double)match_cutoff, 2.99);
    const uint32_t any = order[b];
    const bool minus1 = C.o1[any] != 0, minus2 = C.o2[any] != 0;
    while ((double)m >= min_cluster_cutoff && iter < max_peels) {
        // ---- sums over the current remainder (bu:310-316) ---------------------------------------
        u128 s1 = 0, q1 = 0, s2 = 0, q2 = 0;
        uint32_t fa = 0xFFFFFFFFu;
        for (uint32_t i = threadIdx.x; i < k; i += PEEL_THREADS) {
            if (!alive[b + i]) continue;
            fa = min(fa, i);
            const uint32_t x = order[b + i];
            const int64_t p1 = C.pos1[x], p2 = C.pos2[x];
            s1 += (u128)(__int128)p1; q1 += (u128)(uint64_t)(p1 * p1);
            s2 += (u128)(__int128)p2; q2 += (u128)(uint64_t)(p2 * p2);
        }
        r_s1[threadIdx.x] = s1; r_q1[threadIdx.x] = q1; r_s2[threadIdx.x] = s2; r_q2[threadIdx.x] = q2;
        r_u[threadIdx.x] = fa;
        if (threadIdx.x == 0) sh_kept = 0;
        __syncthreads();
        for (int d = PEEL_THREADS / 2; d > 0; d >>= 1) {
            if (threadIdx.x < d) {
                r_s1[threadIdx.x] += r_s1[threadIdx.x + d]; r_q1[threadIdx.x] += r_q1[threadIdx.x + d];
                r_s2[threadIdx.x] += r_s2[threadIdx.x + d]; r_q2[threadIdx.x] += r_q2[threadIdx.x + d];
                r_u[threadIdx.x] = min(r_u[threadIdx.x], r_u[threadIdx.x + d]);
            }
            __syncthreads();
        }
        const uint32_t first_alive = r_u[0];
        if (threadIdx.x == 0) {
            const double n = (double)m;
            const double E1 = __ddiv_rn(i128_to_double(r_s1[0]), n), E2 = __ddiv_rn(i128_to_double(r_s2[0]), n);
            const double Q1 = __ddiv_rn(u128_to_double(r_q1[0]), n), Q2 = __ddiv_rn(u128_to_double(r_q2[0]), n);
            const double v1 = __dsub_rn(Q1, __dmul_rn(E1, E1)), v2 = __dsub_rn(Q2, __dmul_rn(E2, E2));
            const double sd1 = v1 >= 0 ? fmax(floor_sd, __dsqrt_rn(v1)) : floor_sd;   // bu:317-324
            const double sd2 = v2 >= 0 ? fmax(floor_sd, __dsqrt_rn(v2)) : floor_sd;
            sh_lo1 = __dsub_rn(E1, __dmul_rn(3.0, sd1)); sh_hi1 = __dadd_rn(E1, __dmul_rn(3.0, sd1));
            sh_lo2 = __dsub_rn(E2, __dmul_rn(3.0, sd2)); sh_hi2 = __dadd_rn(E2, __dmul_rn(3.0, sd2));
        }
        __syncthreads();
        // ---- 3-sigma filter (bu:327-331) -> value lists ----------------------------------------
        for (uint32_t i = threadIdx.x; i < k; i += PEEL_THREADS) {
            if (!alive[b + i]) continue;
            const uint32_t x = order[b + i];
            const double p1 = (double)C.pos1[x], p2 = (double)C.pos2[x];
            if (p1 <= sh_hi1 && p1 >= sh_lo1 && p2 <= sh_hi2 && p2 >= sh_lo2) {
                const uint32_t s = atomicAdd(&sh_kept, 1u);
                a1[s] = C.pos1[x];
                a2[s] = C.pos2[x];
            }
        }
        __syncthreads();
        const uint32_t kept = sh_kept;
        int32_t bp1 = minus1 ? 1000000000 : 0, bp2 = minus2 ? 1000000000 : 0;   // bu:305-306
        if (kept > 0) {
            uint32_t P = 2;
            while (P < kept) P <<= 1;
            for (uint32_t i = kept + threadIdx.x; i < P; i += PEEL_THREADS) { a1[i] = INT32_MAX; a2[i] = INT32_MAX; }
            __syncthreads();
            block_bitonic_sort(a1, P);
            block_bitonic_sort(a2, P);
            bp1 = block_call_position(a1, kept, minus1, r_i, r_u);
            bp2 = block_call_position(a2, kept, minus2, r_i, r_u);
        }
        __syncthreads();
        // ---- matching members form bpr (bu:360-374) ------------------------------------------------
        s1 = q1 = s2 = q2 = 0;
        long long ma = 0, mb2 = 0;
        uint32_t nb = 0;
        for (uint32_t i = threadIdx.x; i < k; i += PEEL_THREADS) {
            if (!alive[b + i]) continue;
            const uint32_t x = order[b + i];
            if (bp_match_dev(C.pos1[x], C.pos2[x], minus1, minus2, C.rgap[x], bp1, bp2, match_cutoff)) {
                alive[b + i] = 0;
                o.sub_of[x] = iter;
                nb++;
                const int64_t p1 = C.pos1[x], p2 = C.pos2[x];
                s1 += (u128)(__int128)p1; q1 += (u128)(uint64_t)(p1 * p1);
                s2 += (u128)(__int128)p2; q2 += (u128)(uint64_t)(p2 * p2);
                if (C.flag[x] == 0) { ma += C.mq1[x]; mb2 += C.mq2[x]; } else { ma += C.mq2[x]; mb2 += C.mq1[x]; }
            }
        }
        r_s1[threadIdx.x] = s1; r_q1[threadIdx.x] = q1; r_s2[threadIdx.x] = s2; r_q2[threadIdx.x] = q2;
        r_m1[threadIdx.x] = ma; r_m2[threadIdx.x] = mb2; r_u[threadIdx.x] = nb;
        __syncthreads();
        for (int d = PEEL_THREADS / 2; d > 0; d >>= 1) {
            if (threadIdx.x < d) {
                r_s1[threadIdx.x] += r_s1[threadIdx.x + d]; r_q1[threadIdx.x] += r_q1[threadIdx.x + d];
                r_s2[threadIdx.x] += r_s2[threadIdx.x + d]; r_q2[threadIdx.x] += r_q2[threadIdx.x + d];
                r_m1[threadIdx.x] += r_m1[threadIdx.x + d]; r_m2[threadIdx.x] += r_m2[threadIdx.x + d];
                r_u[threadIdx.x] += r_u[threadIdx.x + d];
            }
            __syncthreads();
        }
        const uint32_t n_match = r_u[0];
        if (threadIdx.x == 0) {
            const size_t row = (size_t)b + iter;
            o.pos1[row] = bp1; o.pos2[row] = bp2; o.nb[row] = (int32_t)n_match;
            o.first_alive[row] = order[b + first_alive];
            double* st = o.stats + 6 * row;
            if (n_match == 0) { for (int z = 0; z < 6; z++) st[z] = 0.0; }
            else {
                const double n = (double)n_match;
                const double E1 = __ddiv_rn(i128_to_double(r_s1[0]), n), E2 = __ddiv_rn(i128_to_double(r_s2[0]), n);
                const double Q1 = __ddiv_rn(u128_to_double(r_q1[0]), n), Q2 = __ddiv_rn(u128_to_double(r_q2[0]), n);
                const double v1 = __dsub_rn(Q1, __dmul_rn(E1, E1)), v2 = __dsub_rn(Q2, __dmul_rn(E2, E2));
                st[0] = E1; st[1] = E2;
                st[2] = v1 >= 0 ? __dsqrt_rn(v1) : -1.0;  // bu:380-387 (-1 marks the `except: 0` branch)
                st[3] = v2 >= 0 ? __dsqrt_rn(v2) : -1.0;
                st[4] = __ddiv_rn((double)r_m1[0], n);
                st[5] = __ddiv_rn((double)r_m2[0], n);
            }
            o.n_sub[c] = iter + 1;
        }
        __syncthreads();
        iter++;
        if (n_match == 0) break;  // bpc2bp returns an empty remainder (bu:375-376)
        m -= n_match;
    }
}

}  // namespace coral

#define LAUNCH_1D(kern, n, ...)                                                          \
    do {                                                                                 \
        if ((n) > 0) {                                                                   \
            kern<<<(unsigned)(((n) + 255) / 256), 256, 0, c->stream>>>(__VA_ARGS__);      \
            c->n_launches++;                                                             \
            CORAL_CUDA(c, cudaGetLastError());                                           \
        }                                                                                \
    } while (0)

// work-buffer slots in ctx->cw
enum { W_CHR1, W_POS1, W_CHR2, W_POS2, W_RGAP, W_MQ1, W_MQ2, W_O1, W_O2, W_FLAG, W_KEYA, W_KEYB, W_IDXA, W_IDXB, W_HEAD,
       W_GID, W_GSTART, W_GFIRST, W_GBIG, W_SEG, W_CSTART, W_LABEL, W_SCR, W_ORDER, W_CPTR, W_ALIVE, W_SORT1, W_SORT2,
       W_ROWI, W_STATS, W_SUBOF, W_NSUB };

// Work buffers for n candidates.
static int cluster_ensure(coral_ctx* c, int64_t n) {
    DevBuf* w = c->cw;
    const size_t n4 = (size_t)n * 4, n8 = (size_t)n * 8;
    const int i4[] = {W_CHR1, W_POS1, W_CHR2, W_POS2, W_RGAP, W_MQ1, W_MQ2, W_IDXA, W_IDXB, W_HEAD, W_GID, W_SEG, W_LABEL, W_SCR,
                      W_ORDER, W_SUBOF};
    for (int k : i4) CORAL_CUDA(c, w[k].ensure(n4 + 16));
    const int i1[] = {W_O1, W_O2, W_FLAG, W_GBIG, W_ALIVE};
    for (int k : i1) CORAL_CUDA(c, w[k].ensure((size_t)n + 16));
    const int i4p[] = {W_GSTART, W_GFIRST, W_CSTART, W_CPTR, W_NSUB};
    for (int k : i4p) CORAL_CUDA(c, w[k].ensure(n4 + 16));
    CORAL_CUDA(c, w[W_KEYA].ensure(n8));
    CORAL_CUDA(c, w[W_KEYB].ensure(n8));
    CORAL_CUDA(c, w[W_SORT1].ensure(2 * n4 + 16));
    CORAL_CUDA(c, w[W_SORT2].ensure(2 * n4 + 16));
    CORAL_CUDA(c, w[W_ROWI].ensure(4 * n4));   // pos1, pos2, nb, first_alive rows
    CORAL_CUDA(c, w[W_STATS].ensure(6 * n8));
    return CORAL_OK;
}

static int cluster_core(coral_ctx* c, int64_t n, coral::CandIn C, double min_cluster_size, int32_t bp_distance_cutoff,
                        double peel_cutoff, int32_t match_cutoff, int32_t max_peels, int64_t* n_clusters);

extern "C" int coral_cluster_call(coral_ctx* c, int64_t n, const int32_t* chr1, const int32_t* pos1, const uint8_t* o1,
                                  const int32_t* chr2, const int32_t* pos2, const uint8_t* o2, const int32_t* rgap,
                                  const uint8_t* flag, const int32_t* mq1, const int32_t* mq2, double min_cluster_size,
                                  int32_t bp_distance_cutoff, double peel_cutoff, int32_t match_cutoff, int32_t max_peels,
                                  int64_t* n_clusters) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c || n < 0) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    c->cl_valid = false; c->cl_n = n; c->cl_n_clusters = 0;
    if (n_clusters) *n_clusters = 0;
    if (c->pass.active) {
        uint64_t k = pass_hash(PASS_H0, "cluster_call", 12);
        const int64_t a[2] = {n, (int64_t)max_peels}; const double d[2] = {min_cluster_size, peel_cutoff};
        const int32_t b[2] = {bp_distance_cutoff, match_cutoff};
        k = pass_hash(pass_hash(pass_hash(k, a, 16), d, 16), b, 8);
        if (n > 0) {
            k = pass_hash(k, chr1, n * 4); k = pass_hash(k, pos1, n * 4); k = pass_hash(k, chr2, n * 4); k = pass_hash(k, pos2, n * 4);
            k = pass_hash(k, o1, n); k = pass_hash(k, o2, n); k = pass_hash(k, rgap, n * 4); k = pass_hash(k, flag, n);
            k = pass_hash(k, mq1, n * 4); k = pass_hash(k, mq2, n * 4);
        }
        pass_call(c, k);
    }
    if (n == 0) { c->cl_valid = true; return CORAL_OK; }
    if (n >= ((int64_t)1 << 31)) { c->err = "too many candidates"; return CORAL_E_ARG; }
    CORAL_TRY(cluster_ensure(c, n));
    DevBuf* w = c->cw;
    CORAL_TRY(h2d_keyed(c, w[W_CHR1].p, chr1, (size_t)n)); CORAL_TRY(h2d_keyed(c, w[W_POS1].p, pos1, (size_t)n));
    CORAL_TRY(h2d_keyed(c, w[W_CHR2].p, chr2, (size_t)n)); CORAL_TRY(h2d_keyed(c, w[W_POS2].p, pos2, (size_t)n));
    CORAL_TRY(h2d_keyed(c, w[W_RGAP].p, rgap, (size_t)n)); CORAL_TRY(h2d_keyed(c, w[W_MQ1].p, mq1, (size_t)n));
    CORAL_TRY(h2d_keyed(c, w[W_MQ2].p, mq2, (size_t)n)); CORAL_TRY(h2d_keyed(c, w[W_O1].p, o1, (size_t)n));
    CORAL_TRY(h2d_keyed(c, w[W_O2].p, o2, (size_t)n)); CORAL_TRY(h2d_keyed(c, w[W_FLAG].p, flag, (size_t)n));
    CandIn C{w[W_CHR1].as<int32_t>(), w[W_POS1].as<int32_t>(), w[W_CHR2].as<int32_t>(), w[W_POS2].as<int32_t>(),
             w[W_RGAP].as<int32_t>(), w[W_MQ1].as<int32_t>(), w[W_MQ2].as<int32_t>(), w[W_O1].as<uint8_t>(),
             w[W_O2].as<uint8_t>(), w[W_FLAG].as<uint8_t>()};
    return cluster_core(c, n, C, min_cluster_size, bp_distance_cutoff, peel_cutoff, match_cutoff, max_peels, n_clusters);
}

// The same call on the candidate table coral_candidates_l / coral_candidates_pair left on the device (no round trip of
// the candidates through the host): what find_breakpoints does with its candidate list (ibg:687-716).
extern "C" int coral_cluster_call_candidates(coral_ctx* c, double min_cluster_size, int32_t bp_distance_cutoff,
                                             double peel_cutoff, int32_t match_cutoff, int32_t max_peels,
                                             int64_t* n_clusters) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->cand.valid) { c->err = "coral_cluster_call_candidates before coral_candidates_l"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    const int64_t n = c->cand.n;
    c->cl_valid = false; c->cl_n = n; c->cl_n_clusters = 0;
    if (n_clusters) *n_clusters = 0;
    if (c->pass.active) {
        uint64_t k = pass_hash(PASS_H0, "cluster_call_candidates", 23);
        const int64_t a[1] = {(int64_t)max_peels}; const double d[2] = {min_cluster_size, peel_cutoff};
        const int32_t b[2] = {bp_distance_cutoff, match_cutoff};
        pass_call(c, pass_hash(pass_hash(pass_hash(k, a, 8), d, 16), b, 8));
    }
    if (n == 0) { c->cl_valid = true; return CORAL_OK; }
    CORAL_TRY(cluster_ensure(c, n));
    CandTable& T = c->cand;
    CandIn C{T.chr1.as<int32_t>(), T.pos1.as<int32_t>(), T.chr2.as<int32_t>(), T.pos2.as<int32_t>(), T.rgap.as<int32_t>(),
             T.mq1.as<int32_t>(), T.mq2.as<int32_t>(), T.o1.as<uint8_t>(), T.o2.as<uint8_t>(), T.flag.as<uint8_t>()};
    return cluster_core(c, n, C, min_cluster_size, bp_distance_cutoff, peel_cutoff, match_cutoff, max_peels, n_clusters);
}

static int cluster_core(coral_ctx* c, int64_t n, coral::CandIn C, double min_cluster_size, int32_t bp_distance_cutoff,
                        double peel_cutoff, int32_t match_cutoff, int32_t max_peels, int64_t* n_clusters) {
    using namespace coral;
    DevBuf* w = c->cw;
    uint64_t* keyA = w[W_KEYA].as<uint64_t>(); uint64_t* keyB = w[W_KEYB].as<uint64_t>();
    uint32_t* idxA = w[W_IDXA].as<uint32_t>(); uint32_t* idxB = w[W_IDXB].as<uint32_t>();
    uint32_t* head = w[W_HEAD].as<uint32_t>(); uint32_t* gid = w[W_GID].as<uint32_t>();
    uint32_t* gstart = w[W_GSTART].as<uint32_t>(); uint32_t* gfirst = w[W_GFIRST].as<uint32_t>();
    uint8_t* gbig = w[W_GBIG].as<uint8_t>(); uint32_t* seg = w[W_SEG].as<uint32_t>();
    uint32_t* cstart = w[W_CSTART].as<uint32_t>(); uint32_t* label = w[W_LABEL].as<uint32_t>();

    c->stage_begin(CORAL_STAGE_CLUSTER);
    // every sort key below is (an id < n) << 32 | 32 payload bits: only the id bits that can be set are sorted on
    int bits_n = 1;
    while (((int64_t)1 << bits_n) < n + 1) bits_n++;
    // groups (first-appearance order is carried by gfirst)
    LAUNCH_1D(k_group_key, n, n, C, keyA, idxA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, idxA, idxB, n, 0, 40));
    LAUNCH_1D(k_key_heads, n, n, keyB, head);
    CORAL_TRY(inclusive_sum_u32(c, head, head, n));
    uint32_t n_groups = 0;
    LAUNCH_1D(k_group_assign, n, n, idxB, head, gid, gstart, gfirst);
    CORAL_TRY(pass_resolve(c, &n_groups, head + n - 1));
    LAUNCH_1D(k_group_big, (int64_t)n_groups, (int64_t)n_groups, gstart, n, min_cluster_size, gbig);
    // runs along pos1 within a group (idxB is grouped by gid already; sort by (gid, pos1))
    LAUNCH_1D(k_seg_pos_key, n, n, idxB, gid, C.pos1, keyA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, idxB, idxA, n, 0, 32 + bits_n));
    LAUNCH_1D(k_gap_heads, n, n, keyB, idxA, gid, gbig, bp_distance_cutoff, head);
    CORAL_TRY(inclusive_sum_u32(c, head, head, n));
    LAUNCH_1D(k_scatter_seg, n, n, idxA, head, seg);
    // components along pos2 within a run
    LAUNCH_1D(k_seg_pos_key, n, n, idxA, seg, C.pos2, keyA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, idxA, idxB, n, 0, 32 + bits_n));
    LAUNCH_1D(k_gap_heads, n, n, keyB, idxB, gid, gbig, bp_distance_cutoff, head);
    CORAL_TRY(inclusive_sum_u32(c, head, head, n));
    uint32_t n_comp = 0;
    LAUNCH_1D(k_scatter_seg, n, n, idxB, head, seg);
    CORAL_TRY(pass_resolve(c, &n_comp, head + n - 1));      // (head is not written again before the next scan)
    // members of each component in input order
    LAUNCH_1D(k_seg_idx_key, n, n, seg, keyA, idxA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, idxA, idxB, n, 0, 32 + bits_n));
    LAUNCH_1D(k_seg_starts, n, n, keyB, cstart, n_comp);
    k_component_cluster<<<n_comp, CL_THREADS, 0, c->stream>>>(cstart, idxB, C, gid, gbig, bp_distance_cutoff, label,
                                                             w[W_SCR].as<uint32_t>());
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    // clusters in reference order, members in input order
    uint32_t* order = w[W_ORDER].as<uint32_t>();
    uint32_t* cptr = w[W_CPTR].as<uint32_t>();
    LAUNCH_1D(k_cluster_key, n, n, gid, gfirst, label, keyA, idxA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, idxA, order, n, 0, 32 + bits_n));
    LAUNCH_1D(k_key_heads, n, n, keyB, head);
    CORAL_TRY(inclusive_sum_u32(c, head, head, n));
    uint32_t n_cl = 0;
    CORAL_TRY(pass_resolve(c, &n_cl, head + n - 1));
    LAUNCH_1D(k_cluster_ptr, n, n, head, cptr, n_cl);
    c->stage_end(CORAL_STAGE_CLUSTER);

    // peel
    c->stage_begin(CORAL_STAGE_BPC2BP);
    int32_t* rowi = w[W_ROWI].as<int32_t>();
    PeelOut o{rowi, rowi + n, rowi + 2 * n, w[W_NSUB].as<int32_t>(), w[W_SUBOF].as<int32_t>(),
              reinterpret_cast<uint32_t*>(rowi + 3 * n), w[W_STATS].as<double>()};
    k_peel<<<n_cl, PEEL_THREADS, 0, c->stream>>>(cptr, order, C, peel_cutoff, match_cutoff, max_peels > 0 ? max_peels : INT32_MAX,
                                                 w[W_ALIVE].as<uint8_t>(), w[W_SORT1].as<int32_t>(), w[W_SORT2].as<int32_t>(), o);
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    c->stage_end(CORAL_STAGE_BPC2BP);
    CORAL_TRY(pass_sync(c));
    c->cl_n_clusters = n_cl;
    c->cl_valid = true;
    if (n_clusters) *n_clusters = n_cl;
    return CORAL_OK;
}

// Operator form of breakpoint_utilities.bp_match (bu:391-416) for n pairs whose chromosomes and orientations already
// agree (the caller compares those: they are strings on its side): pos1 / pos2 of bp1 and bp2, the orientation of each
// end of bp1 (0 '+', 1 '-'), rgap as the reference passes it (any double), one cutoff per coordinate.
extern "C" int coral_bp_match(coral_ctx* c, int64_t n, const int64_t* bp1_pos1, const int64_t* bp1_pos2, const uint8_t* minus1,
                              const uint8_t* minus2, const double* rgap, const int64_t* bp2_pos1, const int64_t* bp2_pos2,
                              int64_t cutoff1, int64_t cutoff2, uint8_t* out) {
    using namespace coral;
    if (!c || n < 0) return CORAL_E_ARG;
    if (n == 0) return CORAL_OK;
    cudaSetDevice(c->device);
    DevBuf& w = c->cw[W_STATS];
    const size_t n8 = (size_t)n * 8;
    CORAL_CUDA(c, w.ensure(5 * n8 + 3 * (size_t)n + 64));
    int64_t* d = w.as<int64_t>();
    uint8_t* u = reinterpret_cast<uint8_t*>(d + 5 * n);
    CORAL_TRY(h2d(c, d, bp1_pos1, (size_t)n)); CORAL_TRY(h2d(c, d + n, bp1_pos2, (size_t)n));
    CORAL_TRY(h2d(c, d + 2 * n, bp2_pos1, (size_t)n)); CORAL_TRY(h2d(c, d + 3 * n, bp2_pos2, (size_t)n));
    CORAL_TRY(h2d(c, d + 4 * n, rgap, (size_t)n));
    CORAL_TRY(h2d(c, u, minus1, (size_t)n)); CORAL_TRY(h2d(c, u + n, minus2, (size_t)n));
    LAUNCH_1D(k_bp_match, n, n, d, d + n, u, u + n, reinterpret_cast<const double*>(d + 4 * n), d + 2 * n, d + 3 * n, cutoff1,
              cutoff2, u + 2 * n);
    CORAL_TRY(d2h(c, out, u + 2 * n, (size_t)n));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    c->cl_valid = false;          // the row-statistics buffer of the last cluster call was reused
    return CORAL_OK;
}

extern "C" int coral_get_clusters(coral_ctx* c, uint32_t* order, uint32_t* cluster_ptr, int32_t* n_sub, int32_t* sub_of,
                                  int32_t* row_pos1, int32_t* row_pos2, int32_t* row_nb, uint32_t* row_first,
                                  double* row_stats) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->cl_valid) { c->err = "coral_get_clusters before coral_cluster_call"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    const size_t n = (size_t)c->cl_n, nc = (size_t)c->cl_n_clusters;
    if (n == 0) { if (cluster_ptr) cluster_ptr[0] = 0; return CORAL_OK; }
    DevBuf* w = c->cw;
    int32_t* rowi = w[W_ROWI].as<int32_t>();
    if (order) CORAL_TRY(d2h_result(c, order, w[W_ORDER].p, n));
    if (cluster_ptr) CORAL_TRY(d2h_result(c, cluster_ptr, w[W_CPTR].p, nc + 1));
    if (n_sub) CORAL_TRY(d2h_result(c, n_sub, w[W_NSUB].p, nc));
    if (sub_of) CORAL_TRY(d2h_result(c, sub_of, w[W_SUBOF].p, n));
    if (row_pos1) CORAL_TRY(d2h_result(c, row_pos1, rowi, n));
    if (row_pos2) CORAL_TRY(d2h_result(c, row_pos2, rowi + n, n));
    if (row_nb) CORAL_TRY(d2h_result(c, row_nb, rowi + 2 * n, n));
    if (row_first) CORAL_TRY(d2h_result(c, row_first, rowi + 3 * n, n));
    if (row_stats) CORAL_TRY(d2h_result(c, row_stats, w[W_STATS].p, 6 * n));
    return pass_sync(c);
}
