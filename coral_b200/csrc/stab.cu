// coral_b200/csrc/stab.cu — concordant-edge read support (reference infer_breakpoint_graph.py:1039-1055).
//
// For the edge (chr, p, '+') -> (chr, p + 1, '-') the reference builds four sets of read NAMES from point
// fetches at p, p + 1, p - d - 1 and p + 1 + d (d = min_bp_match_cutoff_), then
//   ce[9] = names(p) | names(p + 1)
//   ce[8] = len((names(p) & names(p+1) & names(p-d-1) & names(p+1+d)) - rbps)
// with rbps = names of the reads on discordant edges incident to either node.  Here every record emits a
// (edge, record, 4-bit mask) row for the edges whose points it covers, rbps members emit a row with bit 16,
// rows are sorted by (edge, name hash) and the masks OR-ed per name, so two records of one read can jointly
// satisfy the test exactly as the name sets do.
#include <vector>

#include "ctx.h"
#include "dev_util.cuh"
#include "relops.cuh"

namespace coral {

struct Edges {
    const int32_t *tid, *p;
    int32_t n, d;
};

__device__ __forceinline__ int32_t edge_lower_bound(const Edges& E, int32_t tid, int64_t p) {
    int32_t lo = 0, hi = E.n;
    while (lo < hi) {
        int32_t mid = (lo + hi) >> 1;
        bool less = E.tid[mid] < tid || (E.tid[mid] == tid && (int64_t)E.p[mid] < p);
        if (less) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void k_stab_rows(int64_t n, const int32_t* __restrict__ tid, const int32_t* __restrict__ pos,
                            const int32_t* __restrict__ endpos, Edges E, uint32_t* __restrict__ row_edge,
                            uint32_t* __restrict__ row_rec, uint32_t* __restrict__ row_mask, unsigned long long* __restrict__ n_rows,
                            int64_t cap) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t t = tid[i];
    if (t < 0) return;
    const int64_t s = pos[i], e = endpos[i];  // covers x iff s <= x < e
    const int64_t d = E.d;
    for (int32_t k = edge_lower_bound(E, t, s - 1 - d); k < E.n && E.tid[k] == t && (int64_t)E.p[k] <= e + d; k++) {
        const int64_t p = E.p[k];
        uint32_t m = 0;
        if (s <= p && p < e) m |= 1u;
        if (s <= p + 1 && p + 1 < e) m |= 2u;
        if (s <= p - d - 1 && p - d - 1 < e) m |= 4u;
        if (s <= p + 1 + d && p + 1 + d < e) m |= 8u;
        if (!m) continue;
        const unsigned long long r = atomicAdd(n_rows, 1ull);
        if ((int64_t)r < cap) { row_edge[r] = (uint32_t)k; row_rec[r] = (uint32_t)i; row_mask[r] = m; }
    }
}

__global__ void k_stab_rbps(int64_t n_rb, const uint32_t* __restrict__ rb_edge, const uint32_t* __restrict__ rb_rec,
                            uint32_t* __restrict__ row_edge, uint32_t* __restrict__ row_rec, uint32_t* __restrict__ row_mask,
                            int64_t base) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rb) return;
    row_edge[base + i] = rb_edge[i];
    row_rec[base + i] = rb_rec[i];
    row_mask[base + i] = 16u;
}

__global__ void k_stab_iota(int64_t n, uint32_t* p) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = (uint32_t)i;
}

__global__ void k_stab_key(int64_t n, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ row_rec,
                           const uint32_t* __restrict__ row_edge, const uint64_t* __restrict__ name, int which,
                           uint64_t* __restrict__ key) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = perm[i];
    key[i] = which == 2 ? (uint64_t)row_edge[r] : name[row_rec[r]];
}

// one thread per (edge, name) segment head: OR the masks, count full support outside rbps
__global__ void k_stab_count(int64_t n, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ row_edge,
                             const uint32_t* __restrict__ row_rec, const uint32_t* __restrict__ row_mask,
                             const uint64_t* __restrict__ nlo, const uint64_t* __restrict__ nhi,
                             unsigned long long* __restrict__ count) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = perm[i];
    const uint32_t e = row_edge[r];
    const uint64_t lo = nlo[row_rec[r]], hi = nhi[row_rec[r]];
    if (i > 0) {
        const uint32_t q = perm[i - 1];
        if (row_edge[q] == e && nlo[row_rec[q]] == lo && nhi[row_rec[q]] == hi) return;  // not a head
    }
    uint32_t m = 0;
    for (int64_t j = i; j < n; j++) {
        const uint32_t q = perm[j];
        if (row_edge[q] != e || nlo[row_rec[q]] != lo || nhi[row_rec[q]] != hi) break;
        m |= row_mask[q];
    }
    if ((m & 15u) == 15u && !(m & 16u)) atomicAdd(count + e, 1ull);
}

// ---- the same count on explicit (edge, name hash, mask) rows: the multi-GPU combine --------------------------------
__global__ void k_rows_key(int64_t n, const uint32_t* __restrict__ perm, const uint64_t* __restrict__ a, uint64_t* __restrict__ key) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) key[i] = a[perm[i]];
}
__global__ void k_rows_count(int64_t n, const uint32_t* __restrict__ perm, const uint64_t* __restrict__ edge,
                             const uint64_t* __restrict__ lo, const uint64_t* __restrict__ hi, const uint64_t* __restrict__ mask,
                             unsigned long long* __restrict__ count, int64_t n_edges) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = perm[i];
    const uint64_t e = edge[r], l = lo[r], h = hi[r];
    if (i > 0) {
        const uint32_t q = perm[i - 1];
        if (edge[q] == e && lo[q] == l && hi[q] == h) return;  // not a head
    }
    uint64_t m = 0;
    for (int64_t j = i; j < n; j++) {
        const uint32_t q = perm[j];
        if (edge[q] != e || lo[q] != l || hi[q] != h) break;
        m |= mask[q];
    }
    if ((m & 15u) == 15u && !(m & 16u) && (int64_t)e < n_edges) atomicAdd(count + e, 1ull);
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

extern "C" int coral_concordant(coral_ctx* c, const int32_t* tid, const int32_t* p, int32_t n_edges, int64_t n_rbps,
                                const uint32_t* rb_edge, const uint32_t* rb_rec, int64_t* lr_count, int64_t* n_union_rows) {
    using namespace coral;
    if (!c || n_edges < 0 || n_rbps < 0) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_concordant before coral_load_records"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    c->stab_n = 0; c->stab_valid = false;
    if (n_union_rows) *n_union_rows = 0;
    if (n_edges == 0) { c->stab_valid = true; return CORAL_OK; }
    for (int32_t i = 0; i < n_edges; i++) {
        if (i > 0 && (tid[i] < tid[i - 1] || (tid[i] == tid[i - 1] && p[i] <= p[i - 1]))) {
            c->err = "concordant edges must be sorted by (tid, p) and distinct";
            return CORAL_E_ARG;
        }
        if ((int64_t)p[i] - c->prm.min_bp_match_cutoff - 1 < 0) { c->err = "ValueError: start out of range (negative point fetch)"; return CORAL_E_ARG; }
    }
    for (int64_t i = 0; i < n_rbps; i++)
        if (rb_edge[i] >= (uint32_t)n_edges || rb_rec[i] >= (uint64_t)c->nrec) { c->err = "bad rbps row"; return CORAL_E_ARG; }
    RecordTable& rt = c->rt;
    DevBuf* w = c->sw;
    CORAL_CUDA(c, w[0].ensure((size_t)n_edges * 8 + 16));
    int32_t* d_e = w[0].as<int32_t>();
    CORAL_TRY(h2d(c, d_e, tid, (size_t)n_edges));
    CORAL_TRY(h2d(c, d_e + n_edges, p, (size_t)n_edges));
    Edges E{d_e, d_e + n_edges, n_edges, c->prm.min_bp_match_cutoff};
    CORAL_CUDA(c, w[1].ensure((size_t)n_edges * 8 + 16));
    unsigned long long* d_count = w[1].as<unsigned long long>();
    unsigned long long* d_nrows = d_count + n_edges;
    CORAL_CUDA(c, w[2].ensure((size_t)std::max<int64_t>(n_rbps, 1) * 8));
    uint32_t* d_rb = w[2].as<uint32_t>();
    CORAL_TRY(h2d(c, d_rb, rb_edge, (size_t)n_rbps));
    CORAL_TRY(h2d(c, d_rb + n_rbps, rb_rec, (size_t)n_rbps));
    c->stage_begin(CORAL_STAGE_STAB);
    int64_t cap = std::max<int64_t>((int64_t)1 << 16, (int64_t)(w[3].cap / 4) - n_rbps);
    unsigned long long n_rows = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        const size_t tot = (size_t)(cap + n_rbps);
        CORAL_CUDA(c, w[3].ensure(tot * 4)); CORAL_CUDA(c, w[4].ensure(tot * 4)); CORAL_CUDA(c, w[5].ensure(tot * 4));
        CORAL_CUDA(c, cudaMemsetAsync(d_count, 0, (size_t)n_edges * 8 + 8, c->stream));
        LAUNCH_1D(k_stab_rows, c->nrec, c->nrec, rt.tid.as<int32_t>(), rt.pos.as<int32_t>(), rt.endpos.as<int32_t>(), E,
                  w[3].as<uint32_t>(), w[4].as<uint32_t>(), w[5].as<uint32_t>(), d_nrows, cap);
        CORAL_TRY(d2h(c, &n_rows, d_nrows, 1));
        CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
        if ((int64_t)n_rows <= cap) break;
        cap = (int64_t)n_rows + 1024;
    }
    uint32_t* row_edge = w[3].as<uint32_t>(); uint32_t* row_rec = w[4].as<uint32_t>(); uint32_t* row_mask = w[5].as<uint32_t>();
    LAUNCH_1D(k_stab_rbps, n_rbps, n_rbps, d_rb, d_rb + n_rbps, row_edge, row_rec, row_mask, (int64_t)n_rows);
    const int64_t n = (int64_t)n_rows + n_rbps;
    if (n > 0) {
        CORAL_CUDA(c, w[6].ensure((size_t)n * 24 + 64));
        uint64_t* keyA = w[6].as<uint64_t>(); uint64_t* keyB = keyA + n;
        uint32_t* pa = reinterpret_cast<uint32_t*>(keyB + n); uint32_t* pb = pa + n;
        LAUNCH_1D(k_stab_iota, n, n, pa);
        LAUNCH_1D(k_stab_key, n, n, pa, row_rec, row_edge, rt.name_lo.as<uint64_t>(), 0, keyA);
        CORAL_TRY(sort_pairs_u64(c, keyA, keyB, pa, pb, n));
        LAUNCH_1D(k_stab_key, n, n, pb, row_rec, row_edge, rt.name_hi.as<uint64_t>(), 1, keyA);
        CORAL_TRY(sort_pairs_u64(c, keyA, keyB, pb, pa, n));
        LAUNCH_1D(k_stab_key, n, n, pa, row_rec, row_edge, rt.name_lo.as<uint64_t>(), 2, keyA);
        CORAL_TRY(sort_pairs_u64(c, keyA, keyB, pa, pb, n, 0, 32));
        LAUNCH_1D(k_stab_count, n, n, pb, row_edge, row_rec, row_mask, rt.name_lo.as<uint64_t>(), rt.name_hi.as<uint64_t>(), d_count);
    }
    c->stage_end(CORAL_STAGE_STAB);
    std::vector<unsigned long long> h((size_t)n_edges);
    CORAL_TRY(d2h(c, h.data(), d_count, (size_t)n_edges));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int32_t i = 0; i < n_edges; i++) lr_count[i] = (int64_t)h[i];
    c->stab_n = (int64_t)n_rows;
    c->stab_valid = true;
    if (n_union_rows) *n_union_rows = (int64_t)n_rows;
    return CORAL_OK;
}

extern "C" int coral_get_concordant_rows(coral_ctx* c, uint32_t* edge, uint32_t* rec, uint32_t* mask) {
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->stab_valid) { c->err = "coral_get_concordant_rows before coral_concordant"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    const size_t n = (size_t)c->stab_n;
    if (edge) CORAL_TRY(d2h(c, edge, c->sw[3].p, n));
    if (rec) CORAL_TRY(d2h(c, rec, c->sw[4].p, n));
    if (mask) CORAL_TRY(d2h(c, mask, c->sw[5].p, n));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}

// Device addresses of the rows left by the last coral_concordant (edge, rec, mask: uint32 each), so a multi-GPU caller
// can all-gather them (as (edge, name hash, mask)) without a host round trip.
extern "C" int coral_concordant_device_rows(coral_ctx* c, uint64_t* ptrs, int64_t* n_rows) {
    if (!c || !ptrs || !n_rows) return CORAL_E_ARG;
    if (!c->stab_valid) { c->err = "coral_concordant_device_rows before coral_concordant"; return CORAL_E_STATE; }
    for (int k = 0; k < 3; k++) ptrs[k] = (uint64_t)reinterpret_cast<uintptr_t>(c->sw[3 + k].p);
    *n_rows = c->stab_n;
    return CORAL_OK;
}

// ce[8] from explicit rows, all arrays on this ctx's device (uint64 each): per (edge, name hash) the masks are OR-ed;
// a name counts for its edge when bits 0..3 are all set and bit 4 (rbps member, ibg:1047-1053) is not.
extern "C" int coral_count_name_rows(coral_ctx* c, int32_t n_edges, int64_t n, const uint64_t* d_edge, const uint64_t* d_lo,
                                     const uint64_t* d_hi, const uint64_t* d_mask, int64_t* lr_count) {
    using namespace coral;
    if (!c || n_edges < 0 || n < 0 || !lr_count) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    for (int32_t i = 0; i < n_edges; i++) lr_count[i] = 0;
    if (n == 0 || n_edges == 0) return CORAL_OK;
    if (n >= ((int64_t)1 << 32)) { c->err = "too many rows"; return CORAL_E_ARG; }
    DevBuf* w = c->sw;
    CORAL_CUDA(c, w[7].ensure((size_t)n * 24 + 64 + (size_t)n_edges * 8));
    uint64_t* keyA = w[7].as<uint64_t>(); uint64_t* keyB = keyA + n;
    uint32_t* pa = reinterpret_cast<uint32_t*>(keyB + n); uint32_t* pb = pa + n;
    unsigned long long* d_count = reinterpret_cast<unsigned long long*>(pb + n);   // byte offset 24 n: 8-byte aligned
    CORAL_CUDA(c, cudaMemsetAsync(d_count, 0, (size_t)n_edges * 8, c->stream));
    LAUNCH_1D(k_stab_iota, n, n, pa);
    LAUNCH_1D(k_rows_key, n, n, pa, d_lo, keyA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, pa, pb, n));
    LAUNCH_1D(k_rows_key, n, n, pb, d_hi, keyA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, pb, pa, n));
    LAUNCH_1D(k_rows_key, n, n, pa, d_edge, keyA);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, pa, pb, n, 0, 32));
    LAUNCH_1D(k_rows_count, n, n, pb, d_edge, d_lo, d_hi, d_mask, d_count, (int64_t)n_edges);
    std::vector<unsigned long long> h((size_t)n_edges);
    CORAL_TRY(d2h(c, h.data(), d_count, (size_t)n_edges));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int32_t i = 0; i < n_edges; i++) lr_count[i] = (int64_t)h[i];
    return CORAL_OK;
}
