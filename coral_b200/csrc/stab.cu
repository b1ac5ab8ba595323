// coral_b200/csrc/stab.cu — concordant-edge read support (reference infer_breakpoint_graph.py:1039-1055).
//
// For the edge (chr, p, '+') -> (chr, p + 1, '-') the reference builds four sets of read NAMES from point
// fetches at p, p + 1, p - d - 1 and p + 1 + d (d = min_bp_match_cutoff_), then
//   ce[9] = names(p) | names(p + 1)
//   ce[8] = len((names(p) & names(p+1) & names(p-d-1) & names(p+1+d)) - rbps)
// with rbps = names of the reads on discordant edges incident to either node.  Here every record emits a
// (edge, record, 4-bit mask) row for the edges whose points it covers, rbps members emit a row with bit 16,
// rows are bucketed by a 24-bit hash of (edge, name hash) (one 3-pass radix sort), equal (edge, name) triples are
// found by exact comparison inside a bucket and their masks OR-ed, so two records of one read can jointly satisfy
// the test exactly as the name sets do.  The multi-GPU form packs the local rows as one message per rank
// (coral_concordant_pack), the caller all-gathers the messages, and coral_count_name_msgs runs the same count on
// the gathered rows.
#include <vector>

#include "ctx.h"
#include "dev_util.cuh"
#include "msg.cuh"
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

// ---- concordant exchange: (edge, name, mask) rows counted per edge -----------------------------------------
__global__ void k_name_rows(int64_t n, const uint32_t* __restrict__ edge, const uint32_t* __restrict__ rec,
                            const uint32_t* __restrict__ mask, const uint64_t* __restrict__ lo, const uint64_t* __restrict__ hi,
                            NameRow* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = rec[i];
    out[i] = NameRow{edge[i], lo[r], hi[r], mask[i]};
}

struct NameSet {
    const uint8_t* base[17];   // messages of the ranks, then the caller's rbps rows (no header)
    uint64_t row0[18];
    int n;                     // entries in base[]
    int n_msgs;                // the first n_msgs carry a MsgHdr
};

__device__ __forceinline__ NameRow name_row(const NameSet& s, uint64_t g) {
    int r = 0;
    while (r + 1 < s.n && g >= s.row0[r + 1]) r++;
    const uint8_t* b = s.base[r] + (r < s.n_msgs ? sizeof(MsgHdr) : 0);
    return reinterpret_cast<const NameRow*>(b)[g - s.row0[r]];
}

// 32-bit bucket of (edge, name); equal triples share a bucket, the count kernel compares triples exactly
__global__ void k_name_keys(NameSet s, NameRow* __restrict__ rows, uint64_t* __restrict__ key, uint32_t* __restrict__ idx) {
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= s.row0[s.n]) return;
    const NameRow r = name_row(s, g);
    rows[g] = r;
    key[g] = mix64(r.lo ^ mix64(r.hi + 0x9e3779b97f4a7c15ull * (r.edge + 1))) >> 40;   // 24 bits: 3 radix passes
    idx[g] = (uint32_t)g;
}

__global__ void k_name_count(int64_t n, const uint64_t* __restrict__ skey, const uint32_t* __restrict__ sidx,
                             const NameRow* __restrict__ rows, unsigned long long* __restrict__ count, int64_t n_edges) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t k = skey[i];
    const NameRow me = rows[sidx[i]];
    for (int64_t j = i - 1; j >= 0 && skey[j] == k; j--) {     // a head has no equal triple before it in its bucket
        const NameRow q = rows[sidx[j]];
        if (q.edge == me.edge && q.lo == me.lo && q.hi == me.hi) return;
    }
    uint64_t m = me.mask;
    for (int64_t j = i + 1; j < n && skey[j] == k; j++) {
        const NameRow q = rows[sidx[j]];
        if (q.edge == me.edge && q.lo == me.lo && q.hi == me.hi) m |= q.mask;
    }
    if ((m & 15u) == 15u && !(m & 16u) && (int64_t)me.edge < n_edges) atomicAdd(count + me.edge, 1ull);
}

__global__ void k_put_hdr(MsgHdr* h, MsgHdr v) { *h = v; }

}  // namespace coral

#define LAUNCH_1D(kern, n, ...)                                                          \
    do {                                                                                 \
        if ((n) > 0) {                                                                   \
            kern<<<(unsigned)(((n) + 255) / 256), 256, 0, c->stream>>>(__VA_ARGS__);      \
            c->n_launches++;                                                             \
            CORAL_CUDA(c, cudaGetLastError());                                           \
        }                                                                                \
    } while (0)

namespace {
using namespace coral;

// lr_count[e] = names of edge e whose OR-ed mask has bits 0..3 and not bit 4, over the rows of `set`
int count_names(coral_ctx* c, const NameSet& set, int32_t n_edges, int64_t* lr_count) {
    for (int32_t i = 0; i < n_edges; i++) lr_count[i] = 0;
    const int64_t n = (int64_t)set.row0[set.n];
    if (n == 0 || n_edges == 0) return CORAL_OK;
    if (n >= ((int64_t)1 << 32)) { c->err = "too many rows"; return CORAL_E_ARG; }
    DevBuf* w = c->sw;
    CORAL_CUDA(c, w[6].ensure((size_t)n * 56 + 64 + (size_t)n_edges * 8));
    NameRow* rows = w[6].as<NameRow>();
    uint64_t* keyA = reinterpret_cast<uint64_t*>(rows + n); uint64_t* keyB = keyA + n;
    unsigned long long* d_count = reinterpret_cast<unsigned long long*>(keyB + n);
    uint32_t* pa = reinterpret_cast<uint32_t*>(d_count + n_edges); uint32_t* pb = pa + n;
    CORAL_CUDA(c, cudaMemsetAsync(d_count, 0, (size_t)n_edges * 8, c->stream));
    LAUNCH_1D(k_name_keys, n, set, rows, keyA, pa);
    CORAL_TRY(sort_pairs_u64(c, keyA, keyB, pa, pb, n, 0, 24));
    LAUNCH_1D(k_name_count, n, n, keyB, pb, rows, d_count, (int64_t)n_edges);
    CORAL_TRY(d2h_result(c, reinterpret_cast<unsigned long long*>(lr_count), d_count, (size_t)n_edges));   // same 64 bits
    return pass_sync(c);
}

}  // namespace

// lr_count == NULL: only the stab rows are produced (the sharded path counts on the gathered rows).
extern "C" int coral_concordant(coral_ctx* c, const int32_t* tid, const int32_t* p, int32_t n_edges, int64_t n_rbps,
                                const uint32_t* rb_edge, const uint32_t* rb_rec, int64_t* lr_count, int64_t* n_union_rows) {
    using namespace coral;
    if (!c || n_edges < 0 || n_rbps < 0) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_concordant before coral_load_records"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(c));
    cudaSetDevice(c->device);
    c->stab_n = 0; c->stab_valid = false;
    CORAL_TRY(pass_side_join(c, SIDE_STAB));
    if (n_union_rows) *n_union_rows = 0;
    if (c->pass.active) {
        uint64_t k = pass_hash(PASS_H0, lr_count ? "concordant" : "concordant_rows", lr_count ? 10 : 15);
        const int64_t a[2] = {n_edges, n_rbps};
        k = pass_hash(k, a, 16);
        if (n_edges > 0) { k = pass_hash(k, tid, (size_t)n_edges * 4); k = pass_hash(k, p, (size_t)n_edges * 4); }
        if (n_rbps > 0) { k = pass_hash(k, rb_edge, (size_t)n_rbps * 4); k = pass_hash(k, rb_rec, (size_t)n_rbps * 4); }
        pass_call(c, k);
    }
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
    CORAL_TRY(h2d_keyed(c, d_e, tid, (size_t)n_edges));
    CORAL_TRY(h2d_keyed(c, d_e + n_edges, p, (size_t)n_edges));
    Edges E{d_e, d_e + n_edges, n_edges, c->prm.min_bp_match_cutoff};
    CORAL_CUDA(c, w[1].ensure(16));
    unsigned long long* d_nrows = w[1].as<unsigned long long>();
    CORAL_CUDA(c, w[2].ensure((size_t)std::max<int64_t>(n_rbps, 1) * 8));
    uint32_t* d_rb = w[2].as<uint32_t>();
    CORAL_TRY(h2d_keyed(c, d_rb, rb_edge, (size_t)n_rbps));
    CORAL_TRY(h2d_keyed(c, d_rb + n_rbps, rb_rec, (size_t)n_rbps));
    c->stage_begin(CORAL_STAGE_STAB);
    int64_t cap = std::max<int64_t>((int64_t)1 << 16, (int64_t)(w[3].cap / 4) - n_rbps);
    unsigned long long n_rows = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        const size_t tot = (size_t)(cap + n_rbps);
        CORAL_CUDA(c, w[3].ensure(tot * 4)); CORAL_CUDA(c, w[4].ensure(tot * 4)); CORAL_CUDA(c, w[5].ensure(tot * 4));
        CORAL_CUDA(c, cudaMemsetAsync(d_nrows, 0, 8, c->stream));
        LAUNCH_1D(k_stab_rows, c->nrec, c->nrec, rt.tid.as<int32_t>(), rt.pos.as<int32_t>(), rt.endpos.as<int32_t>(), E,
                  w[3].as<uint32_t>(), w[4].as<uint32_t>(), w[5].as<uint32_t>(), d_nrows, cap);
        CORAL_TRY(pass_resolve(c, &n_rows, d_nrows));
        if ((int64_t)n_rows <= cap) break;
        cap = (int64_t)n_rows + 1024;
    }
    uint32_t* row_edge = w[3].as<uint32_t>(); uint32_t* row_rec = w[4].as<uint32_t>(); uint32_t* row_mask = w[5].as<uint32_t>();
    LAUNCH_1D(k_stab_rbps, n_rbps, n_rbps, d_rb, d_rb + n_rbps, row_edge, row_rec, row_mask, (int64_t)n_rows);
    const int64_t n = (int64_t)n_rows + n_rbps;
    c->stab_n = (int64_t)n_rows;
    c->stab_valid = true;
    if (n_union_rows) *n_union_rows = (int64_t)n_rows;
    int rc = CORAL_OK;
    if (lr_count) {
        // the local rows (+ rbps rows, mask 16) as one headerless NameRow array, counted like a gathered set
        CORAL_CUDA(c, w[7].ensure((size_t)std::max<int64_t>(n, 1) * sizeof(NameRow)));
        LAUNCH_1D(k_name_rows, n, n, row_edge, row_rec, row_mask, rt.name_lo.as<uint64_t>(), rt.name_hi.as<uint64_t>(),
                  w[7].as<NameRow>());
        NameSet set;
        set.n = 1; set.n_msgs = 0;
        set.base[0] = w[7].as<uint8_t>();
        set.row0[0] = 0; set.row0[1] = (uint64_t)n;
        rc = count_names(c, set, n_edges, lr_count);
    }
    c->stage_end(CORAL_STAGE_STAB);
    return rc;
}

extern "C" int coral_get_concordant_rows(coral_ctx* c, uint32_t* edge, uint32_t* rec, uint32_t* mask) {
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->stab_valid) { c->err = "coral_get_concordant_rows before coral_concordant"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    const size_t n = (size_t)c->stab_n;
    if (n == 0) return CORAL_OK;
    CORAL_TRY(pass_side_begin(c, SIDE_STAB));  // the rows stay as they are until the next coral_concordant
    if (edge) CORAL_TRY(d2h_result(c, edge, c->sw[3].p, n));
    if (rec) CORAL_TRY(d2h_result(c, rec, c->sw[4].p, n));
    if (mask) CORAL_TRY(d2h_result(c, mask, c->sw[5].p, n));
    return pass_sync(c);
}

// ---- multi-GPU form ----------------------------------------------------------------------------
// The rows of the last coral_concordant as one message: MsgHdr + n NameRow (edge, name hash, mask), on the device.
extern "C" int coral_concordant_pack(coral_ctx* c, uint64_t* d_msg, uint64_t* nbytes) {
    using namespace coral;
    if (!c || !d_msg || !nbytes) return CORAL_E_ARG;
    if (!c->stab_valid) { c->err = "coral_concordant_pack before coral_concordant"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    const int64_t n = c->stab_n;
    const uint64_t total = sizeof(MsgHdr) + (uint64_t)n * sizeof(NameRow);
    DevBuf& msg = c->xmsg[2];
    CORAL_CUDA(c, msg.ensure(total));
    if (c->pass.active) pass_call(c, pass_hash(PASS_H0, "concordant_pack", 15));
    MsgHdr h{};
    h.n_rows = (uint64_t)n; h.kind = MSG_NAMES;
    k_put_hdr<<<1, 1, 0, c->stream>>>(msg.as<MsgHdr>(), h);
    c->n_launches++;
    RecordTable& rt = c->rt;
    LAUNCH_1D(k_name_rows, n, n, c->sw[3].as<uint32_t>(), c->sw[4].as<uint32_t>(), c->sw[5].as<uint32_t>(),
              rt.name_lo.as<uint64_t>(), rt.name_hi.as<uint64_t>(), reinterpret_cast<NameRow*>(msg.as<uint8_t>() + sizeof(MsgHdr)));
    if (!c->pass.active) CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    *d_msg = (uint64_t)reinterpret_cast<uintptr_t>(msg.p);
    *nbytes = total;
    return CORAL_OK;
}

// ce[8] on the gathered messages of all ranks (message r = d_buf[starts[r] : starts[r] + sizes[r]), host arrays) + the
// caller's rbps rows (d_rb_rows: n_rb NameRow on this device, mask 16, ibg:1047-1053).
extern "C" int coral_count_name_msgs(coral_ctx* c, int32_t n_edges, int32_t n_ranks, const uint8_t* d_buf, const uint64_t* starts,
                                     const uint64_t* sizes, const uint64_t* d_rb_rows, int64_t n_rb, int64_t* lr_count) {
    using namespace coral;
    if (!c || n_edges < 0 || !d_buf || !starts || !sizes || n_rb < 0 || !lr_count) return CORAL_E_ARG;
    if (n_ranks < 1 || n_ranks > 16) { c->err = "1..16 ranks per exchange"; return CORAL_E_ARG; }
    cudaSetDevice(c->device);
    NameSet set;
    set.n = n_ranks + (n_rb > 0 ? 1 : 0);
    set.n_msgs = n_ranks;
    set.row0[0] = 0;
    for (int r = 0; r < n_ranks; r++) {
        if ((starts[r] & 15) || sizes[r] < sizeof(MsgHdr) || (sizes[r] - sizeof(MsgHdr)) % sizeof(NameRow)) {
            c->err = "bad name-row message"; return CORAL_E_ARG;
        }
        set.base[r] = d_buf + starts[r];
        set.row0[r + 1] = set.row0[r] + (sizes[r] - sizeof(MsgHdr)) / sizeof(NameRow);
    }
    if (n_rb > 0) {
        set.base[n_ranks] = reinterpret_cast<const uint8_t*>(d_rb_rows);
        set.row0[n_ranks + 1] = set.row0[n_ranks] + (uint64_t)n_rb;
    }
    return count_names(c, set, n_edges, lr_count);
}
