// coral_b200/csrc/paths.cu — concordant-read pass of compute_path_constraints
// (reference infer_breakpoint_graph.py:1295-1321 + path_constraints.py:48-90 alignment_to_path).
//
// The reference walks every record of the amplicon intervals a second time: records with MAPQ >= 20 whose read
// NAME is one of the amplicon's "concordant reads" are turned into the run of sequence edges their reference span
// [reference_start, reference_end] touches; the run is trimmed at both ends (an end edge sharing < min_overlap with
// the alignment goes, then end edges shorter than min_overlap go, never the last edge left) and runs of fewer than
// three edges give no path.  Which run a record has depends only on (tid, pos, endpos) from the record table, so
// here it is one thread per record: two binary searches into the sorted, disjoint sequence edges, the trimming as
// integer compares, a probe of an exact 128-bit name-hash set, and one row (interval, record, first edge, last
// edge) per qualifying (interval, record) pair.  Rows are put in the reference's emission order — interval list
// order, then file order — by one radix sort on (interval << 32 | record).  Spelling a run out as a path (a walk
// over a few nodes of ONE graph per distinct run) stays with the caller.
#include <algorithm>
#include <vector>

#include "ctx.h"
#include "dev_util.cuh"
#include "relops.cuh"

namespace coral {

struct SeqEdges {            // sorted by (tid, l), disjoint: r is then sorted as well inside one tid
    const int32_t *tid, *l, *r;
    int32_t n;
};
struct Ivals {
    const int32_t *tid, *start, *stop;     // fetch(contig, start, stop): pos < stop and endpos > start
    int32_t n;
};
struct NameSet128 {          // open addressing over indices into (key_lo, key_hi); exact: full keys are compared
    const uint64_t *key_lo, *key_hi;
    const uint32_t* slots;
    uint32_t mask;
};
constexpr uint32_t SLOT_EMPTY = 0xffffffffu;

__device__ __forceinline__ uint32_t nameset_home(uint64_t lo, uint64_t hi, uint32_t mask) {
    return (uint32_t)mix64(lo ^ mix64(hi)) & mask;
}

__global__ void k_nameset_keys_from_recs(int64_t n, const uint32_t* __restrict__ rec, const uint64_t* __restrict__ rt_lo,
                                         const uint64_t* __restrict__ rt_hi, uint64_t* __restrict__ key_lo,
                                         uint64_t* __restrict__ key_hi) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = rec[i];
    key_lo[i] = rt_lo[r];
    key_hi[i] = rt_hi[r];
}

__global__ void k_nameset_insert(int64_t n, const uint64_t* __restrict__ key_lo, const uint64_t* __restrict__ key_hi,
                                 uint32_t* __restrict__ slots, uint32_t mask) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t lo = key_lo[i], hi = key_hi[i];
    uint32_t h = nameset_home(lo, hi, mask);
    for (;;) {
        const uint32_t old = atomicCAS(slots + h, SLOT_EMPTY, (uint32_t)i);
        if (old == SLOT_EMPTY) return;
        if (key_lo[old] == lo && key_hi[old] == hi) return;       // the same name through another record
        h = (h + 1) & mask;
    }
}

__device__ __forceinline__ bool nameset_has(const NameSet128& s, uint64_t lo, uint64_t hi) {
    uint32_t h = nameset_home(lo, hi, s.mask);
    for (;;) {
        const uint32_t k = s.slots[h];
        if (k == SLOT_EMPTY) return false;
        if (s.key_lo[k] == lo && s.key_hi[k] == hi) return true;
        h = (h + 1) & s.mask;
    }
}

// first edge index in [b, e) with key[k] >= x  (key sorted ascending on [b, e))
__device__ __forceinline__ int32_t lower_bound_i32(const int32_t* __restrict__ key, int32_t b, int32_t e, int64_t x) {
    while (b < e) {
        const int32_t m = (b + e) >> 1;
        if ((int64_t)key[m] < x) b = m + 1; else e = m;
    }
    return b;
}

// status[0] = records the reference would fail on (reference_end is None: unmapped flag or no CIGAR)
__global__ void k_alignment_runs(int64_t n, const int32_t* __restrict__ tid, const int32_t* __restrict__ pos,
                                 const int32_t* __restrict__ endpos, const uint32_t* __restrict__ flag_mapq,
                                 const uint32_t* __restrict__ ncig, const uint64_t* __restrict__ name_lo,
                                 const uint64_t* __restrict__ name_hi, Ivals iv, SeqEdges se, NameSet128 names,
                                 int32_t min_overlap, uint32_t min_mapq, uint64_t* __restrict__ row_key,
                                 int32_t* __restrict__ row_first, int32_t* __restrict__ row_last,
                                 unsigned long long* __restrict__ n_rows, int64_t cap, int32_t* __restrict__ status) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t t = tid[i];
    if (t < 0) return;
    const uint32_t fm = flag_mapq[i];
    if (((fm >> 16) & 255u) < min_mapq) return;                    // ibg:1308 q >= 20
    const int64_t s = pos[i], e_fetch = endpos[i];
    uint32_t hit = 0;                                              // intervals fetching this record (n <= 32 per call)
    for (int32_t j = 0; j < iv.n; j++)
        if (iv.tid[j] == t && s < (int64_t)iv.stop[j] && e_fetch > (int64_t)iv.start[j]) hit |= 1u << j;
    if (!hit) return;
    if (!nameset_has(names, name_lo[i], name_hi[i])) return;       // ibg:1308 rn in concordant_reads
    if ((fm & 4u) || ncig[i] == 0) { atomicAdd(status, 1); return; }   // reference_end is None -> TypeError in the reference
    const int64_t e = e_fetch;      // reference_end (exclusive) is used as an inclusive bound by interval_overlap (pc:62)
    // edges of this contig
    const int32_t t0 = lower_bound_i32(se.tid, 0, se.n, t), t1 = lower_bound_i32(se.tid, 0, se.n, (int64_t)t + 1);
    int32_t a = lower_bound_i32(se.r, t0, t1, s);                  // first edge with r >= s
    int32_t b = lower_bound_i32(se.l, t0, t1, e + 1) - 1;          // last edge with l <= e
    int32_t cnt = b - a + 1;
    if (cnt <= 2) return;                                          // trimming only removes edges
    const int64_t mo = min_overlap;
#define SHARED_BP(k) (min((int64_t)se.r[k], e) - max((int64_t)se.l[k], s))
#define EDGE_SIZE(k) ((int64_t)se.r[k] - (int64_t)se.l[k] + 1)
    if (cnt > 1 && SHARED_BP(a) < mo) { a++; cnt--; }
    while (cnt > 1 && EDGE_SIZE(a) < mo) { a++; cnt--; }
    if (cnt > 1 && SHARED_BP(b) < mo) { b--; cnt--; }
    while (cnt > 1 && EDGE_SIZE(b) < mo) { b--; cnt--; }
#undef SHARED_BP
#undef EDGE_SIZE
    if (cnt <= 2) return;
    while (hit) {
        const int32_t j = __ffs(hit) - 1;
        hit &= hit - 1;
        const unsigned long long slot = atomicAdd(n_rows, 1ull);
        if ((int64_t)slot < cap) {
            row_key[slot] = ((uint64_t)j << 32) | (uint64_t)i;
            row_first[slot] = a;
            row_last[slot] = b;
        }
    }
}

__global__ void k_path_rows_ordered(int64_t n, const uint64_t* __restrict__ skey, const uint32_t* __restrict__ sidx,
                                    const int32_t* __restrict__ row_first, const int32_t* __restrict__ row_last,
                                    int32_t iv_base, uint32_t* __restrict__ o_rec, int32_t* __restrict__ o_iv,
                                    int32_t* __restrict__ o_first, int32_t* __restrict__ o_last) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t k = skey[i];
    const uint32_t src = sidx[i];
    o_rec[i] = (uint32_t)k;
    o_iv[i] = iv_base + (int32_t)(k >> 32);
    o_first[i] = row_first[src];
    o_last[i] = row_last[src];
}

__global__ void k_iota_u32(int64_t n, uint32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (uint32_t)i;
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

extern "C" int coral_alignment_paths(coral_ctx* c, const int32_t* iv_tid, const int32_t* iv_start, const int32_t* iv_stop,
                                     int32_t n_iv, const int32_t* se_tid, const int32_t* se_l, const int32_t* se_r,
                                     int32_t n_se, const uint32_t* name_rec, int64_t n_name_rec, const uint64_t* name_lo,
                                     const uint64_t* name_hi, int64_t n_name_hash, int32_t min_overlap, int32_t min_mapq,
                                     int64_t* n_rows_out) {
    using namespace coral;
    if (!c || n_iv < 0 || n_se < 0 || n_name_rec < 0 || n_name_hash < 0 || !n_rows_out) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_alignment_paths before coral_load_records"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(c));
    cudaSetDevice(c->device);
    PathTable& pt = c->paths;
    pt.n = 0; pt.valid = false;
    *n_rows_out = 0;
    for (int32_t k = 0; k < n_se; k++) {
        if (se_r[k] < se_l[k] || (k > 0 && (se_tid[k] < se_tid[k - 1] || (se_tid[k] == se_tid[k - 1] && se_l[k] <= se_r[k - 1])))) {
            c->err = "sequence edges must be sorted by (tid, l) and disjoint";
            return CORAL_E_ARG;
        }
    }
    for (int32_t j = 0; j < n_iv; j++)
        if (iv_start[j] < 0) { c->err = "ValueError: start out of range (negative fetch start)"; return CORAL_E_ARG; }
    for (int64_t k = 0; k < n_name_rec; k++)
        if (name_rec[k] >= (uint64_t)c->nrec) { c->err = "bad name record"; return CORAL_E_ARG; }
    const int64_t n_names = n_name_rec + n_name_hash;
    if (n_iv == 0 || n_se < 3 || n_names == 0 || c->nrec == 0) { pt.valid = true; return CORAL_OK; }
    if (n_names >= ((int64_t)1 << 30)) { c->err = "too many names"; return CORAL_E_ARG; }
    RecordTable& rt = c->rt;
    DevBuf* w = c->pw;
    // name set: keys (from records of this ctx and / or given hashes) + open-addressing slots at load <= 0.5
    uint32_t cap_slots = 1024;
    while ((int64_t)cap_slots < 2 * n_names) cap_slots <<= 1;
    CORAL_CUDA(c, w[0].ensure((size_t)n_names * 16));
    CORAL_CUDA(c, w[1].ensure((size_t)cap_slots * 4));
    uint64_t* key_lo = w[0].as<uint64_t>(); uint64_t* key_hi = key_lo + n_names;
    uint32_t* slots = w[1].as<uint32_t>();
    if (n_name_rec > 0) {
        CORAL_CUDA(c, w[2].ensure((size_t)n_name_rec * 4));
        CORAL_TRY(h2d(c, w[2].p, name_rec, (size_t)n_name_rec));
        LAUNCH_1D(k_nameset_keys_from_recs, n_name_rec, n_name_rec, w[2].as<uint32_t>(), rt.name_lo.as<uint64_t>(),
                  rt.name_hi.as<uint64_t>(), key_lo, key_hi);
    }
    if (n_name_hash > 0) {
        // key_hi starts at key_lo + n_names: the given hashes fill the tail of both halves
        CORAL_TRY(h2d(c, key_lo + n_name_rec, name_lo, (size_t)n_name_hash));
        CORAL_TRY(h2d(c, key_hi + n_name_rec, name_hi, (size_t)n_name_hash));
    }
    CORAL_CUDA(c, cudaMemsetAsync(slots, 0xff, (size_t)cap_slots * 4, c->stream));
    LAUNCH_1D(k_nameset_insert, n_names, n_names, key_lo, key_hi, slots, cap_slots - 1);
    NameSet128 names{key_lo, key_hi, slots, cap_slots - 1};
    // sequence edges
    CORAL_CUDA(c, w[3].ensure((size_t)n_se * 12));
    int32_t* d_se = w[3].as<int32_t>();
    CORAL_TRY(h2d(c, d_se, se_tid, (size_t)n_se));
    CORAL_TRY(h2d(c, d_se + n_se, se_l, (size_t)n_se));
    CORAL_TRY(h2d(c, d_se + 2 * (size_t)n_se, se_r, (size_t)n_se));
    SeqEdges se{d_se, d_se + n_se, d_se + 2 * (size_t)n_se, n_se};
    CORAL_CUDA(c, w[4].ensure((size_t)n_iv * 12));
    int32_t* d_iv = w[4].as<int32_t>();
    CORAL_TRY(h2d(c, d_iv, iv_tid, (size_t)n_iv));
    CORAL_TRY(h2d(c, d_iv + n_iv, iv_start, (size_t)n_iv));
    CORAL_TRY(h2d(c, d_iv + 2 * (size_t)n_iv, iv_stop, (size_t)n_iv));
    CORAL_CUDA(c, w[5].ensure(16));
    unsigned long long* d_nrows = w[5].as<unsigned long long>();
    int32_t* d_status = reinterpret_cast<int32_t*>(d_nrows + 1);
    c->stage_begin(CORAL_STAGE_PATHS);
    int64_t total = 0;
    std::vector<int64_t> chunk_rows;
    // intervals are taken 32 at a time (one mask word per record); rows of chunk k precede those of chunk k + 1
    for (int32_t j0 = 0; j0 < n_iv; j0 += 32) {
        const int32_t nj = std::min<int32_t>(32, n_iv - j0);
        Ivals iv{d_iv + j0, d_iv + n_iv + j0, d_iv + 2 * (size_t)n_iv + j0, nj};
        int64_t cap = std::max<int64_t>((int64_t)1 << 16, (int64_t)(w[6].cap / 8));
        unsigned long long n_rows = 0;
        int32_t bad = 0;
        for (int attempt = 0; attempt < 2; attempt++) {
            CORAL_CUDA(c, w[6].ensure((size_t)cap * 8)); CORAL_CUDA(c, w[7].ensure((size_t)cap * 4)); CORAL_CUDA(c, w[8].ensure((size_t)cap * 4));
            CORAL_CUDA(c, cudaMemsetAsync(d_nrows, 0, 16, c->stream));
            LAUNCH_1D(k_alignment_runs, c->nrec, c->nrec, rt.tid.as<int32_t>(), rt.pos.as<int32_t>(), rt.endpos.as<int32_t>(),
                      rt.flag_mapq.as<uint32_t>(), rt.ncig.as<uint32_t>(), rt.name_lo.as<uint64_t>(), rt.name_hi.as<uint64_t>(),
                      iv, se, names, min_overlap, (uint32_t)std::max(min_mapq, 0), w[6].as<uint64_t>(), w[7].as<int32_t>(),
                      w[8].as<int32_t>(), d_nrows, cap, d_status);
            CORAL_TRY(d2h(c, &n_rows, d_nrows, 1));
            CORAL_TRY(d2h(c, &bad, d_status, 1));
            CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
            if ((int64_t)n_rows <= cap) break;
            cap = (int64_t)n_rows + 1024;
        }
        if (bad) {
            c->err = "TypeError: reference_end is None (unmapped or CIGAR-less record with MAPQ >= min_mapq among the concordant reads)";
            return CORAL_E_FORMAT;
        }
        const int64_t n = (int64_t)n_rows;
        if (n >= ((int64_t)1 << 32)) { c->err = "too many rows"; return CORAL_E_ARG; }
        if (n > 0) {
            // order: (interval, record) — the reference's emission order
            CORAL_CUDA(c, w[9].ensure((size_t)n * 8)); CORAL_CUDA(c, w[10].ensure((size_t)n * 8));
            uint32_t* ia = w[10].as<uint32_t>(); uint32_t* ib = ia + n;
            LAUNCH_1D(k_iota_u32, n, n, ia);
            CORAL_TRY(sort_pairs_u64(c, w[6].as<uint64_t>(), w[9].as<uint64_t>(), ia, ib, n, 0, 32 + 5));
            const size_t need = (size_t)(total + n) * 4;
            DevBuf* cols[4] = {&pt.rec, &pt.iv, &pt.first, &pt.last};
            for (DevBuf* col : cols) {
                if (col->cap < need) {               // grow, keeping the rows of earlier chunks
                    DevBuf bigger;
                    CORAL_CUDA(c, bigger.ensure(need * 2));
                    if (total > 0) CORAL_CUDA(c, cudaMemcpyAsync(bigger.p, col->p, (size_t)total * 4, cudaMemcpyDeviceToDevice, c->stream));
                    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
                    col->release();
                    *col = bigger;
                }
            }
            LAUNCH_1D(k_path_rows_ordered, n, n, w[9].as<uint64_t>(), ib, w[7].as<int32_t>(), w[8].as<int32_t>(), j0,
                      pt.rec.as<uint32_t>() + total, pt.iv.as<int32_t>() + total, pt.first.as<int32_t>() + total,
                      pt.last.as<int32_t>() + total);
            total += n;
        }
    }
    c->stage_end(CORAL_STAGE_PATHS);
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    pt.n = total;
    pt.valid = true;
    *n_rows_out = total;
    return CORAL_OK;
}

extern "C" int coral_get_alignment_paths(coral_ctx* c, uint32_t* rec, int32_t* iv, int32_t* first, int32_t* last) {
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->paths.valid) { c->err = "coral_get_alignment_paths before coral_alignment_paths"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    const size_t n = (size_t)c->paths.n;
    if (n == 0) return CORAL_OK;
    if (rec) CORAL_TRY(d2h(c, rec, c->paths.rec.p, n));
    if (iv) CORAL_TRY(d2h(c, iv, c->paths.iv.p, n));
    if (first) CORAL_TRY(d2h(c, first, c->paths.first.p, n));
    if (last) CORAL_TRY(d2h(c, last, c->paths.last.p, n));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}
