// coral_b200/csrc/breakpoints.cu — breakpoint candidate generation.
//
//  * small-deletion rows: reference infer_breakpoint_graph.py:750-762 (records overlapping each amplicon
//    interval, MAPQ >= 20, consecutive aligned blocks more than min_del_len apart).  The gap events were
//    found by the record scan; here they are joined with the interval list and put in emission order
//    (interval order, then file order, then block order).
//  * chimeric candidates: breakpoint_utilities.py:129-186 `alignment2bp_l` with :289-295 `interval2bp`,
//    :11-15 `interval_overlap` and :37-44 `interval_overlap_l`, over the device-resident chimeric table.
#include <vector>

#include <cstring>

#include "ctx.h"
#include "dev_util.cuh"
#include "relops.cuh"
#include "tables.h"

namespace coral {

struct Intervals {
    const int32_t *tid, *start, *end;  // end inclusive (amplicon_intervals[ai][2])
    int32_t n;
};

// ---- small-deletion rows ---------------------------------------------------------------------------
template <bool FILL>
__global__ void k_smalldel_join(int64_t n_ev, const uint32_t* __restrict__ ev_rec, const uint32_t* __restrict__ ev_k,
                                const int32_t* __restrict__ tid, const int32_t* __restrict__ pos,
                                const int32_t* __restrict__ endpos, const uint32_t* __restrict__ flag_mapq, int32_t min_mapq,
                                Intervals I, uint32_t* __restrict__ cnt, const uint32_t* __restrict__ base,
                                uint64_t* __restrict__ key, uint32_t* __restrict__ row_ev) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_ev) return;
    const uint32_t r = ev_rec[e];
    uint32_t c = 0;
    if ((int32_t)(flag_mapq[r] >> 16) >= min_mapq) {
        const int32_t t = tid[r], p = pos[r], q = endpos[r];
        for (int32_t a = 0; a < I.n; a++) {
            // fetch(ai[0], ai[1], ai[2] + 1): pos < stop and endpos > start
            if (I.tid[a] == t && p < I.end[a] + 1 && q > I.start[a]) {
                if (FILL) {
                    key[base[e] + c] = ((uint64_t)a << 48) | ((uint64_t)r << 16) | (uint64_t)(ev_k[e] & 0xffffu);
                    row_ev[base[e] + c] = (uint32_t)e;
                }
                c++;
            }
        }
    }
    if (!FILL) cnt[e] = c;
}

struct SmallDelRows {
    uint32_t* rec;
    int32_t *ai, *next_start, *prev_end, *first_start, *last_end, *mapq;
};

__global__ void k_smalldel_rows(int64_t n_rows, const uint64_t* __restrict__ key, const uint32_t* __restrict__ row_ev,
                                const int32_t* __restrict__ ev_next, const int32_t* __restrict__ ev_prev,
                                const int32_t* __restrict__ fbs, const int32_t* __restrict__ lbe,
                                const uint32_t* __restrict__ flag_mapq, SmallDelRows o) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows) return;
    const uint64_t k = key[i];
    const uint32_t e = row_ev[i];
    const uint32_t r = (uint32_t)((k >> 16) & 0xffffffffu);
    o.rec[i] = r;
    o.ai[i] = (int32_t)(k >> 48);
    o.next_start[i] = ev_next[e];
    o.prev_end[i] = ev_prev[e];
    o.first_start[i] = fbs[r];
    o.last_end[i] = lbe[r];
    o.mapq[i] = (int32_t)(flag_mapq[r] >> 16);
}

// ---- alignment2bp_l ----------------------------------------------------------------------------------
struct ChimView {
    const uint32_t* aln_ptr;
    const int32_t *status, *qs, *qe, *tid, *rs, *re, *mapq;
    const uint8_t* strand;
};
struct CandCols {
    uint32_t* read;
    int32_t *i, *j, *chr1, *pos1, *chr2, *pos2, *rgap, *mq1, *mq2;
    uint8_t *o1, *o2, *flag;
};

// interval_overlap_l(rint, intervals): literal `int1[1] <= int2[2] and int2[1] <= int1[2]` (bu:15)
__device__ __forceinline__ int32_t overlap_first(const Intervals& I, int32_t t, int32_t s, int32_t e) {
    for (int32_t a = 0; a < I.n; a++)
        if (I.tid[a] == t && s <= I.end[a] && I.start[a] <= e) return a;
    return -1;
}

// same-strand firing rule bu:150-161 / :176-185
__device__ __forceinline__ bool fires(const ChimView& C, uint32_t a, uint32_t b, int32_t gap_) {
    if (C.strand[a] != C.strand[b]) return true;
    const int64_t gr = (int64_t)C.qs[b] - C.qe[a];
    const int64_t grr = C.strand[b] == 0 ? (int64_t)C.rs[b] - C.re[a] : (int64_t)C.re[a] - C.rs[b];
    const double lim = fmax((double)gap_, fabs(__dmul_rn((double)gr, 0.2)));
    int64_t d = gr - grr;
    if (d < 0) d = -d;
    return (double)d > lim;
}

// interval2bp(R1 = a, R2 = b, (rn, ia, ib), rgap) + [q[a], q[b]]  (bu:289-295)
__device__ __forceinline__ void emit(const ChimView& C, const int32_t* __restrict__ chr_rank, uint32_t read, uint32_t base,
                                     uint32_t a, uint32_t b, int32_t rgap, uint32_t slot, CandCols o, int32_t* status) {
    const int32_t c1 = chr_rank[C.tid[a]], c2 = chr_rank[C.tid[b]];
    if (c1 < 0 || c2 < 0) atomicAdd(&status[1], 1);  // KeyError in global_names.chr_idx
    const bool keep = (c2 < c1) || (c2 == c1 && C.rs[b] < C.re[a]);
    o.read[slot] = read;
    o.rgap[slot] = rgap;
    if (keep) {
        o.chr1[slot] = C.tid[a]; o.pos1[slot] = C.re[a]; o.o1[slot] = C.strand[a];
        o.chr2[slot] = C.tid[b]; o.pos2[slot] = C.rs[b]; o.o2[slot] = C.strand[b] ^ 1;
        o.i[slot] = (int32_t)(a - base); o.j[slot] = (int32_t)(b - base); o.flag[slot] = 0;
    } else {
        o.chr1[slot] = C.tid[b]; o.pos1[slot] = C.rs[b]; o.o1[slot] = C.strand[b] ^ 1;
        o.chr2[slot] = C.tid[a]; o.pos2[slot] = C.re[a]; o.o2[slot] = C.strand[a];
        o.i[slot] = (int32_t)(b - base); o.j[slot] = (int32_t)(a - base); o.flag[slot] = 1;
    }
    o.mq1[slot] = C.mapq[a];
    o.mq2[slot] = C.mapq[b];
}

// thread per read.  FILL=false counts, FILL=true writes at cand_ptr[read].
template <bool FILL>
__global__ void k_align2bp_l(int64_t n_reads, ChimView C, Intervals I, const int32_t* __restrict__ chr_rank,
                             int32_t min_bp_match_cutoff, int32_t min_mapq, int32_t gap_, int32_t gap_mapq,
                             uint32_t* __restrict__ cnt, const uint32_t* __restrict__ cand_ptr, CandCols o, int32_t* status) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    const uint32_t b0 = C.aln_ptr[r], b1 = C.aln_ptr[r + 1];
    const uint32_t n = b1 - b0;
    uint32_t c = 0;
    const uint32_t slot0 = FILL ? cand_ptr[r] : 0;
    if (n >= 2) {
        // bassigned is consulted only for neighbours: keep it as a 64-bit window when n <= 64, else recompute
        for (uint32_t k = b0; k + 1 < b1; k++) {
            const int32_t gap = C.qs[k + 1] - C.qe[k];
            const int32_t io1 = overlap_first(I, C.tid[k], C.rs[k], C.re[k]);
            const int32_t io2 = overlap_first(I, C.tid[k + 1], C.rs[k + 1], C.re[k + 1]);
            if ((int64_t)gap + min_bp_match_cutoff >= 0 && io1 >= 0 && io1 == io2 && fires(C, k, k + 1, gap_) &&
                C.mapq[k] >= min_mapq && C.mapq[k + 1] >= min_mapq) {
                if (FILL) emit(C, chr_rank, (uint32_t)r, b0, k, k + 1, gap, slot0 + c, o, status);
                c++;
            }
        }
        for (uint32_t k = b0 + 1; k + 1 < b1; k++) {
            if (!(C.mapq[k] < gap_mapq && C.mapq[k - 1] >= min_mapq && C.mapq[k + 1] >= min_mapq)) continue;
            // bassigned[k-1] / bassigned[k]: re-evaluate the adjacent-pair predicate
            bool assigned = false;
            for (uint32_t m = k - 1; m <= k && !assigned; m++) {
                const int32_t gap = C.qs[m + 1] - C.qe[m];
                const int32_t a1 = overlap_first(I, C.tid[m], C.rs[m], C.re[m]);
                const int32_t a2 = overlap_first(I, C.tid[m + 1], C.rs[m + 1], C.re[m + 1]);
                assigned = (int64_t)gap + min_bp_match_cutoff >= 0 && a1 >= 0 && a1 == a2 && fires(C, m, m + 1, gap_) &&
                           C.mapq[m] >= min_mapq && C.mapq[m + 1] >= min_mapq;
            }
            if (assigned) continue;
            const int32_t io1 = overlap_first(I, C.tid[k - 1], C.rs[k - 1], C.re[k - 1]);
            const int32_t io2 = overlap_first(I, C.tid[k + 1], C.rs[k + 1], C.re[k + 1]);
            if (io1 >= 0 && io1 == io2 && fires(C, k - 1, k + 1, gap_)) {
                if (FILL) emit(C, chr_rank, (uint32_t)r, b0, k - 1, k + 1, C.qs[k + 1] - C.qe[k - 1], slot0 + c, o, status);
                c++;
            }
        }
    }
    if (!FILL) cnt[r] = c;
}

// alignment2bp (bu:70-96): two-interval form used by the amplicon-interval search (ibg:433).
// thread per selected read.  FILL=false counts.
__device__ __forceinline__ bool ov1(const Intervals& I, int32_t a, int32_t t, int32_t s, int32_t e) {
    return I.tid[a] == t && s <= I.end[a] && I.start[a] <= e;
}
template <bool FILL>
__global__ void k_align2bp_pair(int64_t n_sel, const uint32_t* __restrict__ sel, ChimView C, Intervals I,
                                const int32_t* __restrict__ chr_rank, int32_t min_bp_match_cutoff, int32_t min_mapq,
                                int32_t gap_mapq, uint32_t* __restrict__ cnt, const uint32_t* __restrict__ cand_ptr, CandCols o,
                                int32_t* status) {
    int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= n_sel) return;
    const uint32_t r = sel[x];
    const uint32_t b0 = C.aln_ptr[r], b1 = C.aln_ptr[r + 1];
    uint32_t c = 0;
    const uint32_t slot0 = FILL ? cand_ptr[x] : 0;
    auto spans = [&](uint32_t a, uint32_t b) {
        return (ov1(I, 0, C.tid[a], C.rs[a], C.re[a]) && ov1(I, 1, C.tid[b], C.rs[b], C.re[b])) ||
               (ov1(I, 0, C.tid[b], C.rs[b], C.re[b]) && ov1(I, 1, C.tid[a], C.rs[a], C.re[a]));
    };
    auto adj = [&](uint32_t k) {
        const int32_t gap = C.qs[k + 1] - C.qe[k];
        return (int64_t)gap + min_bp_match_cutoff >= 0 && C.mapq[k] >= min_mapq && C.mapq[k + 1] >= min_mapq && spans(k, k + 1);
    };
    if (b1 - b0 >= 2) {
        for (uint32_t k = b0; k + 1 < b1; k++)
            if (adj(k)) {
                if (FILL) emit(C, chr_rank, r, b0, k, k + 1, C.qs[k + 1] - C.qe[k], slot0 + c, o, status);
                c++;
            }
        for (uint32_t k = b0 + 1; k + 1 < b1; k++) {
            if (!(C.mapq[k] < gap_mapq && C.mapq[k - 1] >= min_mapq && C.mapq[k + 1] >= min_mapq)) continue;
            if (adj(k - 1) || adj(k)) continue;
            if (spans(k - 1, k + 1)) {
                if (FILL) emit(C, chr_rank, r, b0, k - 1, k + 1, C.qs[k + 1] - C.qe[k - 1], slot0 + c, o, status);
                c++;
            }
        }
    }
    if (!FILL) cnt[x] = c;
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

static uint64_t interval_key(const char* name, const int32_t* tid, const int32_t* start, const int32_t* end, int32_t n) {
    using namespace coral;
    uint64_t k = pass_hash(PASS_H0, name, strlen(name));
    k = pass_hash(k, &n, 4);
    if (n > 0) { k = pass_hash(k, tid, (size_t)n * 4); k = pass_hash(k, start, (size_t)n * 4); k = pass_hash(k, end, (size_t)n * 4); }
    return k;
}

static int upload_intervals(coral_ctx* c, DevBuf& buf, const int32_t* tid, const int32_t* start, const int32_t* end,
                            int32_t n, coral::Intervals* out) {
    using namespace coral;
    CORAL_CUDA(c, buf.ensure((size_t)std::max(n, 1) * 12));
    int32_t* d = buf.as<int32_t>();
    CORAL_TRY(h2d_keyed(c, d, tid, (size_t)n));
    CORAL_TRY(h2d_keyed(c, d + n, start, (size_t)n));
    CORAL_TRY(h2d_keyed(c, d + 2 * (size_t)n, end, (size_t)n));
    *out = Intervals{d, d + n, d + 2 * (size_t)n, n};
    return CORAL_OK;
}

extern "C" int coral_smalldel(coral_ctx* c, const int32_t* tid, const int32_t* start, const int32_t* end_incl,
                              int32_t n_intervals, int64_t* n_rows) {
    using namespace coral;
    if (!c || n_intervals < 0) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_smalldel before coral_load_records"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(c));
    if (n_intervals >= 65536) { c->err = "more than 65535 amplicon intervals"; return CORAL_E_ARG; }
    cudaSetDevice(c->device);
    SmallDelTable& T = c->sdel;
    T.valid = false; T.n = 0;
    CORAL_TRY(pass_side_join(c, SIDE_SDEL));
    const int64_t n_ev = c->gaps.n;
    if (n_rows) *n_rows = 0;
    if (c->pass.active) pass_call(c, interval_key("smalldel", tid, start, end_incl, n_intervals));
    if (n_ev == 0 || n_intervals == 0) { T.valid = true; return CORAL_OK; }
    Intervals I;
    CORAL_TRY(upload_intervals(c, c->s0, tid, start, end_incl, n_intervals, &I));
    RecordTable& rt = c->rt;
    GapEvents& g = c->gaps;
    c->stage_begin(CORAL_STAGE_SMALLDEL);
    CORAL_CUDA(c, c->s1.ensure((size_t)(n_ev + 1) * 8 + 16));
    uint32_t* cnt = c->s1.as<uint32_t>();
    uint32_t* base = cnt + n_ev + 1;
    LAUNCH_1D(k_smalldel_join<false>, n_ev, n_ev, g.rec.as<uint32_t>(), g.k.as<uint32_t>(), rt.tid.as<int32_t>(),
              rt.pos.as<int32_t>(), rt.endpos.as<int32_t>(), rt.flag_mapq.as<uint32_t>(), c->prm.min_mapq, I, cnt,
              (const uint32_t*)nullptr, (uint64_t*)nullptr, (uint32_t*)nullptr);
    CORAL_CUDA(c, cudaMemsetAsync(cnt + n_ev, 0, 4, c->stream));
    CORAL_TRY(exclusive_sum_u32(c, cnt, base, n_ev + 1));
    uint32_t n = 0;
    CORAL_TRY(pass_resolve(c, &n, base + n_ev));
    if (n > 0) {
        CORAL_CUDA(c, c->s2.ensure((size_t)n * 24 + 64));
        uint64_t* key_a = c->s2.as<uint64_t>();
        uint64_t* key_b = key_a + n;
        uint32_t* ev_a = reinterpret_cast<uint32_t*>(key_b + n);
        uint32_t* ev_b = ev_a + n;
        LAUNCH_1D(k_smalldel_join<true>, n_ev, n_ev, g.rec.as<uint32_t>(), g.k.as<uint32_t>(), rt.tid.as<int32_t>(),
                  rt.pos.as<int32_t>(), rt.endpos.as<int32_t>(), rt.flag_mapq.as<uint32_t>(), c->prm.min_mapq, I, cnt, base,
                  key_a, ev_a);
        CORAL_TRY(sort_pairs_u64(c, key_a, key_b, ev_a, ev_b, n));
        DevBuf* cols[] = {&T.rec, &T.ai, &T.next_start, &T.prev_end, &T.first_start, &T.last_end, &T.mapq};
        for (DevBuf* b : cols) CORAL_CUDA(c, b->ensure((size_t)n * 4));
        SmallDelRows o{T.rec.as<uint32_t>(), T.ai.as<int32_t>(), T.next_start.as<int32_t>(), T.prev_end.as<int32_t>(),
                       T.first_start.as<int32_t>(), T.last_end.as<int32_t>(), T.mapq.as<int32_t>()};
        LAUNCH_1D(k_smalldel_rows, (int64_t)n, (int64_t)n, key_b, ev_b, g.next_start.as<int32_t>(), g.prev_end.as<int32_t>(),
                  rt.fbs.as<int32_t>(), rt.lbe.as<int32_t>(), rt.flag_mapq.as<uint32_t>(), o);
    }
    c->stage_end(CORAL_STAGE_SMALLDEL);
    CORAL_TRY(pass_sync(c));
    T.n = n;
    T.valid = true;
    if (n_rows) *n_rows = n;
    return CORAL_OK;
}

extern "C" int coral_get_smalldel(coral_ctx* c, uint32_t* rec, int32_t* ai, int32_t* next_start, int32_t* prev_end,
                                  int32_t* first_start, int32_t* last_end, int32_t* mapq) {
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->sdel.valid) { c->err = "coral_get_smalldel before coral_smalldel"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    SmallDelTable& T = c->sdel;
    const size_t n = (size_t)T.n;
    if (n) CORAL_TRY(pass_side_begin(c, SIDE_SDEL));   // the table stays as it is until the next coral_smalldel
    if (rec) CORAL_TRY(d2h_result(c, rec, T.rec.p, n));
    if (ai) CORAL_TRY(d2h_result(c, ai, T.ai.p, n));
    if (next_start) CORAL_TRY(d2h_result(c, next_start, T.next_start.p, n));
    if (prev_end) CORAL_TRY(d2h_result(c, prev_end, T.prev_end.p, n));
    if (first_start) CORAL_TRY(d2h_result(c, first_start, T.first_start.p, n));
    if (last_end) CORAL_TRY(d2h_result(c, last_end, T.last_end.p, n));
    if (mapq) CORAL_TRY(d2h_result(c, mapq, T.mapq.p, n));
    return pass_sync(c);
}

extern "C" int coral_candidates_l(coral_ctx* c, const int32_t* tid, const int32_t* start, const int32_t* end_incl,
                                  int32_t n_intervals, int64_t* n_cand) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c || n_intervals < 0) return CORAL_E_ARG;
    if (!c->chim.valid) { c->err = "coral_candidates_l before coral_fetch"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    CandTable& T = c->cand;
    T.valid = false; T.n = 0;
    CORAL_TRY(pass_side_join(c, SIDE_CAND));
    if (n_cand) *n_cand = 0;
    ChimericTable& H = c->chim;
    const int64_t nr = H.n_reads;
    if (c->pass.active) pass_call(c, interval_key("candidates_l", tid, start, end_incl, n_intervals));
    if (nr == 0) { T.valid = true; return CORAL_OK; }
    Intervals I;
    CORAL_TRY(upload_intervals(c, c->s0, tid, start, end_incl, n_intervals, &I));
    ChimView V{H.aln_ptr.as<uint32_t>(), H.read_status.as<int32_t>(), H.qs.as<int32_t>(), H.qe.as<int32_t>(),
               H.tid.as<int32_t>(), H.rs.as<int32_t>(), H.re.as<int32_t>(), H.mapq.as<int32_t>(), H.strand.as<uint8_t>()};
    c->stage_begin(CORAL_STAGE_CANDIDATES);
    CORAL_CUDA(c, c->s1.ensure((size_t)(nr + 1) * 8 + 16));
    uint32_t* cnt = c->s1.as<uint32_t>();
    uint32_t* ptr = cnt + nr + 1;
    CORAL_CUDA(c, c->status.ensure(64));
    CORAL_CUDA(c, cudaMemsetAsync(c->status.p, 0, 64, c->stream));
    CandCols none{};
    LAUNCH_1D(k_align2bp_l<false>, nr, nr, V, I, c->d_chr_rank.as<int32_t>(), c->prm.min_bp_match_cutoff, c->prm.min_mapq,
              c->prm.gap, c->prm.gap_mapq, cnt, (const uint32_t*)nullptr, none, c->status.as<int32_t>());
    CORAL_CUDA(c, cudaMemsetAsync(cnt + nr, 0, 4, c->stream));
    CORAL_TRY(exclusive_sum_u32(c, cnt, ptr, nr + 1));
    uint32_t n = 0;
    CORAL_TRY(pass_resolve(c, &n, ptr + nr));
    if (n > 0) {
        DevBuf* c4[] = {&T.read, &T.i, &T.j, &T.chr1, &T.pos1, &T.chr2, &T.pos2, &T.rgap, &T.mq1, &T.mq2};
        for (DevBuf* b : c4) CORAL_CUDA(c, b->ensure((size_t)n * 4));
        DevBuf* c1[] = {&T.o1, &T.o2, &T.flag};
        for (DevBuf* b : c1) CORAL_CUDA(c, b->ensure((size_t)n));
        CandCols o{T.read.as<uint32_t>(), T.i.as<int32_t>(), T.j.as<int32_t>(), T.chr1.as<int32_t>(), T.pos1.as<int32_t>(),
                   T.chr2.as<int32_t>(), T.pos2.as<int32_t>(), T.rgap.as<int32_t>(), T.mq1.as<int32_t>(), T.mq2.as<int32_t>(),
                   T.o1.as<uint8_t>(), T.o2.as<uint8_t>(), T.flag.as<uint8_t>()};
        LAUNCH_1D(k_align2bp_l<true>, nr, nr, V, I, c->d_chr_rank.as<int32_t>(), c->prm.min_bp_match_cutoff, c->prm.min_mapq,
                  c->prm.gap, c->prm.gap_mapq, cnt, ptr, o, c->status.as<int32_t>());
    }
    c->stage_end(CORAL_STAGE_CANDIDATES);
    int32_t st[2] = {0, 0};
    uint64_t st64 = 0;
    CORAL_TRY(pass_resolve(c, &st64, c->status.p));          // two status words
    memcpy(st, &st64, 8);
    if (st[1] != 0) { c->err = "KeyError: contig of a breakpoint is not in global_names.chr_idx"; return CORAL_E_FORMAT; }
    T.n = n;
    T.valid = true;
    if (n_cand) *n_cand = n;
    return CORAL_OK;
}

extern "C" int coral_get_candidates(coral_ctx* c, uint32_t* read, int32_t* i, int32_t* j, int32_t* chr1, int32_t* pos1,
                                    uint8_t* o1, int32_t* chr2, int32_t* pos2, uint8_t* o2, int32_t* rgap, uint8_t* flag,
                                    int32_t* mq1, int32_t* mq2) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->cand.valid) { c->err = "coral_get_candidates before coral_candidates_l"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    CandTable& T = c->cand;
    const size_t n = (size_t)T.n;
    if (n) CORAL_TRY(pass_side_begin(c, SIDE_CAND));   // the table stays as it is until the next coral_candidates_*
    if (read) CORAL_TRY(d2h_result(c, read, T.read.p, n));
    if (i) CORAL_TRY(d2h_result(c, i, T.i.p, n));
    if (j) CORAL_TRY(d2h_result(c, j, T.j.p, n));
    if (chr1) CORAL_TRY(d2h_result(c, chr1, T.chr1.p, n));
    if (pos1) CORAL_TRY(d2h_result(c, pos1, T.pos1.p, n));
    if (o1) CORAL_TRY(d2h_result(c, o1, T.o1.p, n));
    if (chr2) CORAL_TRY(d2h_result(c, chr2, T.chr2.p, n));
    if (pos2) CORAL_TRY(d2h_result(c, pos2, T.pos2.p, n));
    if (o2) CORAL_TRY(d2h_result(c, o2, T.o2.p, n));
    if (rgap) CORAL_TRY(d2h_result(c, rgap, T.rgap.p, n));
    if (flag) CORAL_TRY(d2h_result(c, flag, T.flag.p, n));
    if (mq1) CORAL_TRY(d2h_result(c, mq1, T.mq1.p, n));
    if (mq2) CORAL_TRY(d2h_result(c, mq2, T.mq2.p, n));
    return pass_sync(c);
}

// Replace the device chimeric table with one supplied by the caller (operator forms of alignment2bp*).
extern "C" int coral_set_chimeric(coral_ctx* c, int64_t n_reads, const uint32_t* aln_ptr, const int32_t* qs, const int32_t* qe,
                                  const int32_t* tid, const int32_t* rstart, const int32_t* rend, const uint8_t* strand,
                                  const int32_t* mapq) {
    using namespace coral;
    if (!c || n_reads < 0 || !aln_ptr) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    ChimericTable& T = c->chim;
    T.valid = false;
    CORAL_TRY(pass_side_join(c, SIDE_CHIM));
    const size_t nr = (size_t)n_reads, na = aln_ptr[n_reads];
    const size_t na1 = std::max<size_t>(na, 1), nr1 = std::max<size_t>(nr, 1);
    CORAL_CUDA(c, T.aln_ptr.ensure((nr + 1) * 4)); CORAL_CUDA(c, T.read_status.ensure(nr1 * 4));
    CORAL_CUDA(c, T.read_first_rec.ensure(nr1 * 4)); CORAL_CUDA(c, T.read_len.ensure(nr1 * 8));
    DevBuf* a4[] = {&T.qs, &T.qe, &T.tid, &T.rs, &T.re, &T.mapq};
    for (DevBuf* b : a4) CORAL_CUDA(c, b->ensure(na1 * 4));
    CORAL_CUDA(c, T.strand.ensure(na1)); CORAL_CUDA(c, T.nm.ensure(na1 * 8));
    CORAL_TRY(h2d(c, T.aln_ptr.p, aln_ptr, nr + 1));
    CORAL_CUDA(c, cudaMemsetAsync(T.read_status.p, 0, nr1 * 4, c->stream));
    CORAL_TRY(h2d(c, T.qs.p, qs, na)); CORAL_TRY(h2d(c, T.qe.p, qe, na)); CORAL_TRY(h2d(c, T.tid.p, tid, na));
    CORAL_TRY(h2d(c, T.rs.p, rstart, na)); CORAL_TRY(h2d(c, T.re.p, rend, na)); CORAL_TRY(h2d(c, T.strand.p, strand, na));
    CORAL_TRY(h2d(c, T.mapq.p, mapq, na));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    const int32_t nref = (int32_t)c->ref_names.size();
    for (size_t k = 0; k < na; k++)
        if (tid[k] < 0 || tid[k] >= nref) { c->err = "alignment tid outside the reference table"; return CORAL_E_ARG; }
    T.n_reads = n_reads; T.n_aln = (int64_t)na; T.valid = true;
    return CORAL_OK;
}

extern "C" int coral_candidates_pair(coral_ctx* c, const uint32_t* reads, int64_t n_sel, const int32_t* tid2,
                                     const int32_t* start2, const int32_t* end2, int64_t* n_cand) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c || n_sel < 0) return CORAL_E_ARG;
    if (!c->chim.valid) { c->err = "coral_candidates_pair before coral_fetch"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    CandTable& T = c->cand;
    T.valid = false; T.n = 0;
    CORAL_TRY(pass_side_join(c, SIDE_CAND));
    if (n_cand) *n_cand = 0;
    ChimericTable& H = c->chim;
    if (c->pass.active) {
        uint64_t k = interval_key("candidates_pair", tid2, start2, end2, 2);
        k = pass_hash(k, &n_sel, 8);
        if (n_sel > 0) k = pass_hash(k, reads, (size_t)n_sel * 4);
        pass_call(c, k);
    }
    if (n_sel == 0) { T.valid = true; return CORAL_OK; }
    for (int64_t k = 0; k < n_sel; k++) if ((int64_t)reads[k] >= H.n_reads) { c->err = "read index out of range"; return CORAL_E_ARG; }
    Intervals I;
    CORAL_TRY(upload_intervals(c, c->s0, tid2, start2, end2, 2, &I));
    ChimView V{H.aln_ptr.as<uint32_t>(), H.read_status.as<int32_t>(), H.qs.as<int32_t>(), H.qe.as<int32_t>(),
               H.tid.as<int32_t>(), H.rs.as<int32_t>(), H.re.as<int32_t>(), H.mapq.as<int32_t>(), H.strand.as<uint8_t>()};
    c->stage_begin(CORAL_STAGE_CANDIDATES);
    CORAL_CUDA(c, c->s1.ensure((size_t)(n_sel + 1) * 12 + 16));
    uint32_t* cnt = c->s1.as<uint32_t>();
    uint32_t* ptr = cnt + n_sel + 1;
    uint32_t* d_sel = ptr + n_sel + 1;
    CORAL_TRY(h2d_keyed(c, d_sel, reads, (size_t)n_sel));
    CORAL_CUDA(c, c->status.ensure(64));
    CORAL_CUDA(c, cudaMemsetAsync(c->status.p, 0, 64, c->stream));
    CandCols none{};
    LAUNCH_1D(k_align2bp_pair<false>, n_sel, n_sel, d_sel, V, I, c->d_chr_rank.as<int32_t>(), c->prm.min_bp_match_cutoff,
              c->prm.min_mapq, c->prm.gap_mapq, cnt, (const uint32_t*)nullptr, none, c->status.as<int32_t>());
    CORAL_CUDA(c, cudaMemsetAsync(cnt + n_sel, 0, 4, c->stream));
    CORAL_TRY(exclusive_sum_u32(c, cnt, ptr, n_sel + 1));
    uint32_t n = 0;
    CORAL_TRY(d2h(c, &n, ptr + n_sel, 1));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    if (n > 0) {
        DevBuf* c4[] = {&T.read, &T.i, &T.j, &T.chr1, &T.pos1, &T.chr2, &T.pos2, &T.rgap, &T.mq1, &T.mq2};
        for (DevBuf* b : c4) CORAL_CUDA(c, b->ensure((size_t)n * 4));
        DevBuf* c1[] = {&T.o1, &T.o2, &T.flag};
        for (DevBuf* b : c1) CORAL_CUDA(c, b->ensure((size_t)n));
        CandCols o{T.read.as<uint32_t>(), T.i.as<int32_t>(), T.j.as<int32_t>(), T.chr1.as<int32_t>(), T.pos1.as<int32_t>(),
                   T.chr2.as<int32_t>(), T.pos2.as<int32_t>(), T.rgap.as<int32_t>(), T.mq1.as<int32_t>(), T.mq2.as<int32_t>(),
                   T.o1.as<uint8_t>(), T.o2.as<uint8_t>(), T.flag.as<uint8_t>()};
        LAUNCH_1D(k_align2bp_pair<true>, n_sel, n_sel, d_sel, V, I, c->d_chr_rank.as<int32_t>(), c->prm.min_bp_match_cutoff,
                  c->prm.min_mapq, c->prm.gap_mapq, cnt, ptr, o, c->status.as<int32_t>());
    }
    c->stage_end(CORAL_STAGE_CANDIDATES);
    int32_t st[2] = {0, 0};
    CORAL_TRY(d2h(c, st, c->status.p, 2));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    if (st[1] != 0) { c->err = "KeyError: contig of a breakpoint is not in global_names.chr_idx"; return CORAL_E_FORMAT; }
    T.n = n; T.valid = true;
    if (n_cand) *n_cand = n;
    return CORAL_OK;
}
