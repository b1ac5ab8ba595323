// coral_b200/csrc/chimeric_merge.cu — distributed `fetch` of the multi-GPU engine.
//
// After the row exchange every rank holds ALL gathered rows (exchange.cu), but parses only the reads whose name hash
// falls into its residue class (option "fetch_part_world" / "fetch_part_rank", fetch.cu) — 1/N of the SA entries instead
// of all of them on every rank.  The N partial chimeric tables are then packed (coral_pack_chimeric), all-gathered
// (xchg.cu) and merged (coral_merge_chimeric) into the table the 1-GPU fetch would have built:
//   * reads in the order of their first SA-bearing record (the reference's dict insertion order, ibg:145-151):
//     read_first_rec is a row index of the gathered rows, which are in global file order, so it is the sort key and is
//     unique per read;
//   * each read's alignments unchanged (already in their stable (qs, qe) order, cp:263).
#include <algorithm>
#include <memory>
#include <vector>

#include "ctx.h"
#include "msg.cuh"
#include "relops.cuh"

namespace coral {

struct ChimRead {             // one read of a partial table
    uint32_t first_rec;
    int32_t status;
    int64_t read_len;
    uint32_t n_aln, aln_off;  // its alignments inside the same message
};
static_assert(sizeof(ChimRead) == 24, "ChimRead is 24 bytes on the wire");
struct ChimAln {
    int32_t qs, qe, tid, rs, re, mapq;
    int64_t nm;
    uint32_t strand, pad;
};
static_assert(sizeof(ChimAln) == 40, "ChimAln is 40 bytes on the wire");
enum { MSG_CHIM = 4, MSG_SDEL = 5 };

struct SdelRow {
    uint32_t rec;
    int32_t ai, next_start, prev_end, first_start, last_end, mapq;
    uint32_t pad;
    uint64_t lo, hi;
};
static_assert(sizeof(SdelRow) == 48, "SdelRow is 48 bytes on the wire");

__global__ void k_pack_sdel(int64_t n, const uint32_t* __restrict__ rec, const int32_t* __restrict__ ai,
                            const int32_t* __restrict__ next_start, const int32_t* __restrict__ prev_end,
                            const int32_t* __restrict__ first_start, const int32_t* __restrict__ last_end,
                            const int32_t* __restrict__ mapq, const uint64_t* __restrict__ lo, const uint64_t* __restrict__ hi,
                            SdelRow* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    SdelRow r;
    r.rec = rec[i]; r.ai = ai[i]; r.next_start = next_start[i]; r.prev_end = prev_end[i]; r.first_start = first_start[i];
    r.last_end = last_end[i]; r.mapq = mapq[i]; r.pad = 0; r.lo = lo[r.rec]; r.hi = hi[r.rec];
    out[i] = r;
}

struct ChimColsC {
    const uint32_t *read_first_rec, *aln_ptr;
    const int64_t *read_len, *nm;
    const int32_t *read_status, *qs, *qe, *tid, *rs, *re, *mapq;
    const uint8_t* strand;
};
struct ChimColsW {
    uint32_t *read_first_rec, *aln_ptr;
    int64_t *read_len, *nm;
    int32_t *read_status, *qs, *qe, *tid, *rs, *re, *mapq;
    uint8_t* strand;
};

__global__ void k_put_chim_hdr(MsgHdr* h, MsgHdr v) { *h = v; }

// (the first thread also writes the header and zeroes the padding behind the reads and behind the alignments)
__global__ void k_pack_chim(int64_t n_reads, int64_t n_aln, ChimColsC t, ChimRead* __restrict__ reads, ChimAln* __restrict__ alns,
                            MsgHdr* __restrict__ hdr_out, MsgHdr hdr, uint32_t pad_reads, uint32_t pad_alns) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        *hdr_out = hdr;
        uint8_t* p1 = reinterpret_cast<uint8_t*>(reads + n_reads);
        for (uint32_t k = 0; k < pad_reads; k++) p1[k] = 0;
        uint8_t* p2 = reinterpret_cast<uint8_t*>(alns + n_aln);
        for (uint32_t k = 0; k < pad_alns; k++) p2[k] = 0;
    }
    if (i < n_reads) {
        ChimRead r;
        r.first_rec = t.read_first_rec[i]; r.status = t.read_status[i]; r.read_len = t.read_len[i];
        r.aln_off = t.aln_ptr[i]; r.n_aln = t.aln_ptr[i + 1] - t.aln_ptr[i];
        reads[i] = r;
    }
    if (i < n_aln) {
        ChimAln a;
        a.qs = t.qs[i]; a.qe = t.qe[i]; a.tid = t.tid[i]; a.rs = t.rs[i]; a.re = t.re[i]; a.mapq = t.mapq[i];
        a.nm = t.nm[i]; a.strand = t.strand[i]; a.pad = 0;
        alns[i] = a;
    }
}

struct ChimSet {
    const uint8_t* base[CORAL_XCHG_MAX_RANKS];
    uint64_t row0[CORAL_XCHG_MAX_RANKS + 1];     // prefix sums of the reads per message
    int n;
};
__device__ __forceinline__ const ChimRead* chim_read(const ChimSet& s, uint64_t g, int* rank) {
    int r = 0;
    while (r + 1 < s.n && g >= s.row0[r + 1]) r++;
    *rank = r;
    return reinterpret_cast<const ChimRead*>(s.base[r] + sizeof(MsgHdr)) + (g - s.row0[r]);
}
__device__ __forceinline__ const ChimAln* chim_alns(const ChimSet& s, int r) {
    const uint64_t n_reads = s.row0[r + 1] - s.row0[r];
    return reinterpret_cast<const ChimAln*>(s.base[r] + sizeof(MsgHdr) + ((n_reads * sizeof(ChimRead) + 15) & ~(uint64_t)15));
}

__global__ void k_chim_keys(ChimSet s, uint64_t* __restrict__ key, uint32_t* __restrict__ idx) {
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= s.row0[s.n]) return;
    int r;
    key[g] = chim_read(s, g, &r)->first_rec;
    idx[g] = (uint32_t)g;
}
__global__ void k_chim_counts(ChimSet s, const uint32_t* __restrict__ order, uint32_t* __restrict__ cnt) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= s.row0[s.n]) return;
    int r;
    cnt[j] = chim_read(s, order[j], &r)->n_aln;
}
__global__ void k_chim_fill(ChimSet s, const uint32_t* __restrict__ order, uint32_t total_aln, ChimColsW o) {
    const uint64_t n = s.row0[s.n];
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int r;
    const ChimRead* rd = chim_read(s, order[j], &r);
    o.read_first_rec[j] = rd->first_rec;
    o.read_status[j] = rd->status;
    o.read_len[j] = rd->read_len;
    if (j == n - 1) o.aln_ptr[n] = total_aln;
    const ChimAln* a = chim_alns(s, r) + rd->aln_off;
    const uint32_t base = o.aln_ptr[j];
    for (uint32_t k = 0; k < rd->n_aln; k++) {
        const ChimAln v = a[k];
        const uint32_t d = base + k;
        o.qs[d] = v.qs; o.qe[d] = v.qe; o.tid[d] = v.tid; o.rs[d] = v.rs; o.re[d] = v.re; o.mapq[d] = v.mapq;
        o.nm[d] = v.nm; o.strand[d] = (uint8_t)v.strand;
    }
}

__global__ void k_chim_hdrs(ChimSet s, uint64_t* __restrict__ out) {
    const int t = threadIdx.x;
    const int k = t / 8, w = t % 8;
    if (k < s.n) out[t] = reinterpret_cast<const uint64_t*>(s.base[k])[w];
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

extern "C" {

// The chimeric table coral_fetch left on this ctx as one message (MsgHdr + reads + alignments) in this ctx's HBM.
int coral_pack_chimeric(coral_ctx* c, uint64_t* d_msg, uint64_t* nbytes) {
    using namespace coral;
    if (!c || !d_msg || !nbytes) return CORAL_E_ARG;
    if (!c->chim.valid) { c->err = "coral_pack_chimeric before coral_fetch"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    PassLane lane__(c);
    if (c->pass.active) pass_call(c, pass_hash(PASS_H0, "pack_chimeric", 13));
    ChimericTable& T = c->chim;
    const uint64_t nr = (uint64_t)T.n_reads, na = (uint64_t)T.n_aln;
    const uint64_t reads_bytes = (nr * sizeof(ChimRead) + 15) & ~(uint64_t)15;
    const uint64_t total = (sizeof(MsgHdr) + reads_bytes + na * sizeof(ChimAln) + 15) & ~(uint64_t)15;
    DevBuf& msg = c->xmsg[0];
    CORAL_CUDA(c, msg.ensure(total + 16));
    MsgHdr h{};
    h.n_rows = nr; h.text_bytes = na; h.kind = MSG_CHIM;
    ChimColsC t{T.read_first_rec.as<uint32_t>(), T.aln_ptr.as<uint32_t>(), T.read_len.as<int64_t>(), T.nm.as<int64_t>(),
                T.read_status.as<int32_t>(), T.qs.as<int32_t>(), T.qe.as<int32_t>(), T.tid.as<int32_t>(), T.rs.as<int32_t>(),
                T.re.as<int32_t>(), T.mapq.as<int32_t>(), T.strand.as<uint8_t>()};
    const int64_t nthreads = (int64_t)std::max<uint64_t>(std::max(nr, na), 1);      // at least one thread: it writes the header
    const uint64_t alns_bytes = total - sizeof(MsgHdr) - reads_bytes;
    LAUNCH_1D(k_pack_chim, nthreads, (int64_t)nr, (int64_t)na, t, reinterpret_cast<ChimRead*>(msg.as<uint8_t>() + sizeof(MsgHdr)),
              reinterpret_cast<ChimAln*>(msg.as<uint8_t>() + sizeof(MsgHdr) + reads_bytes), msg.as<MsgHdr>(), h,
              (uint32_t)(reads_bytes - nr * sizeof(ChimRead)), (uint32_t)(alns_bytes - na * sizeof(ChimAln)));
    if (!c->pass.active) CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    *d_msg = (uint64_t)reinterpret_cast<uintptr_t>(msg.p);
    *nbytes = total;
    return CORAL_OK;
}

// The gathered partial tables of all ranks (message r = d_buf[starts[r] : ...), on this device) -> this ctx's chimeric
// table, as the 1-GPU coral_fetch over the whole file would have left it; coral_get_chimeric / coral_candidates_* /
// coral_cluster_call* / coral_cn_segment_index then run unchanged.
int coral_merge_chimeric(coral_ctx* c, int32_t n_ranks, const uint8_t* d_buf, const uint64_t* starts, int64_t* n_reads,
                         int64_t* n_alignments) {
    using namespace coral;
    if (!c || !d_buf || !starts) return CORAL_E_ARG;
    if (n_ranks < 1 || n_ranks > CORAL_XCHG_MAX_RANKS) { c->err = "1..16 ranks per exchange"; return CORAL_E_ARG; }
    cudaSetDevice(c->device);
    PassLane lane__(c);
    if (c->pass.active) {
        uint64_t k = pass_hash(PASS_H0, "merge_chimeric", 14);
        k = pass_hash(k, &n_ranks, 4);
        pass_call(c, pass_hash(k, starts, (size_t)n_ranks * 8));
    }
    ChimSet set;
    set.n = n_ranks;
    for (int r = 0; r < n_ranks; r++) {
        if (starts[r] & 15) { c->err = "message starts must be 16-byte aligned"; return CORAL_E_ARG; }
        set.base[r] = d_buf + starts[r];
    }
    HdrPtrs hp;
    hp.n = n_ranks;
    for (int r = 0; r < n_ranks; r++) hp.base[r] = set.base[r];
    std::vector<MsgHdr> hdrs((size_t)n_ranks);
    CORAL_TRY(pass_resolve_hdrs(c, hp, reinterpret_cast<uint64_t*>(hdrs.data())));
    uint64_t R = 0, A = 0;
    set.row0[0] = 0;
    for (int r = 0; r < n_ranks; r++) {
        if (hdrs[r].kind != (uint64_t)MSG_CHIM) { c->err = "unexpected message kind in the gathered chimeric tables"; return CORAL_E_FORMAT; }
        R += hdrs[r].n_rows; A += hdrs[r].text_bytes;
        set.row0[r + 1] = R;
    }
    if (R >= ((uint64_t)1 << 32) || A >= ((uint64_t)1 << 32)) { c->err = "too many gathered reads / alignments"; return CORAL_E_ARG; }
    ChimericTable& T = c->chim;
    T.valid = false;
    CORAL_TRY(pass_side_join(c, SIDE_CHIM));
    const size_t nr = std::max<uint64_t>(R, 1), na = std::max<uint64_t>(A, 1);
    CORAL_CUDA(c, T.read_first_rec.ensure(nr * 4));
    CORAL_CUDA(c, T.read_len.ensure(nr * 8));
    CORAL_CUDA(c, T.read_status.ensure(nr * 4));
    CORAL_CUDA(c, T.aln_ptr.ensure((nr + 1) * 4));
    DevBuf* a4[] = {&T.qs, &T.qe, &T.tid, &T.rs, &T.re, &T.mapq};
    for (DevBuf* b : a4) CORAL_CUDA(c, b->ensure(na * 4));
    CORAL_CUDA(c, T.strand.ensure(na));
    CORAL_CUDA(c, T.nm.ensure(na * 8));
    if (R > 0) {
        CORAL_CUDA(c, c->s4.ensure(R * 16));
        CORAL_CUDA(c, c->s5.ensure(R * 12 + 16));
        uint64_t* key_a = c->s4.as<uint64_t>(); uint64_t* key_b = key_a + R;
        uint32_t* idx = c->s5.as<uint32_t>(); uint32_t* order = idx + R; uint32_t* cnt = order + R;
        LAUNCH_1D(k_chim_keys, (int64_t)R, set, key_a, idx);
        int key_bits = 1;                      // keys are row indices of the gathered rows (< this ctx's row count)
        while (key_bits < 32 && ((uint64_t)1 << key_bits) < (uint64_t)std::max<int64_t>(c->nrec, 2)) key_bits++;
        CORAL_TRY(sort_pairs_u64(c, key_a, key_b, idx, order, (int64_t)R, 0, key_bits));
        LAUNCH_1D(k_chim_counts, (int64_t)R, set, order, cnt);
        CORAL_TRY(exclusive_sum_u32(c, cnt, T.aln_ptr.as<uint32_t>(), (int64_t)R));
        ChimColsW o{T.read_first_rec.as<uint32_t>(), T.aln_ptr.as<uint32_t>(), T.read_len.as<int64_t>(), T.nm.as<int64_t>(),
                    T.read_status.as<int32_t>(), T.qs.as<int32_t>(), T.qe.as<int32_t>(), T.tid.as<int32_t>(), T.rs.as<int32_t>(),
                    T.re.as<int32_t>(), T.mapq.as<int32_t>(), T.strand.as<uint8_t>()};
        LAUNCH_1D(k_chim_fill, (int64_t)R, set, order, (uint32_t)A, o);
    } else {
        CORAL_CUDA(c, cudaMemsetAsync(T.aln_ptr.p, 0, 4, c->stream));
    }
    if (!c->pass.active) CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    T.n_reads = (int64_t)R;
    T.n_aln = (int64_t)A;
    T.valid = true;
    if (n_reads) *n_reads = T.n_reads;
    if (n_alignments) *n_alignments = T.n_aln;
    return CORAL_OK;
}

// The rows of the last coral_smalldel as one message (MsgHdr + n rows of 48 bytes: record, interval, next_start,
// prev_end, first_start, last_end, MAPQ, 128-bit name hash), this shard's part of find_smalldel_breakpoints' record
// loop (ibg:750-762).  The caller gathers the messages and merges them in (interval, rank, row) order = the
// reference's emission order.
int coral_smalldel_pack(coral_ctx* c, uint64_t* d_msg, uint64_t* nbytes) {
    using namespace coral;
    if (!c || !d_msg || !nbytes) return CORAL_E_ARG;
    if (!c->sdel.valid) { c->err = "coral_smalldel_pack before coral_smalldel"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    if (c->pass.active) pass_call(c, pass_hash(PASS_H0, "smalldel_pack", 13));
    SmallDelTable& T = c->sdel;
    const uint64_t n = (uint64_t)T.n;
    const uint64_t total = sizeof(MsgHdr) + n * sizeof(SdelRow);
    DevBuf& msg = c->xmsg[3];
    CORAL_CUDA(c, msg.ensure(total));
    MsgHdr h{};
    h.n_rows = n; h.kind = MSG_SDEL;
    k_put_chim_hdr<<<1, 1, 0, c->stream>>>(msg.as<MsgHdr>(), h);
    c->n_launches++;
    LAUNCH_1D(k_pack_sdel, (int64_t)n, (int64_t)n, T.rec.as<uint32_t>(), T.ai.as<int32_t>(), T.next_start.as<int32_t>(),
              T.prev_end.as<int32_t>(), T.first_start.as<int32_t>(), T.last_end.as<int32_t>(), T.mapq.as<int32_t>(),
              c->rt.name_lo.as<uint64_t>(), c->rt.name_hi.as<uint64_t>(), reinterpret_cast<SdelRow*>(msg.as<uint8_t>() + sizeof(MsgHdr)));
    if (!c->pass.active) CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    *d_msg = (uint64_t)reinterpret_cast<uintptr_t>(msg.p);
    *nbytes = total;
    return CORAL_OK;
}

// Copy of device bytes the caller got from this library (a gathered exchange buffer, a packed message) to a host array:
// now, or — inside a pipelined pass — with the pass's deferred result copies (valid when coral_pass_end returns).
int coral_read_device(coral_ctx* c, uint64_t d_src, uint64_t nbytes, uint8_t* out, int32_t lane) {
    using namespace coral;
    if (!c || (nbytes > 0 && (!d_src || !out))) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    std::unique_ptr<PassLane> guard;
    if (lane) guard.reset(new PassLane(c));
    if (c->pass.active) pass_call(c, pass_hash(pass_hash(PASS_H0, "read_device", 11), &nbytes, 8));
    CORAL_TRY(d2h_result(c, out, reinterpret_cast<const void*>((uintptr_t)d_src), (size_t)nbytes));
    return pass_sync(c);
}

}  // extern "C"
