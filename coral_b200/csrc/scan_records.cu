// coral_b200/csrc/scan_records.cu — K0: the one pass over raw BAM record bytes.
//
// Replaces, for every record, the htslib/pysam decode work that the reference triggers from
// infer_breakpoint_graph.py: bam_endpos for fetch(region) (:751,:1031,:1043), get_blocks() (:756),
// infer_read_length() (:1031), the per-base loop of count_coverage (:131,:1033), query_name /
// flag / query_length / mapping_quality / get_tag('SA:Z:') / get_cigar_stats()[0][-1] (:141-154).
// Everything downstream works on the SoA table this kernel writes; QUAL bytes are never touched.
//
// Work decomposition: one warp per record, grid-stride.  CIGAR ops are read lane-strided (coalesced),
// reference offsets come from a warp prefix sum, gap events are emitted by ballot; SEQ is scanned
// with 16-byte vector loads and a SWAR nibble test; the aux TLV is walked warp-uniformly with a
// lane-parallel strlen for Z fields.
#include "ctx.h"
#include "dev_util.cuh"

namespace coral {

struct ScanCols {
    int32_t *tid, *pos, *endpos, *fbs, *lbe, *nm;
    uint32_t *flag_mapq, *lseq, *ncig, *infer, *mbases, *nonacgt, *sa_off, *sa_len, *ngap;
    uint64_t *name_lo, *name_hi;
};
struct ScanEvents {
    uint32_t *rec, *k;
    int32_t *next_start, *prev_end;
    unsigned long long* counter;
    int64_t capacity;
};

__device__ __forceinline__ bool op_is_m(uint32_t op) { return op == 0 || op == 7 || op == 8; }
__device__ __forceinline__ bool op_ref(uint32_t op) { return op == 0 || op == 2 || op == 3 || op == 7 || op == 8; }
__device__ __forceinline__ bool op_qh(uint32_t op) { return op == 0 || op == 1 || op == 4 || op == 5 || op == 7 || op == 8; }

constexpr int SCAN_WARPS = 8;

__global__ void __launch_bounds__(SCAN_WARPS * 32)
k_scan_records(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off, int64_t n, int32_t min_del_len,
               ScanCols o, ScanEvents ev, int32_t* __restrict__ status) {
    const uint32_t lane = lane_id();
    const uint32_t lt_mask = (1u << lane) - 1u;
    int64_t rec = (int64_t)blockIdx.x * SCAN_WARPS + (threadIdx.x >> 5);
    const int64_t stride = (int64_t)gridDim.x * SCAN_WARPS;
    for (; rec < n; rec += stride) {
        const uint8_t* r = recs + off[rec];
        // ---- 36-byte core (record start is 16-byte aligned) ------------------------------------
        uint32_t cw = lane < 9 ? __ldg(reinterpret_cast<const uint32_t*>(r) + lane) : 0u;
        const uint32_t block_size = __shfl_sync(FULL_MASK, cw, 0);
        const int32_t tid = (int32_t)__shfl_sync(FULL_MASK, cw, 1);
        const int32_t pos = (int32_t)__shfl_sync(FULL_MASK, cw, 2);
        const uint32_t w3 = __shfl_sync(FULL_MASK, cw, 3);
        const uint32_t w4 = __shfl_sync(FULL_MASK, cw, 4);
        const uint32_t l_seq = __shfl_sync(FULL_MASK, cw, 5);
        const uint32_t l_name = w3 & 255u, mapq = (w3 >> 8) & 255u;
        const uint32_t n_cig = w4 & 0xffffu, flag = w4 >> 16;

        // ---- read-name hash ----------------------------------------------------------------------
        const uint8_t* name = r + 36;
        const uint32_t nl = l_name ? l_name - 1 : 0;
        uint64_t h1 = 0, h2 = 0;
        for (uint32_t i = lane; i < nl; i += 32) {
            uint32_t b = name[i];
            h1 += hterm1(i, b);
            h2 += hterm2(i, b);
        }
        h1 = warp_sum_u64(h1) + mix64(nl);
        h2 = warp_sum_u64(h2) + mix64(nl ^ 0xABCDull);

        // ---- CIGAR ---------------------------------------------------------------------------------
        const uint8_t* cig = name + l_name;
        int32_t refpos = pos;       // reference coordinate at the start of this batch of 32 ops
        int32_t carry_mend = -1;    // end of the last aligned block seen so far (-1: none)
        int32_t first_start = -1;
        uint32_t gcount = 0;
        uint32_t qh = 0, mb = 0;
        for (uint32_t base = 0; base < n_cig; base += 32) {
            const uint32_t k = base + lane;
            uint32_t v = (k < n_cig) ? ld_u32_any(cig + 4ull * k) : 0xFu;  // op 15 = nothing
            const uint32_t op = v & 15u, len = v >> 4;
            const bool is_m = op_is_m(op);
            const int32_t rlen = op_ref(op) ? (int32_t)len : 0;
            const int32_t incl = warp_incl_sum_i32(rlen);
            const int32_t start = refpos + incl - rlen;
            const int32_t end = start + rlen;
            const int32_t mend = is_m ? end : -1;
            int32_t pmax = warp_incl_max_i32(mend);
            int32_t prev = __shfl_up_sync(FULL_MASK, pmax, 1);
            if (lane == 0) prev = -1;
            prev = max(prev, carry_mend);
            const bool is_ev = is_m && prev >= 0 && (start - prev) > min_del_len;
            const uint32_t evb = __ballot_sync(FULL_MASK, is_ev);
            if (evb) {
                unsigned long long slot0 = 0;
                if (lane == 0) slot0 = atomicAdd(ev.counter, (unsigned long long)__popc(evb));
                slot0 = __shfl_sync(FULL_MASK, slot0, 0);
                if (is_ev) {
                    const uint32_t rank = __popc(evb & lt_mask);
                    const unsigned long long s = slot0 + rank;
                    if ((int64_t)s < ev.capacity) {
                        ev.rec[s] = (uint32_t)rec;
                        ev.k[s] = gcount + rank;
                        ev.next_start[s] = start;
                        ev.prev_end[s] = prev;
                    }
                }
                gcount += __popc(evb);
            }
            const uint32_t mb_ballot = __ballot_sync(FULL_MASK, is_m);
            if (first_start < 0 && mb_ballot) first_start = __shfl_sync(FULL_MASK, start, __ffs(mb_ballot) - 1);
            carry_mend = max(carry_mend, __shfl_sync(FULL_MASK, pmax, 31));
            refpos += __shfl_sync(FULL_MASK, incl, 31);
            if (op_qh(op)) qh += len;
            if (is_m) mb += len;
        }
        qh = (uint32_t)warp_sum_i64(qh);
        mb = (uint32_t)warp_sum_i64(mb);
        int32_t reflen = refpos - pos;
        if ((flag & 4u) || n_cig == 0) reflen = 0;
        const int32_t endpos = pos + (reflen > 0 ? reflen : 1);   // htslib bam_endpos

        // ---- SEQ: count nibbles that are not A/C/G/T ---------------------------------------------
        const uint8_t* seq = cig + 4ull * n_cig;
        const uint32_t sbytes = (l_seq + 1) >> 1;
        uint32_t bad = 0;
        {
            const uintptr_t a0 = reinterpret_cast<uintptr_t>(seq), a1 = a0 + sbytes;
            for (uintptr_t c = (a0 & ~uintptr_t(15)) + 16ull * lane; c < a1; c += 512) {
                uint4 q = __ldg(reinterpret_cast<const uint4*>(c));
                uint64_t x = ((uint64_t)q.y << 32) | q.x, y = ((uint64_t)q.w << 32) | q.z;
                if (c < a0 || c + 16 > a1) {  // edge chunk: neutralise bytes outside [a0, a1) with code 0x11
                    const int lo = c < a0 ? (int)(a0 - c) : 0;
                    const int hi = c + 16 > a1 ? (int)(a1 - c) : 16;
                    uint64_t mx = 0, my = 0;
#pragma unroll
                    for (int b = 0; b < 8; b++) {
                        if (b >= lo && b < hi) mx |= 0xFFull << (8 * b);
                        if (b + 8 >= lo && b + 8 < hi) my |= 0xFFull << (8 * b);
                    }
                    x = (x & mx) | (0x1111111111111111ull & ~mx);
                    y = (y & my) | (0x1111111111111111ull & ~my);
                }
                bad += non_acgt_nibbles(x) + non_acgt_nibbles(y);
            }
            bad = (uint32_t)warp_sum_i64(bad);
            if ((l_seq & 1u) && sbytes) {  // low nibble of the last byte is padding, not a base
                uint32_t lowb = seq[sbytes - 1] & 15u;
                if (!(lowb == 1 || lowb == 2 || lowb == 4 || lowb == 8)) bad -= 1;
            }
        }

        // ---- aux TLV walk (warp-uniform) -----------------------------------------------------------
        const uint8_t* p = seq + sbytes + l_seq;
        const uint8_t* end = r + 4 + block_size;
        uint32_t sa_off = 0, sa_len = 0;
        int32_t nm = 0;
        bool have_sa = false, have_nm = false, bad_aux = false;
        while (p + 3 <= end) {
            const uint32_t t0 = p[0], t1 = p[1], ty = p[2];
            p += 3;
            uint32_t sz = 0;
            if (ty == 'A' || ty == 'c' || ty == 'C') sz = 1;
            else if (ty == 's' || ty == 'S') sz = 2;
            else if (ty == 'i' || ty == 'I' || ty == 'f') sz = 4;
            if (sz) {
                if (t0 == 'N' && t1 == 'M' && !have_nm) {
                    have_nm = true;
                    uint32_t raw = 0;
                    for (uint32_t b = 0; b < sz; b++) raw |= (uint32_t)p[b] << (8 * b);
                    if (ty == 'c') nm = (int32_t)(int8_t)raw;
                    else if (ty == 's') nm = (int32_t)(int16_t)raw;
                    else if (ty == 'f' || ty == 'A') nm = 0;
                    else nm = (int32_t)raw;
                }
                p += sz;
            } else if (ty == 'Z' || ty == 'H') {
                uint32_t len = 0;
                bool found = false;
                while (p + len < end) {
                    const uint8_t* q = p + len + lane;
                    uint32_t b = (q < end) ? *q : 0u;
                    uint32_t zb = __ballot_sync(FULL_MASK, b == 0);
                    if (zb) { len += __ffs(zb) - 1; found = true; break; }
                    len += 32;
                }
                if (!found) { bad_aux = true; break; }
                if (t0 == 'S' && t1 == 'A' && !have_sa) {
                    have_sa = true;
                    sa_off = (uint32_t)(p - r);
                    sa_len = len;
                }
                p += len + 1;
            } else if (ty == 'B') {
                if (p + 5 > end) { bad_aux = true; break; }
                const uint32_t sub = p[0];
                uint32_t cnt = 0;
                for (uint32_t b = 0; b < 4; b++) cnt |= (uint32_t)p[1 + b] << (8 * b);
                uint32_t es = (sub == 'c' || sub == 'C') ? 1 : (sub == 's' || sub == 'S') ? 2 :
                              (sub == 'i' || sub == 'I' || sub == 'f') ? 4 : 0;
                if (!es) { bad_aux = true; break; }
                p += 5 + (size_t)cnt * es;
            } else {
                bad_aux = true;
                break;
            }
        }
        if (p > end) bad_aux = true;

        if (lane == 0) {
            if (bad_aux) atomicAdd(&status[0], 1);
            o.tid[rec] = tid;
            o.pos[rec] = pos;
            o.endpos[rec] = endpos;
            o.fbs[rec] = first_start;
            o.lbe[rec] = carry_mend;
            o.nm[rec] = nm;
            o.flag_mapq[rec] = flag | (mapq << 16);
            o.lseq[rec] = l_seq;
            o.ncig[rec] = n_cig;
            o.infer[rec] = qh;
            o.mbases[rec] = mb;
            o.nonacgt[rec] = bad;
            o.sa_off[rec] = have_sa ? sa_off : 0u;
            o.sa_len[rec] = sa_len;
            o.ngap[rec] = gcount;
            o.name_lo[rec] = h1;
            o.name_hi[rec] = h2;
        }
    }
}

}  // namespace coral

int coral_run_scan(coral_ctx* ctx) {
    using namespace coral;
    const int64_t n = ctx->nrec;
    RecordTable& t = ctx->rt;
    const size_t n4 = (size_t)(n > 0 ? n : 1) * 4, n8 = (size_t)(n > 0 ? n : 1) * 8;
    DevBuf* c4[] = {&t.tid, &t.pos, &t.endpos, &t.flag_mapq, &t.lseq, &t.ncig, &t.infer, &t.mbases, &t.nonacgt,
                    &t.fbs, &t.lbe, &t.sa_off, &t.sa_len, &t.nm, &t.ngap};
    for (DevBuf* b : c4) CORAL_CUDA(ctx, b->ensure(n4));
    CORAL_CUDA(ctx, t.name_lo.ensure(n8));
    CORAL_CUDA(ctx, t.name_hi.ensure(n8));
    CORAL_CUDA(ctx, ctx->status.ensure(64));
    CORAL_CUDA(ctx, ctx->gaps.counter.ensure(8));
    if (ctx->gaps.capacity == 0) ctx->gaps.capacity = std::max<int64_t>(1 << 16, n / 2);
    for (int attempt = 0; attempt < 2; attempt++) {
        GapEvents& g = ctx->gaps;
        CORAL_CUDA(ctx, g.rec.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, g.k.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, g.next_start.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, g.prev_end.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, cudaMemsetAsync(g.counter.p, 0, 8, ctx->stream));
        CORAL_CUDA(ctx, cudaMemsetAsync(ctx->status.p, 0, 64, ctx->stream));
        ScanCols o{t.tid.as<int32_t>(), t.pos.as<int32_t>(), t.endpos.as<int32_t>(), t.fbs.as<int32_t>(),
                   t.lbe.as<int32_t>(), t.nm.as<int32_t>(), t.flag_mapq.as<uint32_t>(), t.lseq.as<uint32_t>(),
                   t.ncig.as<uint32_t>(), t.infer.as<uint32_t>(), t.mbases.as<uint32_t>(), t.nonacgt.as<uint32_t>(),
                   t.sa_off.as<uint32_t>(), t.sa_len.as<uint32_t>(), t.ngap.as<uint32_t>(),
                   t.name_lo.as<uint64_t>(), t.name_hi.as<uint64_t>()};
        ScanEvents e{g.rec.as<uint32_t>(), g.k.as<uint32_t>(), g.next_start.as<int32_t>(), g.prev_end.as<int32_t>(),
                     g.counter.as<unsigned long long>(), g.capacity};
        if (n > 0) {
            int64_t want = (n + SCAN_WARPS - 1) / SCAN_WARPS;
            int blocks = (int)std::min<int64_t>(want, (int64_t)ctx->sm_count * 8);
            ctx->stage_begin(CORAL_STAGE_SCAN);
            k_scan_records<<<blocks, SCAN_WARPS * 32, 0, ctx->stream>>>(
                ctx->d_recs, ctx->d_off.as<uint64_t>(), n, ctx->prm.min_del_len, o, e, ctx->status.as<int32_t>());
            ctx->stage_end(CORAL_STAGE_SCAN);
            ctx->n_launches++;
            CORAL_CUDA(ctx, cudaGetLastError());
        }
        unsigned long long total = 0;
        int32_t st[2] = {0, 0};
        CORAL_CUDA(ctx, cudaMemcpyAsync(&total, g.counter.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CORAL_CUDA(ctx, cudaMemcpyAsync(st, ctx->status.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CORAL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (st[0] != 0) {
            ctx->err = "malformed aux field in " + std::to_string(st[0]) + " record(s)";
            return CORAL_E_FORMAT;
        }
        if ((int64_t)total <= g.capacity) {
            g.n = (int64_t)total;
            return CORAL_OK;
        }
        g.capacity = (int64_t)total + (int64_t)total / 8 + 1024;  // rescan with room for every event
    }
    ctx->err = "gap-event buffer overflow after rescan";
    return CORAL_E_NOMEM;
}
