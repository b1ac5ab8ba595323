// coral_b200/csrc/scan_records.cu — K0: the one pass over raw BAM record bytes.
//
// Replaces, for every record, the htslib/pysam decode work that the reference triggers from
// infer_breakpoint_graph.py: bam_endpos for fetch(region) (:751,:1031,:1043), get_blocks() (:756),
// infer_read_length() (:1031), the per-base loop of count_coverage (:131,:1033), query_name /
// flag / query_length / mapping_quality / get_tag('SA:Z:') / get_cigar_stats()[0][-1] (:141-154).
// Everything downstream works on the SoA table this kernel writes; QUAL bytes are never touched.
//
// B200 shape of the kernel
//   * persistent grid (a multiple of the SM count), warps pull record indices from a global atomic queue,
//     so 1 kb and 120 kb records balance without sorting;
//   * each warp owns a ring of SC_NST shared-memory stages and streams its record through them with 1-D
//     bulk async copies (cp.async.bulk -> UBLKCP, completion on an mbarrier per stage): the copy engine
//     keeps (SC_NST - 1) x SC_CH bytes per warp in flight while the lanes work on the stage that landed.
//     Only [core .. end of SEQ) and the aux block are fetched; QUAL is skipped;
//   * CIGAR: 4 ops per lane from one 16-byte LDS (the op array is not word aligned inside a record: one
//     funnel shift per op), class sums by masked adds, gap events (> min_del_len between aligned blocks)
//     only evaluated exactly on the rare 128-op windows whose D/N total could reach the threshold;
//   * SEQ: 32 nibbles per lane per LDS.128 with a 3-instruction "every nibble is one-hot" test per word; the
//     exact per-nibble count only runs for 16-byte chunks that contain a non-ACGT code.
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

constexpr int SC_WARPS = 8;                 // warps per block
constexpr int SC_NST = 3;                   // stages per warp
constexpr int SC_CH = 2048;                 // payload bytes per staged segment
constexpr int SC_SB = SC_CH + 16;           // stage bytes: 16 B of overlap so no CIGAR op straddles two segments

// segment flags
constexpr uint32_t SF_VALID = 1, SF_FIRST = 2, SF_LAST = 4, SF_AUX = 8, SF_AUXG = 16, SF_BAD = 32;
// CIGAR op classes (ops 0..8 = MIDNSHP=X): M-type {M,=,X}, D/N, query-only {I,S,H}; P consumes nothing.

// ---- mbarrier / bulk copy (sm_90+ PTX; SASS: SYNCS.*, UBLKCP) ---------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(bar),
        "r"(parity)
        : "memory");
}

__device__ __forceinline__ uint32_t lds32(const uint8_t* p) { return *reinterpret_cast<const uint32_t*>(p); }
__device__ __forceinline__ uint4 lds128(const uint8_t* p) { return *reinterpret_cast<const uint4*>(p); }

// Class byte of a raw CIGAR word by ONE PRMT: the op nibble selects a table byte; selector bit 3 means
// "replicate the sign bit of byte (op & 7)", so op 8 ('X') lands on M's entry (top bit set) and reads 0xFF.
// Encoding that makes 0xFF decode as M-type only: bit 0 SET = M-type, bit 1 CLEAR = D/N, bit 2 CLEAR = I/S/H.
// (Only result byte 0 is meaningful; the other selector nibbles are length bits.)
__device__ __forceinline__ uint32_t cigar_class(uint32_t v) {
    //             op:   3 N   2 D   1 I   0 M                   7 =   6 P   5 H   4 S
    const uint32_t lo = 0x04040287u, hi = 0x87060202u;
    return __byte_perm(lo, hi, v);
}
__device__ __forceinline__ bool cls_m(uint32_t c) { return (c & 1u) != 0; }
__device__ __forceinline__ bool cls_dn(uint32_t c) { return (c & 2u) == 0; }
__device__ __forceinline__ bool cls_ish(uint32_t c) { return (c & 4u) == 0; }
// nibbles of w that are one-hot (BAM codes A=1, C=2, G=4, T=8): bit 0 of each such nibble is set in the result
__device__ __forceinline__ uint32_t onehot_nibbles(uint32_t w) {
    const uint32_t w1 = w >> 1;
    const uint32_t p = w ^ w1, q = w & w1, o = w | w1;
    const uint32_t two = q | (q >> 2) | (o & (o >> 2));
    return (p ^ (p >> 2)) & ~two & 0x11111111u;
}
// non-zero iff some nibble of w is not one-hot
__device__ __forceinline__ uint32_t not_onehot_test(uint32_t w) {
    const uint32_t t = w - 0x11111111u;
    return (w & t) | (t & ~w & 0x88888888u);
}

// Aux TLV walk (warp-uniform; p/end may point into shared or global memory).  Finds the first SA:Z and NM.
__device__ __noinline__ bool aux_walk(const uint8_t* p, const uint8_t* end, uint32_t p_rel, uint32_t lane,
                                         uint32_t& sa_off, uint32_t& sa_len, int32_t& nm, bool& have_sa) {
    bool have_nm = false, bad_aux = false;
    const uint8_t* const p_org = p;       // p_rel = offset of p_org from the start of the record
    have_sa = false;
    sa_off = 0; sa_len = 0; nm = 0;
    while (p + 3 <= end) {
        const uint32_t t0 = p[0], t1 = p[1], ty = p[2];
        p += 3;
        uint32_t sz = 0;
        if (ty == 'A' || ty == 'c' || ty == 'C') sz = 1;
        else if (ty == 's' || ty == 'S') sz = 2;
        else if (ty == 'i' || ty == 'I' || ty == 'f') sz = 4;
        if (sz) {
            if (p + sz > end) { bad_aux = true; break; }
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
            if (!found || p + len >= end) { bad_aux = true; break; }
            if (t0 == 'S' && t1 == 'A' && !have_sa) {
                have_sa = true;
                sa_off = (uint32_t)(p - p_org) + p_rel;
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
            if (!es || (uint64_t)cnt * es > (uint64_t)(end - p)) { bad_aux = true; break; }
            p += 5 + (size_t)cnt * es;
        } else {
            bad_aux = true;
            break;
        }
    }
    if (p > end) bad_aux = true;
    return bad_aux;
}

// Per-warp control block in shared memory.  Lane 0 is the warp's producer: it owns the generator state,
// pulls record indices from the global queue, issues the bulk copies and publishes one descriptor per
// stage; all lanes consume.  Keeping this state out of registers is what lets 32 warps live on an SM.
struct SegDesc { uint32_t flags_len, gbeg, rec, pad; };                 // flags | len << 8
struct WarpCtl {
    SegDesc desc[SC_NST];
    // generator (lane 0 only)
    uint64_t g_base;
    uint32_t g_rec, g_cov, g_r1_end, g_a0, g_a1, g_state;                // state | auxseg << 8 | auxglob << 9
    // record constants (written when the first segment is decoded)
    uint32_t pos, l_seq, n_cig, c0, s1, rec_size, unmapped, rec_bad;
};
constexpr size_t SC_SMEM_TOTAL = (size_t)SC_WARPS * SC_NST * SC_SB + (size_t)SC_WARPS * SC_NST * 8 +
                                 (size_t)SC_WARPS * sizeof(WarpCtl);

enum { G_NEW = 0, G_R1 = 1, G_AUX = 2, G_DONE = 3 };

// Lane 0: produce the next segment of the warp's record stream into stage s.
// `ahead` = record index fetched from the queue one grab early (the atomic's latency is off the critical path).
__device__ __noinline__ void gen_next(WarpCtl* ctl, int s, uint32_t bar, uint32_t dst, const uint8_t* __restrict__ recs,
                                         const uint64_t* __restrict__ off, uint32_t n, unsigned int* queue,
                                         uint32_t& ahead) {
    uint32_t st = ctl->g_state;
    uint32_t flags = 0, beg = 0, end = 0;
    if ((st & 255u) == G_R1) {
        const uint32_t cov = ctl->g_cov, r1_end = ctl->g_r1_end;
        beg = cov - 16;
        end = min(cov + (uint32_t)SC_CH, r1_end);
        ctl->g_cov = end;
        flags = SF_VALID;
        if (end >= r1_end) {
            if (st & 256u) ctl->g_state = (st & ~255u) | G_AUX;
            else { flags |= SF_LAST | ((st & 512u) ? SF_AUXG : 0u); ctl->g_state = G_NEW; }
        }
    } else if ((st & 255u) == G_AUX) {
        beg = ctl->g_a0; end = ctl->g_a1;
        flags = SF_VALID | SF_LAST | SF_AUX;
        ctl->g_state = G_NEW;
    } else if ((st & 255u) == G_NEW) {
        const uint32_t idx = ahead;
        if (idx >= n) {
            ctl->g_state = G_DONE;
        } else {
            ahead = atomicAdd(queue, 1u);
            const uint64_t o0 = __ldg(off + idx), o1 = __ldg(off + idx + 1);
            const uint32_t len16 = (uint32_t)(o1 - o0);
            const uint4 ca = __ldg(reinterpret_cast<const uint4*>(recs + o0));
            const uint2 cb = __ldg(reinterpret_cast<const uint2*>(recs + o0 + 16));
            const uint32_t block_size = ca.x, w3 = ca.w, w4 = cb.x, l_seq = cb.y;
            const uint64_t seq_end = 36ull + (w3 & 255u) + 4ull * (w4 & 0xffffu) + ((uint64_t)l_seq + 1) / 2;
            const uint64_t aux_beg = seq_end + l_seq, rec_size = 4ull + block_size;
            const uint32_t e0 = min((uint32_t)SC_CH, len16);
            ctl->g_base = o0;
            ctl->g_rec = idx;
            flags = SF_VALID | SF_FIRST;
            end = e0;
            if (aux_beg > rec_size || rec_size > len16) {
                flags |= SF_BAD | SF_LAST;           // framing inconsistent with the core: reported, not decoded
            } else if (rec_size <= e0) {
                flags |= SF_LAST | (aux_beg < rec_size ? SF_AUX : 0u);
            } else {
                const uint32_t r1_end = min(((uint32_t)seq_end + 15u) & ~15u, len16);
                const uint32_t a0 = (uint32_t)aux_beg & ~15u, a1 = ((uint32_t)rec_size + 15u) & ~15u;
                const bool has_aux = aux_beg < rec_size;
                const bool auxseg = has_aux && (a1 - a0) <= (uint32_t)SC_SB;
                const bool auxglob = has_aux && !auxseg;
                ctl->g_r1_end = r1_end; ctl->g_cov = e0; ctl->g_a0 = a0; ctl->g_a1 = a1;
                uint32_t ns = G_NEW;
                if (e0 >= r1_end) {
                    if (auxseg) ns = G_AUX;
                    else flags |= SF_LAST | (auxglob ? SF_AUXG : 0u);
                } else {
                    ns = G_R1;
                }
                ctl->g_state = ns | (auxseg ? 256u : 0u) | (auxglob ? 512u : 0u);
            }
        }
    }
    SegDesc d;
    d.flags_len = flags | ((end - beg) << 8);
    d.gbeg = beg;
    d.rec = ctl->g_rec;
    d.pad = 0;
    *reinterpret_cast<uint4*>(&ctl->desc[s]) = *reinterpret_cast<const uint4*>(&d);
    if (flags & SF_VALID) {
        mbar_expect_tx(bar, end - beg);
        bulk_g2s(dst, recs + ctl->g_base + beg, end - beg, bar);
    }
}

__global__ void __launch_bounds__(SC_WARPS * 32, 4)
k_scan_records(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off, uint32_t n, int32_t min_del_len,
               ScanCols o, ScanEvents ev, int32_t* __restrict__ status, unsigned int* __restrict__ queue) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint8_t* const ring = smem + (size_t)warp * SC_NST * SC_SB;
    const uint32_t bar0 = smem_u32(smem + (size_t)SC_WARPS * SC_NST * SC_SB) + warp * SC_NST * 8;
    WarpCtl* const ctl = reinterpret_cast<WarpCtl*>(smem + (size_t)SC_WARPS * SC_NST * SC_SB + (size_t)SC_WARPS * SC_NST * 8) + warp;
    uint32_t ahead = 0;          // lane 0: next record index, fetched one grab early
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < SC_NST; s++) mbar_init(bar0 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        ctl->g_state = G_NEW;
        ctl->g_rec = 0;
        ctl->g_base = 0;
        ahead = atomicAdd(queue, 1u);
#pragma unroll 1
        for (int s = 0; s < SC_NST; s++)
            gen_next(ctl, s, bar0 + 8 * s, smem_u32(ring + (size_t)s * SC_SB), recs, off, n, queue, ahead);
    }
    __syncwarp();

    // ---- consumer state: the record being decoded -----------------------------------------------------------
    uint32_t cig_next = 0, seq_next = 0;
    uint32_t msum = 0, dnsum = 0, ishsum = 0, bad = 0;      // lane partial sums
    uint32_t carried = 0;                                    // D/N bases since the last aligned block
    bool seen_m = false;
    int32_t first_start = -1;
    uint32_t gcount = 0;
    uint32_t parity = 0;

    {
#pragma unroll 1
        for (int s = 0;; s = (s + 1 == SC_NST) ? 0 : s + 1) {
            const uint4 dq = *reinterpret_cast<const uint4*>(&ctl->desc[s]);
            const uint32_t fl = dq.x & 255u;
            if (!(fl & SF_VALID)) break;
            mbar_wait(bar0 + 8 * s, (parity >> s) & 1u);
            parity ^= 1u << s;
            const uint8_t* buf = ring + (size_t)s * SC_SB;
            const uint32_t gbeg = dq.y, seg_end = gbeg + (dq.x >> 8);
            const uint32_t rec = dq.z;

            if (fl & SF_FIRST) {
                const uint32_t cw = lane < 9 ? lds32(buf + 4 * lane) : 0u;
                const uint32_t block_size = __shfl_sync(FULL_MASK, cw, 0);
                const uint32_t tid = __shfl_sync(FULL_MASK, cw, 1);
                const uint32_t pos_ = __shfl_sync(FULL_MASK, cw, 2);
                const uint32_t w3 = __shfl_sync(FULL_MASK, cw, 3), w4 = __shfl_sync(FULL_MASK, cw, 4);
                uint32_t l_seq = __shfl_sync(FULL_MASK, cw, 5);
                const uint32_t l_name = w3 & 255u;
                uint32_t n_cig = w4 & 0xffffu;
                const uint32_t flag = w4 >> 16;
                const bool rec_bad = (fl & SF_BAD) != 0;
                const uint32_t c0 = 36 + l_name;
                if (rec_bad) { n_cig = 0; l_seq = 0; }
                const uint32_t s0 = c0 + 4 * n_cig;
                cig_next = 0; seq_next = s0;
                msum = dnsum = ishsum = bad = 0;
                carried = 0; seen_m = false; first_start = -1; gcount = 0;
                // read-name hash
                const uint32_t nl = (l_name && !rec_bad) ? l_name - 1 : 0;
                uint64_t h1 = 0, h2 = 0;
                for (uint32_t i = lane; i < nl; i += 32) {
                    const uint32_t b = buf[36 + i];
                    h1 += hterm1(i, b);
                    h2 += hterm2(i, b);
                }
                h1 = warp_sum_u64(h1) + mix64(nl);
                h2 = warp_sum_u64(h2) + mix64(nl ^ 0xABCDull);
                if (lane == 0) {
                    ctl->pos = pos_; ctl->l_seq = l_seq; ctl->n_cig = n_cig; ctl->c0 = c0;
                    ctl->s1 = s0 + ((l_seq + 1) >> 1); ctl->rec_size = 4 + block_size;
                    ctl->unmapped = (flag & 4u); ctl->rec_bad = rec_bad ? 1u : 0u;
                    o.tid[rec] = (int32_t)tid;
                    o.pos[rec] = (int32_t)pos_;
                    o.flag_mapq[rec] = flag | (((w3 >> 8) & 255u) << 16);
                    o.lseq[rec] = l_seq;
                    o.ncig[rec] = n_cig;
                    o.name_lo[rec] = h1;
                    o.name_hi[rec] = h2;
                    o.sa_off[rec] = 0; o.sa_len[rec] = 0; o.nm[rec] = 0;     // overwritten by the aux walk
                    if (rec_bad) atomicAdd(&status[1], 1);
                }
                __syncwarp();
            }
            const uint32_t n_cig = ctl->n_cig, c0 = ctl->c0;
            const int32_t pos = (int32_t)ctl->pos;

            // ---- CIGAR ops fully inside this segment -----------------------------------------------------
            if (cig_next < n_cig && seg_end >= c0 + 4) {
                const uint32_t i_hi = min(n_cig, (seg_end - c0) >> 2);
                if (i_hi > cig_next) {
                    const uint32_t sh = (c0 & 3u) * 8u;
                    const int32_t wc = ((int32_t)(c0 & ~3u) - (int32_t)gbeg) >> 2;   // word of op 0 in buf (may be < 0)
                    const int32_t w_first = wc + (int32_t)cig_next;
                    const int32_t w_end = wc + (int32_t)i_hi;                          // first word slot past the last op
                    for (int32_t w0 = w_first & ~3; w0 < w_end; w0 += 128) {
                        const int32_t mw = w0 + 4 * (int32_t)lane;
                        uint4 q = make_uint4(0, 0, 0, 0);
                        if (mw * 4 + 16 <= SC_SB) q = lds128(buf + 4 * mw);
                        uint32_t nx = __shfl_down_sync(FULL_MASK, q.x, 1);
                        if (lane == 31) nx = (mw * 4 + 20 <= SC_SB) ? lds32(buf + 4 * mw + 16) : 0u;
                        uint32_t v[4];
                        v[0] = __funnelshift_r(q.x, q.y, sh);
                        v[1] = __funnelshift_r(q.y, q.z, sh);
                        v[2] = __funnelshift_r(q.z, q.w, sh);
                        v[3] = __funnelshift_r(q.w, nx, sh);
                        if (w0 < w_first || w0 + 128 > w_end) {    // edge window: neutralise slots outside [cig_next, i_hi)
#pragma unroll
                            for (int j = 0; j < 4; j++)
                                if (mw + j < w_first || mw + j >= w_end) v[j] = 6u;   // 0P: consumes nothing
                        }
                        uint32_t cls[4], m_it = 0, dn_it = 0, dn_after = 0;
                        bool lane_m = false;
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            cls[j] = cigar_class(v[j]);
                            const uint32_t len = v[j] >> 4;
                            if (cls_m(cls[j])) { m_it += len; dn_after = 0; lane_m = true; }
                            if (cls_dn(cls[j])) { dn_it += len; dn_after += len; }
                            if (cls_ish(cls[j])) ishsum += len;
                        }
                        const uint32_t dn_tot = __reduce_add_sync(FULL_MASK, dn_it);
                        const uint32_t mball = __ballot_sync(FULL_MASK, lane_m);
                        // exact gap events: only when the D/N bases of this window (+ the open gap) could exceed the cut
                        if (mball && (int64_t)carried + dn_tot > (int64_t)min_del_len) {
                            const uint32_t run_r = __reduce_add_sync(FULL_MASK, msum + dnsum);
                            const int32_t base_ref = pos + (int32_t)run_r;
                            const int32_t lane_tot = (int32_t)(m_it + dn_it);
                            const int32_t incl = warp_incl_sum_i32(lane_tot);
                            int32_t cur = base_ref + incl - lane_tot;
                            int32_t st[4], last_m_end = -1;
#pragma unroll
                            for (int j = 0; j < 4; j++) {
                                st[j] = cur;
                                if (cls_m(cls[j]) || cls_dn(cls[j])) cur += (int32_t)(v[j] >> 4);
                                if (cls_m(cls[j])) last_m_end = cur;
                            }
                            const uint32_t before = mball & lt_mask;
                            const int src = before ? 31 - __clz(before) : 0;
                            int32_t prev = __shfl_sync(FULL_MASK, last_m_end, src);
                            bool have_prev = before != 0;
                            if (!have_prev) { prev = base_ref - (int32_t)carried; have_prev = seen_m; }
                            bool evj[4];
                            int32_t evprev[4];
                            uint32_t nbefore = 0, ntot = 0;
#pragma unroll
                            for (int j = 0; j < 4; j++) {
                                evj[j] = false; evprev[j] = 0;
                                if (cls_m(cls[j])) {
                                    evj[j] = have_prev && (st[j] - prev) > min_del_len;
                                    evprev[j] = prev;
                                    prev = st[j] + (int32_t)(v[j] >> 4);
                                    have_prev = true;
                                }
                                const uint32_t b = __ballot_sync(FULL_MASK, evj[j]);
                                nbefore += __popc(b & lt_mask);
                                ntot += __popc(b);
                            }
                            if (ntot) {
                                unsigned long long slot0 = 0;
                                if (lane == 0) slot0 = atomicAdd(ev.counter, (unsigned long long)ntot);
                                slot0 = __shfl_sync(FULL_MASK, slot0, 0);
                                uint32_t r = nbefore;
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    if (evj[j]) {
                                        const unsigned long long sl = slot0 + r;
                                        if ((int64_t)sl < ev.capacity) {
                                            ev.rec[sl] = rec;
                                            ev.k[sl] = gcount + r;
                                            ev.next_start[sl] = st[j];
                                            ev.prev_end[sl] = evprev[j];
                                        }
                                        r++;
                                    }
                                }
                                gcount += ntot;
                            }
                        }
                        if (mball) {
                            if (!seen_m) {      // first aligned block of the record: D/N bases in front of it
                                const int fl_ = __ffs(mball) - 1;
                                uint32_t lead = 0;
                                if ((int)lane < fl_) lead = dn_it;
                                else if ((int)lane == fl_) {
                                    bool hit = false;
#pragma unroll
                                    for (int j = 0; j < 4; j++) {
                                        if (cls_m(cls[j])) hit = true;
                                        if (!hit && cls_dn(cls[j])) lead += v[j] >> 4;
                                    }
                                }
                                lead = __reduce_add_sync(FULL_MASK, lead);
                                first_start = pos + (int32_t)(carried + lead);
                                seen_m = true;
                            }
                            const int ll = 31 - __clz(mball);
                            const uint32_t tr = ((int)lane > ll) ? dn_it : ((int)lane == ll ? dn_after : 0u);
                            carried = __reduce_add_sync(FULL_MASK, tr);
                        } else {
                            carried += dn_tot;
                        }
                        msum += m_it;
                        dnsum += dn_it;
                    }
                    cig_next = i_hi;
                }
            }

            // ---- SEQ bytes inside this segment ---------------------------------------------------------------
            {
                const uint32_t s1 = ctl->s1;
                const uint32_t b_hi = min(s1, seg_end);
                if (b_hi > seq_next) {
                    const uint32_t o_lo = seq_next - gbeg, o_hi = b_hi - gbeg;
                    const int32_t pad_o = ((ctl->l_seq & 1u) && b_hi == s1) ? (int32_t)(o_hi - 1) : -1;
                    for (uint32_t oo = (o_lo & ~15u) + 16u * lane; oo < o_hi; oo += 512u) {
                        const uint4 q = lds128(buf + oo);
                        uint32_t w[4] = {q.x, q.y, q.z, q.w};
                        if (oo < o_lo || oo + 16 > o_hi || (pad_o >> 4) == (int32_t)(oo >> 4)) {
#pragma unroll
                            for (int k = 0; k < 4; k++) {
                                const int32_t wb = (int32_t)oo + 4 * k;
                                const int32_t lo = max(0, min(4, (int32_t)o_lo - wb));
                                const int32_t hi = max(0, min(4, (int32_t)o_hi - wb));
                                const uint32_t m = hi > lo ? ((0xFFFFFFFFu >> (8 * (4 - hi))) & (0xFFFFFFFFu << (8 * lo))) : 0u;
                                w[k] = (w[k] & m) | (0x11111111u & ~m);
                                if (pad_o >= wb && pad_o < wb + 4) {     // odd l_seq: the last low nibble is padding
                                    const uint32_t shp = (uint32_t)(pad_o - wb) * 8u;
                                    w[k] = (w[k] & ~(0xFu << shp)) | (0x1u << shp);
                                }
                            }
                        }
                        const uint32_t acc = not_onehot_test(w[0]) | not_onehot_test(w[1]) | not_onehot_test(w[2]) |
                                             not_onehot_test(w[3]);
                        if (acc) {
                            bad += 32u - __popc(onehot_nibbles(w[0])) - __popc(onehot_nibbles(w[1])) -
                                   __popc(onehot_nibbles(w[2])) - __popc(onehot_nibbles(w[3]));
                        }
                    }
                    seq_next = b_hi;
                }
            }

            // ---- aux block -------------------------------------------------------------------------------------
            if (fl & (SF_AUX | SF_AUXG)) {
                const uint32_t rec_size = ctl->rec_size, aux_beg = ctl->s1 + ctl->l_seq;
                uint32_t sa_off, sa_len;
                int32_t nm;
                bool have_sa, bad_aux;
                if (fl & SF_AUX) {
                    bad_aux = aux_walk(buf + (aux_beg - gbeg), buf + (rec_size - gbeg), aux_beg, lane, sa_off, sa_len, nm, have_sa);
                } else {                              // aux longer than a stage: walk it in global memory
                    const uint8_t* r0 = recs + __ldg(off + rec);
                    bad_aux = aux_walk(r0 + aux_beg, r0 + rec_size, aux_beg, lane, sa_off, sa_len, nm, have_sa);
                }
                if (lane == 0) {
                    if (bad_aux) atomicAdd(&status[0], 1);
                    o.sa_off[rec] = have_sa ? sa_off : 0u;
                    o.sa_len[rec] = sa_len;
                    o.nm[rec] = nm;
                }
            }
            if (fl & SF_LAST) {
                const uint32_t m_tot = __reduce_add_sync(FULL_MASK, msum);
                const uint32_t dn_all = __reduce_add_sync(FULL_MASK, dnsum);
                const uint32_t ish_tot = __reduce_add_sync(FULL_MASK, ishsum);
                const uint32_t bad_tot = __reduce_add_sync(FULL_MASK, bad);
                int32_t reflen = (int32_t)(m_tot + dn_all);
                const int32_t lbe = seen_m ? pos + reflen - (int32_t)carried : -1;
                if (ctl->unmapped || n_cig == 0) reflen = 0;
                const int32_t endpos = pos + (reflen > 0 ? reflen : 1);   // htslib bam_endpos
                if (lane == 0) {
                    o.endpos[rec] = endpos;
                    o.fbs[rec] = first_start;
                    o.lbe[rec] = lbe;
                    o.infer[rec] = m_tot + ish_tot;
                    o.mbases[rec] = m_tot;
                    o.nonacgt[rec] = bad_tot;
                    o.ngap[rec] = gcount;
                }
            }
            __syncwarp();          // every lane is done with the stage before the copy engine overwrites it
            if (lane == 0)
                gen_next(ctl, s, bar0 + 8 * s, smem_u32(ring + (size_t)s * SC_SB), recs, off, n, queue, ahead);
            __syncwarp();
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
    CORAL_CUDA(ctx, ctx->gaps.counter.ensure(16));
    if (ctx->gaps.capacity == 0) ctx->gaps.capacity = std::max<int64_t>(1 << 16, n / 2);
    static bool attr_set = false;
    if (!attr_set) {
        CORAL_CUDA(ctx, cudaFuncSetAttribute(k_scan_records, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SC_SMEM_TOTAL));
        attr_set = true;
    }
    for (int attempt = 0; attempt < 2; attempt++) {
        GapEvents& g = ctx->gaps;
        CORAL_CUDA(ctx, g.rec.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, g.k.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, g.next_start.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, g.prev_end.ensure((size_t)g.capacity * 4));
        CORAL_CUDA(ctx, cudaMemsetAsync(g.counter.p, 0, 16, ctx->stream));     // [0] event count, [1] record queue
        CORAL_CUDA(ctx, cudaMemsetAsync(ctx->status.p, 0, 64, ctx->stream));
        ScanCols o{t.tid.as<int32_t>(), t.pos.as<int32_t>(), t.endpos.as<int32_t>(), t.fbs.as<int32_t>(),
                   t.lbe.as<int32_t>(), t.nm.as<int32_t>(), t.flag_mapq.as<uint32_t>(), t.lseq.as<uint32_t>(),
                   t.ncig.as<uint32_t>(), t.infer.as<uint32_t>(), t.mbases.as<uint32_t>(), t.nonacgt.as<uint32_t>(),
                   t.sa_off.as<uint32_t>(), t.sa_len.as<uint32_t>(), t.ngap.as<uint32_t>(),
                   t.name_lo.as<uint64_t>(), t.name_hi.as<uint64_t>()};
        ScanEvents e{g.rec.as<uint32_t>(), g.k.as<uint32_t>(), g.next_start.as<int32_t>(), g.prev_end.as<int32_t>(),
                     g.counter.as<unsigned long long>(), g.capacity};
        if (n > 0) {
            int per_sm = 0;
            CORAL_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_scan_records, SC_WARPS * 32, SC_SMEM_TOTAL));
            if (per_sm < 1) per_sm = 1;
            const int64_t want = (n + SC_WARPS - 1) / SC_WARPS;
            const int blocks = (int)std::min<int64_t>(want, (int64_t)ctx->sm_count * per_sm);   // persistent: one wave
            ctx->stage_begin(CORAL_STAGE_SCAN);
            k_scan_records<<<blocks, SC_WARPS * 32, SC_SMEM_TOTAL, ctx->stream>>>(
                ctx->d_recs, ctx->d_off.as<uint64_t>(), (uint32_t)n, ctx->prm.min_del_len, o, e, ctx->status.as<int32_t>(),
                reinterpret_cast<unsigned int*>(g.counter.as<unsigned long long>() + 1));
            ctx->stage_end(CORAL_STAGE_SCAN);
            ctx->n_launches++;
            CORAL_CUDA(ctx, cudaGetLastError());
        }
        unsigned long long total = 0;
        int32_t st[2] = {0, 0};
        CORAL_CUDA(ctx, cudaMemcpyAsync(&total, g.counter.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CORAL_CUDA(ctx, cudaMemcpyAsync(st, ctx->status.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CORAL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (st[1] != 0) {
            ctx->err = "record framing inconsistent with the record core in " + std::to_string(st[1]) + " record(s)";
            return CORAL_E_FORMAT;
        }
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
