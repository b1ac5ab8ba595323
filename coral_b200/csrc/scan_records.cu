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
#include "relops.cuh"

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
#ifndef SC_NST
#define SC_NST 2                            // stages per warp
#endif
#ifndef SC_CH
#define SC_CH 3072                          // payload bytes per staged segment
#endif
constexpr int SC_SB = SC_CH + 16;           // stage bytes: 16 B of overlap so no CIGAR op straddles two segments

// segment flags
constexpr uint32_t SF_VALID = 1, SF_FIRST = 2, SF_LAST = 4, SF_AUX = 8, SF_AUXG = 16, SF_BAD = 32;

// ---- shared-memory access by 32-bit shared-window address (no generic-pointer arithmetic in the hot loops) --
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds8(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// 32-bit little-endian load from an arbitrarily aligned shared address (stage buffers keep 16 B of slack)
__device__ __forceinline__ uint32_t lds32_any(uint32_t a) {
    const uint32_t lo = lds32(a & ~3u), hi = lds32((a & ~3u) + 4);
    return __funnelshift_r(lo, hi, (a & 3u) * 8u);
}

// ---- mbarrier / bulk copy (sm_90+ PTX; SASS: SYNCS.*, UBLKCP) ---------------------------------------------
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

// CIGAR op classes (ops 0..8 = MIDNSHP=X): M-type {M,=,X}, D/N, query-only {I,S,H}; P consumes nothing.
// One PRMT per class turns the op code into a clean 0/1 multiplier: selector nibble 0 = op & 7 (X = 8 lands on
// M's entry), selector nibbles 1..3 = 6 (P's entry, 0 in every table) so the upper result bytes are zero.  The
// masked adds then run as IMADs on the FMA pipe while the ALU pipe does the selects.
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}
//                                   op:   3 N 2 D 1 I 0 M           7 = 6 P 5 H 4 S
#define CIG_TM_LO 0x00000001u          //   0   0   0   1
#define CIG_TM_HI 0x01000000u          //                              1   0   0   0
#define CIG_TN_LO 0x01010100u          //   1   1   1   0        non-M-type (everything but M, =, X)
#define CIG_TN_HI 0x00010101u          //                              0   1   1   1
#ifndef SC_NM_TABLE
#define SC_NM_TABLE 1
#endif
#define CIG_TD_LO 0x01010000u          //   1   1   0   0
#define CIG_TD_HI 0x00000000u
#define CIG_TQ_LO 0x00000100u          //   0   0   1   0
#define CIG_TQ_HI 0x00000101u          //                              0   0   1   1

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

// Rare paths kept out of line so the compiler cannot if-convert them into the hot loops.
__device__ __noinline__ uint32_t seq_bad_count(uint4 q) {     // non-ACGT nibbles among the 32 of one 16-byte chunk
    return 32u - __popc(onehot_nibbles(q.x)) - __popc(onehot_nibbles(q.y)) - __popc(onehot_nibbles(q.z)) -
           __popc(onehot_nibbles(q.w));
}
__device__ __noinline__ uint4 cigar_mask_edge(uint4 v, int32_t mw, int32_t w_first, int32_t w_end) {
    const uint32_t span = (uint32_t)(w_end - w_first);       // slots outside [w_first, w_end) become 0P: consume nothing
    if ((uint32_t)(mw - w_first) >= span) v.x = 6u;
    if ((uint32_t)(mw + 1 - w_first) >= span) v.y = 6u;
    if ((uint32_t)(mw + 2 - w_first) >= span) v.z = 6u;
    if ((uint32_t)(mw + 3 - w_first) >= span) v.w = 6u;
    return v;
}
__device__ __forceinline__ uint32_t redux_add_volatile(uint32_t x) {   // not speculated above its guard
    uint32_t r;
    asm volatile("redux.sync.add.u32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(x));
    return r;
}

// value-size class of an aux type byte: 1/2/4 = fixed size, 0x80 = NUL-terminated (Z, H), 0x81 = array (B), 0 = invalid
__constant__ uint8_t c_aux_size[128] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,
    //  A     B     C                             H     I
    0, 1, 0x81, 1, 0, 0, 0, 0, 0x80, 4, 0, 0, 0, 0, 0, 0,
    //        S                                         Z
    0, 0, 0, 2, 0, 0, 0, 0, 0, 0, 0x80, 0, 0, 0, 0, 0,
    //           c           f        i
    0, 0, 0, 1, 0, 0, 4, 0, 0, 4, 0, 0, 0, 0, 0, 0,
    //        s
    0, 0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};

struct AuxOut { uint32_t sa_off, sa_len; int32_t nm; uint32_t have_sa, bad; };

// Aux TLV walk over a staged copy (warp-uniform).  p/end are shared-window addresses; p_rel = record offset of p.
__device__ __noinline__ AuxOut aux_walk_smem(uint32_t p, uint32_t end, uint32_t p_rel, uint32_t lane) {
    AuxOut r{0, 0, 0, 0, 0};
    const uint32_t p_org = p;
    bool have_nm = false;
    while (p + 3 <= end) {
        const uint32_t hd = lds32_any(p);                 // tag[2], type, first value byte
        const uint32_t tag = hd & 0xffffu, ty = (hd >> 16) & 255u;
        p += 3;
        const uint32_t sz = ty < 128 ? c_aux_size[ty] : 0u;
        if (sz == 0) { r.bad = 1; break; }
        if (sz < 0x80) {
            if (p + sz > end) { r.bad = 1; break; }
            if (tag == (('M' << 8) | 'N') && !have_nm) {
                have_nm = true;
                const uint32_t raw = lds32_any(p);
                if (ty == 'c') r.nm = (int32_t)(int8_t)raw;
                else if (ty == 'C') r.nm = (int32_t)(raw & 255u);
                else if (ty == 's') r.nm = (int32_t)(int16_t)raw;
                else if (ty == 'S') r.nm = (int32_t)(raw & 0xffffu);
                else if (ty == 'f' || ty == 'A') r.nm = 0;
                else r.nm = (int32_t)raw;
            }
            p += sz;
        } else if (sz == 0x80) {
            uint32_t len = 0;
            bool found = false;
            while (p + len < end) {
                const uint32_t q = p + len + lane;
                const uint32_t b = (q < end) ? lds8(q) : 1u;
                const uint32_t zb = __ballot_sync(FULL_MASK, b == 0);
                if (zb) { len += __ffs(zb) - 1; found = true; break; }
                len += 32;
            }
            if (!found) { r.bad = 1; break; }
            if (tag == (('A' << 8) | 'S') && !r.have_sa) {
                r.have_sa = 1;
                r.sa_off = (p - p_org) + p_rel;
                r.sa_len = len;
            }
            p += len + 1;
        } else {
            if (p + 5 > end) { r.bad = 1; break; }
            const uint32_t sub = lds8(p);
            const uint32_t cnt = lds32_any(p + 1);
            const uint32_t es = sub < 128 ? c_aux_size[sub] : 0u;
            if (es == 0 || es >= 0x80 || (uint64_t)cnt * es > (uint64_t)(end - p - 5)) { r.bad = 1; break; }
            p += 5 + cnt * es;
        }
    }
    if (p > end) r.bad = 1;
    return r;
}

// The same walk in global memory, for the rare record whose aux block is longer than a stage.
__device__ __noinline__ AuxOut aux_walk_global(const uint8_t* p, const uint8_t* end, uint32_t p_rel, uint32_t lane) {
    AuxOut r{0, 0, 0, 0, 0};
    const uint8_t* const p_org = p;
    bool have_nm = false;
    while (p + 3 <= end) {
        const uint32_t t0 = p[0], t1 = p[1], ty = p[2];
        p += 3;
        const uint32_t sz = ty < 128 ? c_aux_size[ty] : 0u;
        if (sz == 0) { r.bad = 1; break; }
        if (sz < 0x80) {
            if (p + sz > end) { r.bad = 1; break; }
            if (t0 == 'N' && t1 == 'M' && !have_nm) {
                have_nm = true;
                uint32_t raw = 0;
                for (uint32_t b = 0; b < sz; b++) raw |= (uint32_t)p[b] << (8 * b);
                if (ty == 'c') r.nm = (int32_t)(int8_t)raw;
                else if (ty == 's') r.nm = (int32_t)(int16_t)raw;
                else if (ty == 'f' || ty == 'A') r.nm = 0;
                else r.nm = (int32_t)raw;
            }
            p += sz;
        } else if (sz == 0x80) {
            uint32_t len = 0;
            bool found = false;
            while (p + len < end) {
                const uint8_t* q = p + len + lane;
                const uint32_t b = (q < end) ? *q : 1u;
                const uint32_t zb = __ballot_sync(FULL_MASK, b == 0);
                if (zb) { len += __ffs(zb) - 1; found = true; break; }
                len += 32;
            }
            if (!found) { r.bad = 1; break; }
            if (t0 == 'S' && t1 == 'A' && !r.have_sa) {
                r.have_sa = 1;
                r.sa_off = (uint32_t)(p - p_org) + p_rel;
                r.sa_len = len;
            }
            p += len + 1;
        } else {
            if (p + 5 > end) { r.bad = 1; break; }
            const uint32_t sub = p[0];
            uint32_t cnt = 0;
            for (uint32_t b = 0; b < 4; b++) cnt |= (uint32_t)p[1 + b] << (8 * b);
            const uint32_t es = sub < 128 ? c_aux_size[sub] : 0u;
            if (es == 0 || es >= 0x80 || (uint64_t)cnt * es > (uint64_t)(end - p - 5)) { r.bad = 1; break; }
            p += 5 + (size_t)cnt * es;
        }
    }
    if (p > end) r.bad = 1;
    return r;
}

// Per-warp control block in shared memory (addressed by shared-window offsets).  Lane 0 is the warp's producer:
// it owns the generator state, pulls record indices from the global queue, issues the bulk copies and publishes
// one descriptor per stage; all lanes consume.  Keeping this state out of registers is what lets 32 warps live
// on an SM.
constexpr uint32_t WC_DESC = 0;                          // SC_NST x {flags | len << 8, gbeg, rec, stage address}
constexpr uint32_t WC_BAR = WC_DESC + 16 * SC_NST;       // SC_NST mbarriers (8 B each)
constexpr uint32_t WC_GEN = WC_BAR + ((8 * SC_NST + 15) & ~15);   // {base lo, base hi, rec, cov} {r1_end, a0, a1, state}
constexpr uint32_t WC_REC = WC_GEN + 32;                 // {pos, l_seq, n_cig, c0} {s1, rec_size, unmapped, -}
constexpr uint32_t WC_BYTES = WC_REC + 32;
constexpr size_t SC_RING_BYTES = (size_t)SC_WARPS * SC_NST * SC_SB;
constexpr size_t SC_SMEM_TOTAL = SC_RING_BYTES + (size_t)SC_WARPS * WC_BYTES;

enum { G_NEW = 0, G_R1 = 1, G_AUX = 2, G_DONE = 3 };

// Lane 0: produce the next segment of the warp's record stream into stage s.
// `ahead` = record index fetched from the queue one grab early (the atomic's latency is off the critical path).
// `order` (nullable): the queue hands out positions of this permutation instead of record indices — records sorted by
// decreasing size, so that the longest records (a 120 kb ONT read takes a warp > 100 us) start first and the kernel's
// tail is made of the shortest ones.
__device__ __noinline__ uint32_t gen_next(uint32_t ctl, uint32_t s, const uint8_t* __restrict__ recs,
                                          const uint64_t* __restrict__ off, uint32_t n, unsigned int* queue, uint32_t ahead,
                                          const uint32_t* __restrict__ order) {
    const uint32_t bar = ctl + WC_BAR + 8 * s;
    const uint32_t dst = lds32(ctl + WC_DESC + 16 * s + 12);
    uint4 g0 = lds128(ctl + WC_GEN), g1 = lds128(ctl + WC_GEN + 16);     // base lo/hi, rec, cov | r1_end, a0, a1, state
    const uint32_t st = g1.w;
    uint32_t flags = 0, beg = 0, end = 0;
    if ((st & 255u) == G_R1) {
        beg = g0.w - 16;
        end = min(g0.w + (uint32_t)SC_CH, g1.x);
        g0.w = end;
        flags = SF_VALID;
        if (end >= g1.x) {
            if (st & 256u) g1.w = (st & ~255u) | G_AUX;
            else { flags |= SF_LAST | ((st & 512u) ? SF_AUXG : 0u); g1.w = G_NEW; }
        }
    } else if ((st & 255u) == G_AUX) {
        beg = g1.y; end = g1.z;
        flags = SF_VALID | SF_LAST | SF_AUX;
        g1.w = G_NEW;
    } else if ((st & 255u) == G_NEW) {
        const uint32_t qpos = ahead;
        if (qpos >= n) {
            g1.w = G_DONE;
        } else {
            ahead = atomicAdd(queue, 1u);
            const uint32_t idx = order ? __ldg(order + qpos) : qpos;
            const uint64_t o0 = __ldg(off + idx), o1 = __ldg(off + idx + 1);
            const uint32_t len16 = (uint32_t)(o1 - o0);
            const uint4 ca = __ldg(reinterpret_cast<const uint4*>(recs + o0));
            const uint2 cb = __ldg(reinterpret_cast<const uint2*>(recs + o0 + 16));
            const uint32_t block_size = ca.x, w3 = ca.w, w4 = cb.x, l_seq = cb.y;
            const uint64_t seq_end = 36ull + (w3 & 255u) + 4ull * (w4 & 0xffffu) + ((uint64_t)l_seq + 1) / 2;
            const uint64_t aux_beg = seq_end + l_seq, rec_size = 4ull + block_size;
            const uint32_t e0 = min((uint32_t)SC_CH, len16);
            g0.x = (uint32_t)o0; g0.y = (uint32_t)(o0 >> 32); g0.z = idx;
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
                g1.x = r1_end; g0.w = e0; g1.y = a0; g1.z = a1;
                uint32_t ns = G_NEW;
                if (e0 >= r1_end) {
                    if (auxseg) ns = G_AUX;
                    else flags |= SF_LAST | (auxglob ? SF_AUXG : 0u);
                } else {
                    ns = G_R1;
                }
                g1.w = ns | (auxseg ? 256u : 0u) | (auxglob ? 512u : 0u);
            }
        }
    }
    sts128(ctl + WC_GEN, g0);
    sts128(ctl + WC_GEN + 16, g1);
    sts128(ctl + WC_DESC + 16 * s, make_uint4(flags | ((end - beg) << 8), beg, g0.z, dst));
    if (flags & SF_VALID) {
        mbar_expect_tx(bar, end - beg);
        bulk_g2s(dst, recs + (((uint64_t)g0.y << 32) | g0.x) + beg, end - beg, bar);
    }
    return ahead;
}

#ifndef SC_MINB
#define SC_MINB 4                           // resident blocks per SM the register budget is set for
#endif
__global__ void __launch_bounds__(SC_WARPS * 32, SC_MINB)
k_scan_records(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off, uint32_t n, int32_t min_del_len,
               ScanCols o, ScanEvents ev, int32_t* __restrict__ status, unsigned int* __restrict__ queue, uint32_t write_aux,
               const uint32_t* __restrict__ order) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t smem0 = smem_u32(smem);
    const uint32_t ctl = smem0 + (uint32_t)SC_RING_BYTES + warp * WC_BYTES;
    uint32_t ahead = 0;          // lane 0: next record index, fetched one grab early
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < SC_NST; s++) {
            mbar_init(ctl + WC_BAR + 8 * s, 1);
            sts32(ctl + WC_DESC + 16 * s + 12, smem0 + (warp * SC_NST + s) * SC_SB);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        sts128(ctl + WC_GEN, make_uint4(0, 0, 0, 0));
        sts128(ctl + WC_GEN + 16, make_uint4(0, 0, 0, G_NEW));
        ahead = atomicAdd(queue, 1u);
#pragma unroll 1
        for (uint32_t s = 0; s < SC_NST; s++)
            ahead = gen_next(ctl, s, recs, off, n, queue, ahead, order);
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
    uint32_t sel_or;                                          // kept in a register so (v & 7) | sel_or is one LOP3
    asm volatile("mov.u32 %0, 0x6660;" : "=r"(sel_or));

#pragma unroll 1
    for (uint32_t s = 0;; s = (s + 1 == SC_NST) ? 0 : s + 1) {
        const uint4 dq = lds128(ctl + WC_DESC + 16 * s);
        const uint32_t fl = dq.x & 255u;
        if (!(fl & SF_VALID)) break;
        mbar_wait(ctl + WC_BAR + 8 * s, (parity >> s) & 1u);
        parity ^= 1u << s;
        const uint32_t buf = dq.w;
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
                const uint32_t b = lds8(buf + 36 + i);
                h1 += hterm1(i, b);
                h2 += hterm2(i, b);
            }
            h1 = warp_sum_u64(h1) + mix64(nl);
            h2 = warp_sum_u64(h2) + mix64(nl ^ 0xABCDull);
            if (lane == 0) {
                sts128(ctl + WC_REC, make_uint4(pos_, l_seq, n_cig, c0));
                sts128(ctl + WC_REC + 16, make_uint4(s0 + ((l_seq + 1) >> 1), 4 + block_size, flag & 4u, 0));
                o.tid[rec] = (int32_t)tid;
                o.pos[rec] = (int32_t)pos_;
                o.flag_mapq[rec] = flag | (((w3 >> 8) & 255u) << 16);
                o.lseq[rec] = l_seq;
                o.ncig[rec] = n_cig;
                o.name_lo[rec] = h1;
                o.name_hi[rec] = h2;
                // overwritten by the aux walk; in the asynchronous mode the pre-scan owns these columns and fetch may
                // already be reading them, so they are left alone
                if (write_aux) { o.sa_off[rec] = 0; o.sa_len[rec] = 0; o.nm[rec] = 0; }
                if (rec_bad) atomicAdd(&status[1], 1);
            }
            __syncwarp();
        }
        const uint4 rc0 = lds128(ctl + WC_REC);           // pos, l_seq, n_cig, c0
        const uint4 rc1 = lds128(ctl + WC_REC + 16);      // s1, rec_size, unmapped
        const int32_t pos = (int32_t)rc0.x;
        const uint32_t n_cig = rc0.z, c0 = rc0.w;

        // ---- CIGAR ops fully inside this segment ---------------------------------------------------------
        if (cig_next < n_cig && seg_end >= c0 + 4) {
            const uint32_t i_hi = min(n_cig, (seg_end - c0) >> 2);
            if (i_hi > cig_next) {
                const uint32_t sh = (c0 & 3u) * 8u;
                const int32_t wc = ((int32_t)(c0 & ~3u) - (int32_t)gbeg) >> 2;   // word of op 0 in buf (may be < 0)
                const int32_t w_first = wc + (int32_t)cig_next;
                const int32_t w_end = wc + (int32_t)i_hi;                          // first word slot past the last op
#pragma unroll 1
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
                        const uint4 mv = cigar_mask_edge(make_uint4(v[0], v[1], v[2], v[3]), mw, w_first, w_end);
                        v[0] = mv.x; v[1] = mv.y; v[2] = mv.z; v[3] = mv.w;
                    }
                    uint32_t bm[4], bd[4], len[4], m_it = 0, dn_it = 0, dn_after = 0;
#if SC_NM_TABLE
                    // variant: the table of the NON-M-type ops {I,D,N,S,H,P} instead of the M-type one.  Like the D/N and
                    // {I,S,H} tables it has 0 for op 0, so the selector is op & 7 alone (its zero upper nibbles pick
                    // entry 0 = clean upper bytes); M-type bases = all bases - non-M bases; the open-gap recurrence
                    // multiplies by the non-M flag directly.
                    uint32_t nmv[4], tot = 0, nm_it = 0;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t sel = v[j] & 7u;
                        nmv[j] = prmt(CIG_TN_LO, CIG_TN_HI, sel);
                        bd[j] = prmt(CIG_TD_LO, CIG_TD_HI, sel);
                        const uint32_t bq = prmt(CIG_TQ_LO, CIG_TQ_HI, sel);
                        len[j] = v[j] >> 4;
                        tot += len[j];
                        nm_it = nmv[j] * len[j] + nm_it;
                        dn_it = bd[j] * len[j] + dn_it;
                        ishsum = bq * len[j] + ishsum;
                        dn_after = bd[j] * len[j] + dn_after * nmv[j];
                    }
                    m_it = tot - nm_it;
#pragma unroll
                    for (int j = 0; j < 4; j++) bm[j] = nmv[j] ^ 1u;          // only the rare paths below read bm[]
                    const bool lane_m = (nmv[0] & nmv[1] & nmv[2] & nmv[3]) == 0;
#else
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t sel = (v[j] & 7u) | sel_or;
                        bm[j] = prmt(CIG_TM_LO, CIG_TM_HI, sel);
                        bd[j] = prmt(CIG_TD_LO, CIG_TD_HI, sel);
                        const uint32_t bq = prmt(CIG_TQ_LO, CIG_TQ_HI, sel);
                        len[j] = v[j] >> 4;
                        m_it = bm[j] * len[j] + m_it;
                        dn_it = bd[j] * len[j] + dn_it;
                        ishsum = bq * len[j] + ishsum;
                        dn_after = bd[j] * len[j] + dn_after * (bm[j] ^ 1u);
                    }
                    const bool lane_m = (bm[0] | bm[1] | bm[2] | bm[3]) != 0;
#endif
                    const uint32_t dn_tot = __reduce_add_sync(FULL_MASK, dn_it);
                    const uint32_t mball = __ballot_sync(FULL_MASK, lane_m);
                    // exact gap events: only when the D/N bases of this window (+ the open gap) could exceed the cut
                    if (mball && (min_del_len < 0 || carried + dn_tot > (uint32_t)min_del_len)) {
                        const uint32_t run_r = redux_add_volatile(msum + dnsum);
                        const int32_t base_ref = pos + (int32_t)run_r;
                        const int32_t lane_tot = (int32_t)(m_it + dn_it);
                        const int32_t incl = warp_incl_sum_i32(lane_tot);
                        int32_t cur = base_ref + incl - lane_tot;
                        int32_t st[4], last_m_end = -1;
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            st[j] = cur;
                            cur += (int32_t)((bm[j] | bd[j]) * len[j]);
                            if (bm[j]) last_m_end = cur;
                        }
                        const uint32_t before = mball & lt_mask;
                        const int src = before ? 31 - __clz(before) : 0;
                        int32_t prev = __shfl_sync(FULL_MASK, last_m_end, src);
                        bool have_prev = before != 0;
                        if (!have_prev) { prev = base_ref - (int32_t)carried; have_prev = seen_m; }
                        uint32_t evmask = 0, nbefore = 0, ntot = 0;
                        int32_t evprev[4];
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            evprev[j] = prev;
                            bool e = false;
                            if (bm[j]) {
                                e = have_prev && (st[j] - prev) > min_del_len;
                                prev = st[j] + (int32_t)len[j];
                                have_prev = true;
                            }
                            if (e) evmask |= 1u << j;
                            const uint32_t b = __ballot_sync(FULL_MASK, e);
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
                                if (evmask & (1u << j)) {
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
                                uint32_t open = 1;
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    open &= bm[j] ^ 1u;
                                    lead += open * bd[j] * len[j];
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

        // ---- SEQ bytes inside this segment -------------------------------------------------------------------
        {
            const uint32_t s1 = rc1.x;
            const uint32_t b_hi = min(s1, seg_end);
            if (b_hi > seq_next) {
                const uint32_t o_lo = seq_next - gbeg, o_hi = b_hi - gbeg;
                // odd l_seq: the low nibble of the last SEQ byte is padding, kept out of the count
                const uint32_t o_body_hi = ((rc0.y & 1u) && b_hi == s1) ? o_hi - 1 : o_hi;
                const uint32_t a_lo = (o_lo + 15u) & ~15u, a_hi = o_body_hi & ~15u;     // whole 16-byte chunks
                if (a_hi > a_lo) {
#pragma unroll 1
                    for (uint32_t oo = a_lo + 16u * lane; oo < a_hi; oo += 512u) {
                        const uint4 q = lds128(buf + oo);
                        const uint32_t acc = not_onehot_test(q.x) | not_onehot_test(q.y) | not_onehot_test(q.z) |
                                             not_onehot_test(q.w);
                        if (acc) bad += seq_bad_count(q);
                    }
                }
                // ragged head / tail (< 16 bytes each, or the whole range when it holds no full chunk): one byte per lane
                {
                    const uint32_t h_hi = a_hi > a_lo ? a_lo : o_body_hi;     // head = [o_lo, h_hi)
                    uint32_t x = o_lo + lane;
                    bool in = x < h_hi;
                    if (!in && a_hi > a_lo) { x = a_hi + (lane - 16u); in = lane >= 16 && x < o_body_hi; }   // tail = [a_hi, o_body_hi)
                    if (in) {
                        const uint32_t b = lds8(buf + x);
                        const uint32_t hi4 = b >> 4, lo4 = b & 15u;
                        bad += (__popc(hi4) != 1) + (__popc(lo4) != 1);
                    }
                    if (o_body_hi != o_hi && lane == 0) {              // the last byte of an odd-length SEQ: high nibble only
                        const uint32_t b = lds8(buf + o_hi - 1);
                        bad += (__popc(b >> 4) != 1);
                    }
                }
                seq_next = b_hi;
            }
        }

        // ---- aux block -------------------------------------------------------------------------------------------
        if ((fl & (SF_AUX | SF_AUXG)) && write_aux) {
            const uint32_t rec_size = rc1.y, aux_beg = rc1.x + rc0.y;
            AuxOut a;
            if (fl & SF_AUX) {
                a = aux_walk_smem(buf + (aux_beg - gbeg), buf + (rec_size - gbeg), aux_beg, lane);
            } else {                              // aux longer than a stage: walk it in global memory
                const uint8_t* r0 = recs + __ldg(off + rec);
                a = aux_walk_global(r0 + aux_beg, r0 + rec_size, aux_beg, lane);
            }
            if (lane == 0) {
                if (a.bad) atomicAdd(&status[0], 1);
                o.sa_off[rec] = a.have_sa ? a.sa_off : 0u;
                o.sa_len[rec] = a.sa_len;
                o.nm[rec] = a.nm;
            }
        }
        if (fl & SF_LAST) {
            const uint32_t m_tot = __reduce_add_sync(FULL_MASK, msum);
            const uint32_t dn_all = __reduce_add_sync(FULL_MASK, dnsum);
            const uint32_t ish_tot = __reduce_add_sync(FULL_MASK, ishsum);
            const uint32_t bad_tot = __reduce_add_sync(FULL_MASK, bad);
            int32_t reflen = (int32_t)(m_tot + dn_all);
            const int32_t lbe = seen_m ? pos + reflen - (int32_t)carried : -1;
            if (rc1.z || n_cig == 0) reflen = 0;
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
        if (lane == 0) ahead = gen_next(ctl, s, recs, off, n, queue, ahead, order);
        __syncwarp();
    }
}

// Pre-scan of the asynchronous mode: the columns `fetch` needs (core fields, name hash, SA / NM location) for every
// record, from the record core, the read name and the aux block only — a few hundred bytes per record — so that
// fetch + the host steps that follow it can run while the full scan is still pulling CIGAR and SEQ bytes.  Same values
// as k_scan_records writes into the same columns.  The records may sit behind PCIe, where every dependent access costs
// microseconds: a warp makes two round trips per record — the first 128 bytes (core + name) as one coalesced load, then
// the aux block as 16-byte loads into a shared-memory stage that the table-driven walk reads — and the SA text is
// copied into a side buffer in HBM (`side`), so that the kernels of fetch never touch the link.
constexpr uint32_t PS_STAGE = 2048;                  // aux bytes staged per warp; longer aux blocks are walked in place

struct SideText {
    uint8_t* text;                // SA strings back to back (16-byte slots), first slot at 64
    unsigned long long* used;     // bump pointer
    uint64_t cap;
    uint64_t* rec_off;            // per record: slot position - sa_off (mod 2^64), so that text + rec_off + sa_off = slot
};

__global__ void __launch_bounds__(256)
k_prescan_records(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off, uint32_t n, ScanCols o, SideText side,
                  int32_t* __restrict__ status) {
    __shared__ __align__(16) uint8_t stage_mem[8][PS_STAGE + 32];
    const uint32_t lane = lane_id();
    const uint32_t stage = smem_u32(stage_mem[threadIdx.x >> 5]);
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t rec = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; rec < n; rec += nw) {
        const uint64_t o0 = __ldg(off + rec), o1 = __ldg(off + rec + 1);
        const uint8_t* r = recs + o0;
        const uint32_t len16 = (uint32_t)(o1 - o0);
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(r);
        // round trip 1: the first 128 bytes (a frame holds at least the 36-byte core)
        const uint32_t w0 = (4 * lane + 4 <= len16) ? __ldg(rw + lane) : 0u;
        const uint32_t block_size = __shfl_sync(FULL_MASK, w0, 0);
        const uint32_t tid = __shfl_sync(FULL_MASK, w0, 1);
        const uint32_t pos_ = __shfl_sync(FULL_MASK, w0, 2);
        const uint32_t w3 = __shfl_sync(FULL_MASK, w0, 3), w4 = __shfl_sync(FULL_MASK, w0, 4);
        uint32_t l_seq = __shfl_sync(FULL_MASK, w0, 5);
        const uint32_t l_name = w3 & 255u;
        uint32_t n_cig = w4 & 0xffffu;
        const uint32_t flag = w4 >> 16;
        const uint64_t seq_end = 36ull + l_name + 4ull * n_cig + ((uint64_t)l_seq + 1) / 2;
        const uint64_t aux_beg = seq_end + l_seq, rec_size = 4ull + block_size;
        const bool rec_bad = aux_beg > rec_size || rec_size > len16;     // the framing test of gen_next
        if (rec_bad) { n_cig = 0; l_seq = 0; }
        const uint32_t nl = (l_name && !rec_bad) ? l_name - 1 : 0;
        uint64_t h1 = 0, h2 = 0;
        for (uint32_t wi = lane; 4 * wi < 36 + nl; wi += 32) {            // name bytes sit at record offsets 36 .. 36 + nl
            const uint32_t w = wi < 32 ? w0 : __ldg(rw + wi);              // (names longer than 92 characters: one more trip)
#pragma unroll
            for (uint32_t b = 0; b < 4; b++) {
                const uint32_t bo = 4 * wi + b;
                if (bo >= 36 && bo < 36 + nl) {
                    const uint32_t ch = (w >> (8 * b)) & 255u;
                    h1 += hterm1(bo - 36, ch);
                    h2 += hterm2(bo - 36, ch);
                }
            }
        }
        h1 = warp_sum_u64(h1) + mix64(nl);
        h2 = warp_sum_u64(h2) + mix64(nl ^ 0xABCDull);
        AuxOut a{0, 0, 0, 0, 0};
        bool staged = false;
        uint32_t a0 = 0;
        if (!rec_bad && aux_beg < rec_size) {
            a0 = (uint32_t)aux_beg & ~15u;
            const uint32_t a1 = ((uint32_t)rec_size + 15u) & ~15u;       // <= len16: frames are 16-byte padded
            if (a1 - a0 <= PS_STAGE) {
                // round trip 2: the aux block, 16 bytes per lane per step
                for (uint32_t c = 16 * lane; c < a1 - a0; c += 512)
                    sts128(stage + c, __ldg(reinterpret_cast<const uint4*>(r + a0 + c)));
                __syncwarp();
                a = aux_walk_smem(stage + ((uint32_t)aux_beg - a0), stage + ((uint32_t)rec_size - a0), (uint32_t)aux_beg, lane);
                staged = true;
            } else {
                a = aux_walk_global(r + aux_beg, r + rec_size, (uint32_t)aux_beg, lane);
            }
        }
        if (a.have_sa && side.text) {          // SA text -> side buffer (16-byte slots)
            unsigned long long slot = 0;
            const uint32_t need = (a.sa_len + 16u) & ~15u;
            if (lane == 0) slot = atomicAdd(side.used, (unsigned long long)need);
            slot = __shfl_sync(FULL_MASK, slot, 0);
            if (slot + need + 64 <= side.cap) {
                for (uint32_t b = lane; b < a.sa_len; b += 32)
                    side.text[slot + b] = staged ? (uint8_t)lds8(stage + (a.sa_off - a0) + b) : r[a.sa_off + b];
                if (lane == 0) side.rec_off[rec] = (uint64_t)slot - (uint64_t)a.sa_off;
            } else if (lane == 0) {
                atomicExch(&status[2], 1);     // side buffer too small: fetch reads the records instead
            }
        }
        __syncwarp();                          // the stage is reused by the next record
        if (lane == 0) {
            o.tid[rec] = (int32_t)tid;
            o.pos[rec] = (int32_t)pos_;
            o.flag_mapq[rec] = flag | (((w3 >> 8) & 255u) << 16);
            o.lseq[rec] = l_seq;
            o.ncig[rec] = n_cig;
            o.name_lo[rec] = h1;
            o.name_hi[rec] = h2;
            o.sa_off[rec] = a.have_sa ? a.sa_off : 0u;
            o.sa_len[rec] = a.sa_len;
            o.nm[rec] = a.nm;
            if (rec_bad) atomicAdd(&status[1], 1);
            if (a.bad) atomicAdd(&status[0], 1);
        }
    }
}

}  // namespace coral

namespace {
using namespace coral;

// size class of every record, larger first: key = 0xFFFF - min(size >> 9, 0xFFFF)
__global__ void k_order_keys(int64_t n, const uint64_t* __restrict__ off, uint64_t big, uint64_t* __restrict__ key, uint32_t* __restrict__ idx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t sz = off[i + 1] - off[i];
    key[i] = big ? (sz >= big ? 0u : 1u) : 0xFFFFu - min(sz >> 9, (uint64_t)0xFFFFu);
    idx[i] = (uint32_t)i;
}

// The permutation the record queue of k_scan_records follows (option "scan_order": 2 = records by decreasing size class,
// stable, the default; 1 = records of more than twice the mean size first, the rest in file order; 0 = file order).
// Built once per load on the ctx's stream (the offsets do not change between rescans).  nullptr = file order.
const uint32_t* scan_order(coral_ctx* ctx) {
    using namespace coral;
    const int64_t n = ctx->nrec;
    if (ctx->scan_order_mode == 0 || n < 2) return nullptr;
    if (ctx->scan_order_valid) return ctx->scan_order.as<uint32_t>() + ctx->scan_order_at;
    if (ctx->scan_order.ensure((size_t)n * 8) != cudaSuccess || ctx->s4.ensure((size_t)n * 16) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    uint64_t* key_a = ctx->s4.as<uint64_t>(); uint64_t* key_b = key_a + n;
    uint32_t* idx = ctx->scan_order.as<uint32_t>(); uint32_t* out = idx + n;
    const uint64_t big = ctx->scan_order_mode == 1 ? 2 * (ctx->nbytes / (uint64_t)n) : 0;
    k_order_keys<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, ctx->d_off.as<uint64_t>(), big, key_a, idx);
    ctx->n_launches++;
    coral_ctx* c = ctx;
    if (sort_pairs_u64(c, key_a, key_b, idx, out, n, 0, big ? 1 : 16) != CORAL_OK) return nullptr;
    ctx->scan_order_at = (size_t)n;
    ctx->scan_order_valid = true;
    return out;
}

int scan_ensure_tables(coral_ctx* ctx) {
    const int64_t n = ctx->nrec;
    RecordTable& t = ctx->rt;
    const size_t n4 = (size_t)(n > 0 ? n : 1) * 4, n8 = (size_t)(n > 0 ? n : 1) * 8;
    DevBuf* c4[] = {&t.tid, &t.pos, &t.endpos, &t.flag_mapq, &t.lseq, &t.ncig, &t.infer, &t.mbases, &t.nonacgt,
                    &t.fbs, &t.lbe, &t.sa_off, &t.sa_len, &t.nm, &t.ngap};
    for (DevBuf* b : c4) CORAL_CUDA(ctx, b->ensure(n4));
    CORAL_CUDA(ctx, t.name_lo.ensure(n8));
    CORAL_CUDA(ctx, t.name_hi.ensure(n8));
    CORAL_CUDA(ctx, ctx->scan_status.ensure(64));
    CORAL_CUDA(ctx, ctx->status.ensure(64));       // the status words of the other entry points
    CORAL_CUDA(ctx, ctx->gaps.counter.ensure(16));
    if (ctx->gaps.capacity == 0) ctx->gaps.capacity = std::max<int64_t>(1 << 16, n / 2);
    GapEvents& g = ctx->gaps;
    CORAL_CUDA(ctx, g.rec.ensure((size_t)g.capacity * 4));
    CORAL_CUDA(ctx, g.k.ensure((size_t)g.capacity * 4));
    CORAL_CUDA(ctx, g.next_start.ensure((size_t)g.capacity * 4));
    CORAL_CUDA(ctx, g.prev_end.ensure((size_t)g.capacity * 4));
    static bool attr_set = false;
    if (!attr_set) {
        CORAL_CUDA(ctx, cudaFuncSetAttribute(k_scan_records, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SC_SMEM_TOTAL));
        attr_set = true;
    }
    return CORAL_OK;
}

ScanCols scan_cols(coral_ctx* ctx) {
    RecordTable& t = ctx->rt;
    return ScanCols{t.tid.as<int32_t>(), t.pos.as<int32_t>(), t.endpos.as<int32_t>(), t.fbs.as<int32_t>(),
                    t.lbe.as<int32_t>(), t.nm.as<int32_t>(), t.flag_mapq.as<uint32_t>(), t.lseq.as<uint32_t>(),
                    t.ncig.as<uint32_t>(), t.infer.as<uint32_t>(), t.mbases.as<uint32_t>(), t.nonacgt.as<uint32_t>(),
                    t.sa_off.as<uint32_t>(), t.sa_len.as<uint32_t>(), t.ngap.as<uint32_t>(),
                    t.name_lo.as<uint64_t>(), t.name_hi.as<uint64_t>()};
}

// counters + status of a finished scan -> CORAL_OK, an error, or 1 = gap-event buffer too small (capacity was raised)
int scan_verdict(coral_ctx* ctx, cudaStream_t st) {
    GapEvents& g = ctx->gaps;
    unsigned long long total = 0;
    int32_t sw[2] = {0, 0};
    if (st == ctx->stream) {          // a rescan inside a pipelined pass replays the event count and checks the status
        unsigned long long sw64 = 0;
        CORAL_TRY(pass_resolve(ctx, &total, g.counter.p));
        CORAL_TRY(pass_resolve(ctx, &sw64, ctx->scan_status.p));
        memcpy(sw, &sw64, 8);
    } else {
        CORAL_CUDA(ctx, cudaMemcpyAsync(&total, g.counter.p, 8, cudaMemcpyDeviceToHost, st));
        CORAL_CUDA(ctx, cudaMemcpyAsync(sw, ctx->scan_status.p, 8, cudaMemcpyDeviceToHost, st));
        CORAL_CUDA(ctx, cudaStreamSynchronize(st));
    }
    if (sw[1] != 0) {
        ctx->err = "record framing inconsistent with the record core in " + std::to_string(sw[1]) + " record(s)";
        return CORAL_E_FORMAT;
    }
    if (sw[0] != 0) {
        ctx->err = "malformed aux field in " + std::to_string(sw[0]) + " record(s)";
        return CORAL_E_FORMAT;
    }
    if ((int64_t)total <= g.capacity) {
        g.n = (int64_t)total;
        return CORAL_OK;
    }
    g.capacity = (int64_t)total + (int64_t)total / 8 + 1024;  // rescan with room for every event
    return 1;
}

}  // namespace

// Asynchronous form of the load-time scan (CORAL_RECORDS_HOST_MAPPED_ASYNC): the pre-scan runs to completion, the
// full scan is left running on its own stream with one block per SM (the link, not the SMs, bounds it, and the
// kernels of fetch need room next to it); coral_wait_scan() joins it.
int coral_start_scan_async(coral_ctx* ctx, bool launch_full_scan) {
    using namespace coral;
    const int64_t n = ctx->nrec;
    CORAL_TRY(scan_ensure_tables(ctx));
    if (!ctx->scan_stream) CORAL_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->scan_stream, cudaStreamNonBlocking));
    const uint32_t* ord = scan_order(ctx);       // (built on the ctx's stream, ahead of everything below)
    GapEvents& g = ctx->gaps;
    CORAL_CUDA(ctx, cudaMemsetAsync(g.counter.p, 0, 16, ctx->stream));
    CORAL_CUDA(ctx, cudaMemsetAsync(ctx->scan_status.p, 0, 64, ctx->stream));
    if (n == 0) { g.n = 0; return CORAL_OK; }
    ScanCols o = scan_cols(ctx);
    // side buffer for the SA strings (fetch then never reads the records): 1/32 of the record bytes, at least 1 MB;
    // a file with more SA text than that falls back to reading the records
    const uint64_t side_cap = std::max<uint64_t>((uint64_t)1 << 20, ctx->nbytes / 32);
    CORAL_CUDA(ctx, ctx->sa_side.ensure(side_cap));
    CORAL_CUDA(ctx, ctx->sa_side_off.ensure((size_t)n * 8 + 16));
    unsigned long long* d_used = reinterpret_cast<unsigned long long*>(ctx->sa_side_off.as<uint64_t>() + n);
    const unsigned long long first_slot = 64;
    CORAL_CUDA(ctx, cudaMemcpyAsync(d_used, &first_slot, 8, cudaMemcpyHostToDevice, ctx->stream));
    SideText side{ctx->sa_side.as<uint8_t>(), d_used, side_cap, ctx->sa_side_off.as<uint64_t>()};
    const int pre_blocks = (int)std::min<int64_t>((n + 7) / 8, (int64_t)ctx->sm_count * 8);
    k_prescan_records<<<pre_blocks, 256, 0, ctx->stream>>>(ctx->d_recs, ctx->d_off.as<uint64_t>(), (uint32_t)n, o, side,
                                                          ctx->scan_status.as<int32_t>());
    ctx->n_launches++;
    CORAL_CUDA(ctx, cudaGetLastError());
    int32_t sw[3] = {0, 0, 0};
    CORAL_CUDA(ctx, cudaMemcpyAsync(sw, ctx->scan_status.p, 12, cudaMemcpyDeviceToHost, ctx->stream));
    CORAL_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->sa_side_valid = sw[2] == 0;
    if (!ctx->sa_side_valid) CORAL_CUDA(ctx, cudaMemsetAsync(ctx->scan_status.as<int32_t>() + 2, 0, 4, ctx->stream));
    if (sw[1] != 0) {
        ctx->err = "record framing inconsistent with the record core in " + std::to_string(sw[1]) + " record(s)";
        return CORAL_E_FORMAT;
    }
    if (sw[0] != 0) {
        ctx->err = "malformed aux field in " + std::to_string(sw[0]) + " record(s)";
        return CORAL_E_FORMAT;
    }
    if (!launch_full_scan) {            // CORAL_RECORDS_HOST_MAPPED_LAZY: the full scan runs when an entry point asks for it
        ctx->scan_lazy = true;
        return CORAL_OK;
    }
    ScanEvents e{g.rec.as<uint32_t>(), g.k.as<uint32_t>(), g.next_start.as<int32_t>(), g.prev_end.as<int32_t>(),
                 g.counter.as<unsigned long long>(), g.capacity};
    const int64_t want = (n + SC_WARPS - 1) / SC_WARPS;
    const int blocks = (int)std::min<int64_t>(want, (int64_t)ctx->sm_count);
    if (!ctx->ev_beg[CORAL_STAGE_SCAN]) { cudaEventCreate(&ctx->ev_beg[CORAL_STAGE_SCAN]); cudaEventCreate(&ctx->ev_end[CORAL_STAGE_SCAN]); }
    cudaEventRecord(ctx->ev_beg[CORAL_STAGE_SCAN], ctx->scan_stream);
    k_scan_records<<<blocks, SC_WARPS * 32, SC_SMEM_TOTAL, ctx->scan_stream>>>(
        ctx->d_recs, ctx->d_off.as<uint64_t>(), (uint32_t)n, ctx->prm.min_del_len, o, e, ctx->scan_status.as<int32_t>(),
        reinterpret_cast<unsigned int*>(g.counter.as<unsigned long long>() + 1), 0u, ord);
    cudaEventRecord(ctx->ev_end[CORAL_STAGE_SCAN], ctx->scan_stream);
    ctx->ev_valid[CORAL_STAGE_SCAN] = true;
    ctx->n_launches++;
    CORAL_CUDA(ctx, cudaGetLastError());
    ctx->scan_pending = true;
    coral_mem::scan_started();
    return CORAL_OK;
}

// Rescan inside a pipelined pass (device-resident records): a pre-scan (record core, read name, aux block: ~150 B per
// record) fills the columns coral_fetch / coral_candidates_* / coral_cluster_call read, then the full scan runs on its own
// stream — all SMs but a few block slots, which is where the small kernels of those entry points run MEANWHILE on the
// ctx's stream.  Entry points that need the scan's CIGAR / SEQ columns join it with a stream dependency
// (coral_wait_scan).  Same kernels, same columns, same values as coral_run_scan.
int coral_start_scan_overlapped(coral_ctx* ctx) {
    using namespace coral;
    const int64_t n = ctx->nrec;
    if (n == 0) return coral_run_scan(ctx);
    ctx->scan_lazy = false;
    ctx->sa_side_valid = false;
    if (ctx->scan_pending) {
        ctx->scan_pending = false;
        cudaStreamSynchronize(ctx->scan_stream);
        coral_mem::scan_joined();
    }
    CORAL_TRY(scan_ensure_tables(ctx));
    if (!ctx->scan_stream) CORAL_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->scan_stream, cudaStreamNonBlocking));
    if (!ctx->ev_fork) CORAL_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    const uint32_t* ord = scan_order(ctx);
    GapEvents& g = ctx->gaps;
    CORAL_CUDA(ctx, cudaMemsetAsync(g.counter.p, 0, 16, ctx->stream));
    CORAL_CUDA(ctx, cudaMemsetAsync(ctx->scan_status.p, 0, 64, ctx->stream));
    ScanCols o = scan_cols(ctx);
    // the pre-scan first, alone on the device (it is latency-bound: two dependent loads per record, and needs every
    // warp slot to finish in tens of microseconds); what runs next on the ctx's stream (fetch ...) reads its columns
    SideText side{nullptr, nullptr, 0, nullptr};
    const int pre_blocks = (int)std::min<int64_t>((n + 7) / 8, (int64_t)ctx->sm_count * 8);
    ctx->stage_begin(CORAL_STAGE_PRESCAN);
    k_prescan_records<<<pre_blocks, 256, 0, ctx->stream>>>(ctx->d_recs, ctx->d_off.as<uint64_t>(), (uint32_t)n, o, side,
                                                          ctx->scan_status.as<int32_t>());
    ctx->stage_end(CORAL_STAGE_PRESCAN);
    ctx->n_launches++;
    CORAL_CUDA(ctx, cudaGetLastError());
    CORAL_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->stream));
    CORAL_CUDA(ctx, cudaStreamWaitEvent(ctx->scan_stream, ctx->ev_fork, 0));
    ScanEvents e{g.rec.as<uint32_t>(), g.k.as<uint32_t>(), g.next_start.as<int32_t>(), g.prev_end.as<int32_t>(),
                 g.counter.as<unsigned long long>(), g.capacity};
    int per_sm = 0;
    CORAL_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_scan_records, SC_WARPS * 32, SC_SMEM_TOTAL));
    if (per_sm < 1) per_sm = 1;
    const int64_t want = (n + SC_WARPS - 1) / SC_WARPS;
    const int64_t full = (int64_t)ctx->sm_count * per_sm;
    // block slots left to the small kernels that run next to the scan: by default one of the per_sm slots of every SM
    const int64_t reserve = per_sm > 1 ? (ctx->pass.scan_reserve >= 0 ? std::min<int64_t>(ctx->pass.scan_reserve, full - ctx->sm_count)
                                                                      : (int64_t)ctx->sm_count) : 0;
    const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(want, full - reserve));
    if (!ctx->ev_beg[CORAL_STAGE_SCAN]) { cudaEventCreate(&ctx->ev_beg[CORAL_STAGE_SCAN]); cudaEventCreate(&ctx->ev_end[CORAL_STAGE_SCAN]); }
    cudaEventRecord(ctx->ev_beg[CORAL_STAGE_SCAN], ctx->scan_stream);
    k_scan_records<<<blocks, SC_WARPS * 32, SC_SMEM_TOTAL, ctx->scan_stream>>>(
        ctx->d_recs, ctx->d_off.as<uint64_t>(), (uint32_t)n, ctx->prm.min_del_len, o, e, ctx->scan_status.as<int32_t>(),
        reinterpret_cast<unsigned int*>(g.counter.as<unsigned long long>() + 1), 0u, ord);
    cudaEventRecord(ctx->ev_end[CORAL_STAGE_SCAN], ctx->scan_stream);
    ctx->ev_valid[CORAL_STAGE_SCAN] = true;
    ctx->n_launches++;
    CORAL_CUDA(ctx, cudaGetLastError());
    ctx->scan_pending = true;
    coral_mem::scan_started();
    return CORAL_OK;
}

extern "C" int coral_wait_scan(coral_ctx* ctx) {
    if (!ctx) return CORAL_E_ARG;
    if (ctx->scan_lazy) {
        cudaSetDevice(ctx->device);
        return coral_run_scan(ctx);
    }
    if (!ctx->scan_pending) return CORAL_OK;
    cudaSetDevice(ctx->device);
    ctx->scan_pending = false;
    int v;
    if (ctx->pass.active) {
        // inside a pipelined pass the join is a dependency between streams, not a host synchronisation: what follows on
        // the ctx's stream waits for the scan kernel; its verdict is resolved like every other size of the pass
        CORAL_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_end[CORAL_STAGE_SCAN], 0));
        v = scan_verdict(ctx, ctx->stream);
    } else {
        v = scan_verdict(ctx, ctx->scan_stream);
    }
    coral_mem::scan_joined();
    if (v == 1) return coral_run_scan(ctx);      // more gap events than the buffer held: synchronous rescan with room
    return v;
}

int coral_run_scan(coral_ctx* ctx) {
    using namespace coral;
    const int64_t n = ctx->nrec;
    ctx->scan_lazy = false;
    ctx->sa_side_valid = false;
    if (ctx->scan_pending) {                    // join an asynchronous scan before starting over
        ctx->scan_pending = false;
        cudaStreamSynchronize(ctx->scan_stream);
        coral_mem::scan_joined();
    }
    for (int attempt = 0; attempt < 2; attempt++) {
        CORAL_TRY(scan_ensure_tables(ctx));
        GapEvents& g = ctx->gaps;
        CORAL_CUDA(ctx, cudaMemsetAsync(g.counter.p, 0, 16, ctx->stream));     // [0] event count, [1] record queue
        CORAL_CUDA(ctx, cudaMemsetAsync(ctx->scan_status.p, 0, 64, ctx->stream));
        ScanCols o = scan_cols(ctx);
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
                ctx->d_recs, ctx->d_off.as<uint64_t>(), (uint32_t)n, ctx->prm.min_del_len, o, e, ctx->scan_status.as<int32_t>(),
                reinterpret_cast<unsigned int*>(g.counter.as<unsigned long long>() + 1), 1u, scan_order(ctx));
            ctx->stage_end(CORAL_STAGE_SCAN);
            ctx->n_launches++;
            CORAL_CUDA(ctx, cudaGetLastError());
        }
        const int v = scan_verdict(ctx, ctx->stream);
        if (v != 1) return v;
    }
    ctx->err = "gap-event buffer overflow after rescan";
    return CORAL_E_NOMEM;
}
