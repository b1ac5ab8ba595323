// coral_b200/csrc/fetch.cu — chimeric-alignment extraction (reference infer_breakpoint_graph.py:139-174
// `fetch` + cigar_parsing.py:232-269 `alignment_from_satags` and the nine cigar2pos* forms :17-215).
//
// Input is the scan table (SA tag location, name hash, flag, l_seq per record).  Steps:
//   1. ordered compaction of SA-bearing records; per record count of ';'-separated entries
//   2. entry table: byte range of every entry in emission order (record order, then entry order)
//   3. parse every entry (thread per entry): rname -> tid, pos, strand, CIGAR signature + numbers, mapq, NM
//   4. stable sort by read-name hash => reads; first-seen de-duplication of identical entry strings
//   5. read_length = l_seq of the first record with flag < 256 carrying the same name (binary-search join)
//   6. per read: (qs, qe, al) by signature, rint by strand, stable insertion sort by (qs, qe)
//   7. reads ordered by their first SA-bearing record (dict insertion order of the reference)
#include <memory>
#include <vector>

#include "ctx.h"
#include "dev_util.cuh"
#include "relops.cuh"
#include "tables.h"

namespace coral {

// ---- 1. ------------------------------------------------------------------------------------------
// part_world > 1 (rows-only ctx of the multi-GPU engine): this rank only parses the reads whose name hash falls into
// its residue class — every rank holds all gathered rows, the reads are split N ways, the parsed tables are merged
// (chimeric_merge.cu)
struct SaFlag {               // evaluated inside the ordered compaction (no flag array, no kernel of its own)
    const uint32_t* sa_off;
    const uint64_t* name_lo;
    uint32_t part_rank, part_world;
    __host__ __device__ uint8_t operator()(uint32_t i) const {
        return sa_off[i] != 0 && (part_world <= 1 || (uint32_t)((name_lo[i] >> 20) % part_world) == part_rank);
    }
};

// warp per SA record: number of entries of s[:-1].split(';')
__global__ void k_count_entries(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off,
                                const uint32_t* __restrict__ sa_rec, int64_t n_sa, const uint32_t* __restrict__ sa_off,
                                const uint32_t* __restrict__ sa_len, uint32_t* __restrict__ n_ent) {
    const uint32_t lane = lane_id();
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_sa) return;
    const uint32_t rec = sa_rec[w];
    const uint8_t* s = recs + off[rec] + sa_off[rec];
    const uint32_t len = sa_len[rec] ? sa_len[rec] - 1 : 0;  // [:-1]
    uint32_t cnt = 0;
    for (uint32_t b = lane; b < len; b += 32) cnt += s[b] == ';';
    cnt = (uint32_t)warp_sum_i64(cnt);
    if (lane == 0) {
        n_ent[w] = cnt + 1;
        if (w == 0) n_ent[n_sa] = 0;       // the scan below runs over n_sa + 1 elements (its last output = the total)
    }
}

// ---- 2. ------------------------------------------------------------------------------------------
__global__ void k_entry_ranges(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off,
                               const uint32_t* __restrict__ sa_rec, int64_t n_sa, const uint32_t* __restrict__ sa_off,
                               const uint32_t* __restrict__ sa_len, const uint32_t* __restrict__ ent_base,
                               uint32_t* __restrict__ e_rec, uint32_t* __restrict__ e_beg, uint32_t* __restrict__ e_end) {
    const uint32_t lane = lane_id();
    const uint32_t lt = (1u << lane) - 1u;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_sa) return;
    const uint32_t rec = sa_rec[w];
    const uint32_t so = sa_off[rec];
    const uint8_t* s = recs + off[rec] + so;
    const uint32_t len = sa_len[rec] ? sa_len[rec] - 1 : 0;
    const uint32_t base = ent_base[w];
    uint32_t seen = 0;
    if (lane == 0) { e_rec[base] = rec; e_beg[base] = so; }
    for (uint32_t b0 = 0; b0 < len; b0 += 32) {
        const uint32_t b = b0 + lane;
        const bool sep = b < len && s[b] == ';';
        const uint32_t m = __ballot_sync(FULL_MASK, sep);
        if (sep) {
            const uint32_t k = seen + __popc(m & lt);
            e_end[base + k] = so + b;
            e_rec[base + k + 1] = rec;
            e_beg[base + k + 1] = so + b + 1;
        }
        seen += __popc(m);
    }
    if (lane == 0) e_end[base + seen] = so + len;
}

// ---- 3. ------------------------------------------------------------------------------------------
// signature codes: index into {"SM","MS","SMS","SMD","MDS","SMDS","SMI","MIS","SMIS"}; 100 = reject
// (no 'S' or no 'M', cp:248-253); 101 = KeyError (signature not in cp:219-229); 102 = malformed.
__device__ __forceinline__ int sig_code(const char* sig, int n) {
    const char* keys[9] = {"SM", "MS", "SMS", "SMD", "MDS", "SMDS", "SMI", "MIS", "SMIS"};
    for (int k = 0; k < 9; k++) {
        int j = 0;
        while (j < n && keys[k][j] && keys[k][j] == sig[j]) j++;
        if (j == n && keys[k][j] == 0) return k;
    }
    return 101;
}

__global__ void k_parse_entries(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off, int64_t n_ent,
                                const uint32_t* __restrict__ e_rec, const uint32_t* __restrict__ e_beg,
                                const uint32_t* __restrict__ e_end, const uint64_t* __restrict__ ref_hash, int32_t n_ref,
                                SaEntry* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_ent) return;
    const uint8_t* s = recs + off[e_rec[i]];
    uint32_t p = e_beg[i];
    const uint32_t end = e_end[i];
    SaEntry e;
    e.rec = e_rec[i];
    e.tid = -1; e.pos1 = 0; e.strand = 1; e.sig = 102; e.mapq = 0; e.nm = 0;
    e.num[0] = e.num[1] = e.num[2] = e.num[3] = 0;
    // entry hash over raw bytes (exact-string de-duplication, ibg:148)
    uint64_t h1 = 0, h2 = 0;
    for (uint32_t b = p; b < end; b++) { h1 += hterm1(b - p, s[b]); h2 += hterm2(b - p, s[b]); }
    e.h1 = h1 + mix64(end - p);
    e.h2 = h2 + mix64((end - p) ^ 0xABCDull);
    // field 0: rname
    uint32_t f0 = p;
    while (p < end && s[p] != ',') p++;
    {
        uint64_t a = 0, b2 = 0;
        for (uint32_t b = f0; b < p; b++) { a += hterm1(b - f0, s[b]); b2 += hterm2(b - f0, s[b]); }
        a += mix64(p - f0); b2 += mix64((p - f0) ^ 0xABCDull);
        for (int32_t t = 0; t < n_ref; t++)
            if (ref_hash[t] == a && ref_hash[n_ref + t] == b2) { e.tid = t; break; }
    }
    bool ok = p < end;
    p++;
    // field 1: pos
    int64_t v = 0; uint32_t nd = 0;
    while (ok && p < end && s[p] != ',') { uint32_t d = s[p] - '0'; if (d > 9) ok = false; v = v * 10 + d; nd++; p++; }
    ok = ok && nd > 0 && p < end; e.pos1 = v; p++;
    // field 2: strand ("+" exactly, anything else behaves as '-', cp:257 / cigar2pos*)
    uint32_t f2 = p;
    while (ok && p < end && s[p] != ',') p++;
    e.strand = (ok && p - f2 == 1 && s[f2] == '+') ? 0 : 1;
    ok = ok && p < end; p++;
    // field 3: CIGAR
    char sig[6]; int ns = 0, nn = 0; bool has_s = false, has_m = false, num_open = false, bad_num = false;
    int64_t cur = 0;
    while (ok && p < end && s[p] != ',') {
        const uint8_t ch = s[p];
        if (ch >= '0' && ch <= '9') { cur = cur * 10 + (ch - '0'); num_open = true; }
        else {
            if (ch == 'S') has_s = true;
            if (ch == 'M') has_m = true;
            if (ns < 5) sig[ns] = (char)ch;
            ns++;
            if (!num_open) bad_num = true;
            if (nn < 4) e.num[nn] = cur;
            nn++; cur = 0; num_open = false;
        }
        p++;
    }
    if (num_open) bad_num = true;  // trailing digits without an op letter
    ok = ok && p < end; p++;
    // field 4: mapq
    v = 0; nd = 0;
    while (ok && p < end && s[p] != ',') { uint32_t d = s[p] - '0'; if (d > 9) ok = false; v = v * 10 + d; nd++; p++; }
    ok = ok && nd > 0; e.mapq = (int32_t)v;
    // last field: NM (float(t[-1]))
    if (ok) {
        uint32_t q = end;
        while (q > e_beg[i] && s[q - 1] != ',') q--;
        v = 0; nd = 0;
        for (uint32_t b = q; b < end; b++) { uint32_t d = s[b] - '0'; if (d > 9) ok = false; v = v * 10 + d; nd++; }
        ok = ok && nd > 0; e.nm = v;
    }
    if (!ok) e.sig = 102;
    else if (!has_s || !has_m) e.sig = 100;
    else if (ns > 4) e.sig = 101;
    else { int c = sig_code(sig, ns); e.sig = (c <= 8 && bad_num) ? 102 : c; }
    out[i] = e;
}

// ---- 4. ------------------------------------------------------------------------------------------
// iota (nullable): also writes the identity permutation (saves a launch before the first sort)
__global__ void k_gather_name(int64_t n, const SaEntry* __restrict__ ent, const uint32_t* __restrict__ perm,
                              const uint64_t* __restrict__ name, uint64_t* __restrict__ key, uint32_t* __restrict__ iota,
                              uint32_t* __restrict__ zero_word) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    key[i] = name[ent[perm ? perm[i] : i].rec];
    if (iota) iota[i] = (uint32_t)i;
    if (zero_word && i == 0) *zero_word = 0;       // k_heads' "mixed" flag of this round
}
__global__ void k_iota(int64_t n, uint32_t* p) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = (uint32_t)i;
}
// head[i] = 1 when sorted position i starts a new read name
// head[i] = entry i (in sorted order) starts a new read name.  The entries were sorted on `key_mask` bits of the low
// hash word only: if two different names share those bits their entries may interleave, which shows as an adjacent
// pair with equal masked keys and different names -> *mixed is set and the caller re-sorts on all 128 bits.
__global__ void k_heads(int64_t n, const SaEntry* __restrict__ ent, const uint32_t* __restrict__ perm,
                        const uint64_t* __restrict__ nlo, const uint64_t* __restrict__ nhi, uint64_t key_mask,
                        uint32_t* __restrict__ head, uint32_t* __restrict__ mixed) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == 0) { head[0] = 1; return; }
    const uint32_t a = ent[perm[i]].rec, b = ent[perm[i - 1]].rec;
    const bool differ = nlo[a] != nlo[b] || nhi[a] != nhi[b];
    head[i] = differ ? 1u : 0u;
    if (differ && ((nlo[a] ^ nlo[b]) & key_mask) == 0) *mixed = 1u;
}
// seg_id (0-based) from inclusive scan of heads; records segment starts; marks duplicates within a segment
__global__ void k_segments(int64_t n, const SaEntry* __restrict__ ent, const uint32_t* __restrict__ perm,
                           const uint32_t* __restrict__ head_incl, uint32_t* __restrict__ seg_start,
                           uint32_t* __restrict__ seg_of) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t sgm = head_incl[i] - 1;
    seg_of[i] = sgm;
    if (i == 0 || head_incl[i - 1] != head_incl[i]) seg_start[sgm] = (uint32_t)i;
}
__global__ void k_mark_dups(int64_t n, const SaEntry* __restrict__ ent, const uint32_t* __restrict__ perm,
                            const uint32_t* __restrict__ seg_of, const uint32_t* __restrict__ seg_start,
                            uint8_t* __restrict__ keep) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const SaEntry& me = ent[perm[i]];
    bool dup = false;
    for (uint32_t j = seg_start[seg_of[i]]; j < (uint32_t)i && !dup; j++) {
        const SaEntry& o = ent[perm[j]];
        dup = (o.h1 == me.h1 && o.h2 == me.h2);
    }
    keep[i] = !dup;
}

// ---- 5. ------------------------------------------------------------------------------------------
// names of the segments: sorted by the masked low word (key_mask != 0; distinct per segment, k_heads checked it), or by
// (hi, lo) after the full LSD lo-then-hi sort (key_mask == 0); binary search join.
__global__ void k_first_primary(int64_t n_rec, const uint32_t* __restrict__ flag_mapq, const uint64_t* __restrict__ nlo,
                                const uint64_t* __restrict__ nhi, int64_t n_seg, const uint64_t* __restrict__ seg_lo,
                                const uint64_t* __restrict__ seg_hi, uint64_t key_mask, uint32_t* __restrict__ first_primary) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rec) return;
    if ((flag_mapq[i] & 0xffffu) >= 256u) return;
    const uint64_t lo_k = nlo[i], hi_k = nhi[i];
    int64_t lo = 0, hi = n_seg;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        bool less = key_mask ? (seg_lo[mid] & key_mask) < (lo_k & key_mask)
                             : (seg_hi[mid] < hi_k || (seg_hi[mid] == hi_k && seg_lo[mid] < lo_k));
        if (less) lo = mid + 1; else hi = mid;
    }
    if (lo < n_seg && seg_hi[lo] == hi_k && seg_lo[lo] == lo_k) atomicMin(&first_primary[lo], (uint32_t)i);
}
__global__ void k_seg_names(int64_t n_seg, const SaEntry* __restrict__ ent, const uint32_t* __restrict__ perm,
                            const uint32_t* __restrict__ seg_start, const uint64_t* __restrict__ nlo,
                            const uint64_t* __restrict__ nhi, uint64_t* __restrict__ seg_lo, uint64_t* __restrict__ seg_hi,
                            uint64_t* __restrict__ seg_first_emit, uint32_t* __restrict__ first_primary,
                            uint32_t* __restrict__ seg_idx) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_seg) return;
    seg_idx[s] = (uint32_t)s;
    const uint32_t e0 = perm[seg_start[s]];  // stable sort => smallest emission index of the segment
    const uint32_t rec = ent[e0].rec;
    seg_lo[s] = nlo[rec];
    seg_hi[s] = nhi[rec];
    seg_first_emit[s] = e0;
    first_primary[s] = 0xFFFFFFFFu;
}

}  // namespace coral

namespace coral {

// ---- 6. ------------------------------------------------------------------------------------------
// cigar2pos{SM,MS,SMS,SMD,MDS,SMDS,SMI,MIS,SMIS} (cp:17-215), n = CIGAR numbers in order.
__device__ __forceinline__ void cigar2pos_dev(int sig, const int64_t* n, bool minus, int64_t rl, int64_t& qs,
                                              int64_t& qe, int64_t& al) {
    switch (sig) {
        case 0: al = n[1];        if (!minus) { qs = n[0]; qe = rl - 1; }            else { qs = 0;    qe = n[1] - 1; }         break;
        case 1: al = n[0];        if (!minus) { qs = 0;    qe = n[0] - 1; }          else { qs = n[1]; qe = rl - 1; }           break;
        case 2: al = n[1];        if (!minus) { qs = n[0]; qe = n[0] + n[1] - 1; }   else { qs = n[2]; qe = n[2] + n[1] - 1; }  break;
        case 3: al = n[1] + n[2]; if (!minus) { qs = n[0]; qe = rl - 1; }            else { qs = 0;    qe = n[1] - 1; }         break;
        case 4: al = n[0] + n[1]; if (!minus) { qs = 0;    qe = n[0] - 1; }          else { qs = n[2]; qe = rl - 1; }           break;
        case 5: al = n[1] + n[2]; if (!minus) { qs = n[0]; qe = rl - n[3] - 1; }     else { qs = n[3]; qe = rl - n[0] - 1; }    break;
        case 6: al = n[1];        if (!minus) { qs = n[0]; qe = rl - 1; }            else { qs = 0;    qe = rl - n[0] - 1; }    break;
        case 7: al = n[0];        if (!minus) { qs = 0;    qe = rl - n[2] - 1; }     else { qs = n[2]; qe = rl - 1; }           break;
        default: al = n[1];       if (!minus) { qs = n[0]; qe = rl - n[3] - 1; }     else { qs = n[3]; qe = rl - n[0] - 1; }    break;
    }
}

struct ChimCols {
    uint32_t *read_first_rec, *aln_ptr;
    int64_t *read_len, *nm;
    int32_t *read_status, *qs, *qe, *tid, *rs, *re, *mapq;
    uint8_t* strand;
};

// thread per segment (read): status + number of alignments
__global__ void k_read_counts(int64_t n_seg, int64_t n_ent, const SaEntry* __restrict__ ent, const uint32_t* __restrict__ perm,
                              const uint32_t* __restrict__ seg_start, const uint8_t* __restrict__ keep,
                              const uint32_t* __restrict__ first_primary, int32_t* __restrict__ seg_status,
                              uint32_t* __restrict__ seg_cnt) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_seg) return;
    const uint32_t b = seg_start[s], e = (s + 1 < n_seg) ? seg_start[s + 1] : (uint32_t)n_ent;
    int32_t st = READ_OK;
    uint32_t cnt = 0;
    if (first_primary[s] == 0xFFFFFFFFu) st = READ_NO_PRIMARY;
    else {
        for (uint32_t i = b; i < e && st == READ_OK; i++) {
            if (!keep[i]) continue;
            const uint8_t sg = ent[perm[i]].sig;
            if (sg == 100) st = READ_REJECTED;
            else if (sg == 101) st = READ_KEYERROR;
            else if (sg == 102 || ent[perm[i]].tid < 0) st = READ_MALFORMED;
            else cnt++;
        }
    }
    seg_status[s] = st;
    seg_cnt[s] = st == READ_OK ? cnt : 0;
}

__global__ void k_gather_u32(int64_t n, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ src, uint32_t* __restrict__ dst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}

// thread per read rank: fill alignments, stable insertion sort by (qs, qe) (cp:263)
__global__ void k_read_fill(int64_t n_seg, int64_t n_ent, const SaEntry* __restrict__ ent, const uint32_t* __restrict__ perm,
                            const uint32_t* __restrict__ seg_start, const uint8_t* __restrict__ keep,
                            const uint32_t* __restrict__ first_primary, const int32_t* __restrict__ seg_status,
                            const uint32_t* __restrict__ order, const uint32_t* __restrict__ lseq, uint32_t total_aln,
                            ChimCols o) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_seg) return;
    const uint32_t s = order[r];
    const uint32_t b = seg_start[s], e = (s + 1 < (uint32_t)n_seg) ? seg_start[s + 1] : (uint32_t)n_ent;
    const int32_t st = seg_status[s];
    o.read_status[r] = st;
    o.read_first_rec[r] = ent[perm[b]].rec;
    const uint32_t fp = first_primary[s];
    const int64_t rl = fp == 0xFFFFFFFFu ? -1 : (int64_t)lseq[fp];
    o.read_len[r] = rl;
    if (r == n_seg - 1) o.aln_ptr[n_seg] = total_aln;
    if (st != READ_OK) return;
    const uint32_t base = o.aln_ptr[r];
    uint32_t k = 0;
    for (uint32_t i = b; i < e; i++) {
        if (!keep[i]) continue;
        const SaEntry& a = ent[perm[i]];
        int64_t qs, qe, al;
        cigar2pos_dev(a.sig, a.num, a.strand != 0, rl, qs, qe, al);
        const int64_t lo = a.pos1 - 1, hi = a.pos1 + al - 2;
        const int32_t rs = (int32_t)(a.strand ? hi : lo), re = (int32_t)(a.strand ? lo : hi);
        // insertion point: after every earlier element with (qs, qe) <= mine (stable)
        uint32_t j = k;
        while (j > 0 && (o.qs[base + j - 1] > qs || (o.qs[base + j - 1] == qs && o.qe[base + j - 1] > qe))) {
            const uint32_t d = base + j, sft = base + j - 1;
            o.qs[d] = o.qs[sft]; o.qe[d] = o.qe[sft]; o.tid[d] = o.tid[sft]; o.rs[d] = o.rs[sft]; o.re[d] = o.re[sft];
            o.strand[d] = o.strand[sft]; o.mapq[d] = o.mapq[sft]; o.nm[d] = o.nm[sft];
            j--;
        }
        const uint32_t d = base + j;
        o.qs[d] = (int32_t)qs; o.qe[d] = (int32_t)qe; o.tid[d] = a.tid; o.rs[d] = rs; o.re[d] = re;
        o.strand[d] = a.strand; o.mapq[d] = a.mapq; o.nm[d] = a.nm;
        k++;
    }
}

// ---- nm_stats (ibg:153-160): records without SA and MAPQ 60 contribute e = NM / query_length -------
constexpr int NM_BLOCK = 256;
__global__ void k_nm_stats(int64_t n, const uint32_t* __restrict__ sa_off, const uint32_t* __restrict__ flag_mapq,
                           const int32_t* __restrict__ tid, const int32_t* __restrict__ nm, const uint32_t* __restrict__ lseq,
                           double* __restrict__ part) {
    __shared__ double s1[NM_BLOCK], s2[NM_BLOCK], s3[NM_BLOCK], s4[NM_BLOCK];
    const int64_t i = (int64_t)blockIdx.x * NM_BLOCK + threadIdx.x;
    double e = 0, cnt = 0, zd = 0;
    if (i < n && tid[i] >= 0 && sa_off[i] == 0 && (flag_mapq[i] >> 16) == 60u) {
        if (lseq[i] == 0) zd = 1;
        else { e = (double)nm[i] / (double)lseq[i]; cnt = 1; }
    }
    s1[threadIdx.x] = e; s2[threadIdx.x] = e * e; s3[threadIdx.x] = cnt; s4[threadIdx.x] = zd;
    __syncthreads();
    for (int d = NM_BLOCK / 2; d > 0; d >>= 1) {
        if (threadIdx.x < d) {
            s1[threadIdx.x] += s1[threadIdx.x + d]; s2[threadIdx.x] += s2[threadIdx.x + d];
            s3[threadIdx.x] += s3[threadIdx.x + d]; s4[threadIdx.x] += s4[threadIdx.x + d];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        part[4 * (size_t)blockIdx.x] = s1[0]; part[4 * (size_t)blockIdx.x + 1] = s2[0];
        part[4 * (size_t)blockIdx.x + 2] = s3[0]; part[4 * (size_t)blockIdx.x + 3] = s4[0];
    }
}

// per-block partial sums of k_nm_stats (4 doubles per block) -> 4 totals, in a fixed order (thread t takes blocks t,
// t + 256, ... in increasing order, then a tree): deterministic, and within rounding of the host's serial sum
__global__ void k_nm_reduce(const double* __restrict__ part, int64_t nblk, double* __restrict__ out) {
    __shared__ double s[4][256];
    double a[4] = {0, 0, 0, 0};
    for (int64_t b = threadIdx.x; b < nblk; b += 256)
        for (int k = 0; k < 4; k++) a[k] += part[4 * b + k];
    for (int k = 0; k < 4; k++) s[k][threadIdx.x] = a[k];
    __syncthreads();
    for (int d = 128; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d)
            for (int k = 0; k < 4; k++) s[k][threadIdx.x] += s[k][threadIdx.x + d];
        __syncthreads();
    }
    if (threadIdx.x < 4) out[threadIdx.x] = s[threadIdx.x][0];
}


}  // namespace coral

// ---- host orchestration ----------------------------------------------------------------------------
#define LAUNCH_1D(kern, n, ...)                                                          \
    do {                                                                                 \
        if ((n) > 0) {                                                                   \
            kern<<<(unsigned)(((n) + 255) / 256), 256, 0, c->stream>>>(__VA_ARGS__);      \
            c->n_launches++;                                                             \
            CORAL_CUDA(c, cudaGetLastError());                                           \
        }                                                                                \
    } while (0)
#define LAUNCH_WARP(kern, nw, ...)                                                       \
    do {                                                                                 \
        if ((nw) > 0) {                                                                  \
            kern<<<(unsigned)(((nw) + 7) / 8), 256, 0, c->stream>>>(__VA_ARGS__);         \
            c->n_launches++;                                                             \
            CORAL_CUDA(c, cudaGetLastError());                                           \
        }                                                                                \
    } while (0)

int coral_run_fetch(coral_ctx* c) {
    using namespace coral;
    if (!c->loaded) { c->err = "coral_fetch before coral_load_records"; return CORAL_E_STATE; }
    if (c->ref_names.empty()) { c->err = "coral_fetch before coral_set_references"; return CORAL_E_STATE; }
    ChimericTable& T = c->chim;
    T.valid = false; T.n_reads = 0; T.n_aln = 0;
    CORAL_TRY(pass_side_join(c, SIDE_CHIM));
    const int64_t n = c->nrec;
    RecordTable& rt = c->rt;
    if (c->pass.active) {
        const int32_t part[2] = {c->fetch_part_rank, c->fetch_part_world};
        pass_call(c, pass_hash(pass_hash(PASS_H0, "fetch", 5), part, 8));
    }
    // where the SA strings are read from: the records, or — after a pre-scan (asynchronous / lazy placements) — its
    // side copy in HBM, addressed through per-record offsets chosen so that base + offset + sa_off is the copy
    const uint8_t* recs = c->sa_side_valid ? c->sa_side.as<uint8_t>() : c->d_recs;
    const uint64_t* off = c->sa_side_valid ? c->sa_side_off.as<uint64_t>() : c->d_off.as<uint64_t>();
    c->stage_begin(CORAL_STAGE_FETCH);

    // nm_stats partials
    // (rows of a rows-only ctx carry no placement, tid -1: they contribute nothing — the shards' sums travel in the
    // headers of the exchange — so the kernel and its copy are skipped there)
    const int64_t nblk = c->rows_only ? 0 : (n + NM_BLOCK - 1) / NM_BLOCK;
    auto part = std::make_shared<std::vector<double>>((size_t)std::max<int64_t>(nblk, 1) * 4, 0.0);
    if (nblk > 0) {
        // partials in a buffer of their own: their copy (a RESULT: inside a pipelined pass it arrives at coral_pass_end)
        // runs on the pass's side stream, off the chain of dependent kernels below
        CORAL_CUDA(c, c->nm_part.ensure((size_t)nblk * 32));
        k_nm_stats<<<(unsigned)nblk, NM_BLOCK, 0, c->stream>>>(n, rt.sa_off.as<uint32_t>(), rt.flag_mapq.as<uint32_t>(),
                                                              rt.tid.as<int32_t>(), rt.nm.as<int32_t>(), rt.lseq.as<uint32_t>(),
                                                              c->nm_part.as<double>());
        c->n_launches++;
        CORAL_CUDA(c, cudaGetLastError());
        if (c->pass.active) {
            CORAL_TRY(pass_side_begin(c, SIDE_CHIM));
            CORAL_TRY(d2h_result(c, part->data(), c->nm_part.p, (size_t)nblk * 4));
            CORAL_TRY(pass_sync(c));
        } else {
            CORAL_TRY(d2h_result(c, part->data(), c->nm_part.p, (size_t)nblk * 4));
        }
    }
    // 1. SA-bearing records in record order
    CORAL_CUDA(c, c->s0.ensure((size_t)std::max<int64_t>(n, 1)));          // flags u8
    CORAL_CUDA(c, c->s1.ensure((size_t)std::max<int64_t>(n, 1) * 4 + 16)); // sa_rec u32 + count
    uint8_t* d_flag = c->s0.as<uint8_t>();
    uint32_t* d_sa_rec = c->s1.as<uint32_t>();
    uint32_t* d_cnt = d_sa_rec + std::max<int64_t>(n, 1);
    (void)d_flag;
    {
        cub::CountingInputIterator<uint32_t> iota(0);
        cub::TransformInputIterator<uint8_t, SaFlag, cub::CountingInputIterator<uint32_t>> flags(
            iota, SaFlag{rt.sa_off.as<uint32_t>(), rt.name_lo.as<uint64_t>(), (uint32_t)c->fetch_part_rank, (uint32_t)c->fetch_part_world});
        if (n > 0) {
            size_t bytes = 0;
            CORAL_CUDA(c, cub::DeviceSelect::Flagged(nullptr, bytes, iota, flags, d_sa_rec, d_cnt, (int)n, c->stream));
            CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
            CORAL_CUDA(c, cub::DeviceSelect::Flagged(c->cub_tmp.p, bytes, iota, flags, d_sa_rec, d_cnt, (int)n, c->stream));
            c->n_launches += 1;
        } else {
            CORAL_CUDA(c, cudaMemsetAsync(d_cnt, 0, 4, c->stream));
        }
    }
    uint32_t n_sa = 0;
    // ... the number of SA-bearing records sizes the launches below
    CORAL_TRY(pass_resolve_lazy(c, &n_sa, d_cnt));
    CORAL_TRY(pass_after(c, [c, part, nblk]() {
        ChimericTable& T = c->chim;
        T.nm_sum = T.nm_sq = 0; T.nm_n = 0; T.nm_zero_div = 0;
        for (int64_t b = 0; b < nblk; b++) {
            T.nm_sum += (*part)[4 * b]; T.nm_sq += (*part)[4 * b + 1];
            T.nm_n += (int64_t)(*part)[4 * b + 2]; T.nm_zero_div += (int64_t)(*part)[4 * b + 3];
        }
        return CORAL_OK;
    }));
    if (n_sa == 0) { CORAL_TRY(pass_flush_checks(c)); T.valid = true; c->stage_end(CORAL_STAGE_FETCH); return CORAL_OK; }

    // entries per SA record -> bases
    CORAL_CUDA(c, c->s2.ensure((size_t)(n_sa + 1) * 8 + 16));
    uint32_t* d_nent = c->s2.as<uint32_t>();
    uint32_t* d_base = d_nent + n_sa + 1;
    LAUNCH_WARP(k_count_entries, n_sa, recs, off, d_sa_rec, (int64_t)n_sa, rt.sa_off.as<uint32_t>(), rt.sa_len.as<uint32_t>(), d_nent);
    CORAL_TRY(exclusive_sum_u32(c, d_nent, d_base, (int64_t)n_sa + 1));
    uint32_t n_ent = 0;
    CORAL_TRY(pass_resolve_lazy(c, &n_ent, d_base + n_sa));

    // 2-3. entry ranges + parse
    CORAL_CUDA(c, c->s3.ensure((size_t)n_ent * 12));
    uint32_t* e_rec = c->s3.as<uint32_t>();
    uint32_t* e_beg = e_rec + n_ent;
    uint32_t* e_end = e_beg + n_ent;
    LAUNCH_WARP(k_entry_ranges, n_sa, recs, off, d_sa_rec, (int64_t)n_sa, rt.sa_off.as<uint32_t>(), rt.sa_len.as<uint32_t>(),
                d_base, e_rec, e_beg, e_end);
    CORAL_CUDA(c, c->s4.ensure((size_t)n_ent * sizeof(SaEntry)));
    SaEntry* ent = c->s4.as<SaEntry>();
    LAUNCH_1D(k_parse_entries, (int64_t)n_ent, recs, off, (int64_t)n_ent, e_rec, e_beg, e_end, c->d_ref_hash.as<uint64_t>(),
              (int32_t)c->ref_names.size(), ent);

    // 4. stable sort of entry indices by name.  One 5-pass sort on 40 bits of the low hash word groups the names unless
    // two names share those bits (p ~ n^2 / 2^41); k_heads detects that exactly and the full 128-bit sort runs instead.
    CORAL_CUDA(c, c->s5.ensure((size_t)n_ent * 24 + 64));
    uint64_t* key_a = c->s5.as<uint64_t>();
    uint64_t* key_b = key_a + n_ent;
    uint32_t* perm_a = reinterpret_cast<uint32_t*>(key_b + n_ent);
    uint32_t* perm_b = perm_a + n_ent;
    CORAL_CUDA(c, c->s6.ensure((size_t)n_ent * 13 + 64 + 16));
    uint32_t* head = c->s6.as<uint32_t>();
    uint32_t* seg_of = head + n_ent;
    uint32_t* seg_start = seg_of + n_ent;
    uint8_t* keep = reinterpret_cast<uint8_t*>(seg_start + n_ent);
    uint32_t* d_mixed = reinterpret_cast<uint32_t*>(c->s6.as<uint8_t>() + (((size_t)n_ent * 13 + 64 + 3) & ~(size_t)3));
    const uint32_t* perm = nullptr;
    uint32_t n_seg = 0;
    uint64_t name_key_mask = 0;
    for (int full = c->force_full_name_sort ? 1 : 0; full < 2; full++) {
        name_key_mask = full ? 0 : ((1ull << 40) - 1);
        LAUNCH_1D(k_gather_name, (int64_t)n_ent, (int64_t)n_ent, ent, (const uint32_t*)nullptr, rt.name_lo.as<uint64_t>(), key_a, perm_a, d_mixed);
        if (!full) {
            CORAL_TRY(sort_pairs_u64(c, key_a, key_b, perm_a, perm_b, n_ent, 0, 40));
            perm = perm_b;
        } else {
            CORAL_TRY(sort_pairs_u64(c, key_a, key_b, perm_a, perm_b, n_ent));
            LAUNCH_1D(k_gather_name, (int64_t)n_ent, (int64_t)n_ent, ent, perm_b, rt.name_hi.as<uint64_t>(), key_a, (uint32_t*)nullptr, (uint32_t*)nullptr);
            CORAL_TRY(sort_pairs_u64(c, key_a, key_b, perm_b, perm_a, n_ent));
            perm = perm_a;
        }
        LAUNCH_1D(k_heads, (int64_t)n_ent, (int64_t)n_ent, ent, perm, rt.name_lo.as<uint64_t>(), rt.name_hi.as<uint64_t>(),
                  full ? ~0ull : name_key_mask, head, d_mixed);
        CORAL_TRY(inclusive_sum_u32(c, head, head, n_ent));
        uint32_t mixed = 0;
        CORAL_TRY(pass_resolve_lazy(c, &n_seg, head + n_ent - 1));
        CORAL_TRY(pass_resolve_lazy(c, &mixed, d_mixed));
        if (!mixed) break;
        CORAL_TRY(pass_flush_checks(c));       // the second round overwrites what the noted checks read
    }
    LAUNCH_1D(k_segments, (int64_t)n_ent, (int64_t)n_ent, ent, perm, head, seg_start, seg_of);
    LAUNCH_1D(k_mark_dups, (int64_t)n_ent, (int64_t)n_ent, ent, perm, seg_of, seg_start, keep);

    // 5. per-read names, first primary
    CORAL_CUDA(c, c->s7.ensure((size_t)n_seg * 48 + 64));
    uint64_t* seg_lo = c->s7.as<uint64_t>();
    uint64_t* seg_hi = seg_lo + n_seg;
    uint64_t* seg_emit = seg_hi + n_seg;
    uint64_t* seg_emit_sorted = seg_emit + n_seg;
    uint32_t* first_primary = reinterpret_cast<uint32_t*>(seg_emit_sorted + n_seg);
    uint32_t* seg_idx = first_primary + n_seg;
    uint32_t* order = seg_idx + n_seg;
    int32_t* seg_status = reinterpret_cast<int32_t*>(order + n_seg);
    LAUNCH_1D(k_seg_names, (int64_t)n_seg, (int64_t)n_seg, ent, perm, seg_start, rt.name_lo.as<uint64_t>(), rt.name_hi.as<uint64_t>(),
              seg_lo, seg_hi, seg_emit, first_primary, seg_idx);
    LAUNCH_1D(k_first_primary, n, n, rt.flag_mapq.as<uint32_t>(), rt.name_lo.as<uint64_t>(), rt.name_hi.as<uint64_t>(), (int64_t)n_seg,
              seg_lo, seg_hi, name_key_mask, first_primary);

    // 6-7. counts, read order, CSR
    CORAL_CUDA(c, T.read_first_rec.ensure((size_t)n_seg * 4));
    CORAL_CUDA(c, T.read_len.ensure((size_t)n_seg * 8));
    CORAL_CUDA(c, T.read_status.ensure((size_t)n_seg * 4));
    CORAL_CUDA(c, T.aln_ptr.ensure((size_t)(n_seg + 1) * 4));
    CORAL_CUDA(c, c->s0.ensure((size_t)n_seg * 8 + 64));
    uint32_t* seg_cnt = c->s0.as<uint32_t>();
    uint32_t* cnt_ranked = seg_cnt + n_seg;
    LAUNCH_1D(k_read_counts, (int64_t)n_seg, (int64_t)n_seg, (int64_t)n_ent, ent, perm, seg_start, keep, first_primary, seg_status, seg_cnt);
    int emit_bits = 1;                                            // keys are entry indices < n_ent
    while (emit_bits < 32 && ((uint64_t)1 << emit_bits) < (uint64_t)n_ent) emit_bits++;
    CORAL_TRY(sort_pairs_u64(c, seg_emit, seg_emit_sorted, seg_idx, order, n_seg, 0, emit_bits));
    LAUNCH_1D(k_gather_u32, (int64_t)n_seg, (int64_t)n_seg, order, seg_cnt, cnt_ranked);
    CORAL_TRY(exclusive_sum_u32(c, cnt_ranked, T.aln_ptr.as<uint32_t>(), n_seg));
    uint32_t last_ptr = 0, last_cnt = 0;
    CORAL_TRY(pass_resolve_lazy(c, &last_ptr, T.aln_ptr.as<uint32_t>() + n_seg - 1));
    CORAL_TRY(pass_resolve_lazy(c, &last_cnt, cnt_ranked + n_seg - 1));
    const uint32_t n_aln = last_ptr + last_cnt;
    const size_t na = std::max<uint32_t>(n_aln, 1);
    DevBuf* a4[] = {&T.qs, &T.qe, &T.tid, &T.rs, &T.re, &T.mapq};
    for (DevBuf* b : a4) CORAL_CUDA(c, b->ensure(na * 4));
    CORAL_CUDA(c, T.strand.ensure(na));
    CORAL_CUDA(c, T.nm.ensure(na * 8));
    ChimCols o{T.read_first_rec.as<uint32_t>(), T.aln_ptr.as<uint32_t>(), T.read_len.as<int64_t>(), T.nm.as<int64_t>(),
               T.read_status.as<int32_t>(), T.qs.as<int32_t>(), T.qe.as<int32_t>(), T.tid.as<int32_t>(), T.rs.as<int32_t>(),
               T.re.as<int32_t>(), T.mapq.as<int32_t>(), T.strand.as<uint8_t>()};
    LAUNCH_1D(k_read_fill, (int64_t)n_seg, (int64_t)n_seg, (int64_t)n_ent, ent, perm, seg_start, keep, first_primary, seg_status, order,
              rt.lseq.as<uint32_t>(), n_aln, o);
    CORAL_TRY(pass_flush_checks(c));           // one kernel for every size this call replayed
    c->stage_end(CORAL_STAGE_FETCH);
    CORAL_TRY(pass_sync(c));
    T.n_reads = n_seg;
    T.n_aln = n_aln;
    T.valid = true;
    return CORAL_OK;
}

// Device form (the packed exchange): the four sums of this ctx's records, reduced on the device; *d_out = 4 doubles
// (sum e, sum e^2, count, zero-division count) in ctx scratch, valid until the next use of s7 on this stream.
int coral_nm_partials_device(coral_ctx* c, double** d_out) {
    using namespace coral;
    const int64_t n = c->nrec;
    const int64_t nblk = (n + NM_BLOCK - 1) / NM_BLOCK;
    RecordTable& rt = c->rt;
    CORAL_CUDA(c, c->s7.ensure((size_t)std::max<int64_t>(nblk, 1) * 32 + 64));
    double* part = c->s7.as<double>();
    double* out = part + (size_t)std::max<int64_t>(nblk, 1) * 4;
    if (n > 0) {
        k_nm_stats<<<(unsigned)nblk, NM_BLOCK, 0, c->stream>>>(n, rt.sa_off.as<uint32_t>(), rt.flag_mapq.as<uint32_t>(),
                                                              rt.tid.as<int32_t>(), rt.nm.as<int32_t>(), rt.lseq.as<uint32_t>(), part);
        c->n_launches++;
        CORAL_CUDA(c, cudaGetLastError());
    }
    k_nm_reduce<<<1, 256, 0, c->stream>>>(part, n > 0 ? nblk : 0, out);
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    *d_out = out;
    return CORAL_OK;
}

// nm_stats partial sums of this ctx's records alone (ibg:153-157); the sharded path all-reduces them.
extern "C" int coral_nm_partials(coral_ctx* c, double* nm_sum, double* nm_sumsq, int64_t* nm_n, int64_t* nm_zero_div) {
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_nm_partials before coral_load_records"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(c));
    cudaSetDevice(c->device);
    const int64_t n = c->nrec;
    const int64_t nblk = (n + NM_BLOCK - 1) / NM_BLOCK;
    RecordTable& rt = c->rt;
    double s1 = 0, s2 = 0;
    int64_t cnt = 0, zd = 0;
    if (n > 0) {
        CORAL_CUDA(c, c->s7.ensure((size_t)nblk * 32));
        k_nm_stats<<<(unsigned)nblk, NM_BLOCK, 0, c->stream>>>(n, rt.sa_off.as<uint32_t>(), rt.flag_mapq.as<uint32_t>(),
                                                              rt.tid.as<int32_t>(), rt.nm.as<int32_t>(), rt.lseq.as<uint32_t>(),
                                                              c->s7.as<double>());
        c->n_launches++;
        CORAL_CUDA(c, cudaGetLastError());
        std::vector<double> part((size_t)nblk * 4, 0.0);
        CORAL_TRY(d2h(c, part.data(), c->s7.p, (size_t)nblk * 4));
        CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int64_t b = 0; b < nblk; b++) {
            s1 += part[4 * b]; s2 += part[4 * b + 1];
            cnt += (int64_t)part[4 * b + 2]; zd += (int64_t)part[4 * b + 3];
        }
    }
    if (nm_sum) *nm_sum = s1;
    if (nm_sumsq) *nm_sumsq = s2;
    if (nm_n) *nm_n = cnt;
    if (nm_zero_div) *nm_zero_div = zd;
    return CORAL_OK;
}

extern "C" int coral_fetch(coral_ctx* c, int64_t* n_reads, int64_t* n_alignments, double* nm_sum, double* nm_sumsq,
                           int64_t* nm_n, int64_t* nm_zero_div) {
    coral::PassLane lane__(c);
    if (!c) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    CORAL_TRY(coral_run_fetch(c));
    if (n_reads) *n_reads = c->chim.n_reads;
    if (n_alignments) *n_alignments = c->chim.n_aln;
    return coral::pass_after(c, [=]() {       // inside a pipelined pass the nm sums arrive at coral_pass_end
        if (nm_sum) *nm_sum = c->chim.nm_sum;
        if (nm_sumsq) *nm_sumsq = c->chim.nm_sq;
        if (nm_n) *nm_n = c->chim.nm_n;
        if (nm_zero_div) *nm_zero_div = c->chim.nm_zero_div;
        return CORAL_OK;
    });
}

extern "C" int coral_get_chimeric(coral_ctx* c, uint32_t* read_first_rec, int64_t* read_len, int32_t* read_status,
                                  uint32_t* aln_ptr, int32_t* qs, int32_t* qe, int32_t* tid, int32_t* rstart, int32_t* rend,
                                  uint8_t* strand, int32_t* mapq, int64_t* nm) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c) return CORAL_E_ARG;
    if (!c->chim.valid) { c->err = "coral_get_chimeric before coral_fetch"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    ChimericTable& T = c->chim;
    const size_t nr = (size_t)T.n_reads, na = (size_t)T.n_aln;
    if (nr == 0) { if (aln_ptr) aln_ptr[0] = 0; return CORAL_OK; }
    CORAL_TRY(pass_side_begin(c, SIDE_CHIM));  // the table stays as it is until the next coral_fetch / merge
    if (read_first_rec) CORAL_TRY(d2h_result(c, read_first_rec, T.read_first_rec.p, nr));
    if (read_len) CORAL_TRY(d2h_result(c, read_len, T.read_len.p, nr));
    if (read_status) CORAL_TRY(d2h_result(c, read_status, T.read_status.p, nr));
    if (aln_ptr) CORAL_TRY(d2h_result(c, aln_ptr, T.aln_ptr.p, nr + 1));
    if (qs) CORAL_TRY(d2h_result(c, qs, T.qs.p, na));
    if (qe) CORAL_TRY(d2h_result(c, qe, T.qe.p, na));
    if (tid) CORAL_TRY(d2h_result(c, tid, T.tid.p, na));
    if (rstart) CORAL_TRY(d2h_result(c, rstart, T.rs.p, na));
    if (rend) CORAL_TRY(d2h_result(c, rend, T.re.p, na));
    if (strand) CORAL_TRY(d2h_result(c, strand, T.strand.p, na));
    if (mapq) CORAL_TRY(d2h_result(c, mapq, T.mapq.p, na));
    if (nm) CORAL_TRY(d2h_result(c, nm, T.nm.p, na));
    return pass_sync(c);
}

// ---- the whole chimeric table in ONE result copy ------------------------------------------------------------------
namespace coral {
struct ColPack {
    const uint8_t* src[12];
    uint64_t dst_off[12], bytes[12];
};
__global__ void k_pack_cols(ColPack p, uint8_t* __restrict__ out) {
    const int seg = blockIdx.y;
    const uint8_t* src = p.src[seg];
    uint8_t* dst = out + p.dst_off[seg];
    const uint64_t nb = p.bytes[seg], nw = nb >> 2;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (uint64_t i = t; i < nw; i += stride) reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src)[i];
    if (t < (nb & 3)) dst[4 * nw + t] = src[4 * nw + t];
}
}  // namespace coral

// coral_get_chimeric with all twelve columns in one caller buffer: column k starts at offsets[k] (64-byte aligned, in the
// order of coral_get_chimeric's arguments).  One packing kernel + ONE device-to-host copy instead of twelve (inside a
// pipelined pass both run on the pass's side stream); out_bytes must hold offsets[11] + 8 * n_alignments.
extern "C" int coral_get_chimeric_packed(coral_ctx* c, uint8_t* out, uint64_t out_bytes, uint64_t* offsets) {
    coral::PassLane lane__(c);
    using namespace coral;
    if (!c || !out || !offsets) return CORAL_E_ARG;
    if (!c->chim.valid) { c->err = "coral_get_chimeric_packed before coral_fetch"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    ChimericTable& T = c->chim;
    const uint64_t nr = (uint64_t)T.n_reads, na = (uint64_t)T.n_aln;
    const uint64_t sizes[12] = {nr * 4, nr * 8, nr * 4, (nr + 1) * 4, na * 4, na * 4, na * 4, na * 4, na * 4, na, na * 4, na * 8};
    const void* srcs[12] = {T.read_first_rec.p, T.read_len.p, T.read_status.p, T.aln_ptr.p, T.qs.p, T.qe.p, T.tid.p, T.rs.p, T.re.p,
                            T.strand.p, T.mapq.p, T.nm.p};
    uint64_t total = 0;
    ColPack p;
    uint64_t most = 0;
    for (int k = 0; k < 12; k++) {
        offsets[k] = total;
        p.src[k] = static_cast<const uint8_t*>(srcs[k]); p.dst_off[k] = total; p.bytes[k] = nr == 0 ? 0 : sizes[k];
        most = std::max(most, sizes[k]);
        total += (sizes[k] + 63) & ~(uint64_t)63;
    }
    if (out_bytes < total) { c->err = "coral_get_chimeric_packed: buffer too small"; return CORAL_E_ARG; }
    if (nr == 0) { memset(out + offsets[3], 0, 4); return CORAL_OK; }
    CORAL_CUDA(c, c->pack_buf.ensure(total));
    CORAL_TRY(pass_side_begin(c, SIDE_CHIM));  // the table stays as it is until the next coral_fetch / merge
    PassState& P = c->pass;
    cudaStream_t st = (P.active && P.side_on) ? P.side_stream : c->stream;
    const unsigned bx = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((most / 4 + 1023) / 1024, 64));
    k_pack_cols<<<dim3(bx, 12), 256, 0, st>>>(p, c->pack_buf.as<uint8_t>());
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    CORAL_TRY(d2h_result(c, out, c->pack_buf.p, (size_t)total));
    return pass_sync(c);
}

// ---- operator form: alignment_from_satags on caller-supplied SA text ---------------------------------
namespace coral {
// thread per read: cigar2pos + rint + stable (qs, qe) insertion sort, over entries [ptr[r], ptr[r+1])
__global__ void k_sa_text_assemble(int64_t n_reads, const uint32_t* __restrict__ ptr, const SaEntry* __restrict__ ent,
                                   const int64_t* __restrict__ read_len, int32_t* __restrict__ status,
                                   int64_t* __restrict__ qs, int64_t* __restrict__ qe, int64_t* __restrict__ al,
                                   int64_t* __restrict__ rs, int64_t* __restrict__ re, uint8_t* __restrict__ strand,
                                   int32_t* __restrict__ mapq, int64_t* __restrict__ nm, uint32_t* __restrict__ src) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    const uint32_t b = ptr[r], e = ptr[r + 1];
    int32_t st = READ_OK;
    for (uint32_t i = b; i < e && st == READ_OK; i++) {
        const uint8_t sg = ent[i].sig;
        if (sg == 100) st = READ_REJECTED; else if (sg == 101) st = READ_KEYERROR; else if (sg == 102) st = READ_MALFORMED;
    }
    status[r] = st;
    if (st != READ_OK) return;
    const int64_t rl = read_len[r];
    uint32_t k = 0;
    for (uint32_t i = b; i < e; i++) {
        const SaEntry& a = ent[i];
        int64_t q0, q1, l;
        cigar2pos_dev(a.sig, a.num, a.strand != 0, rl, q0, q1, l);
        const int64_t lo = a.pos1 - 1, hi = a.pos1 + l - 2;
        uint32_t j = k;
        while (j > 0 && (qs[b + j - 1] > q0 || (qs[b + j - 1] == q0 && qe[b + j - 1] > q1))) {
            const uint32_t d = b + j, s = b + j - 1;
            qs[d] = qs[s]; qe[d] = qe[s]; al[d] = al[s]; rs[d] = rs[s]; re[d] = re[s]; strand[d] = strand[s];
            mapq[d] = mapq[s]; nm[d] = nm[s]; src[d] = src[s];
            j--;
        }
        const uint32_t d = b + j;
        qs[d] = q0; qe[d] = q1; al[d] = l; rs[d] = a.strand ? hi : lo; re[d] = a.strand ? lo : hi; strand[d] = a.strand;
        mapq[d] = a.mapq; nm[d] = a.nm; src[d] = i - b;
        k++;
    }
}
}  // namespace coral

extern "C" int coral_sa_text(coral_ctx* c, const char* text, uint64_t text_len, int64_t n_reads, const uint32_t* ent_ptr,
                             const uint32_t* ent_beg, const uint32_t* ent_end, const int64_t* read_len, int32_t* status,
                             int64_t* qs, int64_t* qe, int64_t* al, int64_t* rstart, int64_t* rend, uint8_t* strand,
                             int32_t* mapq, int64_t* nm, uint32_t* src) {
    using namespace coral;
    if (!c || n_reads < 0 || !ent_ptr) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    if (n_reads == 0) return CORAL_OK;
    const uint32_t n_ent = ent_ptr[n_reads];
    for (uint32_t i = 0; i < n_ent; i++)
        if (ent_beg[i] > ent_end[i] || ent_end[i] > text_len) { c->err = "entry range outside text"; return CORAL_E_ARG; }
    const size_t ne = std::max<uint32_t>(n_ent, 1);
    CORAL_CUDA(c, c->s0.ensure(text_len + 64));
    CORAL_CUDA(c, c->s1.ensure(ne * 12 + 16));
    CORAL_CUDA(c, c->s2.ensure(ne * sizeof(SaEntry)));
    CORAL_CUDA(c, c->s3.ensure((size_t)(n_reads + 1) * 4 + (size_t)n_reads * 12 + 32));
    CORAL_CUDA(c, c->s4.ensure(ne * 64));
    CORAL_TRY(h2d(c, c->s0.p, text, (size_t)text_len));
    uint32_t* e_rec = c->s1.as<uint32_t>(); uint32_t* e_beg = e_rec + ne; uint32_t* e_end = e_beg + ne;
    CORAL_CUDA(c, cudaMemsetAsync(e_rec, 0, ne * 4, c->stream));
    CORAL_TRY(h2d(c, e_beg, ent_beg, (size_t)n_ent));
    CORAL_TRY(h2d(c, e_end, ent_end, (size_t)n_ent));
    uint8_t* p3 = c->s3.as<uint8_t>();
    int64_t* d_rl = reinterpret_cast<int64_t*>(p3);
    uint64_t* d_zero = reinterpret_cast<uint64_t*>(p3 + (size_t)n_reads * 8);
    uint32_t* d_ptr = reinterpret_cast<uint32_t*>(p3 + (size_t)n_reads * 8 + 8);
    int32_t* d_status = reinterpret_cast<int32_t*>(d_ptr + n_reads + 1);
    CORAL_TRY(h2d(c, d_rl, read_len, (size_t)n_reads));
    CORAL_CUDA(c, cudaMemsetAsync(d_zero, 0, 8, c->stream));
    CORAL_TRY(h2d(c, d_ptr, ent_ptr, (size_t)n_reads + 1));
    SaEntry* ent = c->s2.as<SaEntry>();
    LAUNCH_1D(k_parse_entries, (int64_t)n_ent, c->s0.as<uint8_t>(), d_zero, (int64_t)n_ent, e_rec, e_beg, e_end,
              (const uint64_t*)nullptr, 0, ent);
    int64_t* o8 = c->s4.as<int64_t>();
    int64_t *d_qs = o8, *d_qe = o8 + ne, *d_al = o8 + 2 * ne, *d_rs = o8 + 3 * ne, *d_re = o8 + 4 * ne, *d_nm = o8 + 5 * ne;
    int32_t* d_mq = reinterpret_cast<int32_t*>(o8 + 6 * ne);
    uint32_t* d_src = reinterpret_cast<uint32_t*>(d_mq + ne);
    uint8_t* d_strand = reinterpret_cast<uint8_t*>(d_src + ne);
    LAUNCH_1D(k_sa_text_assemble, n_reads, n_reads, d_ptr, ent, d_rl, d_status, d_qs, d_qe, d_al, d_rs, d_re, d_strand, d_mq,
              d_nm, d_src);
    CORAL_TRY(d2h(c, status, d_status, (size_t)n_reads));
    CORAL_TRY(d2h(c, qs, d_qs, (size_t)n_ent)); CORAL_TRY(d2h(c, qe, d_qe, (size_t)n_ent));
    CORAL_TRY(d2h(c, al, d_al, (size_t)n_ent)); CORAL_TRY(d2h(c, rstart, d_rs, (size_t)n_ent));
    CORAL_TRY(d2h(c, rend, d_re, (size_t)n_ent)); CORAL_TRY(d2h(c, nm, d_nm, (size_t)n_ent));
    CORAL_TRY(d2h(c, mapq, d_mq, (size_t)n_ent)); CORAL_TRY(d2h(c, src, d_src, (size_t)n_ent));
    CORAL_TRY(d2h(c, strand, d_strand, (size_t)n_ent));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}
