// sim/simgen.cpp — synthetic long-read alignment-record generator (test/bench data only).
//
// Produces *decompressed BAM alignment records* (SAM/BAM spec layout) for
// reads whose geometry (pieces on the reference) is planned by sim/generate.py.  No aligner and
// no htslib exist in this image, so the records a minimap2/winnowmap run would have produced
// (scripts/align_nanopore_reads.sh:34-36 in the reference) are synthesised directly:
//   * own CIGAR:   [clip] (M | I | D)* [clip]   from an ONT/HiFi error model
//   * SA:Z tag:    rname,pos,strand,[xS]yM[zI|zD][wS],mapq,NM;   (minimap2 dialect, App. B.4)
//   * aux decoys:  NM ms AS nn tp cm s1 de [SA] rl [ML:B:C] to exercise every TLV type.
// Build: g++ -O2 -shared -fPIC -pthread sim/simgen.cpp -o sim/_simgen.so
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <cmath>
#include <vector>
#include <string>
#include <thread>
#include <algorithm>

namespace {

struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed + 0x9E3779B97F4A7C15ull) {}
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    double uni() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
    // geometric on {1,2,...} with mean m (m >= 1)
    uint32_t geom(double mean) {
        if (mean <= 1.0) return 1;
        double p = 1.0 / mean;
        double u = uni();
        if (u <= 0) u = 1e-300;
        double k = std::floor(std::log(u) / std::log(1.0 - p));
        if (k > 1e7) k = 1e7;
        return 1u + (uint32_t)k;
    }
};

inline uint64_t mix(uint64_t a, uint64_t b) {
    Rng r(a * 0xD6E8FEB86659FD93ull ^ (b + 0x632BE59BD9B4E019ull));
    return r.next();
}

}  // namespace

extern "C" {

// One planned alignment piece.  Pieces of one read are contiguous and in read (5'->3') order.
struct SimPiece {
    int32_t read_id;     // index into read arrays
    int32_t tid;         // reference id
    int32_t ref_start;   // 0-based leftmost reference coordinate
    int32_t ref_len;     // reference bases covered (exact, geometry-driven)
    uint8_t strand;      // 0 '+', 1 '-'
    uint8_t mapq;
    uint8_t dropped;     // 1 => no record, no SA entry; still occupies ref_len query bases
    uint8_t pad;
    int32_t forced_beg;  // [forced_beg, forced_end) into forced_off/forced_len (large in-CIGAR deletions)
    int32_t forced_end;
};

struct SimParams {
    double p_mis, p_ins, p_del;  // per reference base event rates
    double indel_mean;           // mean indel length (geometric)
    double p_n;                  // probability a query base is 'N'
    double dense_del_boost;      // multiplies p_del (C3 "dense D ops")
    uint64_t seed;
    int32_t hard_clip_supp;      // 1: supplementary records use H clips (minimap2 default)
    int32_t record_align;        // pad record starts to this many bytes (1 = as on disk)
    int32_t n_threads;
    int32_t name_style;          // 0: "r<id>", 1: 36-char uuid-like, 2 + k: "<letter k><id>" (same length as style 0)
};

}  // extern "C"

namespace {

struct Op { uint32_t len; uint8_t op; };  // op: 0 M, 1 I, 2 D

// Deterministic per-piece CIGAR body (no clips).  Reference length is exact.
void gen_ops(const SimPiece& p, int piece_index, const SimParams& prm, const int32_t* f_off,
             const int32_t* f_len, std::vector<Op>& ops, uint32_t& qlen, uint32_t& nm) {
    ops.clear();
    Rng rng(mix(prm.seed, ((uint64_t)(uint32_t)p.read_id << 20) ^ (uint64_t)piece_index));
    const double pdel = prm.p_del * (prm.dense_del_boost > 0 ? prm.dense_del_boost : 1.0);
    const double pev = prm.p_ins + pdel;
    const int64_t R = p.ref_len;
    int64_t r = 0;
    uint64_t q = 0, edits = 0;
    double mis_acc = rng.uni();
    int fi = p.forced_beg;
    auto push = [&](uint8_t op, uint32_t len) {
        if (len == 0) return;
        if (!ops.empty() && ops.back().op == op) ops.back().len += len; else ops.push_back({len, op});
    };
    while (r < R) {
        int64_t run = (pev > 0) ? (int64_t)rng.geom(1.0 / pev) : (R - r);
        int64_t lim = R - r;
        bool hit_forced = false;
        if (fi < p.forced_end) {
            int64_t d = (int64_t)f_off[fi] - r;
            if (d < 1) d = 1;
            if (d <= run && d < lim) { run = d; hit_forced = true; }
        }
        if (run >= lim) { run = lim; hit_forced = false; }
        push(0, (uint32_t)run);
        r += run; q += run;
        mis_acc += run * prm.p_mis;
        if (r >= R) break;
        if (hit_forced) {
            int64_t dl = f_len[fi++];
            if (dl > R - r - 1) dl = R - r - 1;
            if (dl > 0) { push(2, (uint32_t)dl); r += dl; edits += dl; }
            continue;
        }
        bool ins = rng.uni() * pev < prm.p_ins;
        uint32_t len = rng.geom(prm.indel_mean);
        if (ins) { push(1, len); q += len; edits += len; }
        else {
            int64_t dl = len;
            if (dl > R - r - 1) dl = R - r - 1;
            if (dl > 0) { push(2, (uint32_t)dl); r += dl; edits += dl; }
        }
    }
    // CIGAR must end on M: trailing I possible only if loop ended right after M (it cannot), fine.
    edits += (uint64_t)mis_acc;
    qlen = (uint32_t)q;
    nm = (uint32_t)edits;
}

inline int reg2bin(int64_t beg, int64_t end) {
    --end;
    if (beg >> 14 == end >> 14) return (int)(4681 + (beg >> 14));
    if (beg >> 17 == end >> 17) return (int)(585 + (beg >> 17));
    if (beg >> 20 == end >> 20) return (int)(73 + (beg >> 20));
    if (beg >> 23 == end >> 23) return (int)(9 + (beg >> 23));
    if (beg >> 26 == end >> 26) return (int)(1 + (beg >> 26));
    return 0;
}

struct Buf {
    std::vector<uint8_t> v;
    void u8(uint8_t x) { v.push_back(x); }
    void u16(uint16_t x) { v.push_back(x & 255); v.push_back(x >> 8); }
    void u32(uint32_t x) { for (int i = 0; i < 4; i++) v.push_back((x >> (8 * i)) & 255); }
    void bytes(const void* p, size_t n) { const uint8_t* b = (const uint8_t*)p; v.insert(v.end(), b, b + n); }
    void tag(const char* t, char ty) { v.push_back(t[0]); v.push_back(t[1]); v.push_back((uint8_t)ty); }
    void tag_uint(const char* t, uint32_t x) {  // smallest unsigned type, as htslib writes
        if (x < 256) { tag(t, 'C'); u8((uint8_t)x); }
        else if (x < 65536) { tag(t, 'S'); u16((uint16_t)x); }
        else { tag(t, 'I'); u32(x); }
    }
    void tag_sint(const char* t, int32_t x) {
        if (x >= -128 && x < 128) { tag(t, 'c'); u8((uint8_t)(int8_t)x); }
        else if (x >= -32768 && x < 32768) { tag(t, 's'); u16((uint16_t)(int16_t)x); }
        else { tag(t, 'i'); u32((uint32_t)x); }
    }
};

std::string read_name(int32_t id, const SimParams& prm) {
    char b[64];
    if (prm.name_style == 1) {
        uint64_t a = mix(prm.seed ^ 0xABCDEF, (uint64_t)id), c = mix(a, 77);
        snprintf(b, sizeof b, "%08x-%04x-%04x-%04x-%012llx", (uint32_t)(a >> 32), (uint32_t)(a >> 16) & 0xffff,
                 (uint32_t)a & 0xffff, (uint32_t)(c >> 48) & 0xffff, (unsigned long long)(c & 0xffffffffffffull));
    } else if (prm.name_style >= 2) {      // shards of one file generated separately: a distinct letter per shard
        snprintf(b, sizeof b, "%c%d", (char)('a' + (prm.name_style - 2) % 26), id);
    } else {
        snprintf(b, sizeof b, "r%d", id);
    }
    return b;
}

std::string sa_cigar(uint32_t clip5, uint32_t m, char indel, uint32_t il, uint32_t clip3) {
    char b[96]; int n = 0;
    if (clip5) n += snprintf(b + n, sizeof b - n, "%uS", clip5);
    n += snprintf(b + n, sizeof b - n, "%uM", m);
    if (il) n += snprintf(b + n, sizeof b - n, "%u%c", il, indel);
    if (clip3) n += snprintf(b + n, sizeof b - n, "%uS", clip3);
    return std::string(b, n);
}

struct PieceInfo { uint32_t qlen, nm, n_ops; uint32_t qs; };

}  // namespace

extern "C" {

// Pass 1+2: per-piece CIGAR statistics and per-read layout.
//   out_qlen/out_nm/out_nops [n_pieces]; out_qs [n_pieces] = start of the piece in read coordinates;
//   out_read_len [n_reads].
int sim_plan(const SimPiece* pieces, int64_t n_pieces, int64_t n_reads, const SimParams* prm,
             const int32_t* f_off, const int32_t* f_len, uint32_t* out_qlen, uint32_t* out_nm,
             uint32_t* out_nops, uint32_t* out_qs, uint32_t* out_read_len) {
    int nt = std::max(1, prm->n_threads);
    // piece index within its read
    std::vector<int32_t> pidx(n_pieces);
    for (int64_t i = 0; i < n_pieces; i++)
        pidx[i] = (i > 0 && pieces[i - 1].read_id == pieces[i].read_id) ? pidx[i - 1] + 1 : 0;
    auto work = [&](int t) {
        std::vector<Op> ops;
        for (int64_t i = t; i < n_pieces; i += nt) {
            const SimPiece& p = pieces[i];
            if (p.dropped) { out_qlen[i] = p.ref_len; out_nm[i] = 0; out_nops[i] = 0; continue; }
            uint32_t q, nm;
            gen_ops(p, pidx[i], *prm, f_off, f_len, ops, q, nm);
            out_qlen[i] = q; out_nm[i] = nm; out_nops[i] = (uint32_t)ops.size();
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++) th.emplace_back(work, t);
    for (auto& x : th) x.join();
    for (int64_t r = 0; r < n_reads; r++) out_read_len[r] = 0;
    for (int64_t i = 0; i < n_pieces; i++) {
        out_qs[i] = out_read_len[pieces[i].read_id];
        out_read_len[pieces[i].read_id] += out_qlen[i];
    }
    return 0;
}

// Pass 3: serialise records for the pieces listed in `order` (already coordinate-sorted by the caller).
// Call with out == nullptr to obtain sizes: rec_off[n_order+1] is always filled.
// ref_names: n_ref NUL-terminated strings concatenated; ref_name_off[n_ref].
int64_t sim_write(const SimPiece* pieces, int64_t n_pieces, const SimParams* prm, const int32_t* f_off,
                  const int32_t* f_len, const uint32_t* qlen, const uint32_t* nmv, const uint32_t* qs,
                  const uint32_t* read_len, const int64_t* order, int64_t n_order, const char* ref_names,
                  const int32_t* ref_name_off, uint8_t* out, uint64_t* rec_off) {
    int nt = std::max(1, prm->n_threads);
    std::vector<int32_t> pidx(n_pieces);
    std::vector<int64_t> first(n_pieces), last(n_pieces);
    for (int64_t i = 0; i < n_pieces; i++) {
        bool cont = i > 0 && pieces[i - 1].read_id == pieces[i].read_id;
        pidx[i] = cont ? pidx[i - 1] + 1 : 0;
        first[i] = cont ? first[i - 1] : i;
    }
    for (int64_t i = n_pieces - 1; i >= 0; i--) {
        bool cont = i + 1 < n_pieces && pieces[i + 1].read_id == pieces[i].read_id;
        last[i] = cont ? last[i + 1] : i;
    }
    // primary = longest non-dropped piece of the read (first on ties)
    std::vector<int64_t> primary(n_pieces, -1);
    for (int64_t i = 0; i < n_pieces; i++) {
        if (first[i] != i) continue;
        int64_t best = -1;
        for (int64_t j = i; j <= last[i]; j++)
            if (!pieces[j].dropped && (best < 0 || qlen[j] > qlen[best])) best = j;
        for (int64_t j = i; j <= last[i]; j++) primary[j] = best;
    }
    const int align = std::max(1, prm->record_align);
    std::vector<uint64_t> sizes(n_order, 0);
    for (int pass = (out ? 1 : 0); pass < 2; pass++) {
        auto work = [&](int t) {
            std::vector<Op> ops;
            Buf b;
            for (int64_t oi = t; oi < n_order; oi += nt) {
                int64_t i = order[oi];
                const SimPiece& p = pieces[i];
                uint32_t q, nm;
                gen_ops(p, pidx[i], *prm, f_off, f_len, ops, q, nm);
                const uint32_t L = read_len[p.read_id];
                const bool is_primary = primary[i] == i;
                const bool hard = !is_primary && prm->hard_clip_supp;
                uint32_t c_lo = p.strand ? L - (qs[i] + q) : qs[i];  // clip before body in ref-forward order
                uint32_t c_hi = p.strand ? qs[i] : L - (qs[i] + q);
                std::string name = read_name(p.read_id, *prm);
                uint32_t n_cig = (uint32_t)ops.size() + (c_lo ? 1 : 0) + (c_hi ? 1 : 0);
                uint32_t l_seq = hard ? q : L;
                b.v.clear();
                b.u32(0);  // block_size placeholder
                b.u32((uint32_t)p.tid);
                b.u32((uint32_t)p.ref_start);
                b.u8((uint8_t)(name.size() + 1));
                b.u8(p.mapq);
                b.u16((uint16_t)reg2bin(p.ref_start, (int64_t)p.ref_start + p.ref_len));
                b.u16((uint16_t)n_cig);
                b.u16((uint16_t)((p.strand ? 16 : 0) | (is_primary ? 0 : 2048)));
                b.u32(l_seq);
                b.u32(0xFFFFFFFFu); b.u32(0xFFFFFFFFu); b.u32(0);
                b.bytes(name.c_str(), name.size() + 1);
                const uint8_t clip_op = hard ? 5 : 4;
                if (c_lo) b.u32((c_lo << 4) | clip_op);
                for (const Op& o : ops) b.u32((o.len << 4) | o.op);
                if (c_hi) b.u32((c_hi << 4) | clip_op);
                // SEQ: random ACGT nibbles (1,2,4,8) with optional N (15)
                Rng sr(mix(prm->seed ^ 0x5E0, ((uint64_t)(uint32_t)p.read_id << 20) ^ (uint64_t)pidx[i]));
                size_t nb = (l_seq + 1) / 2;
                size_t seq_at = b.v.size();
                b.v.resize(seq_at + nb);
                static const uint8_t code[4] = {1, 2, 4, 8};
                for (size_t k = 0; k < nb; k += 8) {
                    uint64_t w = sr.next();
                    for (size_t j = 0; j < 8 && k + j < nb; j++) {
                        uint8_t hi = code[(w >> (j * 4)) & 3], lo = code[(w >> (j * 4 + 2)) & 3];
                        b.v[seq_at + k + j] = (uint8_t)((hi << 4) | lo);
                    }
                }
                if (prm->p_n > 0) {
                    double expect = prm->p_n * l_seq;
                    uint32_t cnt = (uint32_t)expect + (sr.uni() < expect - std::floor(expect) ? 1 : 0);
                    for (uint32_t k = 0; k < cnt && l_seq; k++) {
                        uint32_t pos = (uint32_t)(sr.next() % l_seq);
                        uint8_t& by = b.v[seq_at + pos / 2];
                        by = (pos & 1) ? (uint8_t)((by & 0xF0) | 15) : (uint8_t)((by & 0x0F) | 0xF0);
                    }
                }
                if (l_seq & 1) b.v[seq_at + nb - 1] &= 0xF0;
                size_t qual_at = b.v.size();
                b.v.resize(qual_at + l_seq, (uint8_t)20);
                // aux
                b.tag_uint("NM", nm);
                b.tag_uint("ms", q);
                b.tag_uint("AS", q > nm ? q - nm : 0);
                b.tag_uint("nn", 0);
                b.tag("tp", 'A'); b.u8('P');
                b.tag_uint("cm", q / 15);
                b.tag_sint("s1", (int32_t)(q / 3) * ((p.read_id % 97 == 0) ? -1 : 1));
                { b.tag("de", 'f'); float de = q ? (float)nm / (float)q : 0.f; uint32_t u; memcpy(&u, &de, 4); b.u32(u); }
                // SA: all other non-dropped pieces of the read, primary first then read order
                std::string sa;
                auto add_sa = [&](int64_t j) {
                    const SimPiece& o = pieces[j];
                    uint32_t oq = qlen[j], oR = (uint32_t)o.ref_len;
                    uint32_t m = std::min(oq, oR);
                    char indel = oq > oR ? 'I' : 'D';
                    uint32_t il = oq > oR ? oq - oR : oR - oq;
                    uint32_t lo = o.strand ? L - (qs[j] + oq) : qs[j];
                    uint32_t hi = o.strand ? qs[j] : L - (qs[j] + oq);
                    char pre[128];
                    int n = snprintf(pre, sizeof pre, "%s,%d,%c,", ref_names + ref_name_off[o.tid], o.ref_start + 1,
                                     o.strand ? '-' : '+');
                    sa.append(pre, n);
                    sa += sa_cigar(lo, m, indel, il, hi);
                    n = snprintf(pre, sizeof pre, ",%u,%u;", (unsigned)o.mapq, nmv[j]);
                    sa.append(pre, n);
                };
                int64_t pr = primary[i];
                if (pr >= 0 && pr != i) add_sa(pr);
                for (int64_t j = first[i]; j <= last[i]; j++)
                    if (j != i && j != pr && !pieces[j].dropped) add_sa(j);
                if (!sa.empty()) { b.tag("SA", 'Z'); b.bytes(sa.c_str(), sa.size() + 1); }
                b.tag_uint("rl", (uint32_t)(p.read_id % 300));
                if (p.read_id % 101 == 0) {  // B-array decoy
                    b.tag("ML", 'B'); b.u8('C'); b.u32(5); for (int k = 0; k < 5; k++) b.u8((uint8_t)(k * 40));
                }
                uint32_t bs = (uint32_t)b.v.size() - 4;
                memcpy(b.v.data(), &bs, 4);
                if (pass == 0) {
                    sizes[oi] = b.v.size();
                } else {
                    memcpy(out + rec_off[oi], b.v.data(), b.v.size());
                    uint64_t padn = rec_off[oi + 1] - rec_off[oi] - b.v.size();
                    if (padn) memset(out + rec_off[oi] + b.v.size(), 0, padn);
                }
            }
        };
        std::vector<std::thread> th;
        for (int t = 0; t < nt; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
        if (pass == 0) {
            uint64_t o = 0;
            for (int64_t k = 0; k < n_order; k++) {
                rec_off[k] = o;
                o += (sizes[k] + align - 1) / align * align;
            }
            rec_off[n_order] = o;
            if (!out) return (int64_t)o;
        }
    }
    return (int64_t)rec_off[n_order];
}

}  // extern "C"
