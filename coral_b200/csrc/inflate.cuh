// coral_b200/csrc/inflate.cuh — DEFLATE (RFC 1951) decoder core for BGZF members, host + device.
//
// SURVEY.md §8f rank 3 (device-side BAM ingest): the on-disk step before the path.  The reference opens the BAM
// through pysam / htslib (infer_breakpoint_graph.py:65); a BGZF member is a gzip member of at most 64 KB of payload
// with raw DEFLATE inside (SAM/BAM spec §4.1), members are independent, so they inflate in parallel.
//
// The core is a pull decoder: next_op() consumes input bits and returns ONE operation — a literal byte, a match
// (length, distance), a run of stored bytes, end of stream, or an error — and the caller emits it.  The host caller
// (tests, coral_ingest_host_check) emits serially; the device caller (ingest.cu) runs next_op() on lane 0 of a warp
// and lets all 32 lanes do the copy.  Everything here compiles for both sides so that the CPU tests exercise the
// exact decoder the kernel runs (against zlib's output, tests/test_ingest_cpu.py).
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define CORAL_HD __host__ __device__ __forceinline__
#else
#define CORAL_HD inline
#endif

namespace coral {
namespace dfl {

constexpr int FB_LIT = 10;       // bits resolved by the literal/length fast table
constexpr int FB_DST = 8;        // ... by the distance fast table
constexpr int MAX_BITS = 15;

// Per-decoder tables (3.3 KB; shared memory on the device, stack on the host).
// fast[]: index = next FB bits of the stream; entry = symbol | (code length << 9); 0 = longer than FB bits / unused.
// cnt[] / sym[]: canonical Huffman description (codes per length, symbols in code order) for the slow path.
struct Tables {
    uint16_t lit_fast[1 << FB_LIT];
    uint16_t dst_fast[1 << FB_DST];
    uint16_t lit_cnt[MAX_BITS + 1], dst_cnt[MAX_BITS + 1];
    uint16_t lit_sym[288], dst_sym[32];
};

enum Op : uint32_t { OP_LIT = 0, OP_MATCH = 1, OP_STORED = 2, OP_END = 3, OP_ERR = 4 };
enum Mode : uint32_t { M_HEADER = 0, M_STORED = 1, M_CODES = 2, M_DONE = 3 };

struct State {
    const uint8_t* in;
    uint32_t in_len;
    uint32_t pos;          // next input byte to load into acc
    uint64_t acc;          // bit accumulator, LSB first
    uint32_t nbits;
    uint32_t mode;
    uint32_t last;         // BFINAL of the current block
    uint32_t stored_left;
};

struct OpOut { uint32_t op, a, b; };   // LIT: a = byte; MATCH: a = length, b = distance; STORED: a = input offset, b = count

CORAL_HD void init(State& s, const uint8_t* in, uint32_t in_len) {
    s.in = in; s.in_len = in_len; s.pos = 0; s.acc = 0; s.nbits = 0; s.mode = M_HEADER; s.last = 0; s.stored_left = 0;
}

CORAL_HD void refill(State& s) {           // at least 57 bits afterwards; bytes past the end read as zero
    while (s.nbits <= 56) {
        const uint64_t b = s.pos < s.in_len ? s.in[s.pos] : 0u;
        s.pos++;
        s.acc |= b << s.nbits;
        s.nbits += 8;
    }
}
CORAL_HD uint32_t take(State& s, uint32_t n) {   // n <= 32, caller refilled
    const uint32_t v = (uint32_t)(s.acc & ((1ull << n) - 1ull));
    s.acc >>= n;
    s.nbits -= n;
    return v;
}
// input bytes really consumed so far (whole bytes still in the accumulator are given back)
CORAL_HD uint32_t consumed(const State& s) { return s.pos - (s.nbits >> 3); }
CORAL_HD bool overrun(const State& s) { return consumed(s) > s.in_len; }

CORAL_HD uint32_t bitrev(uint32_t v, int n) {
    uint32_t r = 0;
    for (int i = 0; i < n; i++) { r = (r << 1) | (v & 1u); v >>= 1; }
    return r;
}

// Canonical code from code lengths.  -> 0 complete, 1 incomplete (caller decides), -1 over-subscribed.
CORAL_HD int build(const uint8_t* lens, int n, uint16_t* fast, int fb, uint16_t* cnt, uint16_t* sym) {
    for (int i = 0; i <= MAX_BITS; i++) cnt[i] = 0;
    for (int i = 0; i < n; i++) cnt[lens[i]]++;
    for (int i = 0; i < (1 << fb); i++) fast[i] = 0;
    if (cnt[0] == n) return 0;                           // no codes at all: complete by convention, decoding fails later
    int left = 1;
    for (int len = 1; len <= MAX_BITS; len++) {
        left <<= 1;
        left -= cnt[len];
        if (left < 0) return -1;
    }
    uint16_t offs[MAX_BITS + 2];
    offs[1] = 0;
    for (int len = 1; len <= MAX_BITS; len++) offs[len + 1] = (uint16_t)(offs[len] + cnt[len]);
    for (int i = 0; i < n; i++)
        if (lens[i]) sym[offs[lens[i]]++] = (uint16_t)i;
    // fast table: canonical codes in increasing (length, symbol) order
    uint32_t code = 0;
    int idx = 0;
    for (int len = 1; len <= fb; len++) {
        for (int k = 0; k < cnt[len]; k++, idx++, code++) {
            const uint32_t rev = bitrev(code, len);
            const uint16_t e = (uint16_t)(sym[idx] | (len << 9));
            for (uint32_t j = rev; j < (1u << fb); j += (1u << len)) fast[j] = e;
        }
        code <<= 1;
    }
    return left > 0 ? 1 : 0;
}

// One symbol.  -> symbol, or -1 when the bits match no code.
CORAL_HD int decode(State& s, const uint16_t* fast, int fb, const uint16_t* cnt, const uint16_t* sym) {
    const uint16_t e = fast[s.acc & ((1u << fb) - 1u)];
    if (e) {
        const uint32_t len = e >> 9;
        s.acc >>= len;
        s.nbits -= len;
        return e & 511;
    }
    int code = 0, first = 0, index = 0;
    uint64_t a = s.acc;
    for (int len = 1; len <= MAX_BITS; len++) {
        code |= (int)(a & 1u);
        a >>= 1;
        const int count = cnt[len];
        if (code - count < first) {
            s.acc >>= len;
            s.nbits -= len;
            return sym[index + (code - first)];
        }
        index += count;
        first += count;
        first <<= 1;
        code <<= 1;
    }
    return -1;
}

CORAL_HD bool fixed_tables(Tables& t) {
    uint8_t lens[288];
    for (int i = 0; i < 144; i++) lens[i] = 8;
    for (int i = 144; i < 256; i++) lens[i] = 9;
    for (int i = 256; i < 280; i++) lens[i] = 7;
    for (int i = 280; i < 288; i++) lens[i] = 8;
    build(lens, 288, t.lit_fast, FB_LIT, t.lit_cnt, t.lit_sym);
    for (int i = 0; i < 30; i++) lens[i] = 5;
    build(lens, 30, t.dst_fast, FB_DST, t.dst_cnt, t.dst_sym);
    return true;
}

// Dynamic block header (RFC 1951 §3.2.7) -> tables.  The code-length code borrows the distance-table storage.
CORAL_HD bool dynamic_tables(State& s, Tables& t) {
    const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    refill(s);
    const int nlen = (int)take(s, 5) + 257, ndist = (int)take(s, 5) + 1, ncode = (int)take(s, 4) + 4;
    if (nlen > 286 || ndist > 30) return false;
    uint8_t lens[320];
    for (int i = 0; i < 19; i++) lens[i] = 0;
    for (int i = 0; i < ncode; i++) {
        if ((i & 7) == 0) refill(s);
        lens[order[i]] = (uint8_t)take(s, 3);
    }
    if (build(lens, 19, t.dst_fast, 7, t.dst_cnt, t.dst_sym) != 0) return false;   // must be complete
    int i = 0;
    while (i < nlen + ndist) {
        refill(s);
        const int sy = decode(s, t.dst_fast, 7, t.dst_cnt, t.dst_sym);
        if (sy < 0) return false;
        if (sy < 16) { lens[i++] = (uint8_t)sy; continue; }
        int rep, val = 0;
        if (sy == 16) {
            if (i == 0) return false;
            val = lens[i - 1];
            rep = 3 + (int)take(s, 2);
        } else if (sy == 17) {
            rep = 3 + (int)take(s, 3);
        } else {
            rep = 11 + (int)take(s, 7);
        }
        if (i + rep > nlen + ndist) return false;
        while (rep--) lens[i++] = (uint8_t)val;
    }
    if (overrun(s)) return false;
    if (lens[256] == 0) return false;                    // no end-of-block code
    uint8_t dl[32];
    for (int k = 0; k < ndist; k++) dl[k] = lens[nlen + k];
    int r = build(lens, nlen, t.lit_fast, FB_LIT, t.lit_cnt, t.lit_sym);
    if (r < 0 || (r > 0 && nlen != t.lit_cnt[0] + t.lit_cnt[1])) return false;     // incomplete only with a single code
    r = build(dl, ndist, t.dst_fast, FB_DST, t.dst_cnt, t.dst_sym);
    if (r < 0 || (r > 0 && ndist != t.dst_cnt[0] + t.dst_cnt[1])) return false;
    return true;
}

// The next operation of the stream.
CORAL_HD OpOut next_op(State& s, Tables& t) {
    const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    const uint8_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073,
                                4097, 6145, 8193, 12289, 16385, 24577};
    const uint8_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    for (;;) {
        if (s.mode == M_DONE) return OpOut{OP_END, 0, 0};
        if (s.mode == M_HEADER) {
            refill(s);
            s.last = take(s, 1);
            const uint32_t type = take(s, 2);
            if (type == 0) {
                take(s, s.nbits & 7u);                  // to the byte boundary
                refill(s);
                const uint32_t len = take(s, 16), nlen = take(s, 16);
                if ((len ^ 0xffffu) != nlen) return OpOut{OP_ERR, 1, 0};
                const uint32_t at = consumed(s);        // nbits is a multiple of 8 here
                if (at + len > s.in_len) return OpOut{OP_ERR, 2, 0};
                s.pos = at + len; s.acc = 0; s.nbits = 0;
                s.mode = s.last ? M_DONE : M_HEADER;
                if (len) return OpOut{OP_STORED, at, len};
                continue;
            }
            if (type == 1) fixed_tables(t);
            else if (type == 2) { if (!dynamic_tables(s, t)) return OpOut{OP_ERR, 3, 0}; }
            else return OpOut{OP_ERR, 4, 0};
            s.mode = M_CODES;
        }
        // M_CODES
        refill(s);                                      // 57+ bits: code 15 + extra 5 + code 15 + extra 13 = 48
        const int sy = decode(s, t.lit_fast, FB_LIT, t.lit_cnt, t.lit_sym);
        if (sy < 0) return OpOut{OP_ERR, 5, 0};
        if (sy < 256) {
            if (overrun(s)) return OpOut{OP_ERR, 6, 0};
            return OpOut{OP_LIT, (uint32_t)sy, 0};
        }
        if (sy == 256) {
            if (overrun(s)) return OpOut{OP_ERR, 6, 0};
            s.mode = s.last ? M_DONE : M_HEADER;
            continue;
        }
        const int li = sy - 257;
        if (li >= 29) return OpOut{OP_ERR, 7, 0};
        const uint32_t len = lbase[li] + take(s, lext[li]);
        const int ds = decode(s, t.dst_fast, FB_DST, t.dst_cnt, t.dst_sym);
        if (ds < 0 || ds >= 30) return OpOut{OP_ERR, 8, 0};
        const uint32_t dist = dbase[ds] + take(s, dext[ds]);
        if (overrun(s)) return OpOut{OP_ERR, 6, 0};
        return OpOut{OP_MATCH, len, dist};
    }
}

// Serial driver (host tests; also usable by one device thread).  -> bytes written, or -1.
CORAL_HD int64_t inflate_serial(const uint8_t* in, uint32_t in_len, uint8_t* out, uint32_t out_cap, Tables& t) {
    State s;
    init(s, in, in_len);
    uint32_t o = 0;
    for (;;) {
        const OpOut op = next_op(s, t);
        if (op.op == OP_LIT) {
            if (o >= out_cap) return -1;
            out[o++] = (uint8_t)op.a;
        } else if (op.op == OP_MATCH) {
            if (op.b > o || o + op.a > out_cap) return -1;
            for (uint32_t i = 0; i < op.a; i++) out[o + i] = out[o - op.b + i];
            o += op.a;
        } else if (op.op == OP_STORED) {
            if (o + op.b > out_cap) return -1;
            for (uint32_t i = 0; i < op.b; i++) out[o + i] = in[op.a + i];
            o += op.b;
        } else if (op.op == OP_END) {
            return o;
        } else {
            return -1;
        }
    }
}

}  // namespace dfl
}  // namespace coral
