// coral_b200/csrc/ingest.cu — device-side BAM ingest: BGZF inflate + record framing on the
// GPU, so that what crosses PCIe is the compressed file and the host does no zlib work.  Replaces, for callers that
// opt in, the host reader of bam_reader.cpp (reference: pysam.AlignmentFile(path, 'rb'), infer_breakpoint_graph.py:65).
//
// STATUS: the decoder core and every per-element step below are __host__ __device__ and run against zlib / the host
// reader on the CPU (tests/test_ingest_cpu.py); the kernels are verified on the B200 against the host reader and the
// golden fixtures (tests/test_gpu_device_ingest.py) and timed in bench.py (e2e_from_file) — DESIGN.md 4.3.  Opt-in
// (placement="device_ingest"); the default path uses the host reader.
//
//   host    member table (18-byte BGZF headers, sequential but trivial) + BAM header (first members, host run of the
//           same decoder) -> references, offset of the first record
//   H2D     the compressed bytes
//   k_inflate_members    one warp per member: lane 0 pulls operations from the DEFLATE stream (inflate.cuh, Huffman
//                        tables in shared memory), all 32 lanes do the copies -> one contiguous decoded stream in HBM
//   framing              a record's start is only known from the previous record's block_size: a serial chase of
//                        n_records dependent loads.  Done in parallel instead:
//     k_record_candidates  every byte position that passes the necessary conditions of a record core (sizes
//                          consistent, refIDs in range, name NUL-terminated) is a candidate (true starts + rare chance hits)
//     sort, k_link         candidate -> index of the candidate at p + 4 + block_size (binary search), or terminal / dead
//     k_mark / k_jump      pointer doubling from the first record: after ceil(log2 n) rounds exactly the candidates on
//                          the chain from the first record are marked
//     the chain must end exactly at the end of the stream, otherwise (a file with records the conditions are too
//     strict for) one thread chases serially — the result is always the exact framing, speculation only buys time
//   k_frame_records      marked records copied to 16-byte aligned slots (what the bulk copies of the scan need)
//   -> coral_load_records(..., CORAL_RECORDS_DEVICE) on the framed buffer.
#include <cub/cub.cuh>

#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/coral_b200.h"
#include "ctx.h"
#include "dev_util.cuh"
#include "inflate.cuh"

namespace coral {
namespace ing {

using Member = ::IngestMember;

CORAL_HD uint32_t ld32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
CORAL_HD uint32_t ld16(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }

// Necessary conditions for position p of the decoded stream to start an alignment record (SAM/BAM spec §4.2).
CORAL_HD bool plausible_record(const uint8_t* s, uint64_t total, uint64_t p, int32_t n_ref) {
    if (p + 36 > total) return false;
    const uint8_t* r = s + p;
    const uint32_t bs = ld32(r);
    if (bs < 32 || p + 4 + (uint64_t)bs > total) return false;
    const int32_t ref = (int32_t)ld32(r + 4), pos = (int32_t)ld32(r + 8);
    if (ref < -1 || ref >= n_ref || pos < -1) return false;
    const uint32_t l_name = r[12], n_cig = ld16(r + 16);
    const int32_t l_seq = (int32_t)ld32(r + 20);
    const int32_t nref = (int32_t)ld32(r + 24), npos = (int32_t)ld32(r + 28);
    if (l_name < 1 || l_seq < 0 || nref < -1 || nref >= n_ref || npos < -1) return false;
    const uint64_t fixed = 32ull + l_name + 4ull * n_cig + ((uint64_t)l_seq + 1) / 2 + (uint64_t)l_seq;
    if (fixed > bs) return false;
    return r[36 + l_name - 1] == 0;
}

constexpr uint32_t LINK_END = 0xffffffffu;     // resolved to index n (terminal) / n + 1 (dead) by link_one

// candidate i -> index of the candidate that starts where record i ends; n = the stream ends there; n + 1 = nothing does
CORAL_HD uint32_t link_one(const uint8_t* s, uint64_t total, const uint64_t* cand, uint32_t n, uint32_t i) {
    const uint64_t nx = cand[i] + 4 + (uint64_t)ld32(s + cand[i]);
    if (nx == total) return n;
    uint32_t lo = i + 1, hi = n;               // candidates are sorted and nx > cand[i]
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (cand[mid] < nx) lo = mid + 1; else hi = mid;
    }
    return (lo < n && cand[lo] == nx) ? lo : n + 1;
}

// 16-byte aligned frame of a record
CORAL_HD uint64_t frame_len(const uint8_t* s, uint64_t p) { return ((uint64_t)ld32(s + p) + 4 + 15) & ~15ull; }


// ---- kernels ---------------------------------------------------------------------------------------------------
constexpr int INF_WARPS = 8;

// One warp per BGZF member.  Lane 0 owns the bit reader and the Huffman decode (a serial dependency chain).  It decodes
// a RUN: up to INF_LITS literals into the warp's shared-memory buffer, ended by the first operation that is not a
// literal (or by the full buffer).  The run is broadcast once; all lanes store the literals (coalesced) and then emit the
// closing operation, so that match copies and stored runs are coalesced too.  (The first version broadcast and
// synchronised per literal: BAM's QUAL bytes barely compress, so most of a member is literals — 9 GB/s of output,
// DRAM writes 2.4 x the output, profiles/r2_ingest.txt.)
// status[0] counts members that failed (bad stream, or a length different from the member's ISIZE).
constexpr uint32_t INF_LITS = 128;
constexpr uint32_t OP_NONE = 99;               // the literal buffer filled up: nothing closes this run
__global__ void __launch_bounds__(INF_WARPS * 32)
k_inflate_members(const uint8_t* __restrict__ comp, const Member* __restrict__ mem, uint32_t n_mem, uint8_t* out,
                  int32_t* __restrict__ status) {
    __shared__ dfl::Tables tabs[INF_WARPS];
    __shared__ uint8_t lits[INF_WARPS][INF_LITS];
    const uint32_t lane = lane_id();
    dfl::Tables& t = tabs[threadIdx.x >> 5];
    uint8_t* lb = lits[threadIdx.x >> 5];
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t m = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; m < n_mem; m += nw) {
        const Member mm = mem[m];
        const uint8_t* src = comp + mm.in_off;
        uint8_t* dst = out + mm.out_off;
        dfl::State st;
        dfl::init(st, src, mm.in_len);
        uint32_t o = 0;
        bool fail = false;
        // every run emits at least one byte or ends the stream, so the loop is bounded by out_len + blocks
        for (uint32_t guard = 0; guard < 2u * 65536u + 64u; guard++) {
            dfl::OpOut op{OP_NONE, 0, 0};
            uint32_t nl = 0;
            if (lane == 0) {
                for (;;) {
                    op = dfl::next_op(st, t);
                    if (op.op != dfl::OP_LIT) break;
                    lb[nl++] = (uint8_t)op.a;
                    if (nl == INF_LITS) { op.op = OP_NONE; break; }
                }
            }
            __syncwarp();              // lane 0's literals in shared memory are visible to the warp
            nl = __shfl_sync(FULL_MASK, nl, 0);
            const uint32_t kind = __shfl_sync(FULL_MASK, op.op, 0);
            const uint32_t a = __shfl_sync(FULL_MASK, op.a, 0), b = __shfl_sync(FULL_MASK, op.b, 0);
            if (nl) {
                if (o + nl > mm.out_len) { fail = true; break; }
                for (uint32_t i = lane; i < nl; i += 32) dst[o + i] = lb[i];
                o += nl;
                __syncwarp();          // the bytes just written are input of the closing match; lb is reused next round
            }
            if (kind == dfl::OP_MATCH) {
                if (b > o || o + a > mm.out_len) { fail = true; break; }
                const uint8_t* from = dst + o - b;
                if (b >= a) { for (uint32_t i = lane; i < a; i += 32) dst[o + i] = from[i]; }
                else        { for (uint32_t i = lane; i < a; i += 32) dst[o + i] = from[i % b]; }   // run: period b
                o += a;
            } else if (kind == dfl::OP_STORED) {
                if (o + b > mm.out_len) { fail = true; break; }
                for (uint32_t i = lane; i < b; i += 32) dst[o + i] = src[a + i];
                o += b;
            } else if (kind != OP_NONE) {
                fail = kind != dfl::OP_END;
                break;
            }
            __syncwarp();              // the bytes just written are input of later matches
        }
        if ((fail || o != mm.out_len) && lane == 0) atomicAdd(&status[0], 1);
        __syncwarp();
    }
}

// Candidate record starts.  A thread looks at four consecutive positions through three aligned words; a position goes
// to the full test only when its first two fields (block_size, refID) already look like a record.
__global__ void k_record_candidates(const uint8_t* __restrict__ s, uint64_t total, uint64_t start, int32_t n_ref,
                                    uint64_t* __restrict__ cand, uint32_t cap, unsigned int* __restrict__ counter) {
    const uint64_t ngroups = (total + 3) / 4;
    for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t p0 = 4 * g;
        if (p0 + 4 <= start) continue;
        const uint32_t* w = reinterpret_cast<const uint32_t*>(s + p0);      // the stream buffer is 16-byte aligned, 16 B of slack
        const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
#pragma unroll
        for (uint32_t k = 0; k < 4; k++) {
            const uint32_t bs = __funnelshift_r(w0, w1, 8 * k), ref = __funnelshift_r(w1, w2, 8 * k);
            if (bs < 32 || ref + 1u > (uint32_t)n_ref) continue;
            const uint64_t p = p0 + k;
            if (p < start || !plausible_record(s, total, p, n_ref)) continue;
            const unsigned int at = atomicAdd(counter, 1u);
            if (at < cap) cand[at] = p;
        }
    }
}

__global__ void k_link(const uint8_t* __restrict__ s, uint64_t total, const uint64_t* __restrict__ cand, uint32_t n,
                       uint32_t* __restrict__ jump, uint8_t* __restrict__ mark, uint64_t start) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        jump[i] = link_one(s, total, cand, n, i);
        mark[i] = cand[i] == start;
    } else if (i < n + 2) {
        jump[i] = i;                   // terminal (n) and dead (n + 1) point at themselves
        mark[i] = 0;
    }
}
// marked nodes mark their 2^k-th successor, then every node's pointer doubles
__global__ void k_mark(uint32_t n2, const uint32_t* __restrict__ jump, uint8_t* __restrict__ mark) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n2 && mark[i]) mark[jump[i]] = 1;
}
__global__ void k_jump(uint32_t n2, const uint32_t* __restrict__ jin, uint32_t* __restrict__ jout) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n2) jout[i] = jin[jin[i]];
}
__global__ void k_frame_sizes(const uint8_t* __restrict__ s, const uint64_t* __restrict__ src, uint32_t n, uint64_t* __restrict__ len) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) len[i] = frame_len(s, src[i]);
    else if (i == n) len[i] = 0;
}
// one thread: the exact chase (fallback).  count only when src == nullptr.  status[1] = 1 on a truncated record.
__global__ void k_chase_serial(const uint8_t* __restrict__ s, uint64_t total, uint64_t start, uint64_t* __restrict__ src,
                               uint32_t cap, unsigned int* __restrict__ counter, int32_t* __restrict__ status) {
    uint64_t o = start;
    uint32_t n = 0;
    while (o + 4 <= total) {
        const uint32_t bs = ld32(s + o);
        if (bs < 32 || o + 4 + (uint64_t)bs > total) { status[1] = 1; break; }
        if (src && n < cap) src[n] = o;
        n++;
        o += 4 + (uint64_t)bs;
    }
    *counter = n;
}
// warp per record: stream -> 16-byte aligned frame, padding zeroed.  Destination words are aligned; the source is not:
// two aligned source words and a funnel shift per destination word.
__global__ void k_frame_records(const uint8_t* __restrict__ s, const uint64_t* __restrict__ src, const uint64_t* __restrict__ off,
                                uint32_t n, uint8_t* __restrict__ dst) {
    const uint32_t lane = lane_id();
    const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n; r += nw) {
        const uint64_t p = src[r], o0 = off[r], o1 = off[r + 1];
        const uint32_t len = 4 + ld32(s + p);
        const uint32_t words = (uint32_t)((o1 - o0) >> 2), sh = 8 * (uint32_t)(p & 3);
        const uint32_t* sw = reinterpret_cast<const uint32_t*>(s + (p & ~3ull));
        uint32_t* dw = reinterpret_cast<uint32_t*>(dst + o0);
        for (uint32_t j = lane; j < words; j += 32) {
            uint32_t v = 0;
            if (4 * j < len) {
                v = __funnelshift_r(__ldg(sw + j), __ldg(sw + j + 1), sh);
                const uint32_t keep = len - 4 * j;               // bytes of this word that belong to the record
                if (keep < 4) v &= (1u << (8 * keep)) - 1u;
            }
            dw[j] = v;
        }
    }
}


// ---- host side -------------------------------------------------------------------------------------------------
// BGZF member table of a whole file (SAM/BAM spec §4.1): gzip header with the BC extra field, CDATA, CRC32, ISIZE.
int member_table(const uint8_t* raw, uint64_t fsz, std::vector<Member>& mem, uint64_t& total, std::string& err) {
    uint64_t p = 0;
    total = 0;
    mem.clear();
    while (p < fsz) {
        if (p + 18 > fsz || raw[p] != 0x1f || raw[p + 1] != 0x8b || raw[p + 2] != 8 || !(raw[p + 3] & 4)) {
            err = "not a BGZF member at offset " + std::to_string(p);
            return CORAL_E_FORMAT;
        }
        const uint32_t xlen = ld16(raw + p + 10);
        uint64_t q = p + 12;
        const uint64_t xend = p + 12 + xlen;
        int64_t bsize = -1;
        while (q + 4 <= xend && xend <= fsz) {
            const uint32_t slen = ld16(raw + q + 2);
            if (raw[q] == 'B' && raw[q + 1] == 'C' && slen == 2 && q + 6 <= xend) bsize = ld16(raw + q + 4);
            q += 4 + slen;
        }
        if (bsize < 0 || p + (uint64_t)bsize + 1 > fsz || p + (uint64_t)bsize + 1 < xend + 8) {
            err = "bad BGZF BC field at offset " + std::to_string(p);
            return CORAL_E_FORMAT;
        }
        const uint64_t mend = p + (uint64_t)bsize + 1;
        const uint32_t isize = ld32(raw + mend - 4);
        if (isize > 65536u) { err = "BGZF member larger than 64 KB at offset " + std::to_string(p); return CORAL_E_FORMAT; }
        mem.push_back(Member{xend, total, (uint32_t)(mend - 8 - xend), isize});
        total += isize;
        p = mend;
    }
    return CORAL_OK;
}

// BAM header from the first members (decoded on the host with the same core) -> references, offset of the first record.
int parse_header(const uint8_t* raw, const std::vector<Member>& mem, uint64_t total, std::vector<std::string>& names,
                 std::vector<int64_t>& lens, uint64_t& first_record, std::string& err) {
    std::vector<uint8_t> head;
    size_t next = 0;
    dfl::Tables* t = new dfl::Tables();
    auto more = [&]() -> bool {            // decode one more member onto `head`
        while (next < mem.size() && mem[next].out_len == 0) next++;
        if (next >= mem.size()) return false;
        const Member& m = mem[next++];
        const size_t at = head.size();
        head.resize(at + m.out_len);
        return dfl::inflate_serial(raw + m.in_off, m.in_len, head.data() + at, m.out_len, *t) == (int64_t)m.out_len;
    };
    auto need = [&](uint64_t n) -> bool {
        while (head.size() < n) if (!more()) return false;
        return true;
    };
    int rc = CORAL_E_FORMAT;
    do {
        if (total < 12 || !need(12) || memcmp(head.data(), "BAM\1", 4) != 0) { err = "not a BAM stream"; break; }
        uint64_t o = 4;
        const uint32_t l_text = ld32(&head[o]); o += 4;
        if (!need(o + l_text + 4)) { err = "truncated header"; break; }
        o += l_text;
        const uint32_t n_ref = ld32(&head[o]); o += 4;
        bool ok = true;
        names.clear(); lens.clear();
        for (uint32_t i = 0; i < n_ref && ok; i++) {
            if (!need(o + 4)) { ok = false; break; }
            const uint32_t ln = ld32(&head[o]); o += 4;
            if (ln == 0 || !need(o + ln + 4)) { ok = false; break; }
            names.emplace_back((const char*)&head[o], ln - 1); o += ln;
            lens.push_back((int32_t)ld32(&head[o])); o += 4;
        }
        if (!ok) { err = "truncated reference table"; break; }
        first_record = o;
        rc = CORAL_OK;
    } while (0);
    delete t;
    return rc;
}

}  // namespace ing
}  // namespace coral

using namespace coral;
using namespace coral::ing;

namespace {
void set_err(char* err, size_t n, const std::string& s) {
    if (err && n) snprintf(err, n, "%s", s.c_str());
}
__global__ void k_gather_names(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off, const uint32_t* __restrict__ rec,
                               int64_t n, uint8_t* __restrict__ out) {
    const uint32_t lane = lane_id();
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n) return;
    const uint8_t* r = recs + off[rec[w]];
    const uint32_t ln = r[12];
    for (uint32_t j = lane; j < 256; j += 32) out[w * 256 + j] = j < ln ? r[36 + j] : 0;
}
}  // namespace

extern "C" {

int coral_ingest_open(coral_ctx* c, const uint8_t* file, uint64_t n_bytes) {
    if (!c || (!file && n_bytes)) return CORAL_E_ARG;
    c->ing_ready = false;
    CORAL_TRY(member_table(file, n_bytes, c->ing_members, c->ing_total, c->err));
    CORAL_TRY(parse_header(file, c->ing_members, c->ing_total, c->ing_ref_names, c->ing_ref_lens, c->ing_first, c->err));
    if (c->ing_members.size() >= 0xffffffffull) { c->err = "too many BGZF members"; return CORAL_E_ARG; }
    c->ing_file = file;
    c->ing_file_bytes = n_bytes;
    c->ing_ready = true;
    return CORAL_OK;
}
int32_t coral_ingest_n_ref(const coral_ctx* c) { return c ? (int32_t)c->ing_ref_names.size() : -1; }
const char* coral_ingest_ref_name(const coral_ctx* c, int32_t tid) {
    return (c && tid >= 0 && (size_t)tid < c->ing_ref_names.size()) ? c->ing_ref_names[tid].c_str() : nullptr;
}
int64_t coral_ingest_ref_len(const coral_ctx* c, int32_t tid) {
    return (c && tid >= 0 && (size_t)tid < c->ing_ref_lens.size()) ? c->ing_ref_lens[tid] : -1;
}
int coral_ingest_used_serial_chase(const coral_ctx* c) { return c ? c->ing_serial_chase : -1; }

int coral_ingest_run(coral_ctx* c, int64_t* n_records, uint64_t* framed_bytes) {
    if (!c) return CORAL_E_ARG;
    if (!c->ing_ready) { c->err = "coral_ingest_run before coral_ingest_open"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    if (c->scan_pending) CORAL_TRY(coral_wait_scan(c));       // frees below must not race an asynchronous scan
    cudaStream_t st = c->stream;
    const uint64_t total = c->ing_total, first = c->ing_first;
    const uint32_t n_mem = (uint32_t)c->ing_members.size();
    c->ing_serial_chase = 0;
    // compressed bytes and member table -> device; decoded stream with 64 zero bytes behind it
    CORAL_CUDA(c, c->ing_comp.ensure(c->ing_file_bytes + 16));
    CORAL_CUDA(c, cudaMemcpyAsync(c->ing_comp.p, c->ing_file, c->ing_file_bytes, cudaMemcpyHostToDevice, st));
    CORAL_CUDA(c, c->ing_stream.ensure(total + 64));
    CORAL_CUDA(c, cudaMemsetAsync(c->ing_stream.as<uint8_t>() + total, 0, 64, st));
    CORAL_CUDA(c, c->ing_w[0].ensure((size_t)std::max<uint32_t>(n_mem, 1) * sizeof(Member)));
    if (n_mem) CORAL_CUDA(c, cudaMemcpyAsync(c->ing_w[0].p, c->ing_members.data(), (size_t)n_mem * sizeof(Member), cudaMemcpyHostToDevice, st));
    CORAL_CUDA(c, c->status.ensure(64));
    CORAL_CUDA(c, cudaMemsetAsync(c->status.p, 0, 64, st));
    int32_t* d_status = c->status.as<int32_t>();
    unsigned int* d_count = reinterpret_cast<unsigned int*>(d_status + 4);
    const uint8_t* d_stream = c->ing_stream.as<uint8_t>();
    if (n_mem) {
        const int blocks = (int)std::min<int64_t>(((int64_t)n_mem + INF_WARPS - 1) / INF_WARPS, (int64_t)c->sm_count * 8);
        k_inflate_members<<<blocks, INF_WARPS * 32, 0, st>>>(c->ing_comp.as<uint8_t>(), c->ing_w[0].as<Member>(), n_mem,
                                                            c->ing_stream.as<uint8_t>(), d_status);
        c->n_launches++;
        CORAL_CUDA(c, cudaGetLastError());
    }
    int32_t hs[8] = {0};
    CORAL_CUDA(c, cudaMemcpyAsync(hs, d_status, 32, cudaMemcpyDeviceToHost, st));
    CORAL_CUDA(c, cudaStreamSynchronize(st));
    if (hs[0] != 0) { c->err = "inflate failed in " + std::to_string(hs[0]) + " BGZF member(s)"; return CORAL_E_FORMAT; }

    // ---- framing ----
    uint32_t n = 0;
    uint64_t* d_src = nullptr;
    bool serial = false;
    if (first < total) {
        const uint64_t cap64 = total / 48 + 4096;
        if (cap64 >= 0xfffffff0ull) { c->err = "decoded stream too large for one ctx"; return CORAL_E_ARG; }
        const uint32_t cap = (uint32_t)cap64;
        CORAL_CUDA(c, c->ing_w[1].ensure((size_t)cap * 8));
        CORAL_CUDA(c, c->ing_w[2].ensure((size_t)cap * 8));
        k_record_candidates<<<c->sm_count * 16, 256, 0, st>>>(d_stream, total, first, (int32_t)c->ing_ref_names.size(),
                                                             c->ing_w[1].as<uint64_t>(), cap, d_count);
        c->n_launches++;
        CORAL_CUDA(c, cudaGetLastError());
        unsigned int n_cand = 0;
        CORAL_CUDA(c, cudaMemcpyAsync(&n_cand, d_count, 4, cudaMemcpyDeviceToHost, st));
        CORAL_CUDA(c, cudaStreamSynchronize(st));
        serial = n_cand == 0 || n_cand > cap;
        if (!serial) {
            size_t tmp = 0;
            int bits = 1;
            while (bits < 64 && (total >> bits)) bits++;
            CORAL_CUDA(c, cub::DeviceRadixSort::SortKeys(nullptr, tmp, c->ing_w[1].as<uint64_t>(), c->ing_w[2].as<uint64_t>(), (int)n_cand, 0, bits, st));
            CORAL_CUDA(c, c->cub_tmp.ensure(tmp));
            CORAL_CUDA(c, cub::DeviceRadixSort::SortKeys(c->cub_tmp.p, tmp, c->ing_w[1].as<uint64_t>(), c->ing_w[2].as<uint64_t>(), (int)n_cand, 0, bits, st));
            const uint64_t* d_cand = c->ing_w[2].as<uint64_t>();
            const uint32_t n2 = n_cand + 2;
            CORAL_CUDA(c, c->ing_w[3].ensure((size_t)n2 * 4));
            CORAL_CUDA(c, c->ing_w[4].ensure((size_t)n2 * 4));
            CORAL_CUDA(c, c->ing_w[5].ensure((size_t)n2));
            uint32_t* ja = c->ing_w[3].as<uint32_t>();
            uint32_t* jb = c->ing_w[4].as<uint32_t>();
            uint8_t* mark = c->ing_w[5].as<uint8_t>();
            const unsigned gb = (n2 + 255) / 256;
            k_link<<<gb, 256, 0, st>>>(d_stream, total, d_cand, n_cand, ja, mark, first);
            int rounds = 1;
            while ((1ull << rounds) < n2) rounds++;
            for (int r = 0; r <= rounds; r++) {
                k_mark<<<gb, 256, 0, st>>>(n2, ja, mark);
                k_jump<<<gb, 256, 0, st>>>(n2, ja, jb);
                std::swap(ja, jb);
            }
            c->n_launches += 1 + 2 * (rounds + 1);
            CORAL_CUDA(c, cudaGetLastError());
            uint8_t ends[2] = {0, 0};
            CORAL_CUDA(c, cudaMemcpyAsync(ends, mark + n_cand, 2, cudaMemcpyDeviceToHost, st));
            CORAL_CUDA(c, cudaStreamSynchronize(st));
            if (ends[0] != 1 || ends[1] != 0) {
                serial = true;             // the chain from the first record does not end at the end of the stream
            } else {
                CORAL_CUDA(c, c->ing_w[6].ensure((size_t)n_cand * 8));
                d_src = c->ing_w[6].as<uint64_t>();
                CORAL_CUDA(c, cudaMemsetAsync(d_count, 0, 4, st));
                CORAL_CUDA(c, cub::DeviceSelect::Flagged(nullptr, tmp, d_cand, mark, d_src, d_count, (int)n_cand, st));
                CORAL_CUDA(c, c->cub_tmp.ensure(tmp));
                CORAL_CUDA(c, cub::DeviceSelect::Flagged(c->cub_tmp.p, tmp, d_cand, mark, d_src, d_count, (int)n_cand, st));
                CORAL_CUDA(c, cudaMemcpyAsync(&n, d_count, 4, cudaMemcpyDeviceToHost, st));
                CORAL_CUDA(c, cudaStreamSynchronize(st));
            }
        }
        if (serial) {                      // exact chase by one thread: count, then positions
            c->ing_serial_chase = 1;
            k_chase_serial<<<1, 1, 0, st>>>(d_stream, total, first, nullptr, 0, d_count, d_status);
            CORAL_CUDA(c, cudaMemcpyAsync(hs, d_status, 32, cudaMemcpyDeviceToHost, st));
            CORAL_CUDA(c, cudaStreamSynchronize(st));
            if (hs[1] != 0) { c->err = "truncated alignment record"; return CORAL_E_FORMAT; }
            n = (uint32_t)hs[4];
            CORAL_CUDA(c, c->ing_w[6].ensure((size_t)std::max<uint32_t>(n, 1) * 8));
            d_src = c->ing_w[6].as<uint64_t>();
            k_chase_serial<<<1, 1, 0, st>>>(d_stream, total, first, d_src, n, d_count, d_status);
            c->n_launches += 2;
            CORAL_CUDA(c, cudaGetLastError());
        }
    }
    // frames: sizes -> offsets -> copy
    CORAL_CUDA(c, c->ing_w[7].ensure((size_t)(n + 1) * 8));
    CORAL_CUDA(c, c->ing_w[1].ensure((size_t)(n + 1) * 8));
    uint64_t* d_len = c->ing_w[7].as<uint64_t>();
    uint64_t* d_off = c->ing_w[1].as<uint64_t>();
    k_frame_sizes<<<(n + 1 + 255) / 256, 256, 0, st>>>(d_stream, d_src, n, d_len);
    {
        size_t tmp = 0;
        CORAL_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, tmp, d_len, d_off, (int)(n + 1), st));
        CORAL_CUDA(c, c->cub_tmp.ensure(tmp));
        CORAL_CUDA(c, cub::DeviceScan::ExclusiveSum(c->cub_tmp.p, tmp, d_len, d_off, (int)(n + 1), st));
    }
    c->ing_offsets.assign((size_t)n + 1, 0);
    CORAL_CUDA(c, cudaMemcpyAsync(c->ing_offsets.data(), d_off, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, st));
    CORAL_CUDA(c, cudaStreamSynchronize(st));
    const uint64_t nbytes = c->ing_offsets[n];
    CORAL_CUDA(c, c->ing_recs.ensure(nbytes + 64));
    if (n) {
        const int blocks = (int)std::min<int64_t>(((int64_t)n + 7) / 8, (int64_t)c->sm_count * 8);
        k_frame_records<<<blocks, 256, 0, st>>>(d_stream, d_src, d_off, n, c->ing_recs.as<uint8_t>());
    }
    c->n_launches += 2;
    CORAL_CUDA(c, cudaGetLastError());
    CORAL_CUDA(c, cudaMemsetAsync(c->ing_recs.as<uint8_t>() + nbytes, 0, 64, st));
    CORAL_CUDA(c, cudaStreamSynchronize(st));
    c->ing_comp.release();                 // the compressed bytes and the unframed stream are not needed any more
    c->ing_stream.release();
    c->ing_ready = false;
    c->ing_file = nullptr;
    if (n_records) *n_records = n;
    if (framed_bytes) *framed_bytes = nbytes;
    return coral_load_records(c, c->ing_recs.as<uint8_t>(), nbytes, c->ing_offsets.data(), n, nullptr, CORAL_RECORDS_DEVICE);
}

int coral_get_read_names(coral_ctx* c, const uint32_t* rec, int64_t n, uint8_t* out) {
    if (!c || n < 0 || (n > 0 && (!rec || !out))) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_get_read_names before coral_load_records"; return CORAL_E_STATE; }
    if (n == 0) return CORAL_OK;
    for (int64_t i = 0; i < n; i++)
        if ((int64_t)rec[i] >= c->nrec) { c->err = "record index out of range"; return CORAL_E_ARG; }
    cudaSetDevice(c->device);
    CORAL_CUDA(c, c->ing_w[3].ensure((size_t)n * 4));
    CORAL_CUDA(c, c->ing_w[4].ensure((size_t)n * 256));
    CORAL_CUDA(c, cudaMemcpyAsync(c->ing_w[3].p, rec, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    k_gather_names<<<(unsigned)((n * 32 + 255) / 256), 256, 0, c->stream>>>(c->d_recs, c->d_off.as<uint64_t>(), c->ing_w[3].as<uint32_t>(),
                                                                            n, c->ing_w[4].as<uint8_t>());
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    CORAL_CUDA(c, cudaMemcpyAsync(out, c->ing_w[4].p, (size_t)n * 256, cudaMemcpyDeviceToHost, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}

/* CPU run of the same pieces (test hook, no device needed): every member through dfl::inflate_serial, every position
 * through plausible_record, link_one, the mark / jump rounds as plain loops, frames.  `mode` 0 = as the device does
 * (speculative framing, serial chase as fallback), 1 = force the serial chase.  Outputs: framed records (16-byte
 * aligned starts), offsets[n + 1], counts.  -> CORAL_OK / CORAL_E_FORMAT / CORAL_E_NOMEM (caller buffers too small). */
int coral_ingest_host_check(const uint8_t* file, uint64_t n_bytes, int mode, uint8_t* recs_out, uint64_t recs_cap,
                            uint64_t* offsets_out, int64_t rec_cap, int64_t* n_records, uint64_t* framed_bytes,
                            int64_t* n_candidates, int32_t* used_serial, char* err, size_t err_len) {
    std::vector<Member> mem;
    uint64_t total = 0, first = 0;
    std::string e;
    std::vector<std::string> names;
    std::vector<int64_t> lens;
    int rc = member_table(file, n_bytes, mem, total, e);
    if (rc == CORAL_OK) rc = parse_header(file, mem, total, names, lens, first, e);
    if (rc != CORAL_OK) { set_err(err, err_len, e); return rc; }
    std::vector<uint8_t> stream(total + 64, 0);
    {
        dfl::Tables* t = new dfl::Tables();
        int bad = 0;
        for (const Member& m : mem)
            if (dfl::inflate_serial(file + m.in_off, m.in_len, stream.data() + m.out_off, m.out_len, *t) != (int64_t)m.out_len) bad++;
        delete t;
        if (bad) { set_err(err, err_len, "inflate failed in " + std::to_string(bad) + " BGZF member(s)"); return CORAL_E_FORMAT; }
    }
    const uint8_t* s = stream.data();
    std::vector<uint64_t> src;
    bool serial = mode == 1;
    int64_t ncand = 0;
    if (!serial && first < total) {
        std::vector<uint64_t> cand;
        for (uint64_t p = first; p < total; p++)
            if (plausible_record(s, total, p, (int32_t)names.size())) cand.push_back(p);
        ncand = (int64_t)cand.size();
        const uint32_t n = (uint32_t)cand.size(), n2 = n + 2;
        if (n == 0) {
            serial = true;
        } else {
            std::vector<uint32_t> ja(n2), jb(n2);
            std::vector<uint8_t> mark(n2, 0);
            for (uint32_t i = 0; i < n; i++) { ja[i] = link_one(s, total, cand.data(), n, i); mark[i] = cand[i] == first; }
            ja[n] = n; ja[n + 1] = n + 1;
            int rounds = 1;
            while ((1ull << rounds) < n2) rounds++;
            for (int r = 0; r <= rounds; r++) {
                for (uint32_t i = 0; i < n2; i++) if (mark[i]) mark[ja[i]] = 1;
                for (uint32_t i = 0; i < n2; i++) jb[i] = ja[ja[i]];
                ja.swap(jb);
            }
            if (mark[n] != 1 || mark[n + 1] != 0) serial = true;
            else for (uint32_t i = 0; i < n; i++) if (mark[i]) src.push_back(cand[i]);
        }
    }
    if (serial) {
        src.clear();
        uint64_t o = first;
        while (o + 4 <= total) {
            const uint32_t bs = ld32(s + o);
            if (bs < 32 || o + 4 + (uint64_t)bs > total) { set_err(err, err_len, "truncated alignment record"); return CORAL_E_FORMAT; }
            src.push_back(o);
            o += 4 + (uint64_t)bs;
        }
    }
    if (used_serial) *used_serial = serial ? 1 : 0;
    if (n_candidates) *n_candidates = ncand;
    const int64_t n = (int64_t)src.size();
    if (n_records) *n_records = n;
    uint64_t dst = 0;
    for (int64_t i = 0; i < n; i++) dst += frame_len(s, src[i]);
    if (framed_bytes) *framed_bytes = dst;
    if (n > rec_cap || dst > recs_cap) { set_err(err, err_len, "output buffers too small"); return CORAL_E_NOMEM; }
    dst = 0;
    for (int64_t i = 0; i < n; i++) {
        const uint64_t len = 4 + (uint64_t)ld32(s + src[i]), fl = frame_len(s, src[i]);
        offsets_out[i] = dst;
        memcpy(recs_out + dst, s + src[i], len);
        memset(recs_out + dst + len, 0, fl - len);
        dst += fl;
    }
    offsets_out[n] = dst;
    return CORAL_OK;
}

}  // extern "C"
