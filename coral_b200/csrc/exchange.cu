// coral_b200/csrc/exchange.cu — packed messages of the multi-GPU exchanges.
//
// One process per GPU holds a contiguous range of the file.  What crosses ranks is small and is
// packed ON THE DEVICE into one byte message per rank and exchange, so that an exchange is one
// all-gather of bytes (NCCL, issued by the caller) between two kernels:
//
//   fetch (ibg:139-174)        coral_pack_sa_rows     -> all-gather -> coral_pack_extra_rows
//                              (first local primaries of names that are chimeric on ANY rank, + the
//                              nm_stats partial sums) -> all-gather -> coral_unpack_rows into a
//                              rows-only ctx, on which coral_fetch runs unchanged
//   concordant edges           coral_concordant_pack  -> all-gather -> coral_count_name_msgs
//   (ibg:1039-1055)
//
// Message = MsgHdr (64 B) + n_rows fixed-size rows + text bytes, padded to 16 B.  Rows keep file
// order inside a message and messages are consumed in rank order, so the gathered rows are in global
// file order — the order the reference iterates in.
#include <algorithm>
#include <vector>

#include "ctx.h"
#include "msg.cuh"
#include "relops.cuh"

namespace coral {

// ---- pack: flagged records -> rows + text ------------------------------------------------------------------
__global__ void k_flag_has_sa(int64_t n, const uint32_t* __restrict__ sa_off, uint32_t* __restrict__ f) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) f[i] = sa_off[i] != 0 ? 1u : 0u;
}

// Per record: (flagged ? 1 : 0) << 36 | slot bytes (read name with its NUL + SA text) — ONE inclusive scan gives every
// flagged record its row index and the end of its text slot (rows < 2^28, text < 2^36 bytes per message).
constexpr int PACK_SHIFT = 36;
constexpr uint64_t PACK_MASK = ((uint64_t)1 << PACK_SHIFT) - 1;
struct SlotFn {
    const uint32_t* f;
    const uint8_t* recs;
    const uint64_t* off;
    const uint32_t *sa_off, *sa_len;
    __host__ __device__ uint64_t operator()(uint32_t i) const {
        if (!f[i]) return 0;
        return ((uint64_t)1 << PACK_SHIFT) | ((uint64_t)recs[off[i] + 12] + (sa_off[i] ? sa_len[i] : 0u));
    }
};

struct PackIn {
    const uint8_t* recs;
    const uint64_t* off;
    const uint64_t *lo, *hi;
    const uint32_t *flag_mapq, *lseq, *sa_off, *sa_len;
    const int32_t* nm;
};

// one warp per flagged record: lane 0 writes the row, all lanes copy the slot text.  The first thread also writes the
// header (nm sums from d_nm when given: reduced on the device) and zeroes the padding behind the text.
__global__ void k_pack_rows(int64_t n, const uint32_t* __restrict__ f, const uint64_t* __restrict__ incl, PackIn in,
                            MsgHdr* __restrict__ hdr_out, MsgHdr hdr, const double* __restrict__ d_nm,
                            RowMsg* __restrict__ rows, uint8_t* __restrict__ text, uint32_t pad_bytes) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (i == 0 && lane == 0) {
        if (d_nm) { hdr.nm_sum = d_nm[0]; hdr.nm_sq = d_nm[1]; hdr.nm_n = (int64_t)d_nm[2]; hdr.nm_zd = (int64_t)d_nm[3]; }
        *hdr_out = hdr;
        for (uint32_t k = 0; k < pad_bytes; k++) text[hdr.text_bytes + k] = 0;
    }
    if (i >= n || !f[i]) return;
    const uint64_t v = incl[i];
    const uint32_t r = (uint32_t)(v >> PACK_SHIFT) - 1;
    const uint8_t* rec = in.recs + in.off[i];
    const uint32_t name_len = rec[12];
    const uint32_t so = in.sa_off[i];
    const uint32_t sl = so ? in.sa_len[i] : 0u;
    const uint64_t t0 = (v & PACK_MASK) - (name_len + sl);
    if (lane == 0) {
        RowMsg m;
        m.lo = in.lo[i]; m.hi = in.hi[i];
        m.flag_mapq = in.flag_mapq[i]; m.l_seq = in.lseq[i]; m.nm = in.nm[i];
        m.rec = (uint32_t)i; m.name_len = name_len; m.sa_len = sl; m.text_off = t0;
        rows[r] = m;
    }
    for (uint32_t k = lane; k < name_len; k += 32) text[t0 + k] = rec[36 + k];
    for (uint32_t k = lane; k < sl; k += 32) text[t0 + name_len + k] = rec[so + k];
}

// ---- extra primaries: membership of local non-SA primaries in the gathered SA-name set ----------------------
struct MsgSet {               // up to 16 ranks per call chunk is plenty for one node; more ranks loop on the host
    const uint8_t* base[16];
    uint64_t row0[17];        // prefix sums of n_rows
    int n;
};

__device__ __forceinline__ const RowMsg* msg_row(const MsgSet& s, uint64_t g, int* rank_out = nullptr) {
    int r = 0;
    while (r + 1 < s.n && g >= s.row0[r + 1]) r++;
    if (rank_out) *rank_out = r;
    return reinterpret_cast<const RowMsg*>(s.base[r] + sizeof(MsgHdr)) + (g - s.row0[r]);
}

__global__ void k_collect_names(MsgSet s, uint64_t* __restrict__ lo, uint32_t* __restrict__ idx) {
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= s.row0[s.n]) return;
    lo[g] = msg_row(s, g)->lo;
    idx[g] = (uint32_t)g;
}

// position in the lo-sorted gathered names of the FIRST entry equal to (lo, hi), or -1
// (the names are sorted on the low NAME_SORT_BITS bits of the low hash word only — 5 radix passes instead of 8; entries
// that share those bits sit next to each other in arbitrary order and are told apart by the full 128 bits here)
constexpr int NAME_SORT_BITS = 40;
constexpr uint64_t NAME_SORT_MASK = ((uint64_t)1 << NAME_SORT_BITS) - 1;
__device__ __forceinline__ int64_t name_slot(const MsgSet& s, const uint64_t* __restrict__ slo, const uint32_t* __restrict__ sidx,
                                             int64_t m, uint64_t lo, uint64_t hi) {
    const uint64_t key = lo & NAME_SORT_MASK;
    int64_t a = 0, b = m;
    while (a < b) {
        const int64_t mid = (a + b) >> 1;
        if ((slo[mid] & NAME_SORT_MASK) < key) a = mid + 1; else b = mid;
    }
    // the FIRST entry of the run that carries this very name: every prober of a name must pick the same slot
    for (; a < m && (slo[a] & NAME_SORT_MASK) == key; a++)
        if (slo[a] == lo && msg_row(s, sidx[a])->hi == hi) return a;
    return -1;
}

__global__ void k_extra_probe(int64_t n, const uint32_t* __restrict__ flag_mapq, const uint32_t* __restrict__ sa_off,
                              const uint64_t* __restrict__ lo, const uint64_t* __restrict__ hi, MsgSet s,
                              const uint64_t* __restrict__ slo, const uint32_t* __restrict__ sidx, int64_t m,
                              uint32_t* __restrict__ first, int32_t* __restrict__ slot_of) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t p = -1;
    if ((flag_mapq[i] & 0xffffu) < 256u && sa_off[i] == 0) p = name_slot(s, slo, sidx, m, lo[i], hi[i]);
    slot_of[i] = (int32_t)p;
    if (p >= 0) atomicMin(first + p, (uint32_t)i);
}

__global__ void k_extra_flag(int64_t n, const int32_t* __restrict__ slot_of, const uint32_t* __restrict__ first,
                             uint32_t* __restrict__ f) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t p = slot_of[i];
    f[i] = (p >= 0 && first[p] == (uint32_t)i) ? 1u : 0u;
}

}  // namespace coral

namespace coral {

// ---- unpack: gathered messages -> columns of a rows-only ctx --------------------------------------------------
struct UnpackPlan {
    MsgSet sa, ex;
    uint64_t out_base[16];      // first output row of each rank
    uint64_t text_sa[16];       // where each rank's SA-message / extras-message text lands in the ctx text
    uint64_t text_ex[16];
};

struct RowCols {
    uint64_t *off, *lo, *hi;
    uint32_t *flag_mapq, *lseq, *sa_off, *sa_len;
    int32_t* nm;
};

// rows of `other` (same rank) with rec < key
__device__ __forceinline__ uint64_t rows_below(const uint8_t* base, uint64_t n, uint32_t key) {
    const RowMsg* r = reinterpret_cast<const RowMsg*>(base + sizeof(MsgHdr));
    uint64_t a = 0, b = n;
    while (a < b) {
        const uint64_t mid = (a + b) >> 1;
        if (r[mid].rec < key) a = mid + 1; else b = mid;
    }
    return a;
}

__global__ void k_unpack_rows(UnpackPlan p, RowCols o, uint64_t text_total) {
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t n_sa = p.sa.row0[p.sa.n], n_ex = p.ex.row0[p.ex.n];
    if (g == 0) o.off[n_sa + n_ex] = text_total;          // offsets[n]: the end of the text
    if (g >= n_sa + n_ex) return;
    int r;
    const RowMsg* m;
    uint64_t out, toff;
    if (g < n_sa) {
        m = msg_row(p.sa, g, &r);
        out = p.out_base[r] + (g - p.sa.row0[r]) + rows_below(p.ex.base[r], p.ex.row0[r + 1] - p.ex.row0[r], m->rec);
        toff = p.text_sa[r] + m->text_off;
    } else {
        m = msg_row(p.ex, g - n_sa, &r);
        out = p.out_base[r] + (g - n_sa - p.ex.row0[r]) + rows_below(p.sa.base[r], p.sa.row0[r + 1] - p.sa.row0[r], m->rec);
        toff = p.text_ex[r] + m->text_off;
    }
    o.off[out] = toff;
    o.lo[out] = m->lo; o.hi[out] = m->hi;
    o.flag_mapq[out] = m->flag_mapq; o.lseq[out] = m->l_seq; o.nm[out] = m->nm;
    o.sa_off[out] = m->sa_len ? m->name_len : 0u;     // slot = name (with NUL) + SA text; 0 keeps meaning "no SA tag"
    o.sa_len[out] = m->sa_len;
}

// the text of every message -> its place in the rows-only ctx's text buffer (both sides 16-byte aligned): one kernel
// instead of 2 N copy-engine operations between kernels
struct TextCopy {
    const uint4* src[2 * CORAL_XCHG_MAX_RANKS];
    uint4* dst[2 * CORAL_XCHG_MAX_RANKS];
    uint64_t nvec[2 * CORAL_XCHG_MAX_RANKS];
    int n;
};
__global__ void k_copy_text(TextCopy t) {
    const int seg = blockIdx.y;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < t.nvec[seg]; i += stride) t.dst[seg][i] = t.src[seg][i];
}

__global__ void k_collect_hdrs(MsgSet a, MsgSet b, uint64_t* __restrict__ out) {
    const int t = threadIdx.x;
    const int per = (int)(sizeof(MsgHdr) / 8);
    const int k = t / per, w = t % per;
    const uint8_t* base = k < a.n ? a.base[k] : (k - a.n < b.n ? b.base[k - a.n] : nullptr);
    if (base) out[t] = reinterpret_cast<const uint64_t*>(base)[w];
}

}  // namespace coral

// ---- host orchestration ---------------------------------------------------------------------------------------
#define LAUNCH_1D(kern, n, ...)                                                          \
    do {                                                                                 \
        if ((n) > 0) {                                                                   \
            kern<<<(unsigned)(((n) + 255) / 256), 256, 0, c->stream>>>(__VA_ARGS__);      \
            c->n_launches++;                                                             \
            CORAL_CUDA(c, cudaGetLastError());                                           \
        }                                                                                \
    } while (0)

namespace {
using namespace coral;

struct U32ToU64 {
    __host__ __device__ uint64_t operator()(uint32_t v) const { return (uint64_t)v; }
};

int inclusive_sum_u32_to_u64(coral_ctx* c, const uint32_t* in, uint64_t* out, int64_t n) {
    if (n <= 0) return CORAL_OK;
    cub::TransformInputIterator<uint64_t, U32ToU64, const uint32_t*> it(in, U32ToU64());
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(nullptr, bytes, it, out, (int)n, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(c->cub_tmp.p, bytes, it, out, (int)n, c->stream));
    c->n_launches += 1;
    return CORAL_OK;
}

// message of the records flagged in c->s0 (uint32 per record): rows in file order + their text
int pack_flagged(coral_ctx* c, int kind, MsgHdr hdr, DevBuf& msg, uint64_t* d_msg, uint64_t* nbytes, const double* d_nm = nullptr) {
    const int64_t n = c->nrec;
    RecordTable& t = c->rt;
    uint64_t last = 0;
    uint32_t* f = c->s0.as<uint32_t>();
    if (n >= ((int64_t)1 << 28)) { c->err = "more than 2^28 records in one shard's exchange message"; return CORAL_E_ARG; }
    if (n > 0) {
        CORAL_CUDA(c, c->s3.ensure((size_t)n * 8));
        cub::CountingInputIterator<uint32_t> iota(0);
        cub::TransformInputIterator<uint64_t, SlotFn, cub::CountingInputIterator<uint32_t>> it(
            iota, SlotFn{f, c->d_recs, c->d_off.as<uint64_t>(), t.sa_off.as<uint32_t>(), t.sa_len.as<uint32_t>()});
        size_t bytes = 0;
        CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(nullptr, bytes, it, c->s3.as<uint64_t>(), (int)n, c->stream));
        CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
        CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(c->cub_tmp.p, bytes, it, c->s3.as<uint64_t>(), (int)n, c->stream));
        c->n_launches += 1;
        // rows and text bytes shape the message: read back (outside a pass / while recording) or replayed and checked
        CORAL_TRY(pass_resolve(c, &last, c->s3.as<uint64_t>() + (n - 1)));
    }
    const uint64_t n_rows = last >> PACK_SHIFT, text_bytes = last & PACK_MASK;
    const uint64_t body = sizeof(MsgHdr) + n_rows * sizeof(RowMsg);
    const uint64_t total = body + ((text_bytes + 15) & ~(uint64_t)15);
    CORAL_CUDA(c, msg.ensure(total));
    hdr.n_rows = n_rows; hdr.text_bytes = text_bytes; hdr.kind = (uint64_t)kind; hdr.pad = 0;
    PackIn in{c->d_recs, c->d_off.as<uint64_t>(), t.name_lo.as<uint64_t>(), t.name_hi.as<uint64_t>(), t.flag_mapq.as<uint32_t>(),
              t.lseq.as<uint32_t>(), t.sa_off.as<uint32_t>(), t.sa_len.as<uint32_t>(), t.nm.as<int32_t>()};
    const int64_t threads = std::max<int64_t>(n, 1) * 32;      // at least one warp: the header is written by the first thread
    LAUNCH_1D(k_pack_rows, threads, n, f, c->s3.as<uint64_t>(), in, msg.as<MsgHdr>(), hdr, d_nm,
              reinterpret_cast<RowMsg*>(msg.as<uint8_t>() + sizeof(MsgHdr)), msg.as<uint8_t>() + body, (uint32_t)(total - body - text_bytes));
    if (!c->pass.active) CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    if (d_msg) *d_msg = (uint64_t)reinterpret_cast<uintptr_t>(msg.p);
    if (nbytes) *nbytes = total;
    return CORAL_OK;
}

// headers of the gathered messages (device memory) -> host, and the MsgSet the kernels take
int read_msgs(coral_ctx* c, int32_t n_ranks, const uint8_t* d_buf, const uint64_t* starts, int kind, MsgSet* set,
              std::vector<MsgHdr>* hdrs) {
    if (n_ranks < 1 || n_ranks > 16) { c->err = "1..16 ranks per exchange"; return CORAL_E_ARG; }
    set->n = n_ranks;
    for (int r = 0; r < n_ranks; r++) {
        if (starts[r] & 15) { c->err = "message starts must be 16-byte aligned"; return CORAL_E_ARG; }
        set->base[r] = d_buf + starts[r];
    }
    HdrPtrs hp;
    hp.n = n_ranks;
    for (int r = 0; r < n_ranks; r++) hp.base[r] = set->base[r];
    hdrs->resize((size_t)n_ranks);
    static_assert(sizeof(MsgHdr) == 64, "8 words per header");
    CORAL_TRY(pass_resolve_hdrs(c, hp, reinterpret_cast<uint64_t*>(hdrs->data())));
    set->row0[0] = 0;
    for (int r = 0; r < n_ranks; r++) {
        if ((*hdrs)[r].kind != (uint64_t)kind) { c->err = "unexpected message kind in the gathered buffer"; return CORAL_E_FORMAT; }
        set->row0[r + 1] = set->row0[r] + (*hdrs)[r].n_rows;
    }
    return CORAL_OK;
}

}  // namespace

extern "C" {

// Message 1 of the fetch exchange: this shard's SA-bearing records (ibg:145-151), in file order.
int coral_pack_sa_rows(coral_ctx* c, uint64_t* d_msg, uint64_t* nbytes) {
    if (!c) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_pack_sa_rows needs a ctx with records"; return CORAL_E_STATE; }
    // like coral_fetch, the packing kernels only read columns of the pre-scan (core fields, name hash, SA / NM location)
    // and record text: inside a pass whose rescan runs overlapped they start next to the full scan instead of joining it
    if (!(c->pass.active && c->scan_pending)) CORAL_TRY(coral_wait_scan(c));
    cudaSetDevice(c->device);
    coral::PassLane lane__(c);          // part of the chimeric chain: second lane inside a pipelined pass
    if (c->pass.active) pass_call(c, pass_hash(PASS_H0, "pack_sa_rows", 12));
    const int64_t n = c->nrec;
    CORAL_CUDA(c, c->s0.ensure((size_t)std::max<int64_t>(n, 1) * 4));
    LAUNCH_1D(k_flag_has_sa, n, n, c->rt.sa_off.as<uint32_t>(), c->s0.as<uint32_t>());
    MsgHdr h{};
    return pack_flagged(c, MSG_SA, h, c->xmsg[0], d_msg, nbytes);
}

// Message 2: given every rank's message 1 (gathered, on this device), the first local record with flag < 256 of
// each chimeric name that carries no SA tag itself (read_length comes from the first primary in the file,
// ibg:142-143) + this shard's nm_stats partial sums (ibg:153-157) in the header.
int coral_pack_extra_rows(coral_ctx* c, int32_t n_ranks, const uint8_t* d_buf, const uint64_t* starts, uint64_t* d_msg,
                          uint64_t* nbytes) {
    if (!c || !d_buf || !starts) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_pack_extra_rows needs a ctx with records"; return CORAL_E_STATE; }
    if (!(c->pass.active && c->scan_pending)) CORAL_TRY(coral_wait_scan(c));
    cudaSetDevice(c->device);
    coral::PassLane lane__(c);
    if (c->pass.active) {
        uint64_t k = pass_hash(PASS_H0, "pack_extra_rows", 15);
        k = pass_hash(k, &n_ranks, 4);
        pass_call(c, pass_hash(k, starts, (size_t)std::max(n_ranks, 0) * 8));
    }
    MsgSet set;
    std::vector<MsgHdr> hdrs;
    CORAL_TRY(read_msgs(c, n_ranks, d_buf, starts, MSG_SA, &set, &hdrs));
    const int64_t n = c->nrec;
    const int64_t m = (int64_t)set.row0[n_ranks];
    RecordTable& t = c->rt;
    CORAL_CUDA(c, c->s0.ensure((size_t)std::max<int64_t>(n, 1) * 4));
    if (n > 0 && m > 0) {
        CORAL_CUDA(c, c->s4.ensure((size_t)m * 16));     // lo keys in / out
        CORAL_CUDA(c, c->s5.ensure((size_t)m * 12));     // idx in / out, first
        CORAL_CUDA(c, c->s6.ensure((size_t)n * 4));      // slot_of
        uint64_t* lo_in = c->s4.as<uint64_t>(); uint64_t* lo_out = lo_in + m;
        uint32_t* ix_in = c->s5.as<uint32_t>(); uint32_t* ix_out = ix_in + m; uint32_t* first = ix_out + m;
        LAUNCH_1D(k_collect_names, m, set, lo_in, ix_in);
        CORAL_TRY(sort_pairs_u64(c, lo_in, lo_out, ix_in, ix_out, m, 0, NAME_SORT_BITS));
        CORAL_CUDA(c, cudaMemsetAsync(first, 0xFF, (size_t)m * 4, c->stream));
        LAUNCH_1D(k_extra_probe, n, n, t.flag_mapq.as<uint32_t>(), t.sa_off.as<uint32_t>(), t.name_lo.as<uint64_t>(),
                  t.name_hi.as<uint64_t>(), set, lo_out, ix_out, m, first, c->s6.as<int32_t>());
        LAUNCH_1D(k_extra_flag, n, n, c->s6.as<int32_t>(), first, c->s0.as<uint32_t>());
    } else if (n > 0) {
        CORAL_CUDA(c, cudaMemsetAsync(c->s0.p, 0, (size_t)n * 4, c->stream));
    }
    MsgHdr h{};
    // this shard's nm_stats partial sums (ibg:153-157) ride in the header: reduced on the device, written by k_write_hdr
    double* d_nm = nullptr;
    CORAL_TRY(coral_nm_partials_device(c, &d_nm));
    return pack_flagged(c, MSG_EXTRA, h, c->xmsg[1], d_msg, nbytes, d_nm);
}

// Both gathered message sets -> the columns of a rows-only ctx (rows in global file order: rank order, then record
// order inside the rank, the two messages of a rank merged by record index).  nm_* (per rank) may be NULL.
int coral_unpack_rows(coral_ctx* c, int32_t n_ranks, const uint8_t* d_buf1, const uint64_t* starts1, const uint8_t* d_buf2,
                      const uint64_t* starts2, double* nm_sum, double* nm_sumsq, int64_t* nm_n, int64_t* nm_zero_div,
                      int64_t* n_rows_out) {
    if (!c || !d_buf1 || !starts1 || !d_buf2 || !starts2) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    if (c->pass.active) {
        uint64_t k = pass_hash(PASS_H0, "unpack_rows", 11);
        k = pass_hash(k, &n_ranks, 4);
        k = pass_hash(k, starts1, (size_t)std::max(n_ranks, 0) * 8);
        pass_call(c, pass_hash(k, starts2, (size_t)std::max(n_ranks, 0) * 8));
    }
    UnpackPlan p;
    std::vector<MsgHdr> h1, h2;
    CORAL_TRY(read_msgs(c, n_ranks, d_buf1, starts1, MSG_SA, &p.sa, &h1));
    CORAL_TRY(read_msgs(c, n_ranks, d_buf2, starts2, MSG_EXTRA, &p.ex, &h2));
    uint64_t rows = 0, text = 0;
    for (int r = 0; r < n_ranks; r++) {
        p.out_base[r] = rows;
        rows += h1[r].n_rows + h2[r].n_rows;
        p.text_sa[r] = text; text += (h1[r].text_bytes + 15) & ~(uint64_t)15;      // every message's text starts 16-byte aligned
        p.text_ex[r] = text; text += (h2[r].text_bytes + 15) & ~(uint64_t)15;
        if (nm_sum) nm_sum[r] = h2[r].nm_sum;
        if (nm_sumsq) nm_sumsq[r] = h2[r].nm_sq;
        if (nm_n) nm_n[r] = h2[r].nm_n;
        if (nm_zero_div) nm_zero_div[r] = h2[r].nm_zd;
    }
    if (rows >= ((uint64_t)1 << 32) - 65536) { c->err = "too many gathered rows"; return CORAL_E_ARG; }
    c->loaded = false;
    c->rows_only = true;
    const int64_t n = (int64_t)rows;
    CORAL_CUDA(c, c->recs_owned.ensure(text + 64));
    {
        TextCopy tc;
        tc.n = 0;
        uint64_t most = 0;
        for (int r = 0; r < n_ranks; r++) {
            const uint8_t* t1 = p.sa.base[r] + sizeof(MsgHdr) + h1[r].n_rows * sizeof(RowMsg);
            const uint8_t* t2 = p.ex.base[r] + sizeof(MsgHdr) + h2[r].n_rows * sizeof(RowMsg);
            const uint8_t* srcs[2] = {t1, t2};
            const uint64_t bytes[2] = {h1[r].text_bytes, h2[r].text_bytes};
            const uint64_t dsts[2] = {p.text_sa[r], p.text_ex[r]};
            for (int k = 0; k < 2; k++) {
                if (!bytes[k]) continue;
                tc.src[tc.n] = reinterpret_cast<const uint4*>(srcs[k]);      // messages pad their text to 16 bytes
                tc.dst[tc.n] = reinterpret_cast<uint4*>(c->recs_owned.as<uint8_t>() + dsts[k]);
                tc.nvec[tc.n] = (bytes[k] + 15) >> 4;
                most = std::max(most, tc.nvec[tc.n]);
                tc.n++;
            }
        }
        if (tc.n) {
            const unsigned bx = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((most + 1023) / 1024, (uint64_t)c->sm_count));
            k_copy_text<<<dim3(bx, (unsigned)tc.n), 256, 0, c->stream>>>(tc);
            c->n_launches++;
            CORAL_CUDA(c, cudaGetLastError());
        }
    }
    CORAL_CUDA(c, cudaMemsetAsync(c->recs_owned.as<uint8_t>() + text, 0, 64, c->stream));
    c->d_recs = c->recs_owned.as<uint8_t>();
    c->nbytes = text;
    c->nrec = n;
    RecordTable& t = c->rt;
    const size_t n4 = (size_t)std::max<int64_t>(n, 1) * 4, n8 = n4 * 2;
    CORAL_CUDA(c, c->d_off.ensure(n8 + 8));
    DevBuf* b8[] = {&t.name_lo, &t.name_hi};
    for (DevBuf* b : b8) CORAL_CUDA(c, b->ensure(n8));
    DevBuf* b4[] = {&t.flag_mapq, &t.lseq, &t.nm, &t.sa_off, &t.sa_len};
    for (DevBuf* b : b4) CORAL_CUDA(c, b->ensure(n4));
    RowCols o{c->d_off.as<uint64_t>(), t.name_lo.as<uint64_t>(), t.name_hi.as<uint64_t>(), t.flag_mapq.as<uint32_t>(),
              t.lseq.as<uint32_t>(), t.sa_off.as<uint32_t>(), t.sa_len.as<uint32_t>(), t.nm.as<int32_t>()};
    LAUNCH_1D(k_unpack_rows, std::max<int64_t>(n, 1), p, o, text);
    // columns a row does not carry: tid -1 (no placement), the rest 0.  Nothing ever writes them on a rows-only ctx, so a
    // later unpack only fills what a regrown buffer has lost (ten memsets per pass otherwise)
    DevBuf* fill[] = {&t.tid, &t.pos, &t.endpos, &t.ncig, &t.infer, &t.mbases, &t.nonacgt, &t.fbs, &t.lbe, &t.ngap};
    for (int k = 0; k < 10; k++) {
        DevBuf* b = fill[k];
        CORAL_CUDA(c, b->ensure(n4));
        if (c->rows_fill_ptr[k] != b->p || c->rows_fill_bytes[k] < n4) {
            CORAL_CUDA(c, cudaMemsetAsync(b->p, k == 0 ? 0xFF : 0, b->cap, c->stream));
            c->rows_fill_ptr[k] = b->p;
            c->rows_fill_bytes[k] = b->cap;
        }
    }
    c->gaps.n = 0;
    CORAL_CUDA(c, c->status.ensure(64));
    CORAL_CUDA(c, cudaMemsetAsync(c->status.p, 0, 64, c->stream));
    CORAL_CUDA(c, c->gaps.counter.ensure(16));
    if (!c->pass.active) CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    c->loaded = true;
    if (n_rows_out) *n_rows_out = n;
    return CORAL_OK;
}

}  // extern "C"
