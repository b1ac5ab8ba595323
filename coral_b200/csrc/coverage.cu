// coral_b200/csrc/coverage.cu — region coverage from the scan table.
//
// Replaces AlignmentFile.count_coverage(contig, start, stop, quality_threshold=0, 'nofilter') summed
// over its four arrays (reference infer_breakpoint_graph.py:131-132 and :1033-1034) and the record
// count `len([read for read in fetch(...) if read.infer_read_length()])` (:1031-1032).
//
// The reference walks every aligned base in Python/Cython.  Here the sum over a region is closed
// form: a record lying entirely inside a region contributes its aligned-base count (known from the
// scan) minus its non-ACGT aligned bases; only records that straddle a region boundary, or carry
// non-ACGT query bases, re-walk their own CIGAR (and SEQ) in a second, warp-per-pair kernel.
#include "ctx.h"
#include "dev_util.cuh"
#include "relops.cuh"

namespace coral {

struct Regions {
    const int32_t *tid, *start, *stop;
    int32_t n;
};

// lowest region index i with (rtid[i], rstop[i]) > (tid, pos): first region not entirely left of pos
__device__ __forceinline__ int32_t first_region_after(const Regions& R, int32_t tid, int32_t pos) {
    int32_t lo = 0, hi = R.n;
    while (lo < hi) {
        int32_t mid = (lo + hi) >> 1;
        int32_t t = R.tid[mid];
        bool left = (t < tid) || (t == tid && R.stop[mid] <= pos);
        if (left) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// Warp-aggregated atomicAdd of (key, value) pairs: one atomic per distinct key per warp.
__device__ __forceinline__ void warp_agg_add(unsigned long long* acc, int32_t key, unsigned long long val, bool active) {
    uint32_t todo = __ballot_sync(FULL_MASK, active);
    const uint32_t lane = lane_id();
    while (todo) {
        const int leader = __ffs(todo) - 1;
        const int32_t k = __shfl_sync(FULL_MASK, key, leader);
        const bool mine = active && key == k;
        unsigned long long v = mine ? val : 0ull;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL_MASK, v, d);
        if ((int)lane == leader && v) atomicAdd(acc + k, v);
        todo &= ~__ballot_sync(FULL_MASK, mine);
    }
}

__global__ void k_region_cov_fast(int64_t n, const int32_t* __restrict__ tid, const int32_t* __restrict__ pos,
                                  const int32_t* __restrict__ endpos, const uint32_t* __restrict__ flag_mapq,
                                  const uint32_t* __restrict__ lseq, const uint32_t* __restrict__ infer,
                                  const uint32_t* __restrict__ mbases, const uint32_t* __restrict__ nonacgt,
                                  Regions R, unsigned long long* base_sum, unsigned long long* rec_count,
                                  uint32_t* slow_rec, int32_t* slow_reg, unsigned long long* slow_n, int64_t slow_cap) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n;
    int32_t t = -1, p = 0, e = 0, ri = R.n;
    uint32_t fm = 0, ls = 0, inf = 0, mb = 0, bad = 0;
    if (live) {
        t = tid[i]; p = pos[i]; e = endpos[i];
        fm = flag_mapq[i]; ls = lseq[i]; inf = infer[i]; mb = mbases[i]; bad = nonacgt[i];
        if (t >= 0) ri = first_region_after(R, t, p);
    }
    // A record overlaps a (possibly empty) run of consecutive regions; iterate warp-synchronously.
    while (true) {
        const bool ov = live && ri < R.n && R.tid[ri] == t && R.start[ri] < e;
        if (!__any_sync(FULL_MASK, ov)) break;
        const bool inside = ov && p >= R.start[ri] && e <= R.stop[ri];
        const bool clean = inside && bad == 0 && !(fm & 4u);
        warp_agg_add(rec_count, ri, 1ull, ov && inf > 0);
        warp_agg_add(base_sum, ri, (unsigned long long)mb, clean && ls > 0);
        if (ov && !clean && ls > 0 && mb > 0) {
            unsigned long long s = atomicAdd(slow_n, 1ull);
            if ((int64_t)s < slow_cap) { slow_rec[s] = (uint32_t)i; slow_reg[s] = ri; }
        }
        if (ov) ri++;
    }
}

// Exact walk for (record, region) pairs: clipped M/=/X overlap minus non-ACGT aligned query bases.
// One block (8 warps) per pair.  A 30 kb ONT read carries thousands of CIGAR ops and the position of an op depends on
// all ops before it, so a single warp walking the whole CIGAR is a long dependent chain (the first version: 110-145 us
// for a few thousand pairs, latency-bound).  Here the CIGAR is cut into 8 slices: phase 1 sums the reference / query
// advance of each slice (plain per-lane sums, one warp reduction per slice), a prefix over the 8 sums gives every
// slice its start, and phase 2 runs the exact clipped walk only on the slices that overlap the region.
// number of query bases in [q0, q0 + n) whose 4-bit code is not A/C/G/T (1, 2, 4, 8); eight bases per 32-bit word once
// q is on a byte boundary (the order of the nibbles inside a word does not matter for a count)
__device__ __forceinline__ uint32_t non_acgt_in_range(const uint8_t* __restrict__ seq, int64_t q0, int32_t n) {
    uint32_t bad = 0;
    int64_t q = q0;
    const int64_t end = q0 + n;
    if ((q & 1) && q < end) {
        const uint32_t nib = seq[q >> 1] & 15u;
        bad += !(nib == 1 || nib == 2 || nib == 4 || nib == 8);
        q++;
    }
    for (; q + 8 <= end; q += 8)
        bad += non_acgt_nibbles((uint64_t)ld_u32_any(seq + (q >> 1)) | 0x1111111100000000ull);
    for (; q < end; q++) {
        const uint32_t by = seq[q >> 1];
        const uint32_t nib = (q & 1) ? (by & 15u) : (by >> 4);
        bad += !(nib == 1 || nib == 2 || nib == 4 || nib == 8);
    }
    return bad;
}

constexpr int COV_WARPS = 8;
__global__ void __launch_bounds__(COV_WARPS * 32)
k_region_cov_slow(const uint8_t* __restrict__ recs, const uint64_t* __restrict__ off,
                  const uint32_t* __restrict__ slow_rec, const int32_t* __restrict__ slow_reg,
                  int64_t n_pairs, const uint32_t* __restrict__ nonacgt, Regions R,
                  unsigned long long* base_sum) {
    __shared__ int32_t s_ref[COV_WARPS];
    __shared__ long long s_q[COV_WARPS];
    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    for (int64_t w = blockIdx.x; w < n_pairs; w += gridDim.x) {
        const uint32_t rec = slow_rec[w];
        const int32_t ri = slow_reg[w];
        const int32_t rs = R.start[ri], re = R.stop[ri];
        const uint8_t* r = recs + off[rec];
        const int32_t pos = (int32_t)ld_u32_any(r + 8);
        const uint32_t l_name = r[12];
        const uint32_t n_cig = ld_u32_any(r + 16) & 0xffffu;
        const uint8_t* cig = r + 36 + l_name;
        const uint8_t* seq = cig + 4ull * n_cig;
        const bool check_seq = nonacgt[rec] != 0;
        const uint32_t per = ((n_cig + COV_WARPS * 32 - 1) / (COV_WARPS * 32)) * 32;    // ops per slice, a multiple of 32
        const uint32_t b0 = min(warp * per, n_cig), b1 = min(b0 + per, n_cig);
        // phase 1: advance of this slice
        int32_t rsum = 0;
        long long qsum = 0;
        for (uint32_t k = b0 + lane; k < b1; k += 32) {
            const uint32_t v = ld_u32_any(cig + 4ull * k);
            const uint32_t op = v & 15u;
            const int32_t len = (int32_t)(v >> 4);
            const bool is_m = (op == 0 || op == 7 || op == 8);
            rsum += (is_m || op == 2 || op == 3) ? len : 0;
            qsum += (is_m || op == 1 || op == 4) ? len : 0;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            rsum += __shfl_xor_sync(FULL_MASK, rsum, d);
            qsum += __shfl_xor_sync(FULL_MASK, qsum, d);
        }
        if (lane == 0) { s_ref[warp] = rsum; s_q[warp] = qsum; }
        __syncthreads();
        int32_t refpos = pos;
        int64_t qpos = 0;
        for (uint32_t k = 0; k < warp; k++) { refpos += s_ref[k]; qpos += s_q[k]; }
        unsigned long long total = 0;
        // phase 2: only slices that overlap [rs, re)
        if (b0 < b1 && refpos < re && refpos + rsum > rs) {
            for (uint32_t base = b0; base < b1 && refpos < re; base += 32) {
                const uint32_t k = base + lane;
                const uint32_t v = (k < b1) ? ld_u32_any(cig + 4ull * k) : 0xFu;
                const uint32_t op = v & 15u;
                const int32_t len = (int32_t)(v >> 4);
                const bool is_m = (op == 0 || op == 7 || op == 8);
                const int32_t rlen = (is_m || op == 2 || op == 3) ? len : 0;
                const int32_t qlen = (is_m || op == 1 || op == 4) ? len : 0;
                const int32_t ri_incl = warp_incl_sum_i32(rlen);
                int32_t qi_incl = 0;
                if (check_seq) qi_incl = warp_incl_sum_i32(qlen);      // query offsets only matter for the non-ACGT test
                if (is_m) {
                    const int32_t s = refpos + ri_incl - rlen;
                    const int32_t a = max(s, rs), b = min(s + len, re);
                    if (b > a) {
                        total += (unsigned long long)(b - a);
                        if (check_seq) {
                            total -= non_acgt_in_range(seq, qpos + (qi_incl - qlen) + (a - s), b - a);
                        }
                    }
                }
                refpos += __shfl_sync(FULL_MASK, ri_incl, 31);
                if (check_seq) qpos += __shfl_sync(FULL_MASK, qi_incl, 31);
            }
        }
        total = warp_sum_u64(total);
        if (lane == 0 && total) atomicAdd(base_sum + ri, total);
        __syncthreads();          // s_ref / s_q are reused by the next pair
    }
}

}  // namespace coral

int coral_run_region_coverage(coral_ctx* ctx, const int32_t* tid, const int32_t* start, const int32_t* stop,
                              int32_t n, int64_t* base_sum, int64_t* rec_count, coral_xchg* xchg, int xchg_channel) {
    using namespace coral;
    if (!ctx->loaded) { ctx->err = "coral_region_coverage before coral_load_records"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(ctx));
    if (n < 0) { ctx->err = "negative region count"; return CORAL_E_ARG; }
    if (ctx->pass.active) {
        uint64_t k = pass_hash(PASS_H0, xchg ? "region_coverage_x" : "region_coverage", xchg ? 17 : 15);
        k = pass_hash(k, &n, 4);
        if (n > 0) { k = pass_hash(k, tid, (size_t)n * 4); k = pass_hash(k, start, (size_t)n * 4); k = pass_hash(k, stop, (size_t)n * 4); }
        pass_call(ctx, k);
    }
    if (n == 0) return CORAL_OK;
    for (int32_t i = 0; i < n; i++) {
        if (stop[i] < start[i] || tid[i] < 0) { ctx->err = "bad region"; return CORAL_E_ARG; }
        if (i > 0 && (tid[i] < tid[i - 1] || (tid[i] == tid[i - 1] && start[i] < stop[i - 1]))) {
            ctx->err = "regions must be sorted by (tid, start) and disjoint";
            return CORAL_E_ARG;
        }
    }
    cudaStream_t st = ctx->stream;
    const size_t rb = (size_t)n * 4;
    CORAL_CUDA(ctx, ctx->s0.ensure(rb * 3));
    CORAL_CUDA(ctx, ctx->s1.ensure((size_t)n * 16 + 8));
    int32_t* d_reg = ctx->s0.as<int32_t>();
    CORAL_TRY(h2d_keyed(ctx, d_reg, tid, (size_t)n));
    CORAL_TRY(h2d_keyed(ctx, d_reg + n, start, (size_t)n));
    CORAL_TRY(h2d_keyed(ctx, d_reg + 2 * (size_t)n, stop, (size_t)n));
    unsigned long long* d_sum = ctx->s1.as<unsigned long long>();
    unsigned long long* d_cnt = d_sum + n;
    unsigned long long* d_slow_n = d_cnt + n;
    CORAL_CUDA(ctx, cudaMemsetAsync(d_sum, 0, (size_t)n * 16 + 8, st));
    Regions R{d_reg, d_reg + n, d_reg + 2 * (size_t)n, n};
    const int64_t nrec = ctx->nrec;
    int64_t slow_cap = std::max<int64_t>(1 << 16, (int64_t)(ctx->s2.cap / 4));
    RecordTable& t = ctx->rt;
    for (int attempt = 0; attempt < 2 && nrec > 0; attempt++) {
        CORAL_CUDA(ctx, ctx->s2.ensure((size_t)slow_cap * 4));
        CORAL_CUDA(ctx, ctx->s3.ensure((size_t)slow_cap * 4));
        if (attempt) CORAL_CUDA(ctx, cudaMemsetAsync(d_sum, 0, (size_t)n * 16 + 8, st));
        const int threads = 256;
        const int blocks = (int)((nrec + threads - 1) / threads);
        if (attempt == 0) ctx->stage_begin(CORAL_STAGE_COVERAGE);
        k_region_cov_fast<<<blocks, threads, 0, st>>>(
            nrec, t.tid.as<int32_t>(), t.pos.as<int32_t>(), t.endpos.as<int32_t>(), t.flag_mapq.as<uint32_t>(),
            t.lseq.as<uint32_t>(), t.infer.as<uint32_t>(), t.mbases.as<uint32_t>(), t.nonacgt.as<uint32_t>(), R, d_sum,
            d_cnt, ctx->s2.as<uint32_t>(), ctx->s3.as<int32_t>(), d_slow_n, slow_cap);
        ctx->n_launches++;
        CORAL_CUDA(ctx, cudaGetLastError());
        unsigned long long n_slow = 0;
        CORAL_TRY(pass_resolve(ctx, &n_slow, d_slow_n));
        if ((int64_t)n_slow > slow_cap) { slow_cap = (int64_t)n_slow + 1024; continue; }
        if (n_slow) {
            int sblocks = (int)std::min<int64_t>((int64_t)n_slow, (int64_t)ctx->sm_count * 8);
            k_region_cov_slow<<<sblocks, 256, 0, st>>>(ctx->d_recs, ctx->d_off.as<uint64_t>(), ctx->s2.as<uint32_t>(),
                                                       ctx->s3.as<int32_t>(), (int64_t)n_slow, t.nonacgt.as<uint32_t>(),
                                                       R, d_sum);
            ctx->n_launches++;
            CORAL_CUDA(ctx, cudaGetLastError());
        }
        break;
    }
    ctx->stage_end(CORAL_STAGE_COVERAGE);
    // sharded form: a record is owned by exactly one shard, so the sums over the file are the integer sums of the
    // shards' accumulators (both arrays are contiguous), added up in place over peer memory (xchg.cu)
    if (xchg) CORAL_TRY(coral_xchg_sum_device(ctx, xchg, xchg_channel, d_sum, 2 * (int64_t)n));
    // (the sums are below 2^63: the same 64 bits as the caller's int64)
    if (base_sum) CORAL_TRY(d2h_result(ctx, reinterpret_cast<unsigned long long*>(base_sum), d_sum, (size_t)n));
    if (rec_count) CORAL_TRY(d2h_result(ctx, reinterpret_cast<unsigned long long*>(rec_count), d_cnt, (size_t)n));
    return pass_sync(ctx);
}

// ---- allocation-free variant (coral_region_coverage_static) -------------------------------------------------------
// For callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  Everything coral_run_region_coverage keeps in grow-only ctx buffers — the region arrays, the two accumulators,
// the slow-pair counter and the (record, region) list of the exact re-walk — lives in ONE file-scope device array of a
// fixed size instead: every GPU that loads the library holds it (CORAL_STATIC_SCRATCH_BYTES) until the process exits.
// All calls on a device share the array and it keeps the last call's contents: the accumulators and the counter are
// cleared by every call, and two calls whose kernels overlap on the GPU (different ctxs / streams of one device) race.
namespace coral {

constexpr int32_t kStaticMaxRegions = CORAL_STATIC_MAX_REGIONS;
constexpr int64_t kStaticMaxSlow = CORAL_STATIC_MAX_SLOW_PAIRS;
// layout (bytes): [0, 12 R) tid | start | stop (int32 each) ; [12 R, 28 R) base_sum | rec_count (u64 each) ;
//                 [28 R, 28 R + 8) slow counter ; then 248 bytes of padding ; slow_rec u32[S] ; slow_reg i32[S]
constexpr size_t kStaticOffAcc = (size_t)kStaticMaxRegions * 12;
constexpr size_t kStaticOffSlowN = kStaticOffAcc + (size_t)kStaticMaxRegions * 16;
constexpr size_t kStaticOffSlowRec = kStaticOffSlowN + 256;
constexpr size_t kStaticOffSlowReg = kStaticOffSlowRec + (size_t)kStaticMaxSlow * 4;
constexpr size_t kStaticBytes = kStaticOffSlowReg + (size_t)kStaticMaxSlow * 4;
static_assert(kStaticBytes == CORAL_STATIC_SCRATCH_BYTES, "include/coral_b200.h states the size of the static scratch");
static_assert(kStaticOffAcc % 256 == 0 && kStaticOffSlowRec % 256 == 0, "sub-arrays keep the array's alignment");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// In host code `g_static_scratch` is only the host-side stand-in of the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current.
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[64] = {};

static int static_scratch(coral_ctx* ctx, unsigned char** out) {
    if (ctx->device < 0 || ctx->device >= 64) { ctx->err = "device index beyond the static scratch table"; return CORAL_E_ARG; }
    std::lock_guard<std::mutex> g(g_static_mu);
    if (!g_static_ptr[ctx->device]) {
        CORAL_CUDA(ctx, cudaSetDevice(ctx->device));
        void* p = nullptr;
        CORAL_CUDA(ctx, cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) { ctx->err = "cudaGetSymbolAddress returned a null address"; return CORAL_E_CUDA; }
        g_static_ptr[ctx->device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_ptr[ctx->device];
    return CORAL_OK;
}

}  // namespace coral

extern "C" int coral_region_coverage_static(coral_ctx* ctx, const int32_t* tid, const int32_t* start, const int32_t* stop,
                                            int32_t n, int64_t* base_sum, int64_t* rec_count) {
    using namespace coral;
    if (!ctx) return CORAL_E_ARG;
    if (!ctx->loaded || ctx->rows_only) { ctx->err = "coral_region_coverage_static needs loaded records"; return CORAL_E_STATE; }
    if (ctx->scan_pending || ctx->scan_lazy) {
        ctx->err = "coral_region_coverage_static: the record scan is still pending (call coral_wait_scan first)";
        return CORAL_E_STATE;
    }
    if (ctx->pass.active) { ctx->err = "coral_region_coverage_static inside a pipelined pass"; return CORAL_E_STATE; }
    if (n < 0 || (n > 0 && (!tid || !start || !stop))) { ctx->err = "bad region arrays"; return CORAL_E_ARG; }
    if (n > kStaticMaxRegions) {
        char b[160];
        snprintf(b, sizeof b, "coral_region_coverage_static: %d regions exceed the static scratch (at most %d per call)", n,
                 kStaticMaxRegions);
        ctx->err = b;
        return CORAL_E_ARG;
    }
    if (n == 0) return CORAL_OK;
    for (int32_t i = 0; i < n; i++) {
        if (stop[i] < start[i] || tid[i] < 0) { ctx->err = "bad region"; return CORAL_E_ARG; }
        if (i > 0 && (tid[i] < tid[i - 1] || (tid[i] == tid[i - 1] && start[i] < stop[i - 1]))) {
            ctx->err = "regions must be sorted by (tid, start) and disjoint";
            return CORAL_E_ARG;
        }
    }
    CORAL_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    CORAL_TRY(static_scratch(ctx, &base));
    cudaStream_t st = ctx->stream;
    int32_t* d_tid = reinterpret_cast<int32_t*>(base);
    int32_t* d_start = d_tid + kStaticMaxRegions;
    int32_t* d_stop = d_start + kStaticMaxRegions;
    unsigned long long* d_sum = reinterpret_cast<unsigned long long*>(base + kStaticOffAcc);
    unsigned long long* d_cnt = d_sum + kStaticMaxRegions;
    unsigned long long* d_slow_n = reinterpret_cast<unsigned long long*>(base + kStaticOffSlowN);
    uint32_t* d_slow_rec = reinterpret_cast<uint32_t*>(base + kStaticOffSlowRec);
    int32_t* d_slow_reg = reinterpret_cast<int32_t*>(base + kStaticOffSlowReg);
    CORAL_CUDA(ctx, cudaMemcpyAsync(d_tid, tid, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    CORAL_CUDA(ctx, cudaMemcpyAsync(d_start, start, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    CORAL_CUDA(ctx, cudaMemcpyAsync(d_stop, stop, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    // the array keeps the previous call's sums and counter: what must start at zero is cleared here
    CORAL_CUDA(ctx, cudaMemsetAsync(d_sum, 0, kStaticOffSlowN + 8 - kStaticOffAcc, st));
    Regions R{d_tid, d_start, d_stop, n};
    const int64_t nrec = ctx->nrec;
    RecordTable& t = ctx->rt;
    if (nrec > 0) {
        const int threads = 256;
        const int blocks = (int)((nrec + threads - 1) / threads);
        k_region_cov_fast<<<blocks, threads, 0, st>>>(
            nrec, t.tid.as<int32_t>(), t.pos.as<int32_t>(), t.endpos.as<int32_t>(), t.flag_mapq.as<uint32_t>(),
            t.lseq.as<uint32_t>(), t.infer.as<uint32_t>(), t.mbases.as<uint32_t>(), t.nonacgt.as<uint32_t>(), R, d_sum,
            d_cnt, d_slow_rec, d_slow_reg, d_slow_n, kStaticMaxSlow);
        ctx->n_launches++;
        CORAL_CUDA(ctx, cudaGetLastError());
        unsigned long long n_slow = 0;
        CORAL_CUDA(ctx, cudaMemcpyAsync(&n_slow, d_slow_n, 8, cudaMemcpyDeviceToHost, st));
        CORAL_CUDA(ctx, cudaStreamSynchronize(st));
        if ((int64_t)n_slow > kStaticMaxSlow) {
            char b[200];
            snprintf(b, sizeof b, "coral_region_coverage_static: %llu (record, region) pairs need the exact walk, the static "
                     "scratch holds %lld; use coral_region_coverage", n_slow, (long long)kStaticMaxSlow);
            ctx->err = b;
            return CORAL_E_ARG;
        }
        if (n_slow) {
            int sblocks = (int)std::min<int64_t>((int64_t)n_slow, (int64_t)ctx->sm_count * 8);
            k_region_cov_slow<<<sblocks, 256, 0, st>>>(ctx->d_recs, ctx->d_off.as<uint64_t>(), d_slow_rec, d_slow_reg,
                                                       (int64_t)n_slow, t.nonacgt.as<uint32_t>(), R, d_sum);
            ctx->n_launches++;
            CORAL_CUDA(ctx, cudaGetLastError());
        }
    }
    if (base_sum) CORAL_CUDA(ctx, cudaMemcpyAsync(base_sum, d_sum, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    if (rec_count) CORAL_CUDA(ctx, cudaMemcpyAsync(rec_count, d_cnt, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    CORAL_CUDA(ctx, cudaStreamSynchronize(st));
    return CORAL_OK;
}
