// coral_b200/csrc/relops.cuh — small relational helpers (stable radix sort, scan, ordered compaction)
// for the candidate/cluster tables.  CUB device primitives are library code; the kernels that carry
// the reference's semantics are ours and live in the .cu files.
#pragma once
#include <cub/cub.cuh>

#include <cstring>
#include <utility>

#include "ctx.h"

namespace coral {

// Stable sort of (key u64, value u32) pairs on bits [begin_bit, end_bit).
inline int sort_pairs_u64(coral_ctx* c, const uint64_t* k_in, uint64_t* k_out, const uint32_t* v_in, uint32_t* v_out,
                          int64_t n, int begin_bit = 0, int end_bit = 64) {
    if (n <= 0) return CORAL_OK;
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, bytes, k_in, k_out, v_in, v_out, (int)n, begin_bit, end_bit, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, bytes, k_in, k_out, v_in, v_out, (int)n, begin_bit, end_bit, c->stream));
    c->n_launches += 3;
    return CORAL_OK;
}

inline int exclusive_sum_u32(coral_ctx* c, const uint32_t* in, uint32_t* out, int64_t n) {
    if (n <= 0) return CORAL_OK;
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int)n, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceScan::ExclusiveSum(c->cub_tmp.p, bytes, in, out, (int)n, c->stream));
    c->n_launches += 1;
    return CORAL_OK;
}

inline int inclusive_sum_u32(coral_ctx* c, const uint32_t* in, uint32_t* out, int64_t n) {
    if (n <= 0) return CORAL_OK;
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(nullptr, bytes, in, out, (int)n, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceScan::InclusiveSum(c->cub_tmp.p, bytes, in, out, (int)n, c->stream));
    c->n_launches += 1;
    return CORAL_OK;
}

// out[j] = i for every i in [0, n) with flags[i] != 0, in increasing i; *d_count = number selected.
inline int select_flagged_iota(coral_ctx* c, const uint8_t* flags, uint32_t* out, uint32_t* d_count, int64_t n) {
    if (n <= 0) {
        CORAL_CUDA(c, cudaMemsetAsync(d_count, 0, 4, c->stream));
        return CORAL_OK;
    }
    cub::CountingInputIterator<uint32_t> iota(0);
    size_t bytes = 0;
    CORAL_CUDA(c, cub::DeviceSelect::Flagged(nullptr, bytes, iota, flags, out, d_count, (int)n, c->stream));
    CORAL_CUDA(c, c->cub_tmp.ensure(bytes));
    CORAL_CUDA(c, cub::DeviceSelect::Flagged(c->cub_tmp.p, bytes, iota, flags, out, d_count, (int)n, c->stream));
    c->n_launches += 1;
    return CORAL_OK;
}

// ---- pipelined pass helpers (PassState, ctx.h) ---------------------------------------------------------------------
static __global__ void k_pass_check(const void* d, unsigned long long expect, int nbytes, uint32_t* flag) {
    const unsigned long long v = nbytes == 8 ? *static_cast<const unsigned long long*>(d)
                                             : (unsigned long long)*static_cast<const uint32_t*>(d);
    if (v != expect) atomicAdd(flag, 1u);
}

static __global__ void k_copy_words(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, size_t bytes) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 4 <= bytes && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 3) == 0)
        *reinterpret_cast<uint32_t*>(dst + i) = *reinterpret_cast<const uint32_t*>(src + i);
    else
        for (size_t k = i; k < bytes && k < i + 4; k++) dst[k] = src[k];
}

inline uint64_t pass_hash(uint64_t h, const void* p, size_t bytes) {      // FNV-1a over host input arrays
    const uint8_t* b = static_cast<const uint8_t*>(p);
    for (size_t i = 0; i < bytes; i++) { h ^= b[i]; h *= 0x100000001B3ull; }
    return h;
}
constexpr uint64_t PASS_H0 = 0xCBF29CE484222325ull;

// Every entry point that can run inside a pass announces itself with a key over its host inputs.  A replayed pass
// whose call sequence departs from the recorded one continues as a recording pass from that call on.
inline void pass_call(coral_ctx* c, uint64_t key) {
    PassState& P = c->pass;
    if (!P.active) return;
    if (P.replay) {
        if (P.kcur < P.keys.size() && P.keys[P.kcur] == key) { P.kcur++; return; }
        P.replay = false;
        P.vals.resize(P.cur);
        P.keys.resize(P.kcur);
    }
    P.keys.push_back(key);
    P.kcur = P.keys.size();
}

// A device scalar (4 or 8 bytes) that sizes later launches.  Outside a pass and while recording: read back now.
// While replaying: the recorded value, checked on the device.
template <class T>
inline int pass_resolve(coral_ctx* c, T* out, const void* d_src) {
    static_assert(sizeof(T) == 4 || sizeof(T) == 8, "4- or 8-byte scalars");
    PassState& P = c->pass;
    if (P.active && P.replay && P.cur < P.vals.size()) {
        const uint64_t v = P.vals[P.cur++];
        T t;
        memcpy(&t, &v, sizeof(T));
        *out = t;
        k_pass_check<<<1, 1, 0, c->stream>>>(d_src, (unsigned long long)v, (int)sizeof(T), P.d_flag);
        c->n_launches++;
        return CORAL_OK;
    }
    if (P.active && P.replay) {       // the recorded sequence is shorter than this pass: record from here on
        P.replay = false;
        P.vals.resize(P.cur);
        P.keys.resize(P.kcur);
    }
    CORAL_CUDA(c, cudaMemcpyAsync(out, d_src, sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    if (P.active) {
        uint64_t v = 0;
        memcpy(&v, out, sizeof(T));
        P.vals.push_back(v);
        P.cur = P.vals.size();
    }
    return CORAL_OK;
}

// Lazy form for entry points whose resolved scalars stay untouched until the entry point returns: while replaying the
// check is only noted, pass_flush_checks() launches ONE kernel for all noted values (every kernel on the critical path
// of a pass costs microseconds whatever it does).  Must be flushed before the memory behind d_src is reused.
struct PassChecks { const void* d[8]; unsigned long long expect[8]; int nbytes[8]; int n; };
static __global__ void k_pass_check_multi(PassChecks a, uint32_t* flag) {
    const int i = threadIdx.x;
    if (i >= a.n) return;
    const unsigned long long v = a.nbytes[i] == 8 ? *static_cast<const unsigned long long*>(a.d[i])
                                                  : (unsigned long long)*static_cast<const uint32_t*>(a.d[i]);
    if (v != a.expect[i]) atomicAdd(flag, 1u);
}
inline int pass_flush_checks(coral_ctx* c) {
    PassState& P = c->pass;
    for (size_t k = 0; k < P.pending.size(); k += 8) {
        PassChecks a;
        a.n = (int)std::min<size_t>(8, P.pending.size() - k);
        for (int i = 0; i < a.n; i++) { a.d[i] = P.pending[k + i].d; a.expect[i] = P.pending[k + i].expect; a.nbytes[i] = P.pending[k + i].nbytes; }
        k_pass_check_multi<<<1, 8, 0, c->stream>>>(a, P.d_flag);
        c->n_launches++;
    }
    P.pending.clear();
    return CORAL_OK;
}
template <class T>
inline int pass_resolve_lazy(coral_ctx* c, T* out, const void* d_src) {
    PassState& P = c->pass;
    if (P.active && P.replay && P.cur < P.vals.size()) {
        const uint64_t v = P.vals[P.cur++];
        T t;
        memcpy(&t, &v, sizeof(T));
        *out = t;
        P.pending.push_back(PassState::Pending{d_src, (unsigned long long)v, (int)sizeof(T)});
        return CORAL_OK;
    }
    CORAL_TRY(pass_flush_checks(c));
    return pass_resolve(c, out, d_src);
}

// Array form: n (<= 128) 8-byte words, one check kernel for all of them.
struct PassWords { unsigned long long w[128]; };
static __global__ void k_pass_check_n(const unsigned long long* d, PassWords expect, int n, uint32_t* flag) {
    const int i = threadIdx.x;
    if (i < n && d[i] != expect.w[i]) atomicAdd(flag, 1u);
}
inline int pass_resolve_n(coral_ctx* c, uint64_t* out, const void* d_src, int n) {
    if (n <= 0) return CORAL_OK;
    if (n > 128) { c->err = "pass_resolve_n: more than 128 words"; return CORAL_E_ARG; }
    PassState& P = c->pass;
    if (P.active && P.replay && P.cur + (size_t)n <= P.vals.size()) {
        PassWords ex;
        for (int i = 0; i < n; i++) { out[i] = P.vals[P.cur + i]; ex.w[i] = out[i]; }
        P.cur += (size_t)n;
        k_pass_check_n<<<1, 128, 0, c->stream>>>(static_cast<const unsigned long long*>(d_src), ex, n, P.d_flag);
        c->n_launches++;
        return CORAL_OK;
    }
    if (P.active && P.replay) { P.replay = false; P.vals.resize(P.cur); P.keys.resize(P.kcur); }
    CORAL_CUDA(c, cudaMemcpyAsync(out, d_src, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    if (P.active) {
        for (int i = 0; i < n; i++) P.vals.push_back(out[i]);
        P.cur = P.vals.size();
    }
    return CORAL_OK;
}

// The 64-byte headers of up to 16 gathered messages (8 words each) as host values: read back (outside a pass / while
// recording: one gathering kernel + one copy) or replayed and compared IN PLACE by one kernel.
struct HdrPtrs { const uint8_t* base[16]; int n; };
static __global__ void k_hdrs_collect(HdrPtrs h, uint64_t* __restrict__ out) {
    const int t = threadIdx.x, k = t >> 3;
    if (k < h.n) out[t] = reinterpret_cast<const uint64_t*>(h.base[k])[t & 7];
}
static __global__ void k_hdrs_check(HdrPtrs h, PassWords expect, uint32_t* flag) {
    const int t = threadIdx.x, k = t >> 3;
    if (k < h.n && reinterpret_cast<const uint64_t*>(h.base[k])[t & 7] != expect.w[t]) atomicAdd(flag, 1u);
}
inline int pass_resolve_hdrs(coral_ctx* c, const HdrPtrs& h, uint64_t* out) {
    const int n = h.n * 8;
    if (h.n < 1 || h.n > 16) { c->err = "1..16 messages per exchange"; return CORAL_E_ARG; }
    PassState& P = c->pass;
    if (P.active && P.replay && P.cur + (size_t)n <= P.vals.size()) {
        PassWords ex;
        for (int i = 0; i < n; i++) { out[i] = P.vals[P.cur + i]; ex.w[i] = out[i]; }
        P.cur += (size_t)n;
        k_hdrs_check<<<1, 128, 0, c->stream>>>(h, ex, P.d_flag);
        c->n_launches++;
        return CORAL_OK;
    }
    CORAL_CUDA(c, c->s7.ensure((size_t)n * 8));
    k_hdrs_collect<<<1, 128, 0, c->stream>>>(h, c->s7.as<uint64_t>());
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    return pass_resolve_n(c, out, c->s7.p, n);
}

inline uint8_t* pass_stage(coral_ctx* c, size_t bytes) {          // page-locked staging for one deferred copy
    PassState& P = c->pass;
    bytes = (bytes + 63) & ~(size_t)63;
    for (auto& ch : P.pinned)
        if (ch.cap - ch.used >= bytes) { uint8_t* r = ch.p + ch.used; ch.used += bytes; return r; }
    const size_t cap = std::max<size_t>(bytes, (size_t)8 << 20);
    uint8_t* p = nullptr;
    if (cudaHostAlloc(reinterpret_cast<void**>(&p), cap, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    P.pinned.push_back(PassState::Chunk{p, cap, bytes});
    return p;
}

// Copy of a RESULT to the caller's array: now (the caller synchronises) or, inside a pass, at coral_pass_end.
// From here to the next pass_sync() the result copies of a pass run on the pass's side stream, behind what the ctx's
// stream holds now: for tables that nothing overwrites before the pass ends (the chimeric table, the candidate table,
// the small-deletion and concordant rows), so that a dozen copy-engine operations do not sit between two kernels of the
// critical path (measured: 80 us between the end of fetch and the first candidate kernel).
enum { SIDE_CHIM = 1, SIDE_CAND = 2, SIDE_SDEL = 4, SIDE_STAB = 8 };
inline int pass_side_event(coral_ctx* c, cudaEvent_t* out) {
    PassState& P = c->pass;
    if (P.n_side_events == P.side_events.size()) {
        cudaEvent_t e;
        CORAL_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        P.side_events.push_back(e);
    }
    *out = P.side_events[P.n_side_events++];
    return CORAL_OK;
}
// An entry point that rewrites `table` while copies of it may still run on the side stream waits for them first.
inline int pass_side_join(coral_ctx* c, int table) {
    PassState& P = c->pass;
    if (!P.active || !(P.side_mask & table)) return CORAL_OK;
    cudaEvent_t e;
    CORAL_TRY(pass_side_event(c, &e));
    CORAL_CUDA(c, cudaEventRecord(e, P.side_stream));
    CORAL_CUDA(c, cudaStreamWaitEvent(c->stream, e, 0));
    P.side_mask = 0;
    return CORAL_OK;
}
inline int pass_side_begin(coral_ctx* c, int table) {
    PassState& P = c->pass;
    if (!P.active) return CORAL_OK;
    P.side_mask |= table;
    if (!P.side_stream) CORAL_CUDA(c, cudaStreamCreateWithFlags(&P.side_stream, cudaStreamNonBlocking));
    cudaEvent_t e;
    CORAL_TRY(pass_side_event(c, &e));
    CORAL_CUDA(c, cudaEventRecord(e, c->stream));
    CORAL_CUDA(c, cudaStreamWaitEvent(P.side_stream, e, 0));
    P.side_on = P.side_used = true;
    return CORAL_OK;
}

template <class T>
inline int d2h_result(coral_ctx* c, T* dst, const void* src, size_t n) {
    if (n == 0) return CORAL_OK;
    PassState& P = c->pass;
    if (P.active) {
        // a page-locked destination (coral_host_alloc: what the Python binding hands out inside a pass when asked to)
        // takes the copy directly: no staging buffer, no host memcpy at coral_pass_end
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, dst) == cudaSuccess && at.type == cudaMemoryTypeHost) {
            CORAL_CUDA(c, cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, P.side_on ? P.side_stream : c->stream));
            return CORAL_OK;
        }
        cudaGetLastError();
        uint8_t* st = pass_stage(c, n * sizeof(T));
        if (!st) { c->err = "cudaHostAlloc failed for the pass staging buffer"; return CORAL_E_NOMEM; }
        CORAL_CUDA(c, cudaMemcpyAsync(st, src, n * sizeof(T), cudaMemcpyDeviceToHost, P.side_on ? P.side_stream : c->stream));
        P.copies.push_back(PassState::Copy{dst, st, n * sizeof(T), -1});
        return CORAL_OK;
    }
    CORAL_CUDA(c, cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    return CORAL_OK;
}
// The synchronisation that ends an entry point: skipped inside a pass (coral_pass_end synchronises once).
inline int pass_sync(coral_ctx* c) {
    PassState& P = c->pass;
    if (P.active) {
        // the result copies enqueued since the last mark form a group behind one event: coral_pass_end hands a group to
        // the caller's arrays as soon as ITS event has fired, while later stages are still running on the device
        size_t first = P.copies.size();
        while (first > 0 && P.copies[first - 1].group < 0) first--;
        if (first < P.copies.size()) {
            if (P.n_events == P.events.size()) {
                cudaEvent_t e;
                CORAL_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                P.events.push_back(e);
            }
            const int g = (int)P.n_events++;
            CORAL_CUDA(c, cudaEventRecord(P.events[g], P.side_on ? P.side_stream : c->stream));
            for (size_t k = first; k < P.copies.size(); k++) P.copies[k].group = g;
        }
        P.side_on = false;
        return CORAL_OK;
    }
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}
// Host post-processing of results that were just copied: now, or after the pass's synchronisation.
inline int pass_after(coral_ctx* c, std::function<int()> fn) {
    if (c->pass.active) { c->pass.after.push_back(std::move(fn)); return CORAL_OK; }
    return fn();
}

// RAII guard of the entry points of the chimeric chain (coral_fetch, coral_candidates_*, coral_cluster_call*, their
// getters): inside a pass with the second lane enabled they run on the lane's stream with the lane's scratch buffers.
// The lane's stream waits once per pass for everything the main stream was given before the first such call (the
// scan, or its pre-scan); coral_pass_end joins the lane.
struct PassLane {
    coral_ctx* c;
    bool on = false;
    static void swap_sets(coral_ctx* c) {
        DevBuf* m[8] = {&c->s0, &c->s1, &c->s2, &c->s3, &c->s4, &c->s5, &c->s6, &c->s7};
        for (int k = 0; k < 8; k++) std::swap(*m[k], c->lane_s[k]);
        std::swap(c->cub_tmp, c->lane_cub_tmp);
        std::swap(c->status, c->lane_status);
        std::swap(c->stream, c->pass.lane_stream);
    }
    explicit PassLane(coral_ctx* ctx) : c(ctx) {
        if (!c) return;
        PassState& P = c->pass;
        if (!P.active || !P.fork || P.in_lane) return;
        if (!P.lane_stream) {
            if (cudaStreamCreateWithFlags(&P.lane_stream, cudaStreamNonBlocking) != cudaSuccess) return;
            cudaEventCreateWithFlags(&P.ev_lane_begin, cudaEventDisableTiming);
            cudaEventCreateWithFlags(&P.ev_lane_done, cudaEventDisableTiming);
        }
        if (!P.lane_started) {
            cudaEventRecord(P.ev_lane_begin, c->stream);
            cudaStreamWaitEvent(P.lane_stream, P.ev_lane_begin, 0);
            P.lane_started = true;
        }
        swap_sets(c);
        P.in_lane = true;
        on = true;
    }
    ~PassLane() {
        if (!on) return;
        swap_sets(c);
        c->pass.in_lane = false;
    }
};

// Immediate copy (the caller synchronises right after): entry points that do not take part in pipelined passes.
template <class T>
inline int d2h(coral_ctx* c, T* dst, const void* src, size_t n) {
    if (n == 0) return CORAL_OK;
    CORAL_CUDA(c, cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, c->stream));
    return CORAL_OK;
}
// Copy of an INPUT array to the device.  Outside a pass: a plain asynchronous copy.  Inside a pass every input array has
// a slot (its position in the pass) in a device-side cache: a recording pass fills the slot through page-locked staging
// (a copy from pageable memory takes the driver's slow path: tens of microseconds of stream time each), a replayed pass
// — whose inputs are the recorded ones, the call keys say so — copies slot -> destination on the device, so no input
// crosses PCIe again (measured: 0.17 ms of stream time per coral_region_coverage call before, for three small arrays).
// ONLY for inputs that the entry point's call key (pass_call) covers: h2d() below is the form for everything else.
template <class T>
inline int h2d_keyed(coral_ctx* c, void* dst, const T* src, size_t n) {
    if (n == 0) return CORAL_OK;
    PassState& P = c->pass;
    const size_t bytes = n * sizeof(T);
    if (!P.active) {
        CORAL_CUDA(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream));
        return CORAL_OK;
    }
    const size_t slot = P.in_cur++;
    if (P.inputs.size() <= slot) P.inputs.resize(slot + 1);
    PassState::Input& in = P.inputs[slot];
    if (!(P.replay && in.p && in.bytes == bytes)) {
        if (in.cap < bytes) {
            if (in.p) coral_mem::release(in.p, in.cap);
            in.p = nullptr; in.cap = 0;
            size_t got = 0;
            CORAL_CUDA(c, coral_mem::acquire(&in.p, bytes + 256, &got));
            in.cap = got;
        }
        uint8_t* st = pass_stage(c, bytes);
        if (!st) { c->err = "cudaHostAlloc failed for the pass staging buffer"; return CORAL_E_NOMEM; }
        memcpy(st, src, bytes);
        CORAL_CUDA(c, cudaMemcpyAsync(in.p, st, bytes, cudaMemcpyHostToDevice, c->stream));
        in.bytes = bytes;
    }
    // slot -> destination with a kernel, not the copy engine: a copy-engine operation between two kernels of a stream
    // costs a hand-over between the engines on both sides (tens of microseconds of stream time, measured)
    const size_t words = (bytes + 3) / 4;
    k_copy_words<<<(unsigned)((words + 255) / 256), 256, 0, c->stream>>>(static_cast<const uint8_t*>(in.p),
                                                                          static_cast<uint8_t*>(dst), bytes);
    c->n_launches++;
    return CORAL_OK;
}

// Input copy of an entry point without a call key: asynchronous copy, through page-locked staging inside a pass.
template <class T>
inline int h2d(coral_ctx* c, void* dst, const T* src, size_t n) {
    if (n == 0) return CORAL_OK;
    const void* from = src;
    if (c->pass.active) {
        uint8_t* st = pass_stage(c, n * sizeof(T));
        if (!st) { c->err = "cudaHostAlloc failed for the pass staging buffer"; return CORAL_E_NOMEM; }
        memcpy(st, src, n * sizeof(T));
        from = st;
    }
    CORAL_CUDA(c, cudaMemcpyAsync(dst, from, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return CORAL_OK;
}

}  // namespace coral
