// coral_b200/csrc/xchg.cu — the multi-GPU exchanges of SURVEY.md §8e as kernels over peer memory (NVLink / NVSwitch).
//
// One process per GPU; every rank owns one "window" in its own HBM (cudaMalloc + CUDA IPC handle, opened by the
// peers once).  An all-gather of one variable-length message per rank is ONE kernel on the caller's stream and no host
// round trip (k_xchg_push):
//
//   push   every block copies its share of this rank's message into ITS slot of EVERY peer's window with 16-byte stores
//          over NVLink (the rank's own window included); the block that finishes last writes the slot header (size,
//          sequence number) and — after a system-scope fence — releases the sequence number into each peer's flag word;
//   wait   that same block then polls, thread r for rank r, the flag words of the LOCAL window (acquire, system scope)
//          until they carry this exchange's sequence number, and publishes the size of message r.  Bounded: after
//          `timeout` seconds it gives up and reports (a lost peer must not hang the GPU).
//
// After the wait the gathered messages sit in local HBM, message r at a fixed offset, and go straight into the next
// kernel (coral_pack_extra_rows, coral_unpack_rows, coral_merge_chimeric, coral_count_name_msgs ...).  Inside a
// pipelined pass the sizes are replayed from the previous pass and checked on the device (pass_resolve_n), so a whole
// sharded pass is one stream of kernels per rank.  Integer all-reduce (coverage sums) = push + wait + a sum over slots.
//
// Slot reuse: a channel alternates between two slot sets (sequence parity).  Rank A can only deliver exchange s + 2
// of a channel after it has seen every rank's s + 1, which each rank pushes — in stream order — after the kernels that
// consumed s.  Consumers therefore read a slot set on the stream that ran the exchange, before that stream's next
// exchange on the same channel.
#include <algorithm>
#include <memory>
#include <mutex>
#include <string>

#include "ctx.h"
#include "relops.cuh"

struct coral_xchg {
    int device = 0;
    int rank = 0, world = 1;
    uint64_t cap[CORAL_XCHG_CHANNELS] = {};      // slot bytes per channel (64-byte slot header included)
    uint64_t ch_off[CORAL_XCHG_CHANNELS] = {};   // where the channel's 2 x world slots start in a window
    uint64_t bytes = 0;
    uint8_t* local = nullptr;
    uint8_t* peer[CORAL_XCHG_MAX_RANKS] = {};
    bool opened[CORAL_XCHG_MAX_RANKS] = {};
    uint64_t seq[CORAL_XCHG_CHANNELS] = {};      // sequence number of the last exchange per channel (same on every rank)
    uint32_t* d_done = nullptr;                  // [channels] block counters of the push kernel
    uint64_t* d_sizes = nullptr;                 // [channels][MAX_RANKS] sizes of the last gathered messages
    int32_t* d_status = nullptr;                 // [4]: number of timed-out waits
    double timeout_s = 20.0;
    bool connected = false;
    std::string err;
};

namespace coral {

constexpr uint64_t XCHG_FLAGS_BYTES = 4096;      // flags: uint64 [channels][MAX_RANKS] at the start of a window
constexpr uint64_t XCHG_SLOT_HDR = 64;           // slot = [size u64, seq u64, pad] + message

struct PushArgs {
    uint8_t* slot[CORAL_XCHG_MAX_RANKS];         // this rank's slot in every peer's window (current parity)
    uint64_t* flag[CORAL_XCHG_MAX_RANKS];        // this rank's flag word in every peer's window
    int world;
};

// relaxed store at system scope: ordered behind the ONE system-scope fence the publishing thread executes before its
// flag stores (fence + relaxed store = the release pattern; st.release per flag costs a MEMBAR.ALL.SYS per peer)
__device__ __forceinline__ void st_relaxed_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// One kernel per exchange: every block copies its share of the message into this rank's slot of every peer's window;
// the block that finishes last publishes (slot header, then the sequence number into each peer's flag word, release at
// system scope) and then WAITS: thread r polls rank r's flag in the local window (acquire, system scope) until it carries
// this exchange's sequence number and records the size of message r (~0 after the timeout).  The other blocks have left
// by then: one resident block per waiting rank, no second launch.
struct WaitArgs {
    const uint64_t* flags;       // local window: flag words of this channel
    const uint8_t* slots;        // local window: slot set of this exchange
    uint64_t cap, timeout_ns;
    uint64_t* sizes;
    int32_t* status;
};
__global__ void k_xchg_push(const uint4* __restrict__ src, uint64_t nbytes, PushArgs a, uint64_t seq, uint32_t* done, WaitArgs w) {
    __shared__ int s_last;
    const uint64_t nvec = nbytes >> 4;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        const uint4 v = src[i];
        for (int p = 0; p < a.world; p++) reinterpret_cast<uint4*>(a.slot[p] + XCHG_SLOT_HDR)[i] = v;
    }
    __threadfence_system();                       // this thread's remote stores are ordered before what follows
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t prev = atomicAdd(done, 1u);
        s_last = prev == gridDim.x - 1;
        if (s_last) {                             // every block's stores are behind its own fence: publish
            *done = 0;
            __threadfence();
            for (int p = 0; p < a.world; p++) {
                reinterpret_cast<uint64_t*>(a.slot[p])[0] = nbytes;
                reinterpret_cast<uint64_t*>(a.slot[p])[1] = seq;
            }
            __threadfence_system();               // data and headers are visible at system scope before any flag below
            for (int p = 0; p < a.world; p++) st_relaxed_sys(a.flag[p], seq);
        }
    }
    __syncthreads();
    if (!s_last) return;
    const int r = threadIdx.x;
    if (r >= a.world) return;
    const uint64_t t0 = global_timer_ns();
    bool ok = true;
    while (ld_acquire_sys(w.flags + r) < seq) {
        if (global_timer_ns() - t0 > w.timeout_ns) { ok = false; break; }
        __nanosleep(200);
    }
    if (ok) {
        w.sizes[r] = *reinterpret_cast<const volatile uint64_t*>(w.slots + (uint64_t)r * w.cap);
    } else {
        w.sizes[r] = ~0ull;
        atomicAdd(w.status, 1);
    }
}

// acc[i] = sum over ranks of the i-th 8-byte word of each rank's message (integer sums: order does not matter)
__global__ void k_xchg_sum_u64(const uint8_t* slots, uint64_t cap, int world, int64_t n, unsigned long long* __restrict__ acc) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long s = 0;
    for (int r = 0; r < world; r++)
        s += reinterpret_cast<const unsigned long long*>(slots + (uint64_t)r * cap + XCHG_SLOT_HDR)[i];
    acc[i] = s;
}

static int xchg_fail(coral_xchg* x, const char* what, cudaError_t e) {
    x->err = std::string(what) + ": " + cudaGetErrorString(e);
    return CORAL_E_CUDA;
}

// push + wait of one exchange on c->stream; *slots = the gathered slot set in the local window
static int xchg_round(coral_ctx* c, coral_xchg* x, int ch, const void* d_msg, uint64_t nbytes, const uint8_t** slots) {
    if (!x->connected) { c->err = "coral_xchg: not connected"; return CORAL_E_STATE; }
    if (ch < 0 || ch >= CORAL_XCHG_CHANNELS || x->cap[ch] == 0) { c->err = "coral_xchg: no such channel"; return CORAL_E_ARG; }
    if (nbytes & 15) { c->err = "coral_xchg: messages are padded to 16 bytes"; return CORAL_E_ARG; }
    if (reinterpret_cast<uintptr_t>(d_msg) & 15) { c->err = "coral_xchg: message must be 16-byte aligned"; return CORAL_E_ARG; }
    if (c->device != x->device) { c->err = "coral_xchg: ctx and exchange window live on different devices"; return CORAL_E_ARG; }
    if (nbytes + XCHG_SLOT_HDR > x->cap[ch]) {
        char b[200];
        snprintf(b, sizeof b, "coral_xchg: a message of %llu bytes does not fit the %llu-byte slots of channel %d "
                 "(coral_xchg_create slot_bytes)", (unsigned long long)nbytes, (unsigned long long)x->cap[ch], ch);
        c->err = b;
        return CORAL_E_NOMEM;
    }
    const uint64_t seq = ++x->seq[ch];
    const uint64_t set_off = x->ch_off[ch] + (seq & 1) * (uint64_t)x->world * x->cap[ch];
    PushArgs a;
    a.world = x->world;
    for (int p = 0; p < x->world; p++) {
        a.slot[p] = x->peer[p] + set_off + (uint64_t)x->rank * x->cap[ch];
        a.flag[p] = reinterpret_cast<uint64_t*>(x->peer[p]) + (size_t)ch * CORAL_XCHG_MAX_RANKS + x->rank;
    }
    const uint64_t nvec = nbytes >> 4;
    const unsigned blocks = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((nvec + 1023) / 1024, (uint64_t)c->sm_count));
    const uint8_t* set = x->local + set_off;
    WaitArgs w{reinterpret_cast<const uint64_t*>(x->local) + (size_t)ch * CORAL_XCHG_MAX_RANKS, set, x->cap[ch],
               (uint64_t)(x->timeout_s * 1e9), x->d_sizes + (size_t)ch * CORAL_XCHG_MAX_RANKS, x->d_status};
    k_xchg_push<<<blocks, 256, 0, c->stream>>>(static_cast<const uint4*>(d_msg), nbytes, a, seq, x->d_done + ch, w);
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    *slots = set;
    return CORAL_OK;
}

}  // namespace coral

// In-place sum over ranks of n 8-byte words in device memory (used by the sharded coverage entry point).
int coral_xchg_sum_device(coral_ctx* c, coral_xchg* x, int ch, unsigned long long* d_words, int64_t n) {
    using namespace coral;
    if (n <= 0) return CORAL_OK;
    const uint64_t nbytes = ((uint64_t)n * 8 + 15) & ~(uint64_t)15;
    const uint8_t* set = nullptr;
    CORAL_TRY(xchg_round(c, x, ch, d_words, nbytes, &set));
    k_xchg_sum_u64<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(set, x->cap[ch], x->world, n, d_words);
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    return CORAL_OK;
}

extern "C" {

int coral_xchg_create(int device, int32_t rank, int32_t world, const uint64_t* slot_bytes, coral_xchg** out) {
    if (!out || !slot_bytes || world < 1 || world > CORAL_XCHG_MAX_RANKS || rank < 0 || rank >= world) return CORAL_E_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return CORAL_E_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return CORAL_E_CUDA;
    coral_xchg* x = new (std::nothrow) coral_xchg();
    if (!x) return CORAL_E_NOMEM;
    x->device = device; x->rank = rank; x->world = world;
    uint64_t off = coral::XCHG_FLAGS_BYTES;
    for (int ch = 0; ch < CORAL_XCHG_CHANNELS; ch++) {
        x->cap[ch] = slot_bytes[ch] ? ((slot_bytes[ch] + coral::XCHG_SLOT_HDR + 255) & ~(uint64_t)255) : 0;
        x->ch_off[ch] = off;
        off += 2 * (uint64_t)world * x->cap[ch];
    }
    x->bytes = off;
    bool ok = cudaMalloc(reinterpret_cast<void**>(&x->local), x->bytes) == cudaSuccess &&
              cudaMemset(x->local, 0, coral::XCHG_FLAGS_BYTES) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void**>(&x->d_done), CORAL_XCHG_CHANNELS * 4) == cudaSuccess &&
              cudaMemset(x->d_done, 0, CORAL_XCHG_CHANNELS * 4) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void**>(&x->d_sizes), CORAL_XCHG_CHANNELS * CORAL_XCHG_MAX_RANKS * 8) == cudaSuccess &&
              cudaMalloc(reinterpret_cast<void**>(&x->d_status), 16) == cudaSuccess &&
              cudaMemset(x->d_status, 0, 16) == cudaSuccess && cudaDeviceSynchronize() == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        if (x->local) cudaFree(x->local);
        if (x->d_done) cudaFree(x->d_done);
        if (x->d_sizes) cudaFree(x->d_sizes);
        if (x->d_status) cudaFree(x->d_status);
        delete x;
        return CORAL_E_NOMEM;
    }
    x->peer[rank] = x->local;
    if (world == 1) x->connected = true;
    *out = x;
    return CORAL_OK;
}

const char* coral_xchg_last_error(const coral_xchg* x) { return x ? x->err.c_str() : "null exchange"; }

// The handle the peers open (CUDA IPC, 64 bytes) and, for ranks that live in the SAME process (thread ranks: a handle
// cannot be opened by the process that exported it), the window's plain device address.
int coral_xchg_handle(coral_xchg* x, uint8_t* handle, uint64_t* local_ptr) {
    if (!x || !handle) return CORAL_E_ARG;
    cudaSetDevice(x->device);
    cudaIpcMemHandle_t h;
    static_assert(sizeof(h) == CORAL_XCHG_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
    const cudaError_t e = cudaIpcGetMemHandle(&h, x->local);
    if (e != cudaSuccess) return coral::xchg_fail(x, "cudaIpcGetMemHandle", e);
    memcpy(handle, &h, sizeof h);
    if (local_ptr) *local_ptr = (uint64_t)reinterpret_cast<uintptr_t>(x->local);
    return CORAL_OK;
}

// handles: world x 64 bytes, in rank order (the caller all-gathers them once, over any transport);
// same_process_ptrs (nullable): non-zero entry r = rank r's window is that device address in THIS process.
// Every rank must have created its window before any rank connects, and no rank may start an exchange before every
// rank has connected (a host-side barrier of the caller on both sides).
int coral_xchg_connect(coral_xchg* x, const uint8_t* handles, const uint64_t* same_process_ptrs) {
    if (!x || (!handles && !same_process_ptrs)) return CORAL_E_ARG;
    cudaSetDevice(x->device);
    for (int r = 0; r < x->world; r++) {
        if (r == x->rank) continue;
        if (same_process_ptrs && same_process_ptrs[r]) {
            x->peer[r] = reinterpret_cast<uint8_t*>((uintptr_t)same_process_ptrs[r]);
            continue;
        }
        if (!handles) { x->err = "no handle for a peer of another process"; return CORAL_E_ARG; }
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * CORAL_XCHG_HANDLE_BYTES, sizeof h);
        void* p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return coral::xchg_fail(x, "cudaIpcOpenMemHandle (peer window)", e);
        x->peer[r] = static_cast<uint8_t*>(p);
        x->opened[r] = true;
    }
    x->connected = true;
    return CORAL_OK;
}

int coral_xchg_set_timeout(coral_xchg* x, double seconds) {
    if (!x || !(seconds > 0)) return CORAL_E_ARG;
    x->timeout_s = seconds;
    return CORAL_OK;
}

// Number of waits that timed out since the last call (synchronises the device; 0 = every exchange completed).
int coral_xchg_timeouts(coral_xchg* x, int32_t* n) {
    if (!x || !n) return CORAL_E_ARG;
    cudaSetDevice(x->device);
    int32_t v = 0;
    cudaError_t e = cudaMemcpy(&v, x->d_status, 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && v) e = cudaMemset(x->d_status, 0, 4);
    if (e != cudaSuccess) return coral::xchg_fail(x, "reading the exchange status", e);
    *n = v;
    return CORAL_OK;
}

// The caller runs a barrier over all ranks BEFORE this (a peer may still be writing into the window).
void coral_xchg_destroy(coral_xchg* x) {
    if (!x) return;
    cudaSetDevice(x->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < x->world; r++)
        if (x->opened[r]) cudaIpcCloseMemHandle(x->peer[r]);
    cudaFree(x->local);
    cudaFree(x->d_done);
    cudaFree(x->d_sizes);
    cudaFree(x->d_status);
    delete x;
}

int coral_xchg_allgather(coral_ctx* c, coral_xchg* x, int32_t channel, uint64_t d_msg, uint64_t nbytes, int32_t lane,
                         uint64_t* d_buf, uint64_t* starts, uint64_t* sizes) {
    using namespace coral;
    if (!c || !x || !d_buf || !starts || !sizes) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    std::unique_ptr<PassLane> guard;
    if (lane) guard.reset(new PassLane(c));
    if (c->pass.active) {
        uint64_t k = pass_hash(PASS_H0, "xchg_allgather", 14);
        const uint64_t a[3] = {(uint64_t)channel, nbytes, (uint64_t)x->world};
        pass_call(c, pass_hash(k, a, sizeof a));
    }
    const uint8_t* set = nullptr;
    CORAL_TRY(xchg_round(c, x, channel, reinterpret_cast<const void*>((uintptr_t)d_msg), nbytes, &set));
    CORAL_TRY(pass_resolve_n(c, sizes, x->d_sizes + (size_t)channel * CORAL_XCHG_MAX_RANKS, x->world));
    for (int r = 0; r < x->world; r++) {
        if (sizes[r] == ~0ull) {
            c->err = "coral_xchg_allgather: rank " + std::to_string(r) + " did not deliver within the timeout";
            return CORAL_E_STATE;
        }
        if (sizes[r] + XCHG_SLOT_HDR > x->cap[channel]) { c->err = "coral_xchg_allgather: corrupt slot header"; return CORAL_E_FORMAT; }
        starts[r] = (uint64_t)r * x->cap[channel] + XCHG_SLOT_HDR;
    }
    *d_buf = (uint64_t)reinterpret_cast<uintptr_t>(set);
    return CORAL_OK;
}

// In-place integer all-reduce (sum) of n int64 values of the caller (host array): what the sharded engine uses outside
// the coverage entry point (e.g. partial counters).  Synchronises.
int coral_xchg_allreduce_i64(coral_ctx* c, coral_xchg* x, int32_t channel, int64_t* values, int32_t n) {
    using namespace coral;
    if (!c || !x || n < 0 || (n > 0 && !values)) return CORAL_E_ARG;
    if (c->pass.active) { c->err = "coral_xchg_allreduce_i64 inside a pass (use coral_region_coverage_sharded)"; return CORAL_E_STATE; }
    if (n == 0) return CORAL_OK;
    cudaSetDevice(c->device);
    CORAL_CUDA(c, c->s7.ensure((size_t)n * 8 + 16));
    CORAL_CUDA(c, cudaMemsetAsync(c->s7.p, 0, (size_t)n * 8 + 16, c->stream));
    CORAL_CUDA(c, cudaMemcpyAsync(c->s7.p, values, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    CORAL_TRY(coral_xchg_sum_device(c, x, channel, c->s7.as<unsigned long long>(), n));
    CORAL_CUDA(c, cudaMemcpyAsync(values, c->s7.p, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    int32_t bad = 0;
    if (coral_xchg_timeouts(x, &bad) != CORAL_OK || bad) { c->err = "coral_xchg_allreduce_i64: a peer did not deliver within the timeout"; return CORAL_E_STATE; }
    return CORAL_OK;
}

}  // extern "C"
