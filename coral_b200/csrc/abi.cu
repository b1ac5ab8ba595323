// coral_b200/csrc/abi.cu — extern "C" entry points (include/coral_b200.h).
#include <algorithm>
#include <cstring>
#include <new>

#include "ctx.h"
#include "relops.cuh"

extern "C" {

void coral_default_params(coral_params* p) {
    p->min_del_len = 600;
    p->min_bp_match_cutoff = 100;
    p->min_mapq = 20;
    p->gap_mapq = 10;
    p->gap = 100;
    p->bp_distance_cutoff = 2000;
}

int coral_create(int device, coral_ctx** out) {
    if (!out) return CORAL_E_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return CORAL_E_CUDA;  // no CPU fallback
    if (device < 0 || device >= ndev) return CORAL_E_ARG;
    coral_ctx* c = new (std::nothrow) coral_ctx();
    if (!c) return CORAL_E_NOMEM;
    c->device = device;
    coral_default_params(&c->prm);
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete c;
        return CORAL_E_CUDA;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
    *out = c;
    return CORAL_OK;
}

void coral_destroy(coral_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->scan_stream) { cudaStreamSynchronize(c->scan_stream); cudaStreamDestroy(c->scan_stream); }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->scan_pending) coral_mem::scan_joined();
    c->scan_pending = false;
    // (an adopted stream belongs to another ctx that may already be gone: it is not touched here; the caller of
    // coral_adopt_stream destroys this ctx before, or synchronises, the owner — sharded.py keeps the owner alive)
    if (c->stream && c->owns_stream) cudaStreamSynchronize(c->stream);
    if (c->host_registered) cudaHostUnregister(c->host_registered);
    DevBuf* all[] = {&c->d_chr_rank, &c->d_ref_hash, &c->recs_owned, &c->d_off, &c->d_ordinal, &c->status, &c->scan_status, &c->sa_side, &c->sa_side_off, &c->xmsg[0], &c->xmsg[1], &c->xmsg[2], &c->xmsg[3], &c->scan_order, &c->pack_buf, &c->nm_part,
                     &c->s0, &c->s1, &c->s2, &c->s3, &c->s4, &c->s5, &c->s6, &c->s7, &c->cub_tmp,
                     &c->rt.tid, &c->rt.pos, &c->rt.endpos, &c->rt.flag_mapq, &c->rt.lseq, &c->rt.ncig, &c->rt.infer,
                     &c->rt.mbases, &c->rt.nonacgt, &c->rt.fbs, &c->rt.lbe, &c->rt.name_lo, &c->rt.name_hi,
                     &c->rt.sa_off, &c->rt.sa_len, &c->rt.nm, &c->rt.ngap, &c->rt.gap_base,
                     &c->gaps.rec, &c->gaps.k, &c->gaps.next_start, &c->gaps.prev_end, &c->gaps.counter,
                     &c->chim.read_first_rec, &c->chim.read_len, &c->chim.read_status, &c->chim.aln_ptr, &c->chim.qs,
                     &c->chim.qe, &c->chim.tid, &c->chim.rs, &c->chim.re, &c->chim.strand, &c->chim.mapq, &c->chim.nm};
    for (DevBuf* b : all) b->release();
    if (c->pass.d_flag) cudaFree(c->pass.d_flag);
    for (cudaEvent_t e : c->pass.events) cudaEventDestroy(e);
    for (auto& in : c->pass.inputs) if (in.p) coral_mem::release(in.p, in.cap);
    if (c->pass.lane_stream) { cudaStreamSynchronize(c->pass.lane_stream); cudaStreamDestroy(c->pass.lane_stream); }
    if (c->pass.ev_lane_begin) cudaEventDestroy(c->pass.ev_lane_begin);
    if (c->pass.ev_lane_done) cudaEventDestroy(c->pass.ev_lane_done);
    if (c->pass.side_stream) { cudaStreamSynchronize(c->pass.side_stream); cudaStreamDestroy(c->pass.side_stream); }
    for (cudaEvent_t e : c->pass.side_events) cudaEventDestroy(e);
    for (DevBuf& b : c->lane_s) b.release();
    c->lane_cub_tmp.release(); c->lane_status.release();
    for (auto& ch : c->pass.pinned) cudaFreeHost(ch.p);
    for (DevBuf& b : c->sw) b.release();
    for (DevBuf& b : c->cw) b.release();
    for (DevBuf& b : c->pw) b.release();
    for (DevBuf& b : c->ing_w) b.release();
    c->ing_comp.release(); c->ing_stream.release(); c->ing_recs.release();
    DevBuf* tables[] = {&c->paths.rec, &c->paths.iv, &c->paths.first, &c->paths.last,
                        &c->sdel.rec, &c->sdel.ai, &c->sdel.next_start, &c->sdel.prev_end, &c->sdel.first_start, &c->sdel.last_end, &c->sdel.mapq,
                        &c->cand.read, &c->cand.i, &c->cand.j, &c->cand.chr1, &c->cand.pos1, &c->cand.chr2, &c->cand.pos2, &c->cand.rgap,
                        &c->cand.mq1, &c->cand.mq2, &c->cand.o1, &c->cand.o2, &c->cand.flag};
    for (DevBuf* b : tables) b->release();
    coral_release_tables(c);
    for (int i = 0; i < CORAL_N_STAGES; i++) {
        if (c->ev_beg[i]) cudaEventDestroy(c->ev_beg[i]);
        if (c->ev_end[i]) cudaEventDestroy(c->ev_end[i]);
    }
    if (c->owns_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

const char* coral_last_error(const coral_ctx* c) { return c ? c->err.c_str() : "null ctx"; }

int coral_set_params(coral_ctx* c, const coral_params* p) {
    if (!c || !p) return CORAL_E_ARG;
    c->prm = *p;
    return CORAL_OK;
}

int coral_set_option(coral_ctx* c, const char* name, int64_t value) {
    if (!c || !name) return CORAL_E_ARG;
    if (std::string(name) == "full_name_sort") { c->force_full_name_sort = value != 0; return CORAL_OK; }
    if (std::string(name) == "fetch_part_world") { if (value < 1 || value > 16) return CORAL_E_ARG; c->fetch_part_world = (int32_t)value; c->pass.invalidate(); return CORAL_OK; }
    if (std::string(name) == "fetch_part_rank") { if (value < 0 || value > 15) return CORAL_E_ARG; c->fetch_part_rank = (int32_t)value; c->pass.invalidate(); return CORAL_OK; }
    if (std::string(name) == "scan_order") { if (value < 0 || value > 2) return CORAL_E_ARG; c->scan_order_mode = (int)value; c->scan_order_valid = false; return CORAL_OK; }
    if (std::string(name) == "pass_fork") { c->pass.fork = value != 0; return CORAL_OK; }
    if (std::string(name) == "pass_scan_reserve") { c->pass.scan_reserve = value; return CORAL_OK; }
    if (std::string(name) == "pass_overlap_scan") { c->pass.overlap_scan = value != 0; c->pass.invalidate(); return CORAL_OK; }
    c->err = std::string("unknown option: ") + name;
    return CORAL_E_ARG;
}

int coral_set_references(coral_ctx* c, const char* names, int32_t n, const int32_t* chr_rank) {
    if (!c || !names || n < 0 || !chr_rank) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    c->ref_names.clear();
    const char* p = names;
    for (int32_t i = 0; i < n; i++) {
        c->ref_names.emplace_back(p);
        p += c->ref_names.back().size() + 1;
    }
    c->chr_rank_h.assign(chr_rank, chr_rank + n);
    CORAL_CUDA(c, c->d_chr_rank.ensure((size_t)std::max(n, 1) * 4));
    CORAL_CUDA(c, cudaMemcpyAsync(c->d_chr_rank.p, chr_rank, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CORAL_TRY(coral_upload_ref_hashes(c));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}

int coral_load_records(coral_ctx* c, const uint8_t* recs, uint64_t nbytes, const uint64_t* offsets, int64_t n,
                       const uint64_t* global_ordinal, int on_device) {
    if (!c || n < 0 || (n > 0 && (!recs || !offsets))) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    if (c->scan_pending) {                 // a previous asynchronous load is still scanning: join it first
        c->scan_pending = false;
        cudaStreamSynchronize(c->scan_stream);
        coral_mem::scan_joined();
    }
    c->loaded = false;
    c->rows_only = false;
    for (auto& q : c->rows_fill_ptr) q = nullptr;     // the scan rewrites the columns coral_unpack_rows keeps constant
    c->scan_order_valid = false;
    c->sa_side_valid = false;
    c->pass.invalidate();                  // recorded sizes belong to the previous records
    if (c->pass.active) { c->err = "coral_load_records inside a pass"; return CORAL_E_STATE; }
    for (int64_t i = 0; i < n; i++) {
        if (offsets[i] & 15) { c->err = "record offsets must be 16-byte aligned"; return CORAL_E_ARG; }
        if (offsets[i] + 36 > nbytes || offsets[i + 1] < offsets[i] + 36 || offsets[i + 1] > nbytes) {
            c->err = "record offsets out of range";
            return CORAL_E_ARG;
        }
    }
    if (n > 0 && (offsets[n] & 15)) { c->err = "record offsets must be 16-byte aligned (including the end offset)"; return CORAL_E_ARG; }
    if (n >= ((int64_t)1 << 32) - 65536) { c->err = "more than 2^32 - 65536 records per ctx"; return CORAL_E_ARG; }
    if (c->host_registered) {             // a previous CORAL_RECORDS_HOST_MAPPED load page-locked the caller's buffer
        cudaHostUnregister(c->host_registered);
        c->host_registered = nullptr;
    }
    if (on_device == CORAL_RECORDS_DEVICE) {
        if (reinterpret_cast<uintptr_t>(recs) & 15) { c->err = "device record buffer must be 16-byte aligned"; return CORAL_E_ARG; }
        c->d_recs = recs;
    } else if (on_device == CORAL_RECORDS_HOST_MAPPED || on_device == CORAL_RECORDS_HOST_MAPPED_ASYNC ||
               on_device == CORAL_RECORDS_HOST_MAPPED_LAZY) {
        // Records stay in page-locked host memory; every kernel reads them in place over PCIe (the scan only
        // fetches the fields it needs, so QUAL never crosses the link).
        if (reinterpret_cast<uintptr_t>(recs) & 15) { c->err = "host record buffer must be 16-byte aligned"; return CORAL_E_ARG; }
        cudaPointerAttributes at;
        bool mapped = cudaPointerGetAttributes(&at, recs) == cudaSuccess && at.type == cudaMemoryTypeHost &&
                      at.devicePointer != nullptr;
        cudaGetLastError();
        if (!mapped) {
            // the ABI promises the kernels 64 readable bytes past `nbytes` (include/coral_b200.h): page-lock them too, else
            // a record ending just before a page boundary leaves the slack on an unregistered page (device fault)
            if (n > 0) CORAL_CUDA(c, cudaHostRegister(const_cast<uint8_t*>(recs), nbytes + 64, cudaHostRegisterMapped | cudaHostRegisterPortable));
            c->host_registered = const_cast<uint8_t*>(recs);
            if (n > 0) CORAL_CUDA(c, cudaPointerGetAttributes(&at, recs));
        }
        void* dp = nullptr;
        if (n > 0) CORAL_CUDA(c, cudaHostGetDevicePointer(&dp, const_cast<uint8_t*>(recs), 0));
        c->d_recs = static_cast<const uint8_t*>(dp);
    } else if (on_device != CORAL_RECORDS_HOST_COPY) {
        c->err = "unknown record placement";
        return CORAL_E_ARG;
    } else {
        CORAL_CUDA(c, c->recs_owned.ensure(nbytes + 64));
        CORAL_CUDA(c, cudaMemcpyAsync(c->recs_owned.p, recs, nbytes, cudaMemcpyHostToDevice, c->stream));
        CORAL_CUDA(c, cudaMemsetAsync((uint8_t*)c->recs_owned.p + nbytes, 0, 64, c->stream));
        c->d_recs = c->recs_owned.as<uint8_t>();
    }
    c->nbytes = nbytes;
    c->nrec = n;
    CORAL_CUDA(c, c->d_off.ensure((size_t)(n + 1) * 8));
    if (n > 0) CORAL_CUDA(c, cudaMemcpyAsync(c->d_off.p, offsets, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    c->has_ordinal = global_ordinal != nullptr;
    if (global_ordinal && n > 0) {
        CORAL_CUDA(c, c->d_ordinal.ensure((size_t)n * 8));
        CORAL_CUDA(c, cudaMemcpyAsync(c->d_ordinal.p, global_ordinal, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    }
    c->scan_lazy = false;
    if (on_device == CORAL_RECORDS_HOST_MAPPED_ASYNC) CORAL_TRY(coral_start_scan_async(c, true));
    else if (on_device == CORAL_RECORDS_HOST_MAPPED_LAZY) CORAL_TRY(coral_start_scan_async(c, false));
    else CORAL_TRY(coral_run_scan(c));
    c->loaded = true;
    return CORAL_OK;
}

void coral_release_cached_memory(void) { coral_mem::trim(); }

// Page-locked host memory for result arrays: inside a pipelined pass a result copy into such an array is a direct DMA
// (no staging buffer, no host memcpy when the pass ends).
int coral_host_alloc(uint64_t nbytes, void** out) {
    if (!out) return CORAL_E_ARG;
    *out = nullptr;
    return cudaHostAlloc(out, (size_t)std::max<uint64_t>(nbytes, 1), cudaHostAllocPortable) == cudaSuccess ? CORAL_OK : CORAL_E_NOMEM;
}
void coral_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- streams of a ctx (for callers that enqueue their own work — NCCL collectives — between entry points of a pass) --
uint64_t coral_stream_handle(coral_ctx* c, int which) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    if (which == 0) return (uint64_t)reinterpret_cast<uintptr_t>(c->pass.in_lane ? c->pass.lane_stream : c->stream);
    PassState& P = c->pass;
    if (!P.lane_stream) {
        if (cudaStreamCreateWithFlags(&P.lane_stream, cudaStreamNonBlocking) != cudaSuccess) return 0;
        cudaEventCreateWithFlags(&P.ev_lane_begin, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&P.ev_lane_done, cudaEventDisableTiming);
    }
    return (uint64_t)reinterpret_cast<uintptr_t>(P.in_lane ? c->stream : P.lane_stream);
}

int coral_adopt_stream(coral_ctx* c, uint64_t handle) {
    if (!c || !handle) return CORAL_E_ARG;
    if (c->pass.active) { c->err = "coral_adopt_stream inside a pass"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->owns_stream && c->stream) cudaStreamDestroy(c->stream);
    c->stream = reinterpret_cast<cudaStream_t>((uintptr_t)handle);
    c->owns_stream = false;
    return CORAL_OK;
}

// A host value of the CALLER that must be the same in a replayed pass (e.g. the message sizes of the other ranks of an
// exchange): recorded while the pass records, handed back while it replays.
int coral_pass_note(coral_ctx* c, uint64_t value, uint64_t* out, int32_t* replayed) {
    if (!c || !out) return CORAL_E_ARG;
    PassState& P = c->pass;
    if (!P.active) { *out = value; if (replayed) *replayed = 0; return CORAL_OK; }
    if (P.replay && P.cur < P.vals.size()) {
        *out = P.vals[P.cur++];
        if (replayed) *replayed = 1;
        return CORAL_OK;
    }
    if (P.replay) { P.replay = false; P.vals.resize(P.cur); P.keys.resize(P.kcur); }
    P.vals.push_back(value);
    P.cur = P.vals.size();
    *out = value;
    if (replayed) *replayed = 0;
    return CORAL_OK;
}
int coral_pass_replaying(const coral_ctx* c) { return (c && c->pass.active && c->pass.replay) ? 1 : 0; }

// ---- pipelined pass (PassState, ctx.h) -------------------------------------------------------------------------------
int coral_pass_begin(coral_ctx* c) {
    if (!c) return CORAL_E_ARG;
    PassState& P = c->pass;
    if (P.active) { c->err = "coral_pass_begin inside a pass"; return CORAL_E_STATE; }
    // (a ctx without records may open a pass: the rows-only ctx of the multi-GPU engine is filled inside it by
    // coral_unpack_rows; the entry points check what they need themselves)
    cudaSetDevice(c->device);
    if (c->loaded) CORAL_TRY(coral_wait_scan(c));
    if (!P.d_flag) CORAL_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&P.d_flag), 64));
    CORAL_CUDA(c, cudaMemsetAsync(P.d_flag, 0, 64, c->stream));
    P.active = true;
    P.replay = P.complete && !P.vals.empty();
    if (!P.replay) { P.vals.clear(); P.keys.clear(); }
    P.complete = false;
    P.cur = P.kcur = 0;
    P.copies.clear();
    P.after.clear();
    P.n_events = 0;
    P.in_cur = 0;
    P.lane_started = false;
    P.pending.clear();
    P.n_side_events = 0;
    P.side_on = P.side_used = false;
    P.side_mask = 0;
    for (auto& ch : P.pinned) ch.used = 0;
    return CORAL_OK;
}

int coral_pass_end(coral_ctx* c, int32_t* replayed) {
    if (!c) return CORAL_E_ARG;
    PassState& P = c->pass;
    if (!P.active) { c->err = "coral_pass_end outside a pass"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    int join_rc = CORAL_OK;
    if (c->scan_pending) join_rc = coral_wait_scan(c);     // a rescan nothing joined: its verdict belongs to this pass
    if (P.lane_started) {                                  // the second lane joins the ctx's stream
        cudaEventRecord(P.ev_lane_done, P.lane_stream);
        cudaStreamWaitEvent(c->stream, P.ev_lane_done, 0);
        P.lane_started = false;
    }
    coral::pass_flush_checks(c);
    P.active = false;
    uint32_t bad = 0;
    uint32_t* bad_pinned = reinterpret_cast<uint32_t*>(coral::pass_stage(c, 64));
    cudaError_t e = bad_pinned ? cudaMemcpyAsync(bad_pinned, P.d_flag, 4, cudaMemcpyDeviceToHost, c->stream) : cudaErrorMemoryAllocation;
    // groups of result copies whose event has fired go to the caller's arrays now, while the device still works on
    // the later stages (valid only if the pass does not diverge: a divergent pass is reported below and its arrays are
    // garbage either way)
    size_t done = 0;
    if (e == cudaSuccess) {
        for (; done < P.copies.size(); done++) {
            const auto& cp = P.copies[done];
            if (cp.group < 0) break;
            if (cudaEventSynchronize(P.events[cp.group]) != cudaSuccess) break;
            memcpy(cp.dst, cp.src, cp.bytes);
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess && P.side_used) e = cudaStreamSynchronize(P.side_stream);
    if (e == cudaSuccess) bad = *bad_pinned;
    if (e != cudaSuccess) {
        P.invalidate(); P.copies.clear(); P.after.clear();
        c->err = std::string("coral_pass_end: ") + cudaGetErrorString(e);
        return CORAL_E_CUDA;
    }
    const bool was_replay = P.replay;
    if (replayed) *replayed = was_replay ? 1 : 0;
    if (join_rc != CORAL_OK) { P.invalidate(); P.copies.clear(); P.after.clear(); return join_rc; }
    if (bad != 0) {          // the device disagreed with a replayed size: nothing of this pass can be trusted
        P.invalidate(); P.copies.clear(); P.after.clear();
        c->err = "pipelined pass diverged from the recorded one (" + std::to_string(bad) +
                 " size(s) differ): the records behind a CORAL_RECORDS_DEVICE load changed? Run the pass again.";
        return CORAL_E_STATE;
    }
    for (size_t k = done; k < P.copies.size(); k++) memcpy(P.copies[k].dst, P.copies[k].src, P.copies[k].bytes);
    P.copies.clear();
    int rc = CORAL_OK;
    for (auto& fn : P.after) { const int r = fn(); if (r != CORAL_OK && rc == CORAL_OK) rc = r; }
    P.after.clear();
    if (was_replay) { P.vals.resize(P.cur); P.keys.resize(P.kcur); P.n_replayed++; }
    P.complete = rc == CORAL_OK;
    if (!P.complete) P.invalidate();
    return rc;
}

int coral_rescan(coral_ctx* c) {
    if (!c) return CORAL_E_ARG;
    if (!c->loaded || c->rows_only) { c->err = "coral_rescan before coral_load_records"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    if (c->pass.active) {
        coral::pass_call(c, coral::pass_hash(coral::PASS_H0, c->pass.overlap_scan ? "rescan+" : "rescan", 7));
        if (c->pass.overlap_scan) return coral_start_scan_overlapped(c);
    }
    return coral_run_scan(c);
}

float coral_stage_ms(coral_ctx* c, int stage) {
    if (!c || stage < 0 || stage >= CORAL_N_STAGES || !c->ev_valid[stage]) return -1.f;
    cudaSetDevice(c->device);
    float ms = -1.f;
    if (cudaEventSynchronize(c->ev_end[stage]) != cudaSuccess) return -1.f;
    if (cudaEventElapsedTime(&ms, c->ev_beg[stage], c->ev_end[stage]) != cudaSuccess) return -1.f;
    return ms;
}

// All stage times of the most recent runs in one call (out[CORAL_N_STAGES], < 0 where a stage has not run).
int coral_stage_ms_all(coral_ctx* c, float* out) {
    if (!c || !out) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    for (int s = 0; s < CORAL_N_STAGES; s++) {
        out[s] = -1.f;
        if (!c->ev_valid[s]) continue;
        if (cudaEventSynchronize(c->ev_end[s]) != cudaSuccess) continue;
        float ms = -1.f;
        if (cudaEventElapsedTime(&ms, c->ev_beg[s], c->ev_end[s]) == cudaSuccess) out[s] = ms; else cudaGetLastError();
    }
    return CORAL_OK;
}

// Time from one stage event to another (begin: which = 0, end: which = 1), across the ctx's streams: where the stages
// of a pipelined pass sit relative to the scan running next to them.  < 0 when an event is missing.
float coral_stage_offset_ms(coral_ctx* c, int stage_a, int which_a, int stage_b, int which_b) {
    if (!c || stage_a < 0 || stage_a >= CORAL_N_STAGES || stage_b < 0 || stage_b >= CORAL_N_STAGES) return -1.f;
    if (!c->ev_valid[stage_a] || !c->ev_valid[stage_b]) return -1.f;
    cudaSetDevice(c->device);
    cudaEvent_t a = which_a ? c->ev_end[stage_a] : c->ev_beg[stage_a], b = which_b ? c->ev_end[stage_b] : c->ev_beg[stage_b];
    float ms = -1.f;
    if (cudaEventSynchronize(a) != cudaSuccess || cudaEventSynchronize(b) != cudaSuccess) return -1.f;
    if (cudaEventElapsedTime(&ms, a, b) != cudaSuccess) { cudaGetLastError(); return -1.f; }
    return ms;
}
void coral_reset_stage_times(coral_ctx* c) {
    if (c) for (int s = 0; s < CORAL_N_STAGES; s++) c->ev_valid[s] = false;
}

int64_t coral_n_records(const coral_ctx* c) { return c ? c->nrec : -1; }
uint64_t coral_launch_count(const coral_ctx* c) { return c ? c->n_launches : 0; }

int coral_region_coverage(coral_ctx* c, const int32_t* tid, const int32_t* start, const int32_t* stop, int32_t n,
                          int64_t* base_sum, int64_t* rec_count) {
    if (!c) return CORAL_E_ARG;
    if (c->rows_only) { c->err = "coral_region_coverage on a rows-only ctx"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    return coral_run_region_coverage(c, tid, start, stop, n, base_sum, rec_count);
}

// Sharded form: the sums over ALL shards of the file (in-place integer all-reduce of the accumulators over peer
// memory between the kernels and the result copy); every rank calls it with the same regions.
int coral_region_coverage_sharded(coral_ctx* c, coral_xchg* x, int32_t channel, const int32_t* tid, const int32_t* start,
                                  const int32_t* stop, int32_t n, int64_t* base_sum, int64_t* rec_count) {
    if (!c || !x) return CORAL_E_ARG;
    if (c->rows_only) { c->err = "coral_region_coverage_sharded on a rows-only ctx"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    return coral_run_region_coverage(c, tid, start, stop, n, base_sum, rec_count, x, channel);
}

}  // extern "C"
