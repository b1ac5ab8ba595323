// coral_b200/csrc/rows.cu — the multi-GPU exchange boundary of `fetch`.
//
// The reference unions the SA strings of ALL records of a read name in file order and takes
// read_length from the first record with flag < 256 anywhere in the file
// (infer_breakpoint_graph.py:142-151).  When the file is sharded across GPUs the records of one
// chimeric read live on several shards, so every rank exports one small row per relevant record
// (name hash, flag, MAPQ, l_seq, NM, SA text), the rows are all-gathered in rank (= file) order and
// every rank loads the gathered rows into a second, "rows-only" ctx on which coral_fetch runs
// unchanged.  Only coral_fetch, the candidate and the cluster entry points are valid on such a ctx.
#include <algorithm>
#include <vector>

#include "ctx.h"

namespace coral {
__global__ void k_rows_sa(int64_t n, const uint64_t* __restrict__ off, uint32_t* __restrict__ sa_off,
                          uint32_t* __restrict__ sa_len) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t len = off[i + 1] - off[i];
    // the SA text of row i starts 1 byte into its slot, so "sa_off != 0" keeps meaning "has an SA tag"
    sa_off[i] = len > 1 ? 1u : 0u;
    sa_len[i] = len > 1 ? (uint32_t)(len - 1) : 0u;
}
}  // namespace coral

static int h2d(coral_ctx* c, DevBuf& b, const void* src, size_t bytes) {
    CORAL_CUDA(c, b.ensure(std::max<size_t>(bytes, 4)));
    if (bytes) CORAL_CUDA(c, cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
    return CORAL_OK;
}

extern "C" {

int coral_get_record_keys(coral_ctx* c, uint64_t* name_lo, uint64_t* name_hi, uint32_t* sa_off, uint32_t* sa_len,
                          int32_t* nm) {
    if (!c) return CORAL_E_ARG;
    if (!c->loaded) { c->err = "coral_get_record_keys before coral_load_records"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(c));
    cudaSetDevice(c->device);
    const size_t n = (size_t)c->nrec;
    if (n == 0) return CORAL_OK;
    RecordTable& t = c->rt;
    if (name_lo) CORAL_CUDA(c, cudaMemcpyAsync(name_lo, t.name_lo.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
    if (name_hi) CORAL_CUDA(c, cudaMemcpyAsync(name_hi, t.name_hi.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
    if (sa_off) CORAL_CUDA(c, cudaMemcpyAsync(sa_off, t.sa_off.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (sa_len) CORAL_CUDA(c, cudaMemcpyAsync(sa_len, t.sa_len.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (nm) CORAL_CUDA(c, cudaMemcpyAsync(nm, t.nm.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}

int coral_load_rows(coral_ctx* c, int64_t n, const uint8_t* text, uint64_t text_bytes, const uint64_t* text_off,
                    const uint64_t* name_lo, const uint64_t* name_hi, const uint32_t* flag, const uint32_t* mapq,
                    const uint32_t* l_seq, const int32_t* nm) {
    using namespace coral;
    if (!c || n < 0 || (n > 0 && (!text_off || !name_lo || !name_hi || !flag || !mapq || !l_seq || !nm))) return CORAL_E_ARG;
    cudaSetDevice(c->device);
    CORAL_TRY(coral_wait_scan(c));
    c->loaded = false;
    c->rows_only = true;
    for (int64_t i = 0; i < n; i++)
        if (text_off[i + 1] < text_off[i] || text_off[i + 1] > text_bytes) { c->err = "row text offsets out of range"; return CORAL_E_ARG; }
    // slot i of the device text = one pad byte + the SA text of row i (empty slot: the row has no SA tag)
    std::vector<uint8_t> blob;
    std::vector<uint64_t> off((size_t)n + 1, 0);
    std::vector<uint32_t> fm((size_t)std::max<int64_t>(n, 1));
    for (int64_t i = 0; i < n; i++) {
        off[i] = blob.size();
        const uint64_t len = text_off[i + 1] - text_off[i];
        if (len) {
            blob.push_back(0);
            blob.insert(blob.end(), text + text_off[i], text + text_off[i + 1]);
        }
        fm[i] = (flag[i] & 0xffffu) | ((mapq[i] & 255u) << 16);
    }
    off[n] = blob.size();
    blob.resize(blob.size() + 64, 0);
    CORAL_TRY(h2d(c, c->recs_owned, blob.data(), blob.size()));
    c->d_recs = c->recs_owned.as<uint8_t>();
    c->nbytes = off[n];
    c->nrec = n;
    CORAL_TRY(h2d(c, c->d_off, off.data(), ((size_t)n + 1) * 8));
    RecordTable& t = c->rt;
    const size_t n4 = (size_t)n * 4, n8 = (size_t)n * 8;
    CORAL_TRY(h2d(c, t.name_lo, name_lo, n8));
    CORAL_TRY(h2d(c, t.name_hi, name_hi, n8));
    CORAL_TRY(h2d(c, t.flag_mapq, fm.data(), n4));
    CORAL_TRY(h2d(c, t.lseq, l_seq, n4));
    CORAL_TRY(h2d(c, t.nm, nm, n4));
    CORAL_CUDA(c, t.tid.ensure(std::max<size_t>(n4, 4)));
    CORAL_CUDA(c, cudaMemsetAsync(t.tid.p, 0xFF, std::max<size_t>(n4, 4), c->stream));   // tid -1: rows carry no placement
    DevBuf* zero4[] = {&t.pos, &t.endpos, &t.ncig, &t.infer, &t.mbases, &t.nonacgt, &t.fbs, &t.lbe, &t.ngap,
                       &t.sa_off, &t.sa_len};
    for (DevBuf* b : zero4) {
        CORAL_CUDA(c, b->ensure(std::max<size_t>(n4, 4)));
        CORAL_CUDA(c, cudaMemsetAsync(b->p, 0, std::max<size_t>(n4, 4), c->stream));
    }
    if (n > 0) {
        const unsigned blocks = (unsigned)((n + 255) / 256);
        k_rows_sa<<<blocks, 256, 0, c->stream>>>(n, c->d_off.as<uint64_t>(), t.sa_off.as<uint32_t>(), t.sa_len.as<uint32_t>());
        c->n_launches++;
        CORAL_CUDA(c, cudaGetLastError());
    }
    c->gaps.n = 0;
    CORAL_CUDA(c, c->status.ensure(64));            // buffers the record scan would have created
    CORAL_CUDA(c, cudaMemsetAsync(c->status.p, 0, 64, c->stream));
    CORAL_CUDA(c, c->gaps.counter.ensure(16));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    c->loaded = true;
    return CORAL_OK;
}

}  // extern "C"

// ---- device-side views (the packed device-side exchange is in exchange.cu) ---------------------------------------
extern "C" {

// Device addresses of the record buffer, the offsets and the table columns the exchange reads (valid until the next
// load / rescan of this ctx; read-only for the caller).  Order: recs, offsets, name_lo, name_hi, flag_mapq, l_seq, nm,
// sa_off, sa_len.
int coral_device_view(coral_ctx* c, uint64_t* ptrs, int32_t n_ptrs) {
    if (!c || !ptrs || n_ptrs < 9) return CORAL_E_ARG;
    if (!c->loaded) { c->err = "coral_device_view before coral_load_records"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(c));
    RecordTable& t = c->rt;
    const void* v[9] = {c->d_recs, c->d_off.p, t.name_lo.p, t.name_hi.p, t.flag_mapq.p, t.lseq.p, t.nm.p, t.sa_off.p, t.sa_len.p};
    for (int i = 0; i < 9; i++) ptrs[i] = (uint64_t)reinterpret_cast<uintptr_t>(v[i]);
    return CORAL_OK;
}

}  // extern "C"
