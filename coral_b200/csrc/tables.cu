// coral_b200/csrc/tables.cu — host<->device table plumbing for the C ABI.
#include <vector>

#include "ctx.h"
#include "dev_util.cuh"

int coral_upload_ref_hashes(coral_ctx* c) {
    using namespace coral;
    const size_t n = c->ref_names.size();
    std::vector<uint64_t> h(2 * std::max<size_t>(n, 1));
    for (size_t i = 0; i < n; i++) {
        const std::string& s = c->ref_names[i];
        uint64_t a = 0, b = 0;
        for (size_t k = 0; k < s.size(); k++) {
            a += hterm1((uint32_t)k, (uint8_t)s[k]);
            b += hterm2((uint32_t)k, (uint8_t)s[k]);
        }
        h[i] = a + mix64(s.size());
        h[n + i] = b + mix64(s.size() ^ 0xABCDull);
    }
    CORAL_CUDA(c, c->d_ref_hash.ensure(h.size() * 8));
    CORAL_CUDA(c, cudaMemcpyAsync(c->d_ref_hash.p, h.data(), h.size() * 8, cudaMemcpyHostToDevice, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}

void coral_release_tables(coral_ctx*) {}

template <class S, class D>
static int fetch_col(coral_ctx* c, const DevBuf& b, int64_t n, D* out, std::vector<uint8_t>& tmp) {
    if (!out || n == 0) return CORAL_OK;
    tmp.resize((size_t)n * sizeof(S));
    CORAL_CUDA(c, cudaMemcpyAsync(tmp.data(), b.p, (size_t)n * sizeof(S), cudaMemcpyDeviceToHost, c->stream));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    const S* s = reinterpret_cast<const S*>(tmp.data());
    for (int64_t i = 0; i < n; i++) out[i] = (D)s[i];
    return CORAL_OK;
}

extern "C" int coral_get_record_table(coral_ctx* c, int32_t* tid, int32_t* pos, int32_t* endpos, int32_t* flag,
                                      int32_t* mapq, int64_t* l_seq, int32_t* n_cigar, int64_t* infer_len,
                                      int64_t* aligned_bases, int64_t* non_acgt) {
    if (!c) return CORAL_E_ARG;
    if (!c->loaded) { c->err = "no records loaded"; return CORAL_E_STATE; }
    CORAL_TRY(coral_wait_scan(c));
    cudaSetDevice(c->device);
    const int64_t n = c->nrec;
    std::vector<uint8_t> tmp;
    CORAL_TRY((fetch_col<int32_t, int32_t>(c, c->rt.tid, n, tid, tmp)));
    CORAL_TRY((fetch_col<int32_t, int32_t>(c, c->rt.pos, n, pos, tmp)));
    CORAL_TRY((fetch_col<int32_t, int32_t>(c, c->rt.endpos, n, endpos, tmp)));
    if (flag || mapq) {
        std::vector<uint32_t> fm((size_t)std::max<int64_t>(n, 1));
        CORAL_CUDA(c, cudaMemcpyAsync(fm.data(), c->rt.flag_mapq.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
        CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int64_t i = 0; i < n; i++) {
            if (flag) flag[i] = (int32_t)(fm[i] & 0xffffu);
            if (mapq) mapq[i] = (int32_t)(fm[i] >> 16);
        }
    }
    CORAL_TRY((fetch_col<uint32_t, int64_t>(c, c->rt.lseq, n, l_seq, tmp)));
    CORAL_TRY((fetch_col<uint32_t, int32_t>(c, c->rt.ncig, n, n_cigar, tmp)));
    CORAL_TRY((fetch_col<uint32_t, int64_t>(c, c->rt.infer, n, infer_len, tmp)));
    CORAL_TRY((fetch_col<uint32_t, int64_t>(c, c->rt.mbases, n, aligned_bases, tmp)));
    CORAL_TRY((fetch_col<uint32_t, int64_t>(c, c->rt.nonacgt, n, non_acgt, tmp)));
    return CORAL_OK;
}
