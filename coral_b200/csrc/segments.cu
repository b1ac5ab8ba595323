// coral_b200/csrc/segments.cu — hash_alignment_to_seg (reference infer_breakpoint_graph.py:181-210), the lookups.
//
// The reference asks an interval tree, per local alignment of every chimeric read, which CN segment holds the
// lower and the higher end of the alignment (pos2cni, :177-178, on min / max of the '-'-reversed interval).  CN
// segments are half-open [begin, end) and — whenever this entry point is used — do not overlap, so each lookup is
// a binary search on the segment starts of the alignment's contig.  One thread per alignment of the chimeric
// table; the caller builds the Python sets (`rint.append(cniset)`) and chimeric_alignments_seg from the two columns.
#include "ctx.h"
#include "dev_util.cuh"
#include "relops.cuh"

namespace coral {

struct CnSegs {              // sorted by (tid, begin); disjoint inside a tid
    const int32_t *tid, *begin, *end, *idx;
    int32_t n;
};

// index of the segment of `t` holding p, -1 if none, -2 if the contig has no segment at all
__device__ __forceinline__ int32_t cn_segment_at(const CnSegs& S, int32_t t, int32_t p) {
    int32_t lo = 0, hi = S.n;                 // first segment with (tid, begin) > (t, p)
    while (lo < hi) {
        const int32_t mid = (lo + hi) >> 1;
        const int32_t mt = S.tid[mid];
        if (mt < t || (mt == t && S.begin[mid] <= p)) lo = mid + 1; else hi = mid;
    }
    const int32_t k = lo - 1;
    if (k >= 0 && S.tid[k] == t) return S.end[k] > p ? S.idx[k] : -1;
    return (lo < S.n && S.tid[lo] == t) ? -1 : -2;      // before the contig's first segment / contig unknown
}

__global__ void k_cn_segment_index(int64_t n_aln, const int32_t* __restrict__ tid, const int32_t* __restrict__ rs,
                                   const int32_t* __restrict__ re, CnSegs S, int32_t* __restrict__ lo_idx,
                                   int32_t* __restrict__ hi_idx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_aln) return;
    const int32_t t = tid[i], a = rs[i], b = re[i];
    lo_idx[i] = cn_segment_at(S, t, min(a, b));
    hi_idx[i] = cn_segment_at(S, t, max(a, b));
}

}  // namespace coral

extern "C" int coral_cn_segment_index(coral_ctx* c, const int32_t* seg_tid, const int32_t* seg_begin, const int32_t* seg_end,
                                      const int32_t* seg_idx, int32_t n_seg, int32_t* lo_idx, int32_t* hi_idx) {
    using namespace coral;
    if (!c || n_seg < 0 || !lo_idx || !hi_idx) return CORAL_E_ARG;
    if (!c->chim.valid) { c->err = "coral_cn_segment_index before coral_fetch"; return CORAL_E_STATE; }
    cudaSetDevice(c->device);
    for (int32_t k = 0; k < n_seg; k++) {
        if (seg_end[k] < seg_begin[k] || (k > 0 && (seg_tid[k] < seg_tid[k - 1] ||
                                                    (seg_tid[k] == seg_tid[k - 1] && seg_begin[k] < seg_end[k - 1])))) {
            c->err = "CN segments must be sorted by (tid, begin) and must not overlap";
            return CORAL_E_ARG;
        }
    }
    ChimericTable& H = c->chim;
    const int64_t n = H.n_aln;
    if (n == 0) return CORAL_OK;
    const size_t ns = (size_t)std::max(n_seg, 1);
    CORAL_CUDA(c, c->s0.ensure(ns * 16));
    CORAL_CUDA(c, c->s1.ensure((size_t)n * 8));
    int32_t* d_seg = c->s0.as<int32_t>();
    CORAL_TRY(h2d(c, d_seg, seg_tid, (size_t)n_seg));
    CORAL_TRY(h2d(c, d_seg + ns, seg_begin, (size_t)n_seg));
    CORAL_TRY(h2d(c, d_seg + 2 * ns, seg_end, (size_t)n_seg));
    CORAL_TRY(h2d(c, d_seg + 3 * ns, seg_idx, (size_t)n_seg));
    CnSegs S{d_seg, d_seg + ns, d_seg + 2 * ns, d_seg + 3 * ns, n_seg};
    int32_t* d_lo = c->s1.as<int32_t>();
    int32_t* d_hi = d_lo + n;
    k_cn_segment_index<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(n, H.tid.as<int32_t>(), H.rs.as<int32_t>(),
                                                                        H.re.as<int32_t>(), S, d_lo, d_hi);
    c->n_launches++;
    CORAL_CUDA(c, cudaGetLastError());
    CORAL_TRY(d2h(c, lo_idx, d_lo, (size_t)n));
    CORAL_TRY(d2h(c, hi_idx, d_hi, (size_t)n));
    CORAL_CUDA(c, cudaStreamSynchronize(c->stream));
    return CORAL_OK;
}
