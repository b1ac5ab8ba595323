/* oracle/covlib.c — TEST INFRASTRUCTURE (CPU oracle), never linked into the product.
 *
 * Plain-C restatement of the htslib/pysam inner loops that the reference's hot path calls but
 * whose source is NOT under /root/reference (pysam is an un-vendored, unpinned dependency:
 * reference pyproject.toml:32 `pysam >0.1.7`).  Restated from the SAM/BAM specification and
 * pysam's documented behaviour — "parity unpinned" against real pysam:
 *   ora_index           bam_endpos / bam_cigar2rlen               (fetch(region) overlap rule,
 *                                                                  reference call sites
 *                                                                  infer_breakpoint_graph.py:751,1031,1043)
 *   ora_count_coverage  AlignmentFile.count_coverage(quality_threshold=0, 'nofilter')
 *                                                                  (call sites :131, :1033)
 *   ora_blocks          AlignedSegment.get_blocks()               (call site :756)
 * Build: gcc -O2 -shared -fPIC oracle/covlib.c -o oracle/_covlib.so
 */
#include <stdint.h>
#include <string.h>

static inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static inline uint16_t rd16(const uint8_t* p) { uint16_t v; memcpy(&v, p, 2); return v; }

/* Per record: tid, pos, endpos (htslib bam_endpos: pos + max(1, ref-consumed) for mapped records
 * with a CIGAR, pos + 1 otherwise), flag, mapq, l_seq, n_cigar, infer_read_length (M,I,S,=,X,H). */
void ora_index(const uint8_t* buf, const uint64_t* off, int64_t n, int32_t* tid, int32_t* pos, int32_t* endpos,
               int32_t* flag, int32_t* mapq, int64_t* l_seq, int32_t* n_cigar, int64_t* infer_len,
               int64_t* aligned_bases) {
    for (int64_t i = 0; i < n; i++) {
        const uint8_t* r = buf + off[i];
        tid[i] = (int32_t)rd32(r + 4);
        pos[i] = (int32_t)rd32(r + 8);
        uint32_t l_name = r[12];
        mapq[i] = r[13];
        uint32_t nc = rd16(r + 16);
        flag[i] = rd16(r + 18);
        l_seq[i] = rd32(r + 20);
        n_cigar[i] = (int32_t)nc;
        const uint8_t* c = r + 36 + l_name;
        int64_t rlen = 0, qh = 0, mb = 0;
        for (uint32_t k = 0; k < nc; k++) {
            uint32_t v = rd32(c + 4 * k), op = v & 15, len = v >> 4;
            if (op == 0 || op == 2 || op == 3 || op == 7 || op == 8) rlen += len;
            if (op == 0 || op == 1 || op == 4 || op == 5 || op == 7 || op == 8) qh += len;
            if (op == 0 || op == 7 || op == 8) mb += len;
        }
        infer_len[i] = qh;
        aligned_bases[i] = mb;
        if ((flag[i] & 4) || nc == 0) rlen = 0;
        endpos[i] = pos[i] + (int32_t)(rlen > 0 ? rlen : 1);
    }
}

/* Sum over A/C/G/T arrays of count_coverage(contig, start, stop) for the given record indices
 * (already selected by the fetch overlap rule).  Only M/=/X-aligned pairs whose query base is one
 * of A,C,G,T (BAM nibble 1,2,4,8) count; records without SEQ are skipped. */
int64_t ora_count_coverage(const uint8_t* buf, const uint64_t* off, const int64_t* idx, int64_t n_idx,
                           int64_t start, int64_t stop) {
    int64_t total = 0;
    for (int64_t j = 0; j < n_idx; j++) {
        const uint8_t* r = buf + off[idx[j]];
        int64_t rpos = (int32_t)rd32(r + 8);
        uint32_t l_name = r[12], nc = rd16(r + 16), l_seq = rd32(r + 20);
        if (l_seq == 0) continue;
        const uint8_t* c = r + 36 + l_name;
        const uint8_t* seq = c + 4 * (size_t)nc;
        int64_t qpos = 0;
        for (uint32_t k = 0; k < nc; k++) {
            uint32_t v = rd32(c + 4 * k), op = v & 15;
            int64_t len = v >> 4;
            if (op == 0 || op == 7 || op == 8) {
                int64_t a = rpos < start ? start : rpos;
                int64_t b = rpos + len > stop ? stop : rpos + len;
                for (int64_t x = a; x < b; x++) {
                    int64_t q = qpos + (x - rpos);
                    uint8_t by = seq[q >> 1];
                    uint8_t nib = (q & 1) ? (by & 15) : (by >> 4);
                    total += (nib == 1 || nib == 2 || nib == 4 || nib == 8);
                }
                rpos += len; qpos += len;
            } else if (op == 2 || op == 3) {
                rpos += len;
            } else if (op == 1 || op == 4) {
                qpos += len;
            }
        }
    }
    return total;
}

/* get_blocks(): one (start, end) per M/=/X op; D and N advance the reference.  Returns the
 * number of blocks; out (may be NULL) receives 2*n int64. */
int64_t ora_blocks(const uint8_t* buf, uint64_t off, int64_t* out) {
    const uint8_t* r = buf + off;
    int64_t rpos = (int32_t)rd32(r + 8);
    uint32_t l_name = r[12], nc = rd16(r + 16);
    const uint8_t* c = r + 36 + l_name;
    int64_t nb = 0;
    for (uint32_t k = 0; k < nc; k++) {
        uint32_t v = rd32(c + 4 * k), op = v & 15;
        int64_t len = v >> 4;
        if (op == 0 || op == 7 || op == 8) {
            if (out) { out[2 * nb] = rpos; out[2 * nb + 1] = rpos + len; }
            nb++; rpos += len;
        } else if (op == 2 || op == 3) {
            rpos += len;
        }
    }
    return nb;
}
