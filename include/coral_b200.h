/* include/coral_b200.h — C ABI of the B200-native CoRAL reconstruct hot path.
 *
 * The reference has no FFI: its hot path is a Python object protocol
 * (src/infer_breakpoint_graph.py:20-1330 `bam_to_breakpoint_nanopore`, called in fixed order by
 * `reconstruct_graph`, :1349-1386) over pysam.  This header is the boundary a maintainer binds
 * instead (ctypes stub in INTEGRATION.md).  Every entry point names the reference code it replaces.
 *
 * Conventions
 *   - plain C types only; all pointers are HOST pointers unless the name starts with `d_`;
 *   - every call returns 0 on success or a negative CORAL_E_* code; coral_last_error() gives text;
 *   - a ctx is bound to one CUDA device, is not thread-safe, and owns every result table until the
 *     next call of the same entry point or coral_destroy();
 *   - "records" are *decompressed BAM alignment records* (SAM/BAM spec §4.2: block_size, 32-byte core,
 *     read_name, CIGAR u32[], 4-bit SEQ, QUAL, aux TLV), each starting at a 16-byte-aligned offset of
 *     the buffer (coral_bam_read() frames them that way; the padding is what makes 1-D bulk/TMA
 *     copies of a record legal on sm_100a);
 *   - coordinates are 0-based; intervals are [start, stop) unless stated otherwise;
 *   - `tid` is the BAM header reference index.
 */
#ifndef CORAL_B200_H
#define CORAL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CORAL_OK 0
#define CORAL_E_CUDA (-1)      /* CUDA runtime error (text in coral_last_error) */
#define CORAL_E_ARG (-2)       /* bad argument / misaligned record / out-of-range */
#define CORAL_E_STATE (-3)     /* call order violated (e.g. coverage before load) */
#define CORAL_E_FORMAT (-4)    /* malformed record / aux / SA string */
#define CORAL_E_IO (-5)        /* BAM file could not be read */
#define CORAL_E_NOMEM (-6)

typedef struct coral_ctx coral_ctx;

/* Algorithm constants of the reference class (infer_breakpoint_graph.py:25-33 and call-site literals). */
typedef struct coral_params {
    int32_t min_del_len;           /* 600   ibg:33  gap between aligned blocks that makes a small-del event */
    int32_t min_bp_match_cutoff;   /* 100   ibg:28 */
    int32_t min_mapq;              /* 20    ibg:688,754 */
    int32_t gap_mapq;              /* 10    breakpoint_utilities.py:129 */
    int32_t gap;                   /* 100   ibg:688 (gap_) */
    int32_t bp_distance_cutoff;    /* 2000  ibg:31 */
} coral_params;

void coral_default_params(coral_params* p);

/* ---- lifetime --------------------------------------------------------------------------------- */
int coral_create(int device, coral_ctx** out);
void coral_destroy(coral_ctx* ctx);
const char* coral_last_error(const coral_ctx* ctx);
int coral_set_params(coral_ctx* ctx, const coral_params* p);

/* Reference sequence table (BAM header order).  `names` = n NUL-terminated strings back to back.
 * `chr_rank[tid]` = global_names.chr_idx value (global_names.py:13-18) or -1 when the contig is not
 * in that table (the reference raises KeyError in interval2bp, breakpoint_utilities.py:293). */
int coral_set_references(coral_ctx* ctx, const char* names, int32_t n, const int32_t* chr_rank);

/* ---- host-side BAM ingest (replaces pysam.AlignmentFile(path,'rb'), ibg:65) ---------------------
 * Inflates every BGZF member with zlib on `n_threads` host threads and frames the alignment records
 * into one pinned host buffer, 16-byte aligned starts.  The returned pointers stay valid until
 * coral_bam_free().  Header text / reference names are returned through the accessor calls. */
typedef struct coral_bam coral_bam;
int coral_bam_read(const char* path, int n_threads, coral_bam** out, char* err, size_t err_len);
void coral_bam_free(coral_bam* b);
int32_t coral_bam_n_ref(const coral_bam* b);
const char* coral_bam_ref_name(const coral_bam* b, int32_t tid);
int64_t coral_bam_ref_len(const coral_bam* b, int32_t tid);
int64_t coral_bam_n_records(const coral_bam* b);
const uint8_t* coral_bam_records(const coral_bam* b, uint64_t* nbytes);   /* pinned host buffer */
const uint64_t* coral_bam_offsets(const coral_bam* b);                     /* n_records + 1 */

/* ---- load + scan (replaces htslib record decode under every pysam call of the path) -------------
 * Copies the records to the device (unless `on_device` != 0, then `recs` is a device pointer that
 * must outlive the ctx's use of it) and runs the record-scan kernel once: per record core fields,
 * reference end (bam_endpos), infer_read_length, aligned bases, non-ACGT count, name hash, SA/NM tag
 * location, and the in-CIGAR gap events > min_del_len.  `global_ordinal` (nullable) is the record's
 * index in the whole file when this ctx holds one shard. */
int coral_load_records(coral_ctx* ctx, const uint8_t* recs, uint64_t nbytes, const uint64_t* offsets,
                       int64_t n_records, const uint64_t* global_ordinal, int on_device);
int64_t coral_n_records(const coral_ctx* ctx);
/* Number of kernels this library has launched on the ctx so far (bench.py reports it as gpu_launches). */
uint64_t coral_launch_count(const coral_ctx* ctx);

/* Re-run the record-scan kernel on the records already resident on the device (what one "step" of the
 * device-resident benchmark starts with). */
int coral_rescan(coral_ctx* ctx);

/* Device time of the kernels of one stage during its most recent run, from CUDA events recorded on the
 * ctx's own stream around the launches (milliseconds; < 0 if the stage has not run). */
#define CORAL_STAGE_SCAN 0
#define CORAL_STAGE_COVERAGE 1
#define CORAL_STAGE_FETCH 2
#define CORAL_STAGE_SMALLDEL 3
#define CORAL_STAGE_CANDIDATES 4
#define CORAL_STAGE_CLUSTER 5
#define CORAL_STAGE_BPC2BP 6
#define CORAL_STAGE_STAB 7
#define CORAL_N_STAGES 8
float coral_stage_ms(coral_ctx* ctx, int stage);

/* Per-record scan table copied to host arrays (any pointer may be NULL to skip that column). */
int coral_get_record_table(coral_ctx* ctx, int32_t* tid, int32_t* pos, int32_t* endpos, int32_t* flag,
                           int32_t* mapq, int64_t* l_seq, int32_t* n_cigar, int64_t* infer_len,
                           int64_t* aligned_bases, int64_t* non_acgt);

/* ---- coverage --------------------------------------------------------------------------------- *
 * replaces AlignmentFile.count_coverage(contig,start,stop,quality_threshold=0,'nofilter') summed over
 * A,C,G,T (ibg:131, ibg:1033-1034) and the record count of ibg:1031-1032 for n regions [start,stop).
 * Regions must be sorted by (tid,start) and pairwise disjoint.
 *   base_sum[i]  = #(record, refpos) pairs, refpos in region i, M/=/X-aligned, query base in {A,C,G,T}
 *   rec_count[i] = #records overlapping region i with infer_read_length() > 0                      */
int coral_region_coverage(coral_ctx* ctx, const int32_t* tid, const int32_t* start, const int32_t* stop,
                          int32_t n_regions, int64_t* base_sum, int64_t* rec_count);

/* ---- fetch ------------------------------------------------------------------------------------- *
 * replaces bam_to_breakpoint_nanopore.fetch (ibg:139-174) + cigar_parsing.alignment_from_satags
 * (cigar_parsing.py:232-269, cigar2pos* :17-215) over ALL loaded records.
 * Result (kept on the device for coral_candidates and readable with coral_get_chimeric):
 *   reads in the reference's dict order (order of each name's first SA-bearing record);
 *   read_first_rec[r] = index of that record (the caller takes the name string from its own buffer);
 *   read_len[r]       = l_seq of the first record with flag < 256 and the same name (-1: none);
 *   read_status[r]    = 0 ok | 1 no primary (ibg:164-169) | 2 rejected -> ([],[],[]) (cp:248-253)
 *                       | 3 KeyError signature (cp:255) | 4 malformed entry text;
 *   alignments CSR aln_ptr[n_reads+1], stably sorted by (qs, qe): qs, qe, tid, rstart, rend (rint[1], rint[2];
 *   start > end on '-'), strand (0 '+', 1 '-'), mapq, nm (raw last field; the caller divides by qe - qs, cp:268).
 * nm_*: sum of e = NM/l_seq, sum of e*e and count over records without SA and MAPQ 60 (ibg:153-157);
 * nm_zero_div counts such records with l_seq == 0 (the reference raises ZeroDivisionError). */
int coral_fetch(coral_ctx* ctx, int64_t* n_reads, int64_t* n_alignments, double* nm_sum, double* nm_sumsq,
                int64_t* nm_n, int64_t* nm_zero_div);
int coral_get_chimeric(coral_ctx* ctx, uint32_t* read_first_rec, int64_t* read_len, int32_t* read_status,
                       uint32_t* aln_ptr, int32_t* qs, int32_t* qe, int32_t* tid, int32_t* rstart, int32_t* rend,
                       uint8_t* strand, int32_t* mapq, int64_t* nm);

#ifdef __cplusplus
}
#endif
#endif /* CORAL_B200_H */
