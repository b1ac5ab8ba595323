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
/* Device buffers of destroyed (or regrown) ctxs are kept by a per-process caching allocator and handed to the next ctx
 * (opening a ctx per BAM then costs no cudaMalloc / cudaFree, which synchronise the device); this gives them all back
 * to the driver. */
void coral_release_cached_memory(void);
int coral_set_params(coral_ctx* ctx, const coral_params* p);
/* Named switches that do not change any result.  "full_name_sort" = 1: coral_fetch groups SA entries by a full
 * 128-bit name sort instead of the 40-bit sort it normally tries first (the path it falls back to when two read
 * names share those bits; exposed so that tests can cover it).  "pass_overlap_scan" = 0: coral_rescan inside a
 * pipelined pass runs the one-kernel scan on the ctx's stream instead of pre-scan + full scan on a second stream
 * (default 1, see coral_pass_begin).  Unknown names return CORAL_E_ARG. */
int coral_set_option(coral_ctx* ctx, const char* name, int64_t value);

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
 * Makes the records visible to the device and runs the record-scan kernel once.  `on_device` says where
 * `recs` lives and where it stays:
 *   CORAL_RECORDS_HOST_COPY    host buffer, copied to HBM (every later pass runs at HBM speed);
 *   CORAL_RECORDS_DEVICE       device pointer, used in place (must outlive the ctx's use of it);
 *   CORAL_RECORDS_HOST_MAPPED  page-locked host buffer read IN PLACE by the kernels over PCIe: the scan
 *                              only pulls the fields it needs (QUAL, about half of a record, never crosses
 *                              the link).  A pageable buffer is page-locked for the lifetime of the load.
 *                              The buffer must stay valid and extend >= 64 bytes past `nbytes`.
 * Every record starts at a 16-byte-aligned offset and offsets[n_records] is 16-byte aligned too.
 * The scan computes, per record: the core fields,
 * reference end (bam_endpos), infer_read_length, aligned bases, non-ACGT count, name hash, SA/NM tag
 * location, and the in-CIGAR gap events > min_del_len.  `global_ordinal` (nullable) is the record's
 * index in the whole file when this ctx holds one shard. */
#define CORAL_RECORDS_HOST_COPY 0
#define CORAL_RECORDS_DEVICE 1
#define CORAL_RECORDS_HOST_MAPPED 2
/*   CORAL_RECORDS_HOST_MAPPED_ASYNC  as HOST_MAPPED, but the call returns once a pre-scan (record core, read name and
 *                              aux block only: what coral_fetch needs) is done; the full scan keeps running on its
 *                              own stream while the caller already runs coral_fetch / coral_candidates_* /
 *                              coral_cluster_call / coral_cn_segment_index.  Every other entry point joins the scan
 *                              first (coral_wait_scan), where its errors (malformed record -> CORAL_E_FORMAT) surface. */
#define CORAL_RECORDS_HOST_MAPPED_ASYNC 3
/*   CORAL_RECORDS_HOST_MAPPED_LAZY   only the pre-scan runs at load; the full scan runs (synchronously) inside the first
 *                              entry point that needs it, or never — a caller that only wants coral_fetch (the HSR
 *                              mode, hsr.py:21-49) moves a few hundred bytes per record instead of CIGAR and SEQ. */
#define CORAL_RECORDS_HOST_MAPPED_LAZY 4
int coral_wait_scan(coral_ctx* ctx);
int coral_load_records(coral_ctx* ctx, const uint8_t* recs, uint64_t nbytes, const uint64_t* offsets,
                       int64_t n_records, const uint64_t* global_ordinal, int on_device);
int64_t coral_n_records(const coral_ctx* ctx);
/* Number of kernels this library has launched on the ctx so far (bench.py reports it as gpu_launches). */
uint64_t coral_launch_count(const coral_ctx* ctx);

/* Re-run the record-scan kernel on the records already resident on the device (what one "step" of the
 * device-resident benchmark starts with). */
int coral_rescan(coral_ctx* ctx);

/* ---- pipelined pass ---------------------------------------------------------------------------------------------- *
 * Between coral_pass_begin and coral_pass_end the entry points listed below only ENQUEUE their kernels and return: no
 * call synchronises with the device, every output array they were given is filled when coral_pass_end returns (the
 * arrays must stay valid until then), and the whole pass costs one synchronisation.  The scalars that size the
 * launches of a later kernel (number of SA-bearing records, SA entries, chimeric reads, alignments, candidates,
 * clusters, gap events ...) — and the counts these calls return (*n_reads, *n_rows, *n_clusters ...) — are those of
 * the last complete pass over the SAME records with the SAME sequence of calls and host inputs; a one-thread kernel per
 * value checks the device's own value against it.  The first pass over new inputs (and any pass whose calls depart from
 * the recorded sequence, from that call on) reads those scalars back as usual and records them.  coral_pass_end
 * returns CORAL_E_STATE if the device disagreed with a replayed value (only possible when the caller overwrote a
 * CORAL_RECORDS_DEVICE buffer in place): nothing of that pass is delivered and the next pass records again.
 * *replayed (nullable) = 1 when the pass ran without intermediate synchronisation.
 * Entry points that take part: coral_rescan, coral_region_coverage, coral_fetch + coral_get_chimeric, coral_smalldel +
 * coral_get_smalldel, coral_candidates_l + coral_get_candidates, coral_cluster_call(_candidates) + coral_get_clusters,
 * coral_concordant + coral_get_concordant_rows.  Any other entry point may be called inside a pass; it synchronises.
 * Two lanes: inside a pass the chimeric chain (coral_fetch, coral_candidates_*, coral_cluster_call* and their getters)
 * runs on a second stream with its own scratch buffers, next to the coverage / small-deletion / concordant entry points
 * on the ctx's stream (option "pass_fork", default 1); coral_pass_end joins the two.  Option "pass_overlap_scan" = 1
 * additionally splits coral_rescan as the asynchronous load does (pre-scan, then the full scan on a third stream with
 * "pass_scan_reserve" block slots left free; default 0 — measured slower on one GPU).
 * Streams of a process should not share a hardware queue: set CUDA_DEVICE_MAX_CONNECTIONS=32 before CUDA starts.
 * This is how one step of the device-resident benchmark runs the path (reference call order
 * infer_breakpoint_graph.py:1354-1386) as one stream of kernels. */
/* Result arrays in page-locked memory (coral_host_alloc) receive their copies by direct DMA inside a pass: no staging
 * buffer and no host memcpy at coral_pass_end (1.3 MB of result tables per C2 pass: 0.1 ms of the step otherwise). */
int coral_host_alloc(uint64_t nbytes, void** out);
void coral_host_free(void* p);
int coral_pass_begin(coral_ctx* ctx);
int coral_pass_end(coral_ctx* ctx, int32_t* replayed);
/* For callers that enqueue work of their own between the entry points of a pass (the multi-GPU engine: NCCL
 * collectives): the ctx's streams as CUstream / cudaStream_t values (which = 0: the stream the entry points outside the
 * chimeric chain use; 1: the second lane), a way to make a ctx run on a stream it does not own (the rows-only ctx of the
 * multi-GPU engine runs on the second lane of the records ctx), and values of the caller that a replayed pass must
 * see again (the message sizes of the other ranks): recorded while the pass records, handed back while it replays. */
uint64_t coral_stream_handle(coral_ctx* ctx, int which);
int coral_adopt_stream(coral_ctx* ctx, uint64_t stream);
int coral_pass_note(coral_ctx* ctx, uint64_t value, uint64_t* out, int32_t* replayed);
int coral_pass_replaying(const coral_ctx* ctx);

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
#define CORAL_STAGE_PATHS 8
#define CORAL_STAGE_PRESCAN 9   /* the pre-scan of an asynchronous load / of coral_rescan inside a pipelined pass */
#define CORAL_N_STAGES 10
float coral_stage_ms(coral_ctx* ctx, int stage);
int coral_stage_ms_all(coral_ctx* ctx, float* out /* [CORAL_N_STAGES] */);
/* Forget the stage times recorded so far (coral_stage_ms then reports only stages that ran since), and the time from
 * one stage event to another — begin (which = 0) or end (which = 1) — across the ctx's streams: where the stages of a
 * pipelined pass sit relative to the scan that runs next to them. */
void coral_reset_stage_times(coral_ctx* ctx);
float coral_stage_offset_ms(coral_ctx* ctx, int stage_a, int which_a, int stage_b, int which_b);

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

/* Allocation-free variant of coral_region_coverage.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, same results, but the call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine): the region arrays, the accumulators and the (record, region) list of the exact re-walk live in a
 * file-scope `static __device__` array of CORAL_STATIC_SCRATCH_BYTES (csrc/coverage.cu) that every GPU which loads
 * the library holds until the process exits.  Limits: n_regions <= CORAL_STATIC_MAX_REGIONS and at most
 * CORAL_STATIC_MAX_SLOW_PAIRS (record, region) pairs that need the exact walk (records straddling a region boundary
 * or carrying non-ACGT bases); beyond either the call fails with CORAL_E_ARG and a text that says which (use
 * coral_region_coverage then).  The record scan must have been joined (CORAL_E_STATE while an asynchronous scan is
 * pending) and the call is refused inside a pipelined pass.  All calls on one device share the array, which keeps the
 * previous call's contents (each call clears what must start at zero): calls that overlap on the GPU — two ctxs or
 * threads on one device — race and must be serialised by the caller. */
#define CORAL_STATIC_MAX_REGIONS 4096
#define CORAL_STATIC_MAX_SLOW_PAIRS 65536
#define CORAL_STATIC_SCRATCH_BYTES 639232
int coral_region_coverage_static(coral_ctx* ctx, const int32_t* tid, const int32_t* start, const int32_t* stop,
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

/* The same table with all twelve columns in ONE caller buffer: column k (in the order of coral_get_chimeric's arguments)
 * starts at offsets[k] (64-byte aligned); one packing kernel + one device-to-host copy instead of twelve copies.
 * out_bytes >= sum over the columns of their sizes rounded up to 64 (n_reads / n_alignments from coral_fetch). */
int coral_get_chimeric_packed(coral_ctx* ctx, uint8_t* out, uint64_t out_bytes, uint64_t* offsets);

/* ---- multi-GPU exchange of `fetch` (one process per GPU, each holding a contiguous range of the file) --------
 * The reference unions the SA strings of all records of a read name in file order and takes read_length from
 * the first record with flag < 256 anywhere in the file (ibg:142-151).  A rank exports one row per SA-bearing
 * record (and per first local primary of a chimeric name), the rows of all ranks are concatenated in rank
 * (= file) order and loaded into a second, rows-only ctx on which coral_fetch / coral_candidates_* /
 * coral_cluster_call run unchanged; every other entry point returns CORAL_E_STATE on such a ctx.
 *   coral_get_record_keys : per-record 128-bit name hash, SA tag location (sa_off == 0: no SA tag) and NM;
 *   coral_nm_partials     : this ctx's partial sums of ibg:153-157 (all-reduced by the caller);
 *   coral_load_rows       : row i = name hash, flag, MAPQ, l_seq, NM and SA text text[text_off[i] : text_off[i+1])
 *                           (empty: the record has no SA tag). */
int coral_get_record_keys(coral_ctx* ctx, uint64_t* name_lo, uint64_t* name_hi, uint32_t* sa_off, uint32_t* sa_len,
                          int32_t* nm);
int coral_nm_partials(coral_ctx* ctx, double* nm_sum, double* nm_sumsq, int64_t* nm_n, int64_t* nm_zero_div);
int coral_load_rows(coral_ctx* ctx, int64_t n_rows, const uint8_t* text, uint64_t text_bytes, const uint64_t* text_off,
                    const uint64_t* name_lo, const uint64_t* name_hi, const uint32_t* flag, const uint32_t* mapq,
                    const uint32_t* l_seq, const int32_t* nm);

/* Device-side form of the same exchange (what the multi-GPU engine uses; the rows never visit the host).  Every rank
 * packs what it contributes into ONE byte message in its own HBM, the caller all-gathers the messages (NCCL) into a
 * buffer on this ctx's device — message r = d_buf[starts[r] : ...], starts 16-byte aligned, host array — and hands
 * the gathered buffer to the next call:
 *   coral_pack_sa_rows    : message 1 = this shard's SA-bearing records (name hash, flag, MAPQ, l_seq, NM, record
 *                           index, read name + SA text), file order;
 *   coral_pack_extra_rows : given every rank's message 1: message 2 = the first local record with flag < 256 of each
 *                           chimeric name that carries no SA tag itself (ibg:142-143), + this shard's nm_stats
 *                           partial sums (ibg:153-157) in its header;
 *   coral_unpack_rows     : both gathered sets -> a rows-only ctx (rows in global file order); nm_* [n_ranks]
 *                           receive the shards' partial sums (may be NULL).
 * *d_msg / *nbytes: device address and size of the message, valid until the next pack call of the same kind.
 * At most 16 ranks per exchange.  coral_device_view: device addresses of the record buffer, offsets and table
 * columns (recs, offsets, name_lo, name_hi, flag|mapq<<16, l_seq, nm, sa_off, sa_len; valid until the next load /
 * rescan / unpack, read-only). */
int coral_device_view(coral_ctx* ctx, uint64_t* ptrs, int32_t n_ptrs);
int coral_pack_sa_rows(coral_ctx* ctx, uint64_t* d_msg, uint64_t* nbytes);
int coral_pack_extra_rows(coral_ctx* ctx, int32_t n_ranks, const uint8_t* d_buf, const uint64_t* starts, uint64_t* d_msg,
                          uint64_t* nbytes);
int coral_unpack_rows(coral_ctx* ctx, int32_t n_ranks, const uint8_t* d_buf1, const uint64_t* starts1, const uint8_t* d_buf2,
                      const uint64_t* starts2, double* nm_sum, double* nm_sumsq, int64_t* nm_n, int64_t* nm_zero_div,
                      int64_t* n_rows);

/* ---- exchanges over peer memory (NVLink / NVSwitch), csrc/xchg.cu ------------------------------------------------ *
 * The transport of the multi-GPU form of the path (the one real exchange is the SA union / first-primary
 * rule of fetch, ibg:141-151; coverage sums and concordant name sets are combined the same way).  One process per GPU;
 * every rank owns a WINDOW in its own HBM that the peers map once (CUDA IPC).  An all-gather of one variable-length
 * message per rank is one kernel on the ctx's stream — it stores the message into this rank's slot of every peer's
 * window, releases a sequence number and then waits on the local flags — and no host round trip: inside a
 * pipelined pass the message sizes are replayed and checked on the device like every other launch size.
 *   coral_xchg_create  : slot_bytes[CORAL_XCHG_CHANNELS] = largest message per channel (0: channel unused); the window
 *                        holds 2 x world slots per channel.  A message beyond the slot size fails with CORAL_E_NOMEM.
 *   coral_xchg_handle  : the 64-byte IPC handle the peers open (+ the window's device address, for ranks that are
 *                        threads of the SAME process, where a handle cannot be opened);
 *   coral_xchg_connect : handles of all ranks in rank order (the caller all-gathers them once, any transport); a host
 *                        barrier of the caller must separate create / connect / first exchange / destroy on all ranks;
 *   coral_xchg_allgather : message r of the group = d_buf[starts[r] : starts[r] + sizes[r]) in THIS rank's HBM, valid
 *                        until this stream's next exchange but one on the channel; all ranks call the exchanges of a
 *                        channel in the same order.  lane = 1: inside a pipelined pass, run on the ctx's second lane.
 *   coral_xchg_allreduce_i64 : in-place integer sum over ranks of a small host array (synchronises).
 * A wait gives up after the timeout (default 20 s): the call (or the pass) fails with CORAL_E_STATE instead of
 * hanging the GPU; coral_xchg_timeouts reads and clears the count.  At most CORAL_XCHG_MAX_RANKS ranks. */
typedef struct coral_xchg coral_xchg;
#define CORAL_XCHG_MAX_RANKS 16
#define CORAL_XCHG_CHANNELS 8
#define CORAL_XCHG_HANDLE_BYTES 64
int coral_xchg_create(int device, int32_t rank, int32_t world, const uint64_t* slot_bytes, coral_xchg** out);
void coral_xchg_destroy(coral_xchg* x);
const char* coral_xchg_last_error(const coral_xchg* x);
int coral_xchg_handle(coral_xchg* x, uint8_t* handle, uint64_t* local_ptr);
int coral_xchg_connect(coral_xchg* x, const uint8_t* handles, const uint64_t* same_process_ptrs);
int coral_xchg_set_timeout(coral_xchg* x, double seconds);
int coral_xchg_timeouts(coral_xchg* x, int32_t* n);
int coral_xchg_allgather(coral_ctx* ctx, coral_xchg* x, int32_t channel, uint64_t d_msg, uint64_t nbytes, int32_t lane,
                         uint64_t* d_buf, uint64_t* starts, uint64_t* sizes);
int coral_xchg_allreduce_i64(coral_ctx* ctx, coral_xchg* x, int32_t channel, int64_t* values, int32_t n);
/* coral_region_coverage over ALL shards of the file: every rank calls it with the same regions on its own shard; the
 * accumulators are summed in place over peer memory between the kernels and the result copy (a record is owned by
 * exactly one shard, so the sums of ibg:131 and ibg:1031-1034 are additive). */
int coral_region_coverage_sharded(coral_ctx* ctx, coral_xchg* x, int32_t channel, const int32_t* tid, const int32_t* start,
                                  const int32_t* stop, int32_t n_regions, int64_t* base_sum, int64_t* rec_count);

/* Distributed fetch of the multi-GPU engine (csrc/chimeric_merge.cu): with options "fetch_part_world" = N and
 * "fetch_part_rank" = r, coral_fetch on the rows-only ctx parses only the reads whose name hash falls into residue class
 * r (all ranks hold all gathered rows; the SA entries are split N ways instead of parsed N times).
 *   coral_pack_chimeric  : this ctx's chimeric table as one message (valid until the next pack on the ctx);
 *   coral_merge_chimeric : the gathered messages of all ranks -> the table the 1-GPU coral_fetch over the whole file
 *                          leaves (reads in the order of their first SA-bearing record, ibg:145-151; alignments in their
 *                          stable (qs, qe) order, cp:263); coral_get_chimeric / coral_candidates_* / coral_cluster_call*
 *                          then run unchanged;
 *   coral_smalldel_pack  : the rows of the last coral_smalldel (+ 128-bit name hash) as one message (ibg:750-762, sharded);
 *   coral_read_device    : device bytes handed out by this library -> host array, deferred to coral_pass_end inside a
 *                          pipelined pass. */
int coral_pack_chimeric(coral_ctx* ctx, uint64_t* d_msg, uint64_t* nbytes);
int coral_merge_chimeric(coral_ctx* ctx, int32_t n_ranks, const uint8_t* d_buf, const uint64_t* starts, int64_t* n_reads,
                         int64_t* n_alignments);
int coral_smalldel_pack(coral_ctx* ctx, uint64_t* d_msg, uint64_t* nbytes);
int coral_read_device(coral_ctx* ctx, uint64_t d_src, uint64_t nbytes, uint8_t* out, int32_t lane);

/* Operator form of cigar_parsing.alignment_from_satags / cigar2pos* (cigar_parsing.py:17-269) on SA text the
 * caller supplies: `text` holds the entries, read r owns entries [ent_ptr[r], ent_ptr[r+1]) whose bytes are
 * text[ent_beg[k] : ent_end[k]].  Per read: status as in coral_fetch; per entry, in the read's stable
 * (qs, qe) order: qs, qe, al, rint start/end, strand, mapq, raw NM, and src = index of the entry in the input
 * list (the caller takes the rname string from it). */
int coral_sa_text(coral_ctx* ctx, const char* text, uint64_t text_len, int64_t n_reads, const uint32_t* ent_ptr,
                  const uint32_t* ent_beg, const uint32_t* ent_end, const int64_t* read_len, int32_t* status,
                  int64_t* qs, int64_t* qe, int64_t* al, int64_t* rstart, int64_t* rend, uint8_t* strand,
                  int32_t* mapq, int64_t* nm, uint32_t* src);

/* ---- small-deletion rows ------------------------------------------------------------------------- *
 * replaces the record loop of find_smalldel_breakpoints (ibg:750-762): for every amplicon interval
 * [start, end_incl] in LIST order, every record overlapping it (fetch(chr, start, end_incl + 1)) with
 * MAPQ >= min_mapq in file order, every pair of consecutive aligned blocks more than min_del_len apart.
 * Rows come back in that emission order: record index, interval index, next block start, previous block
 * end, first block start, last block end, MAPQ (= large_indel_alignments[name] entries, ibg:760). */
int coral_smalldel(coral_ctx* ctx, const int32_t* tid, const int32_t* start, const int32_t* end_incl,
                   int32_t n_intervals, int64_t* n_rows);
int coral_get_smalldel(coral_ctx* ctx, uint32_t* rec, int32_t* ai, int32_t* next_start, int32_t* prev_end,
                       int32_t* first_start, int32_t* last_end, int32_t* mapq);

/* ---- chimeric breakpoint candidates ---------------------------------------------------------------- *
 * replaces `for r in chimeric_alignments: alignment2bp_l(r, ..., 100, 20, 100, amplicon_intervals)`
 * (ibg:687-688 -> breakpoint_utilities.py:129-186, interval2bp :289-295) over the table left on the device
 * by coral_fetch.  Candidates in reference order.  Row = interval2bp fields: chr1 (tid), pos1, o1, chr2, pos2,
 * o2 (0 '+', 1 '-'), read index (into the coral_fetch read order) with alignment indices (i, j) already
 * swapped when flag == 1, rgap, flag, and the two MAPQs in (first, second) alignment order (bu:148). */
int coral_candidates_l(coral_ctx* ctx, const int32_t* tid, const int32_t* start, const int32_t* end_incl,
                       int32_t n_intervals, int64_t* n_candidates);
int coral_get_candidates(coral_ctx* ctx, uint32_t* read, int32_t* i, int32_t* j, int32_t* chr1, int32_t* pos1,
                         uint8_t* o1, int32_t* chr2, int32_t* pos2, uint8_t* o2, int32_t* rgap, uint8_t* flag,
                         int32_t* mq1, int32_t* mq2);

/* Two-interval form, breakpoint_utilities.alignment2bp (bu:70-96), as called by the amplicon-interval search
 * (ibg:432-434) for a chosen subset of reads: intervals [0] = intrvl1, [1] = intrvl2 (end inclusive).
 * Candidates come back through coral_get_candidates, in the order of `reads`. */
int coral_candidates_pair(coral_ctx* ctx, const uint32_t* reads, int64_t n_reads_selected, const int32_t* tid2,
                          const int32_t* start2, const int32_t* end2, int64_t* n_candidates);
/* Replace the device chimeric table by a caller-supplied one (operator forms of alignment2bp / alignment2bp_l
 * on Python-side chimeric alignments). */
int coral_set_chimeric(coral_ctx* ctx, int64_t n_reads, const uint32_t* aln_ptr, const int32_t* qs, const int32_t* qe,
                       const int32_t* tid, const int32_t* rstart, const int32_t* rend, const uint8_t* strand,
                       const int32_t* mapq);

/* ---- clustering + exact breakpoint calling --------------------------------------------------------- *
 * replaces cluster_bp_list(bp_list, min_cluster_size, bp_distance_cutoff) (breakpoint_utilities.py:252-286)
 * followed, for every cluster with len >= peel_cutoff, by the peel loop
 *   while len(rest) >= peel_cutoff: bp, bpr, stats, rest = bpc2bp(rest, match_cutoff)
 * (ibg:698-716; bpc2bp/bp_match breakpoint_utilities.py:299-416), at most max_peels times (0 = unbounded).
 * Input: n candidates as parallel arrays (chr ids are opaque equal/not-equal labels).
 * Output (coral_get_clusters):
 *   order[n], cluster_ptr[n_clusters+1]: clusters in the reference's output order, members in input order;
 *   n_sub[c]: number of bpc2bp calls made for cluster c; sub_of[i]: the call (0-based) whose bpr received
 *   candidate i, or -1;
 *   row arrays indexed cluster_ptr[c] + k for call k: called pos1/pos2, len(bpr), the candidate that was
 *   bp_cluster[0] of that call (supplies bp[0], bp[2], bp[3], bp[5..8]), and bp_stats_[6].  Only the first n_sub[c]
 *   slots of a cluster are written; the others are undefined. */
int coral_cluster_call(coral_ctx* ctx, int64_t n, const int32_t* chr1, const int32_t* pos1, const uint8_t* o1,
                       const int32_t* chr2, const int32_t* pos2, const uint8_t* o2, const int32_t* rgap,
                       const uint8_t* flag, const int32_t* mq1, const int32_t* mq2, double min_cluster_size,
                       int32_t bp_distance_cutoff, double peel_cutoff, int32_t match_cutoff, int32_t max_peels,
                       int64_t* n_clusters);
/* The same call on the candidate table coral_candidates_l / coral_candidates_pair left on the device (what
 * find_breakpoints does with its candidate list, ibg:687-716): the candidates do not travel through the host. */
int coral_cluster_call_candidates(coral_ctx* ctx, double min_cluster_size, int32_t bp_distance_cutoff, double peel_cutoff,
                                  int32_t match_cutoff, int32_t max_peels, int64_t* n_clusters);
int coral_get_clusters(coral_ctx* ctx, uint32_t* order, uint32_t* cluster_ptr, int32_t* n_sub, int32_t* sub_of,
                       int32_t* row_pos1, int32_t* row_pos2, int32_t* row_nb, uint32_t* row_first, double* row_stats);

/* Operator form of breakpoint_utilities.bp_match(bp1, bp2, rgap, [cutoff1, cutoff2]) (bu:391-416) over n pairs whose
 * chromosomes and orientations already agree (strings on the caller's side): positions of both breakpoints, the
 * orientation of each end of bp1 (0 '+', 1 '-'), rgap as the reference passes it (any double).  out[i] = 1 / 0.
 * Invalidates the result of the last coral_cluster_call (shares its buffers). */
int coral_bp_match(coral_ctx* ctx, int64_t n, const int64_t* bp1_pos1, const int64_t* bp1_pos2, const uint8_t* minus1,
                   const uint8_t* minus2, const double* rgap, const int64_t* bp2_pos1, const int64_t* bp2_pos2,
                   int64_t cutoff1, int64_t cutoff2, uint8_t* out);

/* ---- concordant edges ------------------------------------------------------------------------------ *
 * replaces the concordant-edge loop of assign_cov (ibg:1039-1055) for n edges (tid, p) -> (tid, p+1),
 * sorted by (tid, p).  rbps: rows (edge index, record index) naming, through any record of the read, the
 * reads on discordant edges incident to either node of that edge (ibg:1047-1053).
 *   lr_count[e] = len((names(p) & names(p+1) & names(p-d-1) & names(p+1+d)) - rbps)     (ce[8])
 * coral_get_concordant_rows returns (edge, record, mask) for every record covering one of the four points
 * (mask bit0 = p, bit1 = p+1, bit2 = p-d-1, bit3 = p+1+d); ce[9] = names of rows with mask & 3. */
int coral_concordant(coral_ctx* ctx, const int32_t* tid, const int32_t* p, int32_t n_edges, int64_t n_rbps,
                     const uint32_t* rb_edge, const uint32_t* rb_rec, int64_t* lr_count, int64_t* n_rows);
int coral_get_concordant_rows(coral_ctx* ctx, uint32_t* edge, uint32_t* rec, uint32_t* mask);

/* Multi-GPU form: coral_concordant with lr_count == NULL only produces the rows of this shard;
 * coral_concordant_pack turns them into one message (64-byte header + n rows of 4 x uint64: edge, name hash lo, hi,
 * mask) in this ctx's HBM; the caller all-gathers the messages; coral_count_name_msgs counts ce[8] on the gathered
 * messages (message r = d_buf[starts[r] : starts[r] + sizes[r]), host arrays) + the rbps rows (d_rb_rows: n_rb rows
 * of the same 4 x uint64 layout on this device, mask 16).  Same result on every rank. */
int coral_concordant_pack(coral_ctx* ctx, uint64_t* d_msg, uint64_t* nbytes);
int coral_count_name_msgs(coral_ctx* ctx, int32_t n_edges, int32_t n_ranks, const uint8_t* d_buf, const uint64_t* starts,
                          const uint64_t* sizes, const uint64_t* d_rb_rows, int64_t n_rb, int64_t* lr_count);

/* ---- CN-segment hashing of the chimeric alignments ---------------------------------------------------------- *
 * replaces the lookups of hash_alignment_to_seg (ibg:181-210: pos2cni on the lower and the higher end of every local
 * alignment).  CN segments [begin, end) sorted by (tid, begin), not overlapping, seg_idx = the caller's segment ids.
 * lo_idx / hi_idx (one entry per alignment of the chimeric table, table order): the id of the segment holding that
 * end, -1 when no segment of the contig holds it, -2 when the contig has no CN segment at all (ibg:184-186). */
int coral_cn_segment_index(coral_ctx* ctx, const int32_t* seg_tid, const int32_t* seg_begin, const int32_t* seg_end,
                           const int32_t* seg_idx, int32_t n_seg, int32_t* lo_idx, int32_t* hi_idx);

/* ---- subpath constraints: the concordant-read pass ------------------------------------------------------ *
 * replaces the record loop of compute_path_constraints (ibg:1303-1310: fetch(interval) + `q >= 20 and rn in
 * concordant_reads` + path_constraints.py:48-84 alignment_to_path up to the graph walk).  Intervals are
 * [start, stop) in the caller's list order; sequence edges [l, r] inclusive, sorted by (tid, l) and disjoint (what
 * build_graph + sort_edges produce).  The name set ("concordant_reads") is given through
 * records of this ctx that carry the names (name_rec) and / or as 128-bit name hashes (name_lo, name_hi: the hash of
 * coral_get_record_keys; the multi-GPU form, where the set spans shards); either may be empty.
 * One row per (interval, record) pair whose trimmed run keeps >= 3 edges: (record, interval index, first edge, last
 * edge), in the reference's order — interval list order, then file order.  A record of the set whose reference_end
 * pysam reports as None (unmapped flag or no CIGAR) makes the reference raise TypeError: CORAL_E_FORMAT. */
int coral_alignment_paths(coral_ctx* ctx, const int32_t* iv_tid, const int32_t* iv_start, const int32_t* iv_stop,
                          int32_t n_iv, const int32_t* se_tid, const int32_t* se_l, const int32_t* se_r, int32_t n_se,
                          const uint32_t* name_rec, int64_t n_name_rec, const uint64_t* name_lo, const uint64_t* name_hi,
                          int64_t n_name_hash, int32_t min_overlap, int32_t min_mapq, int64_t* n_rows);
int coral_get_alignment_paths(coral_ctx* ctx, uint32_t* rec, int32_t* iv, int32_t* first, int32_t* last);

/* ---- device-side BAM ingest (replaces pysam.AlignmentFile(path,'rb'), ibg:65, for callers
 * that opt in) -----------------------------------------------------------------------------------------------------
 * The whole .bam file (BGZF bytes) is handed over as it is on disk; members are inflated and the alignment records
 * framed on the GPU (csrc/ingest.cu), so the host does no zlib work and only compressed bytes cross PCIe.
 *   coral_ingest_open : host only — BGZF member table and BAM header; `file` must stay valid until coral_ingest_run.
 *                       The caller then reads the reference table (coral_ingest_n_ref / _ref_name / _ref_len) and
 *                       calls coral_set_references, as after coral_bam_read.
 *   coral_ingest_run  : H2D of the compressed bytes, inflate, framing, then coral_load_records(CORAL_RECORDS_DEVICE)
 *                       on the framed buffer (owned by the ctx) — the ctx is then in the same state as after a load
 *                       of the host reader's records, with identical offsets.
 *   coral_get_read_names : read names of the given records (any placement), 256 bytes per record, NUL padded — what
 *                       a caller without a host copy of the records uses for pysam's query_name.
 *   coral_ingest_host_check : the same decoder and framing steps run on the CPU (test hook; mode 1 forces the serial
 *                       chase).  Verified on the CPU against the zlib reader and on the B200 (tests/test_gpu_device_ingest.py); timing in
 *                       DESIGN.md 4.3. */
int coral_ingest_open(coral_ctx* ctx, const uint8_t* file, uint64_t n_bytes);
int32_t coral_ingest_n_ref(const coral_ctx* ctx);
const char* coral_ingest_ref_name(const coral_ctx* ctx, int32_t tid);
int64_t coral_ingest_ref_len(const coral_ctx* ctx, int32_t tid);
int coral_ingest_run(coral_ctx* ctx, int64_t* n_records, uint64_t* framed_bytes);
int coral_ingest_used_serial_chase(const coral_ctx* ctx);
int coral_get_read_names(coral_ctx* ctx, const uint32_t* rec, int64_t n, uint8_t* out);
int coral_ingest_host_check(const uint8_t* file, uint64_t n_bytes, int mode, uint8_t* recs_out, uint64_t recs_cap,
                            uint64_t* offsets_out, int64_t rec_cap, int64_t* n_records, uint64_t* framed_bytes,
                            int64_t* n_candidates, int32_t* used_serial, char* err, size_t err_len);

#ifdef __cplusplus
}
#endif
#endif /* CORAL_B200_H */
