"""The coverage track of reference ``plot_amplicons.py`` (``graph_vis.plot_graph``, ``:376-412``) — the second caller of
the coverage kernel: per amplified interval, the mean long-read coverage of fixed-size windows,
``sum(count_coverage(chrom, w, w + window_size, quality_threshold, 'nofilter')) / window_size``, window by window
through pysam in the reference, here all windows of the figure in a few ``coral_region_coverage`` calls.

Only the numbers are produced: the rectangles the reference draws are returned as rows, matplotlib is not touched
(the figure, gene track, arcs and file parsing of ``plot_amplicons.py`` are out of scope).

``quality_threshold`` is pysam's BASE-quality threshold (the reference passes ``--min_mapq`` into it, ``:935``; default
0).  With 0 every aligned base counts, which is what the record scan computes; any other value needs the QUAL bytes,
which the scan never reads: ``NotImplementedError``.
"""
import numpy as np


def window_size_of(int_, plot_bounds=None):
    """reference :388-396."""
    ival_len = int_[1] - int_[0]
    if plot_bounds:
        ival_len = plot_bounds[2] - plot_bounds[1]
    if ival_len >= 1000000:
        return 10000
    if ival_len >= 100000:
        return 1000
    return 150


def coverage_windows(intervals_from_graph, sorted_chrs, plot_bounds=None):
    """The windows of reference :378-412 in drawing order -> [(chrom, interval index, window start, window size)]:
    per interval the regular windows from its start, then one more window aligned to its end when the remainder of
    (length + 1) / window_size is not zero (it overlaps the last regular ones)."""
    out = []
    for chrom in sorted_chrs:
        for inti, int_ in enumerate(intervals_from_graph[chrom]):
            if plot_bounds:
                if chrom != plot_bounds[0]:
                    continue
                if not (int_[1] >= plot_bounds[1] and int_[0] <= plot_bounds[2]):
                    continue
            ws = window_size_of(int_, plot_bounds)
            for w in range(int_[0], int_[1], ws):
                out.append((chrom, inti, w, ws))
            w = int_[1] - ((int_[1] - int_[0] + 1) % ws)
            if w < int_[1]:
                out.append((chrom, inti, w, ws))
    return out


def _disjoint_batches(regions):
    """regions [(tid, start, stop, key)] -> batches sorted by (tid, start) with no two regions overlapping (what
    coral_region_coverage takes); first fit, so overlapping tails go to a second batch."""
    batches, last = [], []
    for r in sorted(regions, key=lambda x: (x[0], x[1], x[2])):
        for b in range(len(batches)):
            if last[b][0] != r[0] or last[b][1] <= r[1]:
                batches[b].append(r)
                last[b] = (r[0], r[2])
                break
        else:
            batches.append([r])
            last.append((r[0], r[2]))
    return batches


def windowed_coverage(region_coverage, references, intervals_from_graph, sorted_chrs, plot_bounds=None, quality_threshold=0):
    """-> (rows, max_cov): rows [(chrom, interval index, window start, window size, mean coverage)] in the reference's
    drawing order, max_cov as the reference tracks it (:401-402).
    `region_coverage(tid[], start[], stop[]) -> (base_sum[], rec_count[])` for disjoint regions sorted by (tid, start):
    `Context.region_coverage` of a ctx that holds the records (or `bam_to_breakpoint_nanopore._region_counts`)."""
    if quality_threshold != 0:
        raise NotImplementedError("base-quality filtered coverage (quality_threshold = %r) needs the QUAL bytes, which the "
                                  "record scan does not read" % (quality_threshold,))
    tid_of = {c: i for i, c in enumerate(references)}
    wins = coverage_windows(intervals_from_graph, sorted_chrs, plot_bounds)
    sums = np.zeros(len(wins), np.int64)
    regions = []
    for k, (chrom, _, w, ws) in enumerate(wins):
        if chrom not in tid_of:
            raise KeyError(chrom)                  # pysam: ValueError/KeyError for a contig the BAM does not have
        regions.append((tid_of[chrom], max(w, 0), w + ws, k))
    for batch in _disjoint_batches(regions):
        bs, _ = region_coverage([r[0] for r in batch], [r[1] for r in batch], [r[2] for r in batch])
        for r, v in zip(batch, bs):
            sums[r[3]] = int(v)
    rows, max_cov = [], 0
    for k, (chrom, inti, w, ws) in enumerate(wins):
        cov = int(sums[k]) * 1.0 / ws
        if cov > max_cov:
            max_cov = cov
        rows.append((chrom, inti, w, ws, cov))
    return rows, max_cov
