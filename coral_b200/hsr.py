"""Drop-in for the breakpoint search of reference ``hsr.py`` (``CoRAL.py hsr``: where does an ecDNA integrate into a
chromosome) — the second caller of the hot operators: ``fetch`` (``hsr.py:21-49``) is the same
one pass over the records as ``bam_to_breakpoint_nanopore.fetch`` and runs on the same kernels; the breakpoint list of
``locate_hsrs`` (``:112-149``) is clustered and called (``:151-173``) by ONE device call (cluster_bp_list + the bpc2bp
peel loop, csrc/cluster.cu).  Candidate generation walks the few alignments of each chimeric read with the interval
predicates, as the reference does; the figure (``:175-225``, matplotlib) is not drawn — ``locate_hsrs`` prints the
reference's lines and returns the points it would plot.

Reference quirks kept: the skip-pair test calls ``interval_overlap`` (not ``_l``) on the LIST of ecDNA intervals
(``:143,147``), which is always ``False >= 0`` i.e. true; called breakpoints keep their read tuples as a LIST
(``:172``), so merging a second call into an existing breakpoint (``:167``, ``list |= set``) raises ``TypeError``.
"""
import sys

from . import global_names, gpu_ops
from .breakpoint_utilities import interval2bp, interval_include, interval_overlap_l


def fetch(lr_bamfh, device=0):
    """reference :21-49.  `lr_bamfh`: coral_b200._abi.BamFile (or any record set the engine accepts), or the PATH of a
    .bam file: the file is then inflated and framed on the GPU (placement "device_ingest": the host reader's zlib pass
    is most of what this call costs on a large file).
    -> (read_length of the chimeric reads, chimeric_alignments); reads without a primary record are dropped."""
    from .infer_breakpoint_graph import bam_to_breakpoint_nanopore
    if isinstance(lr_bamfh, str):
        eng = bam_to_breakpoint_nanopore(lr_bamfh, None, device=device, placement="device_ingest")
        eng._nm_stats_required = False
        eng._warn_no_primary = False
        eng.fetch()
        out = eng.read_length, eng.chimeric_alignments
        eng.closebam()
        return out
    slack = int(getattr(lr_bamfh, "rec_slack", 0))
    # only core / name / aux bytes are needed: the lazy placement never runs the CIGAR + SEQ scan for this caller
    eng = bam_to_breakpoint_nanopore(None, None, device=device, records=lr_bamfh,
                                     placement="mapped_lazy" if slack >= 64 and lr_bamfh.n_records else "copy")
    eng._nm_stats_required = False           # hsr.fetch keeps no NM statistics (and must not fail without them)
    eng._warn_no_primary = False
    eng.fetch()
    out = eng.read_length, eng.chimeric_alignments
    eng._ctx.close()
    return out


def candidate_breakpoints(chimeric_alignments, ecdna_intervals, min_mapq=20, gap_mapq=10):
    """reference :112-149: reads with an alignment inside an ecDNA interval; junctions with exactly one side on ecDNA."""
    bp_list = []
    for r, ca in chimeric_alignments.items():
        r_int, rr_int, q_ = ca[0], ca[1], ca[2]
        on_cycle = False
        for interval in ecdna_intervals:
            i = interval_overlap_l(interval, rr_int)
            if i >= 0 and interval_include(rr_int[i], interval):
                on_cycle = True
                break
        if not on_cycle:
            continue
        n = len(rr_int)
        ecdna = [interval_overlap_l(a, ecdna_intervals) >= 0 for a in rr_int]
        used = [0] * max(n - 1, 0)
        for ri in range(n - 1):
            if q_[ri] >= min_mapq and q_[ri + 1] >= min_mapq and ecdna[ri] != ecdna[ri + 1]:
                bp_list.append(interval2bp(rr_int[ri], rr_int[ri + 1], (r, ri, ri + 1),
                                           int(r_int[ri + 1][0]) - int(r_int[ri][1])) + [q_[ri], q_[ri + 1]])
                used[ri] = 1
        for ri in range(1, n - 1):
            # :141-144 — the last clause of the reference is `interval_overlap(a, LIST) >= 0`: always true
            if used[ri - 1] == 0 and used[ri] == 0 and q_[ri] < gap_mapq and q_[ri - 1] >= min_mapq and \
                    q_[ri + 1] >= min_mapq and not ecdna[ri - 1]:
                bp_list.append(interval2bp(rr_int[ri - 1], rr_int[ri + 1], (r, ri - 1, ri + 1),
                                           int(r_int[ri + 1][0]) - int(r_int[ri - 1][1])) + [q_[ri - 1], q_[ri + 1]])
    return bp_list


def called_breakpoints_device(bp_list, min_size, bp_match_cutoff, cluster_cutoff, ctx=None):
    """cluster_bp_list(bp_list, min_size, cluster_cutoff) and, per cluster of at least min_size members, the bpc2bp peel
    loop (reference :151-157) as one coral_cluster_call -> (bp, bpr, bp_stats_) per call, in the reference's order."""
    if not bp_list:
        return
    ctx = ctx or gpu_ops.default_context()
    a = gpu_ops.flatten_bp_list(bp_list)
    r = gpu_ops.cluster_arrays(ctx, a, min_size, cluster_cutoff, min_size, bp_match_cutoff)
    order, ptr, sub_of = r["order"], r["cluster_ptr"], r["sub_of"]
    for c in range(r["n_clusters"]):
        b, e = int(ptr[c]), int(ptr[c + 1])
        if e - b < min_size:
            continue
        members = order[b:e]
        subs = sub_of[members]
        for k in range(int(r["n_sub"][c])):
            row = b + k
            first = bp_list[int(r["row_first"][row])]
            bp = list(first[:-2])
            bp[1], bp[4] = int(r["row_pos1"][row]), int(r["row_pos2"][row])
            if r["row_nb"][row] == 0:
                yield bp, [], [0, 0, 0, 0, 0, 0]
                continue
            yield bp, [bp_list[int(m)][6] for m in members[subs == k].tolist()], gpu_ops.stats_row(r["row_stats"][row])


def refine(called, min_size, bp_match_cutoff):
    """reference :158-173: keep calls with enough distinct read tuples; a call next to an earlier one would be merged
    into it — which the reference cannot do (`list |= set`), so it raises TypeError there as well."""
    bp_refined, bp_stats = [], []
    for bp, bpr, stats in called:
        if len(set(bpr)) < min_size:
            continue
        for old in bp_refined:
            if bp[0] == old[0] and bp[3] == old[3] and bp[2] == old[2] and bp[5] == old[5] and \
                    abs(bp[1] - old[1]) <= bp_match_cutoff and abs(bp[4] - old[4]) < bp_match_cutoff:
                old[-1] |= set(bpr)            # TypeError, as in the reference (old[-1] is a list)
                break
        else:
            bp_refined.append(bp + [bpr])
            bp_stats.append(stats)
    return bp_refined, bp_stats


def read_ecdna_intervals(cycle_fn, bp_match_cutoff):
    if not cycle_fn.endswith(".bed"):          # the reference converts *_cycles.txt first (cycle2bed, out of scope here)
        sys.stderr.write(cycle_fn + "\n")
        sys.stderr.write("Cycles file must be a converted .bed file!\n")
        sys.exit(1)
    ivals, ext = [], []
    with open(cycle_fn) as fp:
        for line in fp:
            if line.startswith("#"):
                continue
            s = line.strip().split()
            ivals.append([s[0], int(s[1]), int(s[2])])
            ext.append([s[0], int(s[1]) - bp_match_cutoff, int(s[2]) + bp_match_cutoff])
    return ivals, ext


def read_cn_segments(cn_seg):
    cns = {}
    with open(cn_seg) as fp:
        for line in fp:
            s = line.strip().split()
            if line.startswith('chromosome'):
                continue
            if cn_seg.endswith(".cns"):
                cn = 2 * (2 ** float(s[4]))
            elif cn_seg.endswith(".bed"):
                cn = float(s[3])
            else:
                sys.stderr.write(cn_seg + "\n")
                sys.stderr.write("Invalid cn_seg file format!\n")
            cns.setdefault(s[0], []).append([int(s[1]), int(s[2]), cn])
    return cns


def integration_points(bp_refined, ecdna_intervals_ext, cns_dict, normal_cov):
    """reference :176-211: breakpoints with one end on (extended) ecDNA and the other in a CN <= 5 segment with at most
    2.5 x normal coverage support -> [(printed line, x in % of the genome, y = support)]."""
    sizes = global_names.chr_sizes
    total = sum(sizes.values())
    start, agg = {}, 0
    for c in sizes:
        agg += sizes[c]
        start[c] = (agg - sizes[c]) * 100.0 / total
    out = []
    for bp in bp_refined:
        on1 = interval_overlap_l([bp[0], bp[1], bp[1]], ecdna_intervals_ext) >= 0
        on2 = interval_overlap_l([bp[3], bp[4], bp[4]], ecdna_intervals_ext) >= 0
        if on1 == on2:
            continue
        c, p = (bp[3], bp[4]) if on1 else (bp[0], bp[1])
        if c not in start:
            continue
        cn = 0.0
        for seg in cns_dict[c]:
            if seg[0] < p < seg[1]:
                cn = seg[2]
                break
        if cn <= 5.0 and len(bp[-1]) <= float(normal_cov) * 2.5:
            out.append(("Breakpoint %s Support =  %d" % (bp[:6], len(bp[-1])), start[c] + p * 100.0 / total, len(bp[-1])))
    return out


def locate_hsrs(args, device=0, lr_bamfh=None, device_ingest=False):
    """reference :52-225 without the figure.  -> dict(bp_refined, bp_stats, points).
    device_ingest: hand the .bam file to the GPU as it is on disk instead of inflating it with zlib on the host."""
    from . import _abi
    ecdna, ecdna_ext = read_ecdna_intervals(args.cycles, args.bp_match_cutoff)
    print("ecDNA intervals:")
    for ival in ecdna:
        print(ival)
    cns = read_cn_segments(args.cn_seg)
    fh = lr_bamfh if lr_bamfh is not None else (None if device_ingest else _abi.BamFile(args.lr_bam))
    gpu_ops.set_default_device(device)
    _, chimeric = fetch(args.lr_bam if fh is None else fh, device=device)
    print("Fetched %d chimeric alignments." % len(chimeric))
    bp_list = candidate_breakpoints(chimeric, ecdna)
    t = float(args.normal_cov) * 0.5
    bp_refined, bp_stats = refine(called_breakpoints_device(bp_list, t, args.bp_match_cutoff, args.bp_match_cutoff_clustering),
                                  t, args.bp_match_cutoff)
    print("Found %d breakpoints connecting ecDNA and chromosomes." % len(bp_refined))
    if lr_bamfh is None and fh is not None:
        fh.close()
    pts = integration_points(bp_refined, ecdna_ext, cns, args.normal_cov)
    for line, _, _ in pts:
        print(line)
    return {"bp_refined": bp_refined, "bp_stats": bp_stats, "points": [[x, y] for _, x, y in pts]}
