"""oracle/coral_oracle.py — TEST INFRASTRUCTURE: CPU restatement of CoRAL's reconstruct hot path.

This is the checker for the CUDA path.  It is imported only by ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s cpu_baseline / ``--impl reference`` legs; the product never routes through it.

Every function cites the reference lines it restates (paths relative to ``/root/reference/src``):
``cp`` = cigar_parsing.py, ``bu`` = breakpoint_utilities.py, ``ibg`` = infer_breakpoint_graph.py.
The restatement is pinned by fixtures under ``tests/golden/`` that were produced by executing the
reference's own code in the build container (``tests/golden/make_golden.py``).  The reference
itself ships no tests or golden vectors, and the pysam layer underneath is restated
from documentation (``oracle/pysam_shim.py``) — that part of parity is unpinned.
"""
import math
import re
from collections import Counter

import numpy as np

CHR_IDX = {('chr%d' % (i + 1)): i for i in range(22)}
CHR_IDX.update({'chrX': 22, 'chrY': 23, 'chrM': 24})  # global_names.py:13-18
FLIP = {'+': '-', '-': '+'}  # global_names.py:9

_OPS = re.compile(r"(\d+)([A-Z=])")

# SURVEY.md App. A.2 / cp:17-215.  For signature -> (al, plus(qs,qe), minus(qs,qe)); n = numbers in order.
_SIG = {
    "SM":   (lambda n, rl: n[1],        lambda n, rl: (n[0], rl - 1),           lambda n, rl: (0, n[1] - 1)),
    "MS":   (lambda n, rl: n[0],        lambda n, rl: (0, n[0] - 1),            lambda n, rl: (n[1], rl - 1)),
    "SMS":  (lambda n, rl: n[1],        lambda n, rl: (n[0], n[0] + n[1] - 1),  lambda n, rl: (n[2], n[2] + n[1] - 1)),
    "SMD":  (lambda n, rl: n[1] + n[2], lambda n, rl: (n[0], rl - 1),           lambda n, rl: (0, n[1] - 1)),
    "MDS":  (lambda n, rl: n[0] + n[1], lambda n, rl: (0, n[0] - 1),            lambda n, rl: (n[2], rl - 1)),
    "SMDS": (lambda n, rl: n[1] + n[2], lambda n, rl: (n[0], rl - n[3] - 1),    lambda n, rl: (n[3], rl - n[0] - 1)),
    "SMI":  (lambda n, rl: n[1],        lambda n, rl: (n[0], rl - 1),           lambda n, rl: (0, rl - n[0] - 1)),
    "MIS":  (lambda n, rl: n[0],        lambda n, rl: (0, rl - n[2] - 1),       lambda n, rl: (n[2], rl - 1)),
    "SMIS": (lambda n, rl: n[1],        lambda n, rl: (n[0], rl - n[3] - 1),    lambda n, rl: (n[3], rl - n[0] - 1)),
}


def cigar2pos(cigar, strand, read_length):
    """cp:17-215 via the dispatch at cp:219-229 / :254-255.  KeyError for foreign signatures."""
    sig = "".join(ch for ch in cigar if not ch.isdigit())
    al_f, plus_f, minus_f = _SIG[sig]
    nums = [int(x) for x, _ in _OPS.findall(cigar)]
    qs, qe = plus_f(nums, read_length) if strand == '+' else minus_f(nums, read_length)
    return qs, qe, al_f(nums, read_length)


def alignment_from_satags(sa_list, read_length):
    """cp:232-269."""
    rows = []
    for entry in sa_list:
        f = entry.split(',')
        if 'S' not in f[3] or 'M' not in f[3]:  # cp:248-253 — whole read rejected, 3-tuple
            return ([], [], [])
        qs, qe, al = cigar2pos(f[3], f[2], read_length)
        p = int(f[1])
        ref = [f[0], p - 1, p + al - 2, '+'] if f[2] == '+' else [f[0], p + al - 2, p - 1, '-']  # cp:257-260
        rows.append(([qs, qe], ref, int(f[4]), float(f[-1])))
    rows.sort(key=lambda r: (r[0][0], r[0][1]))  # stable, cp:263
    return ([r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows],
            [r[3] / (r[0][1] - r[0][0]) for r in rows])  # cp:268


def interval_overlap(a, b):
    """bu:11-15 (literal; degenerates to containment for reversed '-' intervals, App. A.4)."""
    return a[0] == b[0] and int(a[1]) <= int(b[2]) and int(b[1]) <= int(a[2])


def interval_overlap_l(a, lst):
    """bu:37-44: first matching index or -1."""
    for k, b in enumerate(lst):
        if interval_overlap(a, b):
            return k
    return -1


def interval2bp(R1, R2, r=(), rgap=0):
    """bu:289-295."""
    c1, c2 = CHR_IDX[R1[0]], CHR_IDX[R2[0]]
    if c2 < c1 or (c2 == c1 and R2[1] < R1[2]):
        return [R1[0], R1[2], R1[3], R2[0], R2[1], FLIP[R2[3]], r, rgap, 0]
    return [R2[0], R2[1], FLIP[R2[3]], R1[0], R1[2], R1[3], (r[0], r[2], r[1]), rgap, 1]


def _fires(a, b, qa, qb, gap_):
    """Same-strand / opposite-strand firing rule of bu:146-161 and :174-185 (without MAPQ terms)."""
    if a[3] != b[3]:
        return True
    gr = qb[0] - qa[1]
    grr = (b[1] - a[2]) if b[3] == '+' else (a[2] - b[1])
    return abs(gr - grr) > max(gap_, abs(gr * 0.2))


def alignment2bp_l(rn, ca, min_bp_match_cutoff, min_mapq, gap_, intrvls, gap_mapq=10):
    """bu:129-186."""
    q, r, mq = ca[0], ca[1], ca[2]
    n = len(r)
    out, assigned = [], [0] * max(n - 1, 0)
    for i in range(n - 1):
        io1, io2 = interval_overlap_l(r[i], intrvls), interval_overlap_l(r[i + 1], intrvls)
        gap = int(q[i + 1][0]) - int(q[i][1])
        if gap + min_bp_match_cutoff >= 0 and io1 >= 0 and io1 == io2:
            if _fires(r[i], r[i + 1], q[i], q[i + 1], gap_) and mq[i] >= min_mapq and mq[i + 1] >= min_mapq:
                out.append(interval2bp(r[i], r[i + 1], (rn, i, i + 1), gap) + [mq[i], mq[i + 1]])
                assigned[i] = 1
    for i in range(1, n - 1):
        io1, io2 = interval_overlap_l(r[i - 1], intrvls), interval_overlap_l(r[i + 1], intrvls)
        if assigned[i - 1] == 0 and assigned[i] == 0 and mq[i] < gap_mapq and mq[i - 1] >= min_mapq and \
                mq[i + 1] >= min_mapq and io1 >= 0 and io1 == io2:
            if _fires(r[i - 1], r[i + 1], q[i - 1], q[i + 1], gap_):
                out.append(interval2bp(r[i - 1], r[i + 1], (rn, i - 1, i + 1), int(q[i + 1][0]) - int(q[i - 1][1]))
                           + [mq[i - 1], mq[i + 1]])
    return out


def alignment2bp(rn, ca, min_bp_match_cutoff, min_mapq, intrvl1, intrvl2, gap_mapq=10):
    """bu:70-96."""
    q, r, mq = ca[0], ca[1], ca[2]
    n = len(r)
    out, assigned = [], [0] * max(n - 1, 0)

    def spans(x, y):
        return (interval_overlap(x, intrvl1) and interval_overlap(y, intrvl2)) or \
               (interval_overlap(y, intrvl1) and interval_overlap(x, intrvl2))

    for i in range(n - 1):
        gap = int(q[i + 1][0]) - int(q[i][1])
        if gap + min_bp_match_cutoff >= 0 and mq[i] >= min_mapq and mq[i + 1] >= min_mapq and spans(r[i], r[i + 1]):
            out.append(interval2bp(r[i], r[i + 1], (rn, i, i + 1), gap) + [mq[i], mq[i + 1]])
            assigned[i] = 1
    for i in range(1, n - 1):
        if assigned[i - 1] == 0 and assigned[i] == 0 and mq[i] < gap_mapq and mq[i - 1] >= min_mapq and \
                mq[i + 1] >= min_mapq and spans(r[i - 1], r[i + 1]):
            out.append(interval2bp(r[i - 1], r[i + 1], (rn, i - 1, i + 1), int(q[i + 1][0]) - int(q[i - 1][1]))
                       + [mq[i - 1], mq[i + 1]])
    return out


def cluster_bp_list(bp_list, min_cluster_size, bp_distance_cutoff):
    """bu:252-286: group by (chr1, chr2, o1, o2) in first-appearance order, then greedy first-fit."""
    groups = {}
    for k, bp in enumerate(bp_list):
        groups.setdefault((bp[0], bp[3], bp[2], bp[5]), []).append(k)
    clusters = []
    for members in groups.values():
        if len(members) < min_cluster_size:
            clusters.append([bp_list[k] for k in members])
            continue
        local = []
        for k in members:
            bp = bp_list[k]
            home = None
            for cl in local:
                if any(abs(int(bp[1]) - int(m[1])) < bp_distance_cutoff and
                       abs(int(bp[4]) - int(m[4])) < bp_distance_cutoff for m in cl):
                    home = cl
                    break
            if home is None:
                local.append([bp])
            else:
                home.append(bp)
        clusters.extend(local)
    return clusters


def bp_match(bp1, bp2, rgap, cut):
    """bu:391-416."""
    if not (bp1[0] == bp2[0] and bp1[3] == bp2[3] and bp1[2] == bp2[2] and bp1[5] == bp2[5]):
        return False
    d1, d2 = int(bp1[1]) - int(bp2[1]), int(bp1[4]) - int(bp2[4])
    near1, near2 = abs(d1) < cut[0], abs(d2) < cut[1]
    if rgap <= 0:
        return near1 and near2
    slack, used1, used2 = rgap, False, False
    if bp1[2] == '+' and d1 <= -cut[0]:
        slack -= (-d1 - cut[0] + 1); used1 = True
    if bp1[2] == '-' and d1 >= cut[0]:
        slack -= (d1 - cut[0] + 1); used1 = True
    if bp1[5] == '+' and d2 <= -cut[1]:
        slack -= (-d2 - cut[1] + 1); used2 = True
    if bp1[5] == '-' and d2 >= cut[1]:
        slack -= (d2 - cut[1] + 1); used2 = True
    return ((used1 and slack >= 0) or near1) and ((used2 and slack >= 0) or near2)


def _call_position(vals, orient):
    """bu:336-357: unique mode, else median with ceil ('+') / floor ('-') on even counts."""
    top = Counter(vals).most_common(2)
    if len(top) == 1 or top[0][1] > top[1][1]:
        return top[0][0]
    med = np.median(vals)
    if len(vals) % 2 == 1:
        return int(med)
    return int(math.ceil(med)) if orient == '+' else int(math.floor(med))


def bpc2bp(bp_cluster, bp_distance_cutoff):
    """bu:299-388."""
    n = len(bp_cluster)
    bp = bp_cluster[0][:-2]
    bp[1] = 0 if bp[2] == '+' else 1000000000
    bp[4] = 0 if bp[5] == '+' else 1000000000
    floor_sd = bp_distance_cutoff / 2.99
    m1 = sum(b[1] for b in bp_cluster) / (n * 1.0)
    m2 = sum(b[4] for b in bp_cluster) / (n * 1.0)
    e1 = sum(b[1] * b[1] for b in bp_cluster) / (n * 1.0)
    e2 = sum(b[4] * b[4] for b in bp_cluster) / (n * 1.0)

    def sd(e, m):
        v = e - m * m
        return max(floor_sd, math.sqrt(v)) if v >= 0 else floor_sd  # sqrt(<0) -> except -> floor (bu:317-324)

    s1, s2 = sd(e1, m1), sd(e2, m2)
    kept = [b for b in bp_cluster if m1 - 3 * s1 <= b[1] <= m1 + 3 * s1 and m2 - 3 * s2 <= b[4] <= m2 + 3 * s2]
    if kept:
        last = bp_cluster[-1]  # the loop-leaked `bp_` at bu:343/354
        bp[1] = _call_position([b[1] for b in kept], last[2])
        bp[4] = _call_position([b[4] for b in kept], last[5])
    bpr, rest = [], []
    acc = [0, 0, 0, 0, 0, 0]
    for b in bp_cluster:
        if bp_match(b, bp, b[7] * 1.2, [bp_distance_cutoff, bp_distance_cutoff]):
            bpr.append(b[6])
            acc[0] += b[1]; acc[2] += b[1] * b[1]; acc[1] += b[4]; acc[3] += b[4] * b[4]
            if b[-3] == 0:
                acc[4] += b[-2]; acc[5] += b[-1]
            else:
                acc[4] += b[-1]; acc[5] += b[-2]
        else:
            rest.append(b)
    if not bpr:
        return bp, bpr, [0, 0, 0, 0, 0, 0], []
    st = [a / (len(bpr) * 1.0) for a in acc]
    for k in (2, 3):
        v = st[k] - st[k - 2] * st[k - 2]
        st[k] = math.sqrt(v) if v >= 0 else 0  # bu:380-387
    return bp, bpr, st, rest


# ------------------------------------------------------------------------------------------------
# L3: the hot methods of class bam_to_breakpoint_nanopore (ibg:20-1330), restated over the pysam shim.
# ------------------------------------------------------------------------------------------------
class OracleHotPath(object):
    """CPU restatement of the hot methods.  State attributes mirror ibg:22-61 but are per-instance.

    The order-dependent host steps that are NOT on the hot path (hash_alignment_to_seg,
    find_amplicon_intervals, build_graph) come from ``coral_b200.host_steps.HostSteps`` when a full
    pipeline is assembled (tests/); they are pinned separately by the golden stage dumps.
    """
    min_bp_match_cutoff_ = 100          # ibg:28
    max_breakpoint_distance_cutoff = 2000   # ibg:31
    min_del_len = 600                   # ibg:33

    def __init__(self, lr_bamfile, seedfile=None, amplicon_intervals=None):
        from oracle import pysam_shim
        self.lr_bamfh = pysam_shim.AlignmentFile(lr_bamfile, 'rb')  # ibg:65
        self.min_cluster_cutoff = 3
        self.min_bp_cov_factor = 1.0
        self.normal_cov = 0.0
        self.read_length = {}
        self.chimeric_alignments = {}
        self.large_indel_alignments = {}
        self.nm_stats = [0.0, 0.0, 0]
        self.amplicon_intervals = []
        self.amplicon_interval_connections = {}
        self.new_bp_list, self.new_bp_stats, self.new_bp_ccids = [], [], []
        self.lr_graph = []
        if seedfile is not None:  # ibg:66-69
            with open(seedfile) as fp:
                for line in fp:
                    s = line.strip().split()
                    self.amplicon_intervals.append([s[0], int(s[1]), int(s[2]), -1])
        if amplicon_intervals is not None:
            self.amplicon_intervals = [list(a) for a in amplicon_intervals]

    # -- a12: coverage part of read_cns (ibg:129-135) ------------------------------------------------
    def normal_coverage(self, segments, total_int_len):
        """segments: [[chr, start, end_inclusive], ...] chosen at ibg:105-125."""
        nnc = 0
        for c, s, e in segments:
            nnc += sum(sum(a) for a in self.lr_bamfh.count_coverage(c, s, e + 1, quality_threshold=0,
                                                                    read_callback='nofilter'))
        self.normal_cov = nnc * 1.0 / total_int_len
        self.min_cluster_cutoff = max(self.min_cluster_cutoff, self.min_bp_cov_factor * self.normal_cov)
        return nnc

    # -- a1: ibg:139-174 -------------------------------------------------------------------------------
    def fetch(self):
        sa_of = {}
        for read in self.lr_bamfh.fetch():
            name = read.query_name
            if read.flag < 256 and name not in self.read_length:
                self.read_length[name] = read.query_length
            try:
                tag = read.get_tag('SA:Z:')
            except KeyError:
                tag = None
            if tag is not None:
                lst = sa_of.setdefault(name, [])
                for entry in tag[:-1].split(';'):
                    if entry not in lst:
                        lst.append(entry)
            elif read.mapping_quality == 60:
                e = read.get_cigar_stats()[0][-1] / read.query_length
                self.nm_stats[0] += e
                self.nm_stats[1] += e * e
                self.nm_stats[2] += 1
        self.nm_stats[0] /= self.nm_stats[2]
        self.nm_stats[1] = math.sqrt(self.nm_stats[1] / self.nm_stats[2] - self.nm_stats[0] ** 2)
        self.chimeric_alignments = {}
        for name, lst in sa_of.items():
            if name in self.read_length:  # ibg:162-173: reads without a primary are dropped
                self.chimeric_alignments[name] = alignment_from_satags(lst, self.read_length[name])

    # -- a9: ibg:326-340 -------------------------------------------------------------------------------
    def addbp(self, bp_, bpr_, bp_stats_, ccid):
        for k, bp in enumerate(self.new_bp_list):
            if bp[0] == bp_[0] and bp[3] == bp_[3] and bp[2] == bp_[2] and bp[5] == bp_[5] and \
                    abs(bp[1] - bp_[1]) < 200 and abs(bp[4] - bp_[4]) < 200:
                bp[-1] |= set(bpr_)
                return k
        self.new_bp_list.append(bp_ + [bpr_])
        self.new_bp_ccids.append(ccid)
        self.new_bp_stats.append(bp_stats_)
        return len(self.new_bp_list) - 1

    def _call_clusters(self, candidates):
        """Shared tail of ibg:691-718 and ibg:775-802: cluster -> peel with bpc2bp -> accept -> addbp."""
        cutoff = self.min_cluster_cutoff
        for c in cluster_bp_list(candidates, cutoff, self.max_breakpoint_distance_cutoff):
            if len(c) < cutoff:
                continue
            sub, rest = 0, c
            while len(rest) >= cutoff:
                bp, bpr, stats, rest = bpc2bp(rest, self.min_bp_match_cutoff_)
                support = len(set(bpr))
                if (sub == 0 and support >= cutoff) or support >= max(self.normal_cov * self.min_bp_cov_factor, 3.0):
                    io1 = interval_overlap_l([bp[0], bp[1], bp[1]], self.amplicon_intervals)
                    io2 = interval_overlap_l([bp[3], bp[4], bp[4]], self.amplicon_intervals)
                    if io1 >= 0 and io2 >= 0:
                        assert self.amplicon_intervals[io1][3] == self.amplicon_intervals[io2][3]
                        k = self.addbp(bp, set(bpr), stats, self.amplicon_intervals[io1][3])
                        self.amplicon_interval_connections.setdefault((min(io1, io2), max(io1, io2)), set()).add(k)
                sub += 1

    # -- a3: ibg:721-802 (nm_filter is hard-wired False, ibg:40) ----------------------------------------
    def smalldel_candidates(self):
        for ai in self.amplicon_intervals:
            for read in self.lr_bamfh.fetch(ai[0], ai[1], ai[2] + 1):
                mq = read.mapping_quality
                if mq < 20:
                    continue
                blocks = read.get_blocks()
                for k in range(len(blocks) - 1):
                    if abs(blocks[k + 1][0] - blocks[k][1]) > self.min_del_len:
                        self.large_indel_alignments.setdefault(read.query_name, []).append(
                            [ai[0], blocks[k + 1][0], blocks[k][1], blocks[0][0], blocks[-1][1], mq])
        cands = []
        for name, rows in self.large_indel_alignments.items():
            for k, row in enumerate(rows):
                hi, lo = row[1], row[2]
                if lo > hi:      # ibg:769-771 (aliased swap collapses both to row[1]; unreachable for real CIGARs)
                    lo = hi
                cands.append([row[0], hi, '-', row[0], lo, '+', (name, k, k), 0, 0, -1, -1])
        return cands

    def find_smalldel_breakpoints(self):
        self._call_clusters(self.smalldel_candidates())

    # -- a4: ibg:676-718 -------------------------------------------------------------------------------
    def chimeric_candidates(self):
        cands = []
        for name, ca in self.chimeric_alignments.items():
            cands += alignment2bp_l(name, ca, self.min_bp_match_cutoff_, 20, 100, self.amplicon_intervals)
        return cands

    def find_breakpoints(self):
        self._call_clusters(self.chimeric_candidates())

    # -- a10/a11: ibg:1019-1056 ------------------------------------------------------------------------
    def assign_cov(self):
        fh = self.lr_bamfh
        d = self.min_bp_match_cutoff_
        for g in self.lr_graph:
            for se in g.sequence_edges:
                if se[5] == -1:
                    se[5] = len([r for r in fh.fetch(se[0], se[1], se[2] + 1) if r.infer_read_length()])
                    se[6] = sum(sum(a) for a in fh.count_coverage(se[0], se[1], se[2] + 1, quality_threshold=0,
                                                                  read_callback='nofilter'))
        for g in self.lr_graph:
            for ce in g.concordant_edges:
                def names(c, p):
                    return set(r.query_name for r in fh.fetch(contig=c, start=p, stop=p + 1))
                rls, rrs = names(ce[0], ce[1]), names(ce[3], ce[4])
                rls1, rrs1 = names(ce[0], ce[1] - d - 1), names(ce[3], ce[4] + d)
                rbps = set()
                for node in ((ce[0], ce[1], ce[2]), (ce[3], ce[4], ce[5])):
                    for bpi in g.nodes[node][2]:
                        for r in g.discordant_edges[bpi][10]:
                            rbps.add(r[0])
                ce[9] = rls | rrs
                ce[8] = len((rls & rrs & rls1 & rrs1) - rbps)


    # -- §8f-2: concordant-read pass of compute_path_constraints (ibg:1303-1310) --------------------------
    def _concordant_read_runs(self, amplicon_idx, intervals, concordant_reads):
        """Every record of the amplicon's intervals (list order, file order) with MAPQ >= 20 and a name in
        `concordant_reads` -> its run of sequence edges (path_constraints.py:60-84); distinct runs with their record
        counts in first-seen order.  The graph walk that spells a run out as a path is shared host code."""
        g = self.lr_graph[amplicon_idx]
        runs = {}
        for aint in intervals:
            for read in self.lr_bamfh.fetch(aint[0], aint[1], aint[2] + 1):
                if read.mapq >= 20 and read.query_name in concordant_reads:
                    ids = alignment_run(g, [read.reference_name, read.reference_start, read.reference_end])
                    if ids:
                        key = (ids[0], ids[-1])
                        runs[key] = runs.get(key, 0) + 1
        return [(a, b, n) for (a, b), n in runs.items()]


def alignment_run(g, rint, min_overlap=500):
    """path_constraints.py:60-84: sequence edges touched by [chr, start, end], by start; the first / last edge is
    dropped when the alignment shares < min_overlap with it, then edges shorter than min_overlap are dropped from
    that end (never the last one left); runs of fewer than three edges give no path."""
    lst = [i for i, sseg in enumerate(g.sequence_edges) if interval_overlap(rint, sseg)]
    if len(lst) == 0:
        return []
    lst = sorted(lst, key=lambda i: g.sequence_edges[i][1])

    def shared(i):
        return min(g.sequence_edges[i][2], rint[2]) - max(g.sequence_edges[i][1], rint[1])
    for end in (0, -1):
        if len(lst) > 1 and shared(lst[end]) < min_overlap:
            del lst[end]
        while len(lst) > 1 and g.sequence_edges[lst[end]][7] < min_overlap:
            del lst[end]
    return lst if len(lst) > 2 else []


# ---- remaining interval helpers used by the host steps (bu:18-34, :54-67) ---------------------------------
def interval_adjacent(a, b):
    """bu:25-34."""
    if a[0] != b[0]:
        return False
    return (b[1] == a[2] + 1) if a[1] <= b[1] else (a[1] == b[2] + 1)


def interval_exclusive(a, lst):
    """bu:54-67."""
    covered, parts = set(), [[x for x in a]]
    for k in range(len(lst)):
        for j in range(len(parts))[::-1]:
            p = parts[j]
            if interval_overlap(p, lst[k]):
                covered.add(k)
                del parts[j]
                if p[1] < lst[k][1]:
                    parts.append([p[0], p[1], lst[k][1] - 1, -1])
                if p[2] > lst[k][2]:
                    parts.append([p[0], lst[k][2] + 1, p[2], -1])
    return covered, parts
