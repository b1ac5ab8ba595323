"""Order-dependent HOST steps of ``reconstruct_graph`` that are not on the hot path:
CN-segment bookkeeping, ``hash_alignment_to_seg``, ``find_amplicon_intervals`` / ``find_interval_i`` (BFS),
``addbp`` and ``build_graph``.  They are tiny, sequential and full of order-dependent list surgery, so they
stay Python — restated here (not imported from the reference, which is absent on the GPU box) with the
reference's quirks kept on purpose.  The data-parallel operators they call (alignment2bp, cluster_bp_list,
bpc2bp) come from ``self._ops`` — the CUDA-backed drop-ins in the product, the oracle's restatement in tests.

Citations: ``ibg`` = reference ``src/infer_breakpoint_graph.py``.
"""
import bisect
import contextlib
import gc
import logging
import sys
import time
from collections import namedtuple

import numpy as np

from . import global_names
from .breakpoint_graph import BreakpointGraph as _LocalBreakpointGraph


def reference_graph_class():
    """The reference's own ``breakpoint_graph.BreakpointGraph`` (reference ``breakpoint_graph.py:83-765``) when that
    module is importable under its bare name — i.e. when the drop-in runs inside the reference (``src/`` on
    ``sys.path``, cvxopt installed): the graphs handed to ``compute_cn_lr`` (``infer_breakpoint_graph.py:1388-1389``) and
    to ``cycle_decomposition.reconstruct_cycles`` are then instances of the class those functions were written for.
    None when it is not importable (the GPU box of this repo: no reference, no cvxopt)."""
    try:
        import breakpoint_graph as ref_bg           # bare name: the reference's module layout (ibg:10-14)
    except Exception:                                # noqa  (ImportError, or cvxopt missing inside it)
        return None
    cls = getattr(ref_bg, "BreakpointGraph", None)
    return cls if cls is not None and hasattr(cls, "compute_cn_lr") else None
from .path_steps import PathConstraintSteps

_Hit = namedtuple("_Hit", ["begin", "end", "data"])


class SegmentIndex(object):
    """Half-open interval lookup with the two calls the reference makes on ``intervaltree.IntervalTree``
    (ibg:88-91 ``tree[a:b] = idx``; ibg:178 ``tree[pos]`` -> set of objects carrying ``.data``)."""

    def __init__(self):
        self._b, self._iv, self._span = [], [], 0

    def __setitem__(self, key, data):
        b, e = key.start, key.stop
        if b >= e:
            raise ValueError("IntervalTree: Null Interval objects not allowed in IntervalTree: Interval(%r, %r, %r)" % (b, e, data))
        k = bisect.bisect_right(self._b, b)
        self._b.insert(k, b)
        self._iv.insert(k, _Hit(b, e, data))
        self._span = max(self._span, e - b)

    def __getitem__(self, pos):
        out = set()
        k = bisect.bisect_right(self._b, pos) - 1
        while k >= 0 and self._b[k] >= pos - self._span:
            if self._iv[k].end > pos:
                out.add(self._iv[k])
            k -= 1
        return out


@contextlib.contextmanager
def gc_paused():
    """The host steps build hundreds of thousands of small acyclic lists; the cyclic collector finds nothing in
    them and costs more than the building itself, so it is paused while they are built."""
    on = gc.isenabled()
    gc.disable()
    try:
        yield
    finally:
        if on:
            gc.enable()


def _t():
    return "#TIME " + '%.4f\t' % (time.time() - global_names.TSTART)


class HostSteps(PathConstraintSteps):
    # class constants of the reference (ibg:25-33)
    max_seq_len = 2000000
    cn_gain = 5.0
    min_bp_match_cutoff_ = 100
    interval_delta = 100000
    max_breakpoint_distance_cutoff = 2000
    min_del_len = 600
    nm_filter = False
    graph_cls = None            # container class of lr_graph; None = the reference's BreakpointGraph when importable

    def _init_state(self):
        """Per-instance containers (the reference keeps them as class attributes, ibg:22-61; App. A.10)."""
        self.lr_graph = []
        self.min_bp_cov_factor = 1.0
        self.min_cluster_cutoff = 3
        self.read_length = {}
        self.chimeric_alignments = {}
        self.chimeric_alignments_seg = {}
        self._read_seg_pairs = {}    # read -> [(chromosome, CN-segment index)] of its alignments (find_interval_i's cache)
        self.large_indel_alignments = {}
        self.nm_stats = [0.0, 0.0, 0]
        self.amplicon_intervals = []
        self.amplicon_interval_connections = {}
        self.cns_intervals = []
        self.cns_intervals_by_chr = {}
        self.log2_cn = []
        self.cns_tree = {}
        self.normal_cov = 0.0
        self.ccid2id = {}
        self.new_bp_list = []
        self.new_bp_stats = []
        self.new_bp_ccids = []
        self.source_edges = []
        self.source_edge_ccids = []
        self.path_constraints = {}
        self.longest_path_constraints = {}
        self.cycles = {}
        self.cycle_weights = {}
        self.path_constraints_satisfied = {}

    def _read_seeds(self, seedfile):
        with open(seedfile, 'r') as fp:  # ibg:66-69
            for line in fp:
                s = line.strip().split()
                self.amplicon_intervals.append([s[0], int(s[1]), int(s[2]), -1])

    # ---- read_cns: file parsing and segment choice (ibg:79-127); the coverage call is the hot part -----
    def _parse_cns(self, cns):
        self.cns_intervals, self.log2_cn = [], []
        idx = 0
        with open(cns, 'r') as fp:
            for line in fp:
                s = line.strip().split()
                if s[0] == "chromosome":
                    continue
                self.cns_intervals.append([s[0], int(s[1]), int(s[2]) - 1])
                if s[0] not in self.cns_tree:
                    self.cns_tree[s[0]] = SegmentIndex()
                    self.cns_intervals_by_chr[s[0]] = []
                    idx = 0
                self.cns_tree[s[0]][int(s[1]): int(s[2])] = idx
                idx += 1
                if cns.endswith(".cns"):
                    self.cns_intervals_by_chr[s[0]].append([s[0], int(s[1]), int(s[2]) - 1, 2 * (2 ** float(s[4]))])
                    self.log2_cn.append(float(s[4]))
                elif cns.endswith(".bed"):
                    self.cns_intervals_by_chr[s[0]].append([s[0], int(s[1]), int(s[2]) - 1, float(s[3])])
                    self.log2_cn.append(np.log2(float(s[3]) / 2.0))
                else:
                    sys.stderr.write(cns + "\n")
                    sys.stderr.write("Invalid cn_seg file format!\n")

    def _median_cn_segments(self):
        """Segments around the median log2 CN until >= 10 Mb (ibg:105-125; IndexError when there are too few)."""
        order = np.argsort(self.log2_cn)
        im = int(len(order) / 2.4)
        ip = im + 1
        chosen = [self.cns_intervals[order[ip]], self.cns_intervals[order[im]]]
        total = sum(s[2] - s[1] + 1 for s in chosen)
        i = 1
        while total < 10000000:
            for k in (ip + i, im - i):
                seg = self.cns_intervals[order[k]]
                chosen.append(seg)
                total += seg[2] - seg[1] + 1
            i += 1
        return chosen, total

    def pos2cni(self, chr, pos):
        return self.cns_tree[chr][pos]

    # ---- ibg:181-210 ------------------------------------------------------------------------------------
    def _cni_tables(self):
        """Per chromosome (begins, ends, data) arrays of the CN segments when they do not overlap, else None
        (then every lookup goes through pos2cni, which keeps the reference's assertion on double hits)."""
        tabs = {}
        for chrom, tree in self.cns_tree.items():
            if not isinstance(tree, SegmentIndex):
                return None
            b = np.asarray(tree._b, dtype=np.int64)
            e = np.asarray([h.end for h in tree._iv], dtype=np.int64)
            if len(b) > 1 and bool((e[:-1] > b[1:]).any()):
                return None
            tabs[chrom] = (b, e, np.asarray([h.data for h in tree._iv], dtype=np.int64))
        return tabs

    def _hash_alignment_to_seg_scalar(self):
        for read, ca in self.chimeric_alignments.items():
            for rint in ca[1]:
                if rint[0] not in self.cns_tree:
                    rint.append(set([-1]))
                    continue
                lo = self.pos2cni(rint[0], min(rint[1], rint[2]))
                hi = self.pos2cni(rint[0], max(rint[1], rint[2]))
                assert len(lo) <= 1 and len(hi) <= 1
                cniset = set([list(lo)[0].data if len(lo) == 1 else -1, list(hi)[0].data if len(hi) == 1 else -1])
                if len(cniset) > 1 and -1 in cniset:
                    cniset.remove(-1)
                rint.append(cniset)
                by_seg = self.chimeric_alignments_seg.setdefault(rint[0], {})
                for cni in cniset:
                    if cni != -1:
                        by_seg.setdefault(cni, []).append(read)

    def _cn_segment_ends(self, tabs):
        """(lo, hi) lists over the alignments of chimeric_alignments in dict order: CN-segment index holding the lower /
        higher end, -1 = none, -2 = chromosome without CN segments.  numpy form (the test oracle's engine); the product
        overrides it with the device lookup (csrc/segments.cu)."""
        flat = [rint for ca in self.chimeric_alignments.values() for rint in ca[1]]
        n = len(flat)
        chrom = [r[0] for r in flat]
        p1 = np.fromiter((r[1] for r in flat), dtype=np.int64, count=n)
        p2 = np.fromiter((r[2] for r in flat), dtype=np.int64, count=n)
        plo, phi = np.minimum(p1, p2), np.maximum(p1, p2)
        lo = np.full(n, -2, dtype=np.int64)          # -2: chromosome without CN segments
        hi = np.full(n, -2, dtype=np.int64)
        names, inv = np.unique(np.asarray(chrom, dtype=object), return_inverse=True) if n else ([], None)
        for ci, c in enumerate(names):
            if c not in tabs:
                continue
            b, e, d = tabs[c]
            sel = np.nonzero(inv == ci)[0]
            for pos, out in ((plo[sel], lo), (phi[sel], hi)):
                k = np.searchsorted(b, pos, side="right") - 1
                kk = np.maximum(k, 0)
                out[sel] = np.where((k >= 0) & (e[kk] > pos), d[kk], -1)
        return lo.tolist(), hi.tolist()

    def hash_alignment_to_seg(self):
        """CN-segment indices of both ends of every alignment.  Lookups are batched per chromosome (binary search
        over the non-overlapping segments); the sets and chimeric_alignments_seg are then filled in read order,
        exactly as the per-alignment loop of the reference does."""
        self._read_seg_pairs = {}
        tabs = self._cni_tables()
        if tabs is None:
            return self._hash_alignment_to_seg_scalar()
        lo, hi = self._cn_segment_ends(tabs)
        seg = self.chimeric_alignments_seg
        k = 0
        for read, ca in self.chimeric_alignments.items():
            for rint in ca[1]:
                a, z = lo[k], hi[k]
                k += 1
                if a == -2:
                    rint.append(set([-1]))
                    continue
                cniset = set([a, z])
                if len(cniset) > 1 and -1 in cniset:
                    cniset.remove(-1)
                rint.append(cniset)
                by_seg = seg.setdefault(rint[0], {})
                for cni in cniset:
                    if cni != -1:
                        by_seg.setdefault(cni, []).append(read)

    # ---- ibg:326-340 ------------------------------------------------------------------------------------
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

    # ---- ibg:864-1005 -----------------------------------------------------------------------------------
    def build_graph(self):
        """Split amplicon intervals at breakpoint ends lying strictly inside them; one graph per component."""
        splits = {}

        def add_split(ai, pos, orient, who, which):
            cut = (pos, pos + 1, who, which, '+') if orient == '+' else (pos - 1, pos, who, which, '-')
            splits.setdefault(ai, []).append(cut)

        for bpi, bp in enumerate(self.new_bp_list):
            for ai, seg in enumerate(self.amplicon_intervals):
                if bp[0] == seg[0] and seg[1] < bp[1] < seg[2] and bp[2] in '+-':
                    add_split(ai, bp[1], bp[2], bpi, 1)
                if bp[3] == seg[0] and seg[1] < bp[4] < seg[2] and bp[5] in '+-':
                    add_split(ai, bp[4], bp[5], bpi, 4)
        for srci, srce in enumerate(self.source_edges):
            for ai, seg in enumerate(self.amplicon_intervals):
                if srce[3] == seg[0] and seg[1] < srce[4] < seg[2] and srce[5] in '+-':
                    add_split(ai, srce[4], srce[5], len(self.new_bp_list) + srci, 4)

        # the container: `graph_cls` of the instance / class when set, else the reference's class when importable, else
        # the local one with the same row layouts (coral_b200/breakpoint_graph.py)
        graph_cls = getattr(self, "graph_cls", None) or reference_graph_class() or _LocalBreakpointGraph
        next_id = 1
        for seg in self.amplicon_intervals:
            if seg[3] not in self.ccid2id:
                self.ccid2id[seg[3]] = next_id
                next_id += 1
        for _ in range(len(self.ccid2id)):
            self.lr_graph.append(graph_cls())

        for ai in splits.keys():
            cuts = splits[ai]
            cuts.sort(key=lambda c: c[0])
            seg = self.amplicon_intervals[ai]
            g = self.lr_graph[self.ccid2id[seg[3]] - 1]
            chrom = seg[0]
            left = seg[1]
            prev_first = None
            for cut in cuts:
                if prev_first is not None and not cut[0] > prev_first:
                    prev_first = cut[0]   # duplicate split position: skipped, but it still becomes "previous" (ibg:956)
                    continue_left = True
                else:
                    continue_left = False
                if not continue_left:
                    g.add_node((chrom, left, '-'))
                    g.add_node((chrom, cut[0], '+'))
                    g.add_node((chrom, cut[1], '-'))
                    g.add_sequence_edge(chrom, left, cut[0])
                    g.add_concordant_edge(chrom, cut[0], '+', chrom, cut[1], '-')
                    prev_first = cut[0]
                left = cut[1]
            g.add_node((chrom, cuts[-1][1], '-'))
            g.add_node((chrom, seg[2], '+'))
            g.add_sequence_edge(chrom, cuts[-1][1], seg[2])
        for ai, seg in enumerate(self.amplicon_intervals):
            if ai not in splits:
                g = self.lr_graph[self.ccid2id[seg[3]] - 1]
                g.add_node((seg[0], seg[1], '-'))
                g.add_node((seg[0], seg[2], '+'))
                g.add_sequence_edge(seg[0], seg[1], seg[2])
        for g in self.lr_graph:
            g.sort_edges()

        for seg in self.amplicon_intervals:
            g = self.lr_graph[self.ccid2id[seg[3]] - 1]
            g.amplicon_intervals.append([seg[0], seg[1], seg[2]])
            g.add_endnode((seg[0], seg[1], '-'))
            g.add_endnode((seg[0], seg[2], '+'))

        for bpi, bp in enumerate(self.new_bp_list):
            io1 = self._ops.interval_overlap_l([bp[0], bp[1], bp[1]], self.amplicon_intervals)
            io2 = self._ops.interval_overlap_l([bp[3], bp[4], bp[4]], self.amplicon_intervals)
            assert self.amplicon_intervals[io1][3] == self.amplicon_intervals[io2][3]
            ccid = self.amplicon_intervals[io1][3]
            if ccid != self.new_bp_ccids[bpi]:
                self.new_bp_ccids[bpi] = ccid
            self.lr_graph[self.ccid2id[ccid] - 1].add_discordant_edge(bp[0], bp[1], bp[2], bp[3], bp[4], bp[5],
                                                                     lr_count=len(bp[-1]), reads=bp[-1])
        for srci, srce in enumerate(self.source_edges):
            self.lr_graph[self.ccid2id[self.source_edge_ccids[srci]] - 1].add_source_edge(srce[3], srce[4], srce[5])

    # ---- ibg:213-323 ------------------------------------------------------------------------------------
    def _no_nm_filter(self):
        """nm_filter = True routes the reference through the edit-distance filtered candidate forms (ibg:427, :682, :728).
        The reference never sets it (ibg:40; ibg:1351 is commented out) and this class does not build that route: say so
        instead of silently returning the unfiltered result."""
        if self.nm_filter:
            raise NotImplementedError("nm_filter = True is not supported (dead in the reference, ibg:40 / :1351); the "
                                      "alignment2bp_nm / alignment2bp_nm_l forms exist in breakpoint_utilities")

    def find_amplicon_intervals(self):
        self._no_nm_filter()
        ops = self._ops
        d = self.interval_delta
        for iv in self.amplicon_intervals:  # seeds -> CN-segment bounds, padded when a segment exists there
            c = iv[0]
            lcni = list(self.pos2cni(c, iv[1]))[0].data
            rcni = list(self.pos2cni(c, iv[2]))[0].data
            left = self.cns_intervals_by_chr[c][lcni][1]
            iv[1] = left - d if len(list(self.pos2cni(c, left - d))) > 0 else left
            right = self.cns_intervals_by_chr[c][rcni][2]
            iv[2] = right + d if len(list(self.pos2cni(c, right + d))) > 0 else right

        ccid = 0
        for ai in range(len(self.amplicon_intervals)):   # range fixed before the BFS appends new intervals
            if self.amplicon_intervals[ai][3] == -1:
                self.find_interval_i(ai, ccid)
                ccid += 1

        ivs = self.amplicon_intervals
        conns = self.amplicon_interval_connections
        sidx = sorted(range(len(ivs)), key=lambda i: (global_names.chr_idx[ivs[i][0]], ivs[i][1]))
        siv = [ivs[i] for i in sidx]
        runs, start = [], 0
        for k in range(len(siv) - 1):
            if not (ops.interval_adjacent(siv[k + 1], siv[k]) or ops.interval_overlap(siv[k], siv[k + 1])):
                if k > start:
                    runs.append([start, k])
                start = k + 1
        if len(siv) > 0 and start < len(siv) - 1:
            runs.append([start, len(siv) - 1])
        for a, b in runs[::-1]:
            siv[a][2] = siv[b][2]                      # quirk kept: end of the LAST interval of the run
            for k in range(a + 1, b + 1):
                if siv[k][3] != siv[a][3]:
                    old = siv[k][3]
                    for x in siv:
                        if x[3] == old:
                            x[3] = siv[a][3]
            cmap = {c: c for c in conns.keys()}
            for k in range(a + 1, b + 1):
                keep, gone = sidx[a], sidx[k]
                for c in cmap.keys():
                    if gone == cmap[c][0]:
                        cmap[c] = (keep, cmap[c][1])
                    if gone == cmap[c][1]:
                        cmap[c] = (cmap[c][0], keep)
                    if cmap[c][1] < cmap[c][0]:
                        cmap[c] = (cmap[c][1], cmap[c][0])
            for c in cmap.keys():
                tgt = cmap[c]
                if c != tgt:
                    if tgt not in conns:
                        conns[tgt] = conns[c]
                    else:
                        conns[tgt] |= conns[c]
                    del conns[c]
                    if tgt[0] == tgt[1]:
                        del conns[tgt]
            for k in range(b, a, -1):
                del siv[k]
                del sidx[k]

        self.amplicon_intervals = list(siv)
        where = {sidx[i]: i for i in range(len(sidx))}
        remap = {c: (min(where[c[0]], where[c[1]]), max(where[c[0]], where[c[1]])) for c in conns.keys()}
        self.amplicon_interval_connections = {remap[c]: conns[c] for c in conns.keys()}
        seen = np.zeros(len(self.amplicon_intervals))
        for ai in range(len(self.amplicon_intervals)):
            root = self.amplicon_intervals[ai][3]
            if seen[ai] == 0:
                queue = [ai]
                while len(queue) > 0:
                    cur = queue.pop(0)
                    seen[cur] = 1
                    if self.amplicon_intervals[cur][3] != root:
                        self.amplicon_intervals[cur][3] = root
                    for (a1, a2) in self.amplicon_interval_connections.keys():
                        if a1 == cur and seen[a2] == 0:
                            queue.append(a2)
                        elif a2 == cur and seen[a1] == 0:
                            queue.append(a1)

    # ---- ibg:343-673 ------------------------------------------------------------------------------------
    def _new_bounds(self, chrom_l, chrom_r, seg_l, seg_r, pos_l, pos_r, lir, tail, quirk_l=False, first_cn_guard=False):
        """Padding rules shared by the four interval-refinement sites (ibg:505-523, 534-553, 574-593, 600-619).
        seg_l/seg_r: CN-segment indices of the run's first/last breakpoint; pos_l/pos_r: their coordinates;
        lir: right end used when the right flank is amplified (differs between the in-loop and tail sites)."""
        by = self.cns_intervals_by_chr
        d, half = self.interval_delta, int(self.max_seq_len / 2)
        amp_l = by[chrom_l][seg_l][3] >= self.cn_gain
        amp_r = by[chrom_r][seg_r][3] >= self.cn_gain
        if not amp_l:
            l = max(pos_l - d, by[chrom_l][0][1])
        else:
            l = max(by[chrom_l][seg_l][1] - d, by[chrom_l][0][1])
        if not amp_r:
            r = min(pos_r + d, by[chrom_r][-1][2])
        else:
            r = min(lir + d, by[chrom_r][-1][2])
        if first_cn_guard:
            if by[chrom_l][seg_l][3] and pos_l - half > l:      # ibg:516 (CN truthiness test kept)
                l = pos_l - half
        elif pos_l - half > l:
            l = (pos_l - half > l) if quirk_l else pos_l - half  # ibg:547 assigns the BOOLEAN (kept on purpose)
        if pos_r + half < r:
            r = pos_r + half
        return l, r

    def _pair_candidates(self, reads, intrvl1, intrvl2):
        """alignment2bp over `reads` in order (ibg:432-434); the product overrides this with one batched launch."""
        out = []
        for r in reads:
            out += self._ops.alignment2bp(r, self.chimeric_alignments[r], self.min_bp_match_cutoff_, 20, intrvl1, intrvl2)
        return out

    def _called_breakpoints(self, cands):
        """(bp, bpr, bp_stats) of every sub-cluster peeled off every cluster of `cands`, in the reference's order
        (ibg:436-446: cluster_bp_list, then bpc2bp on the remainder while it keeps min_cluster_cutoff members).
        The CUDA engine overrides this with one device call per candidate list."""
        ops = self._ops
        for c in ops.cluster_bp_list(cands, self.min_cluster_cutoff, self.max_breakpoint_distance_cutoff):
            if len(c) >= self.min_cluster_cutoff:
                rest = c
                while len(rest) >= self.min_cluster_cutoff:
                    bp, bpr, bp_stats_, rest = ops.bpc2bp(rest, self.min_bp_match_cutoff_)
                    yield bp, bpr, bp_stats_

    def find_interval_i(self, ai, ccid):
        ops = self._ops
        by = self.cns_intervals_by_chr
        queue = [ai]
        while len(queue) > 0:
            ai_ = queue.pop(0)
            cur = self.amplicon_intervals[ai_]
            chrom, s, e = cur[0], cur[1], cur[2]
            if cur[3] == -1:
                cur[3] = ccid
            try:
                si = list(self.pos2cni(chrom, s))[0].data
                ei = list(self.pos2cni(chrom, e))[0].data
            except:  # noqa: E722  (ibg:366: any failure skips this interval)
                continue

            # reads hashed to the CN segments of this interval that also map outside it (ibg:369-384)
            # (same insertion order of chromosomes into d1 as the reference's nested loops — the loop over d1.keys()
            # below depends on it; the (chromosome, segment) pairs of a read, without the -1 entries that never insert
            # anything, are spelled out once per read and reused by every interval that meets the read)
            d1 = {}
            seg_here = self.chimeric_alignments_seg[chrom]
            pairs_of = self._read_seg_pairs
            ca_all = self.chimeric_alignments
            for i in range(si, ei + 1):
                if i in seg_here:
                    for r in seg_here[i]:
                        pr = pairs_of.get(r)
                        if pr is None:
                            pr = pairs_of[r] = [(int_[0], i_) for int_ in ca_all[r][1] for i_ in int_[-1] if i_ != -1]
                        for c_r, i_ in pr:
                            if c_r != chrom or i_ <= si or i_ >= ei:
                                dc = d1.get(c_r)
                                if dc is None:
                                    dc = d1[c_r] = {}
                                st = dc.get(i_)
                                if st is None:
                                    dc[i_] = {r}
                                else:
                                    st.add(r)
            for c_ in list(d1.keys()):   # ibg:386-398
                for segi in [k for k in d1[c_].keys() if len(d1[c_][k]) < self.min_cluster_cutoff]:
                    del d1[c_][segi]
                if len(d1[c_]) == 0:
                    del d1[c_]

            refined, refined_conn = [], []
            for c_ in d1.keys():
                bins = sorted(d1[c_].keys())
                runs, nir, lasti = [], set(), 0
                for i in range(len(bins) - 1):          # ibg:408-419
                    nil = by[c_][bins[i + 1]][1]
                    lir = by[c_][bins[i]][2]
                    nir |= d1[c_][bins[i]]
                    if bins[i + 1] - bins[i] > 2 or nil - lir > self.max_seq_len:
                        runs.append([c_, bins[lasti], bins[i], nir])
                        lasti = i + 1
                        nir = set()
                nir |= d1[c_][bins[-1]]
                runs.append([c_, bins[lasti], bins[-1], nir])

                for nint_ in runs:
                    ns = by[nint_[0]][nint_[1]][1]
                    ne = by[nint_[0]][nint_[2]][2]
                    # NOTE: iteration order of a set of read names (ibg:432) — process-dependent in the reference too
                    cands = self._pair_candidates(list(nint_[-1]), [nint_[0], ns, ne], self.amplicon_intervals[ai_])
                    kept_bp = []
                    num_subcluster = 0     # never incremented in the reference (ibg:442-457)
                    for bp, bpr, bp_stats_ in self._called_breakpoints(cands):
                        if (num_subcluster == 0 and len(set(bpr)) >= self.min_cluster_cutoff) or \
                                (len(set(bpr)) >= max(self.normal_cov * self.min_bp_cov_factor, 3.0)):
                            bpi = self.addbp(bp, set(bpr), bp_stats_, ccid)
                            if bpi not in kept_bp:
                                kept_bp.append(bpi)

                    segs, segs_ = [], []
                    if len(kept_bp) > 0:
                        target = [nint_[0], ns, ne]
                        for bpi in kept_bp:
                            bp = self.new_bp_list[bpi][:6]
                            e1, e2 = [bp[0], bp[1], bp[1]], [bp[3], bp[4], bp[4]]
                            try:
                                if ops.interval_overlap(e1, self.amplicon_intervals[ai_]) and ops.interval_overlap(e2, target):
                                    segs.append([list(self.pos2cni(bp[3], bp[4]))[0].data, bp[4], bpi])
                                elif ops.interval_overlap(e2, self.amplicon_intervals[ai_]) and ops.interval_overlap(e1, target):
                                    segs.append([list(self.pos2cni(bp[0], bp[1]))[0].data, bp[1], bpi])
                                else:
                                    logging.warning(_t() + "\t\tExact breakpoint outside amplicon interval.")
                                    logging.warning(_t() + "\t\tBreakpoint %s." % bp)
                                    logging.warning(_t() + "\t\tCurrent interval %s." % self.amplicon_intervals[ai_])
                                    logging.warning(_t() + "\t\tNew interval %s." % target)
                                    o1, o2 = ops.interval_overlap(e1, target), ops.interval_overlap(e2, target)
                                    k1 = list(self.pos2cni(bp[0], bp[1]))[0].data
                                    if o1 and o2:
                                        segs.append([k1, bp[1], bpi])
                                        segs.append([list(self.pos2cni(bp[3], bp[4]))[0].data, bp[4], bpi])
                                    elif o1:
                                        segs.append([k1, bp[1], bpi])
                                        segs_.append([bp[3], list(self.pos2cni(bp[3], bp[4]))[0].data, bp[4], bpi])
                                    elif o2:
                                        segs_.append([bp[0], k1, bp[1], bpi])
                                        segs.append([list(self.pos2cni(bp[3], bp[4]))[0].data, bp[4], bpi])
                                    else:
                                        segs_.append([bp[0], k1, bp[1], bpi])
                                        segs_.append([bp[3], list(self.pos2cni(bp[3], bp[4]))[0].data, bp[4], bpi])
                            except:  # noqa: E722  (ibg:490)
                                pass
                        segs = sorted(segs, key=lambda x: (x[0], x[1]))
                        segs_ = sorted(segs_, key=lambda x: (global_names.chr_idx[x[0]], x[1], x[2]))
                        half = self.max_seq_len / 2
                        # breakpoints landing in the target interval (ibg:494-561)
                        lasti = 0
                        for i in range(len(segs) - 1):
                            nil = by[c_][segs[i + 1][0]][1]
                            ncn = by[c_][segs[i + 1][0]][3]
                            lir = by[c_][segs[i][0]][2]
                            lcn = by[c_][segs[i][0]][3]
                            amp = (ncn >= self.cn_gain) or (lcn >= self.cn_gain)
                            if (segs[i + 1][0] - segs[i][0] > 2) or (nil - lir > half) or \
                                    (segs[i + 1][1] - segs[i][1] > self.max_seq_len) or \
                                    (not amp and nil - lir > 2 * self.interval_delta) or \
                                    (not amp and segs[i + 1][1] - segs[i][1] > 3 * self.interval_delta):
                                l, r = self._new_bounds(c_, c_, segs[lasti][0], segs[i][0], segs[lasti][1], segs[i][1], lir,
                                                        False, first_cn_guard=True)
                                if len(list(self.pos2cni(c_, l))) == 0:
                                    l = by[c_][segs[lasti][0]][1]
                                if len(list(self.pos2cni(c_, r))) == 0:
                                    r = lir
                                refined.append([c_, l, r, -1])
                                refined_conn.append([segs[k][2] for k in range(lasti, i + 1)])
                                lasti = i + 1
                        if len(segs) > 0:
                            l, r = self._new_bounds(c_, c_, segs[lasti][0], segs[-1][0], segs[lasti][1], segs[-1][1],
                                                    by[c_][segs[-1][0]][2], True, quirk_l=True)
                            if len(list(self.pos2cni(c_, l))) == 0:
                                l = by[c_][segs[lasti][0]][1]
                            if len(list(self.pos2cni(c_, r))) == 0:
                                r = by[c_][segs[-1][0]][2]
                            refined.append([c_, l, r, -1])
                            refined_conn.append([segs[k][2] for k in range(lasti, len(segs))])
                        # breakpoint ends outside both intervals (ibg:562-623); their connections are dropped
                        lasti = 0
                        for i in range(len(segs_) - 1):
                            a, b = segs_[i], segs_[i + 1]
                            nil = by[b[0]][b[1]][1]
                            ncn = by[b[0]][b[1]][3]
                            lir = by[a[0]][a[1]][2]
                            lcn = by[a[0]][a[1]][3]
                            amp = (ncn >= self.cn_gain) or (lcn >= self.cn_gain)
                            if (b[0] != a[0]) or (b[1] - a[1] > 2) or (nil - lir > half) or (b[2] - a[2] > self.max_seq_len) or \
                                    (not amp and nil - lir > 2 * self.interval_delta) or \
                                    (not amp and b[2] - a[2] > 3 * self.interval_delta):
                                f = segs_[lasti]
                                l, r = self._new_bounds(f[0], a[0], f[1], a[1], f[2], a[2], lir, False)
                                if len(list(self.pos2cni(f[0], l))) == 0:
                                    l = by[f[0]][f[1]][1]
                                if len(list(self.pos2cni(a[0], r))) == 0:
                                    r = lir
                                refined.append([f[0], l, r, -1])
                                refined_conn.append([])
                                lasti = i + 1
                        if len(segs_) > 0:
                            f, z = segs_[lasti], segs_[-1]
                            l, r = self._new_bounds(f[0], z[0], f[1], z[1], f[2], z[2], by[z[0]][z[1]][2], True)
                            if len(list(self.pos2cni(f[0], l))) == 0:
                                l = by[f[0]][f[1]][1]
                            if len(list(self.pos2cni(f[0], r))) == 0:
                                r = by[f[0]][z[1]][2]
                            refined.append([f[0], l, r, -1])
                            refined_conn.append([])

            # enqueue (ibg:627-673)
            for ni in range(len(refined)):
                hit, pieces = ops.interval_exclusive(refined[ni], self.amplicon_intervals)
                if len(pieces) == 0:
                    for bpi in refined_conn[ni]:
                        bp = self.new_bp_list[bpi][:6]
                        for ei_ in hit:
                            conn = (min(ai_, ei_), max(ai_, ei_))
                            if ei_ != ai_ and ops.interval_overlap([bp[0], bp[1], bp[1]], self.amplicon_intervals[ei_]) or \
                                    ops.interval_overlap([bp[3], bp[4], bp[4]], self.amplicon_intervals[ei_]):
                                self.amplicon_interval_connections.setdefault(conn, set()).add(bpi)
                    for ei_ in hit:
                        if ei_ != ai_ and self.amplicon_intervals[ei_][3] < 0:
                            queue.append(ei_)
                else:
                    for piece in pieces:
                        nai = len(self.amplicon_intervals)
                        self.amplicon_intervals.append(piece)
                        self.amplicon_interval_connections[(ai_, nai)] = set([])
                        if len(hit) == 0:
                            for bpi in refined_conn[ni]:
                                self.amplicon_interval_connections[(ai_, nai)].add(bpi)
                        else:
                            for bpi in refined_conn[ni]:
                                bp = self.new_bp_list[bpi][:6]
                                for ei_ in hit:
                                    conn = (min(ai_, ei_), max(ai_, ei_))
                                    if ops.interval_overlap([bp[0], bp[1], bp[1]], self.amplicon_intervals[ei_]) or \
                                            ops.interval_overlap([bp[3], bp[4], bp[4]], self.amplicon_intervals[ei_]):
                                        self.amplicon_interval_connections.setdefault(conn, set()).add(bpi)
                                    else:
                                        self.amplicon_interval_connections[(ai_, nai)].add(bpi)
                        queue.append(nai)
