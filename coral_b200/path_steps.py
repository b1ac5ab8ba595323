"""``compute_path_constraints`` of reference ``bam_to_breakpoint_nanopore`` (``infer_breakpoint_graph.py:1059-1323``):
the step after the hot path.  Reads that support breakpoint edges, and reads that merely run
through several sequence edges, become subpath constraints of their amplicon's graph.

Two halves:

* reads on discordant edges (``:1067-1293``): a few thousand reads, each a handful of list entries — host logic that
  walks one graph per read with ``path_constraints.py``;
* concordant reads (``:1295-1321``): a second pass over EVERY record of the amplicon intervals (the reference's
  ``fetch`` + ``alignment_to_path`` per record).  That pass is the data-parallel one: ``_concordant_read_runs`` is the
  hook — ``csrc/paths.cu`` on the GPU engine, the pysam-shim loop in the CPU oracle — and returns the distinct runs
  of sequence edges with their record counts, in first-seen order.

The engine mixes this class in through ``HostSteps``; it relies on the containers of the earlier stages
(``lr_graph``, ``chimeric_alignments``, ``large_indel_alignments``, ``amplicon_intervals``, ``ccid2id``).
"""
import logging

from . import path_constraints as pc


class PathConstraintSteps(object):

    # ---- bookkeeping of ibg:1280-1291 / 1311-1321 -------------------------------------------------------------
    def _add_subpath(self, amplicon_idx, path, amp_label, count=1):
        """A path (or its reverse) already on the list gains `count` supporting reads, else it is appended."""
        if not (len(path) > 5 and pc.valid_path(self.lr_graph[amplicon_idx], path)):
            return
        paths, counts, labels = self.path_constraints[amplicon_idx]
        where = self._subpath_index.setdefault(amplicon_idx, {})       # tuple(path) -> position, in step with `paths`
        for cand in (path, path[::-1]):
            k = where.get(tuple(cand))
            if k is not None:
                counts[k] += count
                return
        where[tuple(path)] = len(paths)
        paths.append(path)
        counts.append(count)
        labels.append(amp_label)

    @staticmethod
    def _blocks(entries):
        """entries sorted by min(i, j): maximal runs in which each breakpoint starts on the alignment the previous one
        ended on (ibg:1101-1109)."""
        blocks = [[0]]
        for k in range(1, len(entries)):
            if min(entries[k][0], entries[k][1]) == max(entries[k - 1][0], entries[k - 1][1]):
                blocks[-1].append(k)
            else:
                blocks.append([k])
        return blocks

    def _overlapping_local_alignments(self, rn):
        q = self.chimeric_alignments[rn][0]     # ibg:1113-1117
        return any(q[k + 1][0] - q[k][1] < -self.min_bp_match_cutoff_ for k in range(len(q) - 1))

    def _block_paths(self, g, rints, entries, blocks):
        for block in blocks:
            bps = [entries[k][2] for k in block]
            if len(set(bps)) < len(bps):        # the same breakpoint twice in one block (ibg:1129)
                continue
            yield pc.chimeric_alignment_to_path(g, rints, [entries[k][:2] for k in block], bps)

    # ---- the five read classes of ibg:1086-1279 ------------------------------------------------------------------
    def _paths_of_bp_read(self, g, rn, split, sdel):
        """split: [[i, j, discordant edge]] from chimeric read tuples (i != j); sdel: the same from in-CIGAR
        deletions (i == j = index into large_indel_alignments[rn])."""
        if len(split) == 1 and not sdel:
            rints = [a[:4] for a in self.chimeric_alignments[rn][1]]
            return [pc.chimeric_alignment_to_path_i(g, rints, split[0][0], split[0][1], split[0][2])]
        if len(split) > 1 and not sdel:
            split = sorted(split, key=lambda it: min(it[0], it[1]))
            if self._overlapping_local_alignments(rn):
                return []
            rints = [a[:4] for a in self.chimeric_alignments[rn][1]]
            return list(self._block_paths(g, rints, split, self._blocks(split)))
        if not split and len(sdel) == 1:
            return self._paths_single_deletion(g, rn, sdel)
        if not split:
            return self._paths_many_deletions(g, rn, sdel)
        return self._paths_split_and_deletions(g, rn, split, sdel)

    def _paths_single_deletion(self, g, rn, sdel):
        row = self.large_indel_alignments[rn][0]     # ibg:1139: always the read's FIRST row
        c, nxt, prv, first, last = row[0], row[1], row[2], row[3], row[4]
        if first < last:
            if not prv < nxt:
                return []
            two = [[c, first, prv, '+'], [c, nxt, last, '+']]
            return [pc.chimeric_alignment_to_path_i(g, two, 1, 0, sdel[0][2])]
        if not prv > nxt:
            return []
        two = [[c, first, prv, '-'], [c, nxt, last, '-']]
        return [pc.chimeric_alignment_to_path_i(g, two, 0, 1, sdel[0][2])]

    def _paths_many_deletions(self, g, rn, sdel):
        rows = self.large_indel_alignments[rn]        # ibg:1172-1206
        spans = set((r[0], min(r[3], r[4]), max(r[3], r[4])) for r in rows)
        if len(spans) > 1 or len(rows) <= 1:
            return []
        pieces = [[r[0], min(r[3], r[4]), max(r[3], r[4]), '+'] for r in rows]
        for k, r in enumerate(sorted(rows, key=lambda it: min(it[1], it[2]))):
            pieces.append([r[0], min(r[3], r[4]), max(r[3], r[4]), '+'])
            pieces[k][2] = min(r[1], r[2])
            pieces[k + 1][1] = max(r[1], r[2])
        sdel = sorted(sdel, key=lambda it: it[0])
        blocks = [[]]
        for k, it in enumerate(sdel):
            if k == 0 or it[0] == sdel[k - 1][0] + 1:
                blocks[-1].append(k)
            else:
                blocks.append([k])
        out = []
        for block in blocks:
            bps = [sdel[k][2] for k in block]
            if len(set(bps)) < len(bps):
                continue
            out.append(pc.chimeric_alignment_to_path(g, pieces, [[sdel[k][0], sdel[k][0] + 1] for k in block], bps))
        return out

    def _paths_split_and_deletions(self, g, rn, split, sdel):
        rints = [a[:4] for a in self.chimeric_alignments[rn][1]]      # ibg:1207-1279
        rows = self.large_indel_alignments[rn]
        inside = []
        for r in rows:       # each deletion must fall strictly inside one local alignment
            lo, hi = min(r[1], r[2]), max(r[1], r[2])
            for ri, a in enumerate(rints):
                if r[0] == a[0] and lo > min(a[1], a[2]) and hi < max(a[1], a[2]):
                    inside.append(ri)
                    break
            else:
                return []
        split = [list(it) for it in split]
        for rsi, ri in enumerate(inside):      # cut alignment ri at deletion rsi; later indices shift by one
            lo, hi = min(rows[rsi][1], rows[rsi][2]), max(rows[rsi][1], rows[rsi][2])
            rints.insert(ri, rints[ri][:])
            fwd = rints[ri][3] == '+'
            rints[ri][2], rints[ri + 1][1] = (lo, hi) if fwd else (hi, lo)
            for it in split:
                if it[0] >= ri and it[1] >= ri:
                    it[0] += 1
                    it[1] += 1
            for it in sdel:
                if it[0] == rsi:
                    split.append([ri + 1, ri, it[2]] if fwd else [ri, ri + 1, it[2]])
        split.sort(key=lambda it: min(it[0], it[1]))
        blocks = self._blocks(split)
        if self._overlapping_local_alignments(rn):
            return []
        return list(self._block_paths(g, rints, split, blocks))

    # ---- ibg:1059-1323 ---------------------------------------------------------------------------------------------
    def compute_path_constraints(self):
        self._subpath_index = {}
        for amplicon_idx, g in enumerate(self.lr_graph):
            self.path_constraints[amplicon_idx] = [[], [], []]
            self.longest_path_constraints[amplicon_idx] = [[], [], []]
            bp_reads = {}
            for di, de in enumerate(g.discordant_edges):
                for r_ in de[10]:
                    slot = bp_reads.setdefault(r_[0], [[], []])
                    slot[1 if r_[1] == r_[2] else 0].append([r_[1], r_[2], di])
            logging.debug("There are %d reads in total covering at least one breakpoint in amplicon %d."
                          % (len(bp_reads), amplicon_idx + 1))
            for rn, (split, sdel) in bp_reads.items():
                for path in self._paths_of_bp_read(g, rn, split, sdel):
                    self._add_subpath(amplicon_idx, path, amplicon_idx)
            logging.debug("There are %d distinct subpaths due to reads involving breakpoints in amplicon %d."
                          % (len(self.path_constraints[amplicon_idx][0]), amplicon_idx + 1))
            # reads that only run through concordant edges (ibg:1295-1321)
            concordant_reads = set()
            for ce in g.concordant_edges:
                for rn in ce[9]:
                    if rn not in self.large_indel_alignments and rn not in self.chimeric_alignments:
                        concordant_reads.add(rn)
            logging.debug("There are %d concordant reads within amplicon intervals in amplicon %d."
                          % (len(concordant_reads), amplicon_idx + 1))
            intervals = [a for a in self.amplicon_intervals if self.ccid2id[a[3]] - 1 == amplicon_idx]
            for first, last, count in self._concordant_read_runs(amplicon_idx, intervals, concordant_reads):
                self._add_subpath(amplicon_idx, pc.sequence_run_to_path(g, first, last), amplicon_idx, count)
            logging.debug("There are %d distinct subpaths in total in amplicon %d."
                          % (len(self.path_constraints[amplicon_idx][0]), amplicon_idx + 1))

    # _concordant_read_runs(amplicon_idx, intervals, concordant_reads) is provided by the engine (no default here: the
    # product's is the GPU pass of infer_breakpoint_graph.py, the test oracle's is the pysam-shim loop):
    # [(first sequence edge, last sequence edge, number of records)] in first-seen order over the records of
    # `intervals` (list order, then file order) with MAPQ >= 20 whose name is in `concordant_reads`
    # (ibg:1303-1310 + path_constraints.py:60-84).
