"""Drop-in for the hot methods of reference ``infer_breakpoint_graph.bam_to_breakpoint_nanopore``
(``:20-1330``): same class name, attributes, method names and call order
(``read_cns -> fetch -> hash_alignment_to_seg -> find_amplicon_intervals -> find_smalldel_breakpoints ->
find_breakpoints -> build_graph -> assign_cov``, reference ``reconstruct_graph`` ``:1354-1386``).

The hot methods run on the GPU through the C ABI (``_abi.Context``); this file only converts result tables
into the reference's Python containers.  Without a CUDA device the constructor raises — there is no CPU path.
"""
import functools
import logging
import math
import time

import numpy as np

from . import _abi, breakpoint_utilities, global_names, gpu_ops
from .breakpoint_graph import output_breakpoint_graph_lr, output_breakpoint_info_lr
from .host_steps import HostSteps, _t, gc_paused


def _gc_paused(fn):
    @functools.wraps(fn)
    def run(*a, **k):
        with gc_paused():
            return fn(*a, **k)
    return run


class bam_to_breakpoint_nanopore(HostSteps):
    _ops = breakpoint_utilities

    _auto_mapped = "mapped_async"   # what placement="auto" means for a buffer with slack (the sharded engine joins at load)
    _normal_pending = None      # (class default: _init_state assigns the properties before __init__ sets it)

    def __init__(self, lr_bamfile, seedfile, device=0, n_threads=0, records=None, placement="auto"):
        """`records` (optional): an object with .references, .rec (uint8 host buffer of framed records), .rec_off and
        .query_name(i) — lets a caller that already holds decompressed records skip the BAM file.
        `placement`: "copy" (records copied to HBM), "mapped" (kernels read the page-locked host buffer in place; the
        scan runs once per BAM, so only the fields it needs cross PCIe), "mapped_async" (as "mapped", and the
        constructor returns after a pre-scan of the fields `fetch` needs: `fetch` and `hash_alignment_to_seg` then run
        while the full scan is still pulling CIGAR / SEQ bytes over the link; `read_cns`'s coverage sum is evaluated
        when `normal_cov` / `min_cluster_cutoff` are first read; 172 vs 200 ms per C2 pass),
        "mapped_lazy" (pre-scan only; the full scan runs when a method needs it), "auto" (mapped_async when the
        buffer has the 64 bytes of slack the mapped modes need, else copy) or "device_ingest" (`lr_bamfile` is
        inflated and framed on the GPU, the records never exist in host memory: csrc/ingest.cu)."""
        self._init_state()
        self._normal_pending = None              # read_cns's coverage call, deferred while the scan is still running
        if placement == "device_ingest":
            # SURVEY 8f rank 3 (csrc/ingest.cu): the file's BGZF bytes go to the GPU as they are on disk and are
            # inflated and framed there; no host copy of the records exists, read names come back through the ctx
            self._ctx = _abi.Context(device)
            gpu_ops.set_default_device(device)
            self.lr_bamfh = _abi.DeviceBam(lr_bamfile, self._ctx)
            self._ctx.set_references(self.lr_bamfh.references)
            self.lr_bamfh.run()
            self._scan_async = False
        else:
            self.lr_bamfh = records if records is not None else _abi.BamFile(lr_bamfile, n_threads)   # replaces pysam (ibg:65)
            self._ctx = _abi.Context(device)
            gpu_ops.set_default_device(device)
            self._ctx.set_references(self.lr_bamfh.references)
            rec, rec_off = self.lr_bamfh.rec, self.lr_bamfh.rec_off
            slack = int(getattr(self.lr_bamfh, "rec_slack", 0))
            if placement == "auto":
                nb = int(rec_off[-1]) if len(rec_off) else 0
                placement = self._auto_mapped if nb > 0 and getattr(rec, "nbytes", 0) + slack >= nb + 64 else "copy"
            self._scan_async = placement in ("mapped_async", "mapped_lazy")
            self._ctx.load_records(rec, rec_off, placement=placement, slack=slack)
        self._tid = {c: i for i, c in enumerate(self.lr_bamfh.references)}
        self._name_rec = {}     # read name -> index of one record carrying it (for concordant-edge rbps)
        self._read_names = []   # chimeric read order of the device table
        self._read_index = {}
        self._nm_stats_required = True   # hsr.fetch (the second caller of this pass) keeps no NM statistics
        self._warn_no_primary = True
        self._aln_ptr = np.zeros(1, np.uint32)
        self._read_kept = np.zeros(0, bool)
        self._ce_recs = None    # (names, a record of each) over all concordant-edge name sets of the last assign_cov
        self._ce_rec_of = None
        if seedfile is not None:
            self._read_seeds(seedfile)

    # ---- a12: ibg:75-136 ---------------------------------------------------------------------------------
    def read_cns(self, cns):
        self._parse_cns(cns)
        segs, total = self._median_cn_segments()
        order = sorted(range(len(segs)), key=lambda k: (self._tid[segs[k][0]], segs[k][1]))
        self._normal_pending = ([segs[k] for k in order], total, self.min_bp_cov_factor)   # the factor as of this call
        if not self._scan_async:
            self._resolve_normal_cov()

    def _resolve_normal_cov(self):
        """ibg:129-135.  With an asynchronous scan this runs when normal_cov / min_cluster_cutoff are first read (the
        coverage sum needs the whole scan; fetch and hash_alignment_to_seg, which the reference calls next, do not)."""
        pending, self._normal_pending = self._normal_pending, None
        if pending is None:
            return
        segs, total, factor = pending
        nnc = self._coverage_sum(segs)
        self._normal_cov = nnc * 1.0 / total
        logging.info(_t() + "LR normal cov = %f." % (self._normal_cov))
        self._min_cluster_cutoff = max(self._min_cluster_cutoff, factor * self._normal_cov)

    @property
    def normal_cov(self):
        if self._normal_pending is not None:
            self._resolve_normal_cov()
        return self._normal_cov

    @normal_cov.setter
    def normal_cov(self, v):
        if self._normal_pending is not None:      # the reference's "last assignment wins": read_cns's own assignment first
            self._resolve_normal_cov()
        self._normal_cov = v

    @property
    def min_cluster_cutoff(self):
        if self._normal_pending is not None:
            self._resolve_normal_cov()
        return self._min_cluster_cutoff

    @min_cluster_cutoff.setter
    def min_cluster_cutoff(self, v):
        if self._normal_pending is not None:
            self._resolve_normal_cov()
        self._min_cluster_cutoff = v

    def _coverage_sum(self, segs):
        """Sum of count_coverage over inclusive segments; overlapping requests are issued in disjoint batches."""
        total = 0
        pending = list(segs)
        while pending:
            batch, rest, last = [], [], None
            for s in pending:
                if last is not None and self._tid[s[0]] == self._tid[last[0]] and s[1] <= last[2]:
                    rest.append(s)
                else:
                    batch.append(s)
                    last = s
            bs, _ = self._region_counts([self._tid[s[0]] for s in batch], [s[1] for s in batch], [s[2] + 1 for s in batch])
            total += int(bs.sum())
            pending = rest
        return total

    # ---- a1/a2: ibg:139-174 ------------------------------------------------------------------------------
    def _run_fetch(self):
        """Device fetch -> (table dict, record indices -> read names).  The sharded engine overrides this with the
        exchange of SURVEY.md §8e (rows all-gathered, fetch on the gathered rows)."""
        return self._ctx.fetch(), self._names_of

    @property
    def _chim_ctx(self):
        """The ctx that holds the chimeric table (the shard's own ctx, or the gathered-rows ctx when sharded)."""
        return self._ctx

    @_gc_paused
    def fetch(self):
        f, names_of = self._run_fetch()
        if self._nm_stats_required:
            if f["nm_zero_div"] or f["nm_n"] == 0:
                raise ZeroDivisionError("division by zero")              # ibg:154 / ibg:159
            mean = f["nm_sum"] / f["nm_n"]
            self.nm_stats = [mean, math.sqrt(f["nm_sumsq"] / f["nm_n"] - mean ** 2), f["nm_n"]]
        refs = self.lr_bamfh.references
        n_reads = f["n_reads"]
        ap = f["aln_ptr"].tolist() if n_reads else [0]
        logging.info(_t() + "Fetched %d chimeric reads." % n_reads)
        # flat Python lists built once (the per-read work below is slicing only)
        d = (f["qe"].astype(np.int64) - f["qs"].astype(np.int64))
        zero = np.nonzero(d == 0)[0]
        nmn_all = (f["nm"].astype(np.float64) / np.where(d == 0, 1, d)).tolist()          # cp:268
        qint_all = [[a_, b_] for a_, b_ in zip(f["qs"].tolist(), f["qe"].tolist())]
        rint_all = [[refs[t_], a_, b_, '-' if s_ else '+'] for t_, a_, b_, s_ in
                    zip(f["tid"].tolist(), f["rstart"].tolist(), f["rend"].tolist(), f["strand"].tolist())]
        mapq_all = f["mapq"].tolist()
        first_rec = f["read_first_rec"].astype(np.int64) if n_reads else np.zeros(0, np.int64)
        names = names_of(first_rec)
        self._aln_ptr = f["aln_ptr"] if n_reads else np.zeros(1, np.uint32)
        self._read_kept = np.zeros(n_reads, bool)     # reads whose alignments entered chimeric_alignments (table order)
        status = f["read_status"].tolist() if n_reads else []
        rlen = f["read_len"].tolist() if n_reads else []
        recs = first_rec.tolist()
        self._read_names = names
        for r in range(n_reads):
            name = names[r]
            self._read_index[name] = r
            self._name_rec[name] = recs[r]
            st = status[r]
            if st == 1:      # ibg:164-169
                if self._warn_no_primary:
                    logging.warning(_t() + "Found chimeric read without primary alignment.")
                    logging.warning(_t() + "\tRead name: %s; Read length: N/A." % name)
                continue
            self.read_length[name] = rlen[r]
            if st == 2:      # cp:248-253
                logging.warning(_t() + "Found chimeric alignment without match or soft clips.")
                self.chimeric_alignments[name] = ([], [], [])
                continue
            if st == 3:
                raise KeyError("CIGAR signature of an SA entry of read %s is not in cigar2pos_ops" % name)
            if st != 0:
                raise ValueError("malformed SA entry on read %s" % name)
            a, b = ap[r], ap[r + 1]
            if len(zero) and np.searchsorted(zero, a) != np.searchsorted(zero, b):
                raise ZeroDivisionError("float division by zero")      # cp:268 with qe == qs
            self.chimeric_alignments[name] = (qint_all[a:b], rint_all[a:b], mapq_all[a:b], nmn_all[a:b])
            self._read_kept[r] = True
        logging.info(_t() + "Computed alignment intervals on all chimeric reads.")

    # ---- shared tail of ibg:691-718 and ibg:775-802 -------------------------------------------------------
    def _call_clusters(self, cand, tuples):
        """cand: dict of candidate arrays (chr ids = BAM tids); tuples[k] = the (name, i, j) of candidate k."""
        n = len(tuples)
        if n == 0:
            return
        refs = self.lr_bamfh.references
        cutoff = self.min_cluster_cutoff
        r = gpu_ops.cluster_arrays(self._ctx, cand, cutoff, self.max_breakpoint_distance_cutoff, cutoff,
                                   self.min_bp_match_cutoff_)
        order, ptr, sub_of = r["order"], r["cluster_ptr"], r["sub_of"]
        accept2 = max(self.normal_cov * self.min_bp_cov_factor, 3.0)
        for c in range(r["n_clusters"]):
            b, e = int(ptr[c]), int(ptr[c + 1])
            if e - b < cutoff:
                continue
            members = order[b:e]
            subs = sub_of[members]
            for k in range(int(r["n_sub"][c])):
                row = b + k
                bpr = [tuples[int(m)] for m in members[subs == k]]
                support = len(set(bpr))
                if (k == 0 and support >= cutoff) or support >= accept2:
                    f0 = int(r["row_first"][row])
                    bp = [refs[cand["chr1"][f0]], int(r["row_pos1"][row]), '+-'[cand["o1"][f0]], refs[cand["chr2"][f0]],
                          int(r["row_pos2"][row]), '+-'[cand["o2"][f0]], tuples[f0], int(cand["rgap"][f0]), int(cand["flag"][f0])]
                    io1 = breakpoint_utilities.interval_overlap_l([bp[0], bp[1], bp[1]], self.amplicon_intervals)
                    io2 = breakpoint_utilities.interval_overlap_l([bp[3], bp[4], bp[4]], self.amplicon_intervals)
                    if io1 >= 0 and io2 >= 0:
                        assert self.amplicon_intervals[io1][3] == self.amplicon_intervals[io2][3]
                        bpi = self.addbp(bp, set(bpr), gpu_ops.stats_row(r["row_stats"][row]), self.amplicon_intervals[io1][3])
                        self.amplicon_interval_connections.setdefault((min(io1, io2), max(io1, io2)), set()).add(bpi)

    def _interval_arrays(self):
        ivs = self.amplicon_intervals
        return [self._tid[a[0]] for a in ivs], [a[1] for a in ivs], [a[2] for a in ivs]

    # ---- a3: ibg:721-802 ------------------------------------------------------------------------------------
    @_gc_paused
    def find_smalldel_breakpoints(self):
        self._no_nm_filter()
        t, s, e = self._interval_arrays()
        self._ctx.set_params(min_del_len=self.min_del_len, min_bp_match_cutoff=self.min_bp_match_cutoff_)
        lia = self.large_indel_alignments
        for (name, ai, ns, pe, fs, le, mq) in self._smalldel_rows(t, s, e):
            lia.setdefault(name, []).append([self.amplicon_intervals[ai][0], ns, pe, fs, le, mq])
        logging.info(_t() + "Fetched %d reads with large indels in CIGAR." % (len(lia)))
        self._call_clusters(*self._smalldel_candidates())

    def _smalldel_rows(self, t, s, e):
        """Rows (name, interval index, next_start, prev_end, first_start, last_end, mapq) in the reference's
        emission order: interval list order, file order, block order (ibg:750-762)."""
        rows = self._ctx.smalldel(t, s, e)
        if rows["n"] == 0:
            return []
        urec, inv = np.unique(rows["rec"].astype(np.int64), return_inverse=True)
        unames = self._names_of(urec)
        for nm, rc in zip(unames, urec.tolist()):
            self._name_rec.setdefault(nm, rc)
        names = [unames[k] for k in inv.tolist()]
        return list(zip(names, *(rows[k].tolist() for k in ("ai", "next_start", "prev_end", "first_start", "last_end", "mapq"))))

    def _smalldel_candidates(self):
        tuples, chrs, p1, p2 = [], [], [], []
        for name, lst in self.large_indel_alignments.items():          # ibg:765-772
            for idx, row in enumerate(lst):
                hi, lo = row[1], row[2]
                if lo > hi:
                    lo = hi
                tuples.append((name, idx, idx)); chrs.append(self._tid[row[0]]); p1.append(hi); p2.append(lo)
        n = len(tuples)
        cand = {"chr1": np.array(chrs, np.int32), "pos1": np.array(p1, np.int32), "o1": np.ones(n, np.uint8),
                "chr2": np.array(chrs, np.int32), "pos2": np.array(p2, np.int32), "o2": np.zeros(n, np.uint8),
                "rgap": np.zeros(n, np.int32), "flag": np.zeros(n, np.uint8), "mq1": np.full(n, -1, np.int32),
                "mq2": np.full(n, -1, np.int32)}
        return cand, tuples

    # ---- a4/a5: ibg:676-718 ---------------------------------------------------------------------------------
    @_gc_paused
    def find_breakpoints(self):
        self._no_nm_filter()
        t, s, e = self._interval_arrays()
        self._ctx.set_params(min_del_len=self.min_del_len, min_bp_match_cutoff=self.min_bp_match_cutoff_)
        c = self._chim_ctx.candidates_l(t, s, e)
        names = self._read_names
        rd, ii, jj = c["read"].tolist(), c["i"].tolist(), c["j"].tolist()
        tuples = [(names[rd[k]], ii[k], jj[k]) for k in range(c["n"])]
        self._call_clusters(c, tuples)

    def _pair_candidates(self, reads, intrvl1, intrvl2):
        """One batched launch of alignment2bp over the device chimeric table (ibg:432-434).  The candidates stay a
        column table (chr ids = BAM tids); _called_breakpoints clusters and calls them in one device call."""
        idx = [self._read_index[r] for r in reads]
        return self._chim_ctx.candidates_pair(idx, [self._tid[intrvl1[0]], self._tid[intrvl2[0]]], [intrvl1[1], intrvl2[1]],
                                              [intrvl1[2], intrvl2[2]])

    def _called_breakpoints(self, cand):
        """cluster_bp_list + the bpc2bp peel loop of find_interval_i (ibg:436-446) as one coral_cluster_call."""
        if cand["n"] == 0:
            return
        refs = self.lr_bamfh.references
        names = self._read_names
        cutoff = self.min_cluster_cutoff
        r = gpu_ops.cluster_arrays(self._ctx, cand, cutoff, self.max_breakpoint_distance_cutoff, cutoff,
                                   self.min_bp_match_cutoff_)
        order, ptr, sub_of = r["order"], r["cluster_ptr"], r["sub_of"]
        tups = list(zip([names[x] for x in cand["read"].tolist()], cand["i"].tolist(), cand["j"].tolist()))
        tup = tups.__getitem__
        for c in range(r["n_clusters"]):
            b, e = int(ptr[c]), int(ptr[c + 1])
            if e - b < cutoff:
                continue
            members = order[b:e]
            subs = sub_of[members]
            for k in range(int(r["n_sub"][c])):
                row = b + k
                f0 = int(r["row_first"][row])
                bp = [refs[cand["chr1"][f0]], int(r["row_pos1"][row]), '+-'[cand["o1"][f0]], refs[cand["chr2"][f0]],
                      int(r["row_pos2"][row]), '+-'[cand["o2"][f0]], tup(f0), int(cand["rgap"][f0]), int(cand["flag"][f0])]
                if r["row_nb"][row] == 0:        # bpc2bp with no member near the called position (bu:352-353)
                    yield bp, [], [0, 0, 0, 0, 0, 0]
                    continue
                yield bp, [tups[m] for m in members[subs == k].tolist()], gpu_ops.stats_row(r["row_stats"][row])

    def _cn_segment_ends(self, tabs):
        """The lookups of hash_alignment_to_seg (ibg:181-210) on the device chimeric table (coral_cn_segment_index)."""
        rows = sorted((self._tid[c], b, e, d) for c, (bs, es, ds) in tabs.items() if c in self._tid
                      for b, e, d in zip(bs.tolist(), es.tolist(), ds.tolist()))
        n_aln = int(self._aln_ptr[-1])
        lo, hi = self._chim_ctx.cn_segment_index([r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows],
                                                 [r[3] for r in rows], n_aln)
        per_read = np.diff(self._aln_ptr.astype(np.int64))
        keep = np.repeat(self._read_kept, per_read)
        n_py = sum(len(ca[1]) for ca in self.chimeric_alignments.values())
        if int(keep.sum()) != n_py:
            raise RuntimeError("chimeric_alignments no longer matches the device table fetch() built")
        return lo[keep].tolist(), hi[keep].tolist()

    @_gc_paused
    def hash_alignment_to_seg(self):
        return HostSteps.hash_alignment_to_seg(self)

    @_gc_paused
    def find_amplicon_intervals(self):
        return HostSteps.find_amplicon_intervals(self)

    # ---- a10/a11: ibg:1019-1056 -----------------------------------------------------------------------------
    def _edge_queries(self):
        """Device queries of assign_cov: sequence-edge regions in disjoint sorted batches, concordant-edge points
        (sorted, de-duplicated) with their owners, and the rbps rows (edge index, a record of the read)."""
        todo = []
        for gi, g in enumerate(self.lr_graph):
            for si, se in enumerate(g.sequence_edges):
                if se[5] == -1:
                    todo.append((self._tid[se[0]], se[1], se[2] + 1, gi, si))
        todo.sort()
        batches = [[]]
        for row in todo:   # disjoint by construction (App. A.14); split defensively otherwise
            if batches[-1] and batches[-1][-1][0] == row[0] and row[1] < batches[-1][-1][2]:
                batches.append([])
            batches[-1].append(row)
        edges = []
        for gi, g in enumerate(self.lr_graph):
            for ci, ce in enumerate(g.concordant_edges):
                edges.append((self._tid[ce[0]], ce[1], gi, ci))
        edges.sort()
        uniq, owners = [], []
        for (t, p, gi, ci) in edges:
            if uniq and uniq[-1] == (t, p):
                owners[-1].append((gi, ci))
            else:
                uniq.append((t, p)); owners.append([(gi, ci)])
        rbps = []
        for k, own in enumerate(owners):
            gi, ci = own[0]
            g = self.lr_graph[gi]
            ce = g.concordant_edges[ci]
            seen = set()
            for node in ((ce[0], ce[1], ce[2]), (ce[3], ce[4], ce[5])):
                for bpi in g.nodes[node][2]:
                    for r in g.discordant_edges[bpi][10]:
                        seen.add(r[0])
            rbps.append(seen)
        return batches, uniq, owners, rbps

    def _names_of(self, recs):
        """Read names of local records, in bulk when the BAM handle can (one pass over the host buffer)."""
        fh = self.lr_bamfh
        if hasattr(fh, "query_names"):
            return fh.query_names(recs)
        return [fh.query_name(int(r)) for r in recs]

    def _rbps_rows(self, rbps):
        """rbps name sets -> (edge index, a local record of that read) rows for coral_concordant."""
        name_rec = self._name_rec
        rb_edge, rb_rec = [], []
        for k, names in enumerate(rbps):
            hit = [name_rec[nm] for nm in names if nm in name_rec]   # a read with no loaded record cannot cover any point
            rb_rec.extend(hit)
            rb_edge.extend([k] * len(hit))
        return rb_edge, rb_rec

    def _region_counts(self, tid, start, stop):
        """(base_sum, rec_count) of disjoint sorted regions; the sharded engine all-reduces the shard sums."""
        return self._ctx.region_coverage(tid, start, stop)

    def _concordant_support(self, uniq, rbps):
        """Per unique concordant edge: ce[8] and the name set ce[9] (ibg:1043-1055)."""
        rb_edge, rb_rec = self._rbps_rows(rbps)
        cnt, rows = self._ctx.concordant([u[0] for u in uniq], [u[1] for u in uniq], rb_edge, rb_rec)
        sel = (rows["mask"] & 3) != 0
        r_edge, r_rec = rows["edge"][sel], rows["rec"][sel]
        o = np.argsort(r_edge, kind="stable")
        r_edge, r_rec = r_edge[o], r_rec[o]
        bounds = np.searchsorted(r_edge, np.arange(len(uniq) + 1))
        urec, inv = np.unique(r_rec, return_inverse=True)
        unames = self._names_of(urec)
        self._ce_recs = (unames, urec)          # a record per name of every ce[9]: compute_path_constraints' name set
        self._ce_rec_of = None
        row_names = [unames[i] for i in inv.tolist()]
        out = [set(row_names[bounds[k]:bounds[k + 1]]) for k in range(len(uniq))]
        return [int(x) for x in cnt], out

    @_gc_paused
    def assign_cov(self):
        batches, uniq, owners, rbps = self._edge_queries()
        for batch in batches:
            if not batch:
                continue
            bs, rc = self._region_counts([r[0] for r in batch], [r[1] for r in batch], [r[2] for r in batch])
            for k, (_, _, _, gi, si) in enumerate(batch):
                self.lr_graph[gi].sequence_edges[si][5] = int(rc[k])
                self.lr_graph[gi].sequence_edges[si][6] = int(bs[k])
        if not uniq:
            return
        self._ctx.set_params(min_del_len=self.min_del_len, min_bp_match_cutoff=self.min_bp_match_cutoff_)
        cnt, names_of = self._concordant_support(uniq, rbps)
        for k, own in enumerate(owners):
            for (gi, ci) in own:
                ce = self.lr_graph[gi].concordant_edges[ci]
                ce[9] = set(names_of[k]) if len(own) > 1 else names_of[k]
                ce[8] = cnt[k]

    # ---- §8f-2: concordant-read pass of compute_path_constraints (ibg:1303-1321) -----------------------------------
    def _sorted_sequence_edges(self, g):
        """Sequence edges of one graph as device columns sorted by (tid, l) + the map back to graph edge ids."""
        se = g.sequence_edges
        perm = sorted(range(len(se)), key=lambda k: (self._tid[se[k][0]], se[k][1]))
        return ([self._tid[se[k][0]] for k in perm], [se[k][1] for k in perm], [se[k][2] for k in perm],
                np.array(perm, np.int64))

    def _name_set_records(self, concordant_reads):
        """A local record per read name of the set (names come from the ce[9] sets assign_cov built)."""
        if self._ce_rec_of is None:
            unames, urec = self._ce_recs if self._ce_recs is not None else ([], np.zeros(0, np.int64))
            self._ce_rec_of = dict(zip(unames, urec.tolist()))
        rec_of, other = self._ce_rec_of, self._name_rec
        return [rec_of[nm] if nm in rec_of else other[nm] for nm in concordant_reads]

    @staticmethod
    def _distinct_runs(first, last):
        """Rows in emission order -> [(first, last, count)] in first-seen order."""
        if len(first) == 0:
            return []
        key = (first.astype(np.int64) << 32) | last.astype(np.int64)
        uk, idx, cnt = np.unique(key, return_index=True, return_counts=True)
        o = np.argsort(idx, kind="stable")
        return [(int(uk[k] >> 32), int(uk[k] & 0xffffffff), int(cnt[k])) for k in o]

    def _path_rows(self, g, intervals, **name_set):
        t, l, r, perm = self._sorted_sequence_edges(g)
        rows = self._ctx.alignment_paths([self._tid[a[0]] for a in intervals], [a[1] for a in intervals],
                                         [a[2] + 1 for a in intervals], t, l, r, **name_set)
        return rows["iv"], perm[rows["first"]], perm[rows["last"]]

    def _concordant_read_runs(self, amplicon_idx, intervals, concordant_reads):
        if not intervals or not concordant_reads:
            return []
        _, first, last = self._path_rows(self.lr_graph[amplicon_idx], intervals,
                                         name_rec=self._name_set_records(concordant_reads))
        return self._distinct_runs(first, last)

    @_gc_paused
    def compute_path_constraints(self):
        return HostSteps.compute_path_constraints(self)

    def closebam(self):
        self._ctx.close()            # first: joins a scan that may still be reading the record buffer
        self.lr_bamfh.close()

    def __del__(self):
        try:                         # an engine dropped without closebam(): the ctx goes before the records it reads
            self._ctx.close()
        except Exception:
            pass


BamToBreakpointNanopore = bam_to_breakpoint_nanopore   # name used by BASELINE.json's north_star

edge_type_to_index = {'s': 0, 'c': 1, 'd': 2}          # reference :17


def print_complete_message():
    """Reference :1398-1399 (called by CoRAL.py:30 after the run)."""
    logging.info(_t() + "Total runtime.")


@_gc_paused
def reconstruct_graph(args):
    """Reference :1333-1395 up to (not including) compute_cn_lr, which stays on the reference's cvxopt code.
    The cyclic collector is paused for the whole pass (see host_steps.gc_paused)."""
    global_names.TSTART = time.time()
    b2bn = bam_to_breakpoint_nanopore(args.lr_bam, args.cnv_seed)
    b2bn.min_bp_cov_factor = args.min_bp_support
    logging.info(_t() + "Opened LR bam files.")
    b2bn.read_cns(args.cn_seg)
    logging.info(_t() + "Completed parsing CN segment files.")
    b2bn.fetch()
    logging.info(_t() + "Completed fetching reads containing breakpoints.")
    b2bn.hash_alignment_to_seg()
    logging.info(_t() + "Completed hashing chimeric reads to CN segments.")
    b2bn.find_amplicon_intervals()
    logging.info(_t() + "Completed finding amplicon intervals.")
    b2bn.find_smalldel_breakpoints()
    logging.info(_t() + "Completed finding small del breakpoints.")
    b2bn.find_breakpoints()
    logging.info(_t() + "Completed finding all discordant breakpoints.")
    b2bn.build_graph()
    logging.info(_t() + "Breakpoint graph built for all amplicons.")
    if getattr(args, "output_bp", False):
        for gi, g in enumerate(b2bn.lr_graph):
            stats = []
            for de in g.discordant_edges:
                for bpi, bp_ in enumerate(b2bn.new_bp_list):
                    if de[:6] == bp_[:6]:
                        stats.append(b2bn.new_bp_stats[bpi])
                        break
            output_breakpoint_info_lr(g, args.output_prefix + "_amplicon" + str(gi + 1) + "_breakpoints.txt", stats)
    else:
        b2bn.assign_cov()
        logging.info(_t() + "Fetched read coverage for all sequence and concordant edges.")
        for gi, g in enumerate(b2bn.lr_graph):
            if hasattr(g, "compute_cn_lr"):          # the reference's own class (host_steps.reference_graph_class)
                g.compute_cn_lr(b2bn.normal_cov)     # reference :1388-1389, cvxopt
            else:
                logging.warning(_t() + "compute_cn_lr unavailable (the reference's breakpoint_graph / cvxopt is not "
                                "importable here): amplicon %d graph is written with CN 0.0." % (gi + 1))
            output_breakpoint_graph_lr(g, args.output_prefix + "_amplicon" + str(gi + 1) + "_graph.txt")
    return b2bn
