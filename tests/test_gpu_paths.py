"""GPU parity of the concordant-read pass of compute_path_constraints (csrc/paths.cu, coral_alignment_paths) against
the oracle's restatement of infer_breakpoint_graph.py:1303-1310 + path_constraints.py:60-84, through the C ABI:
seeded datasets with random sequence-edge partitions (edges shorter than min_overlap included), more than 32
intervals, overlapping intervals, the name set given by records and by hashes, hand-made records around every
trimming threshold, and the argument / error behaviour."""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


class _G(object):
    def __init__(self, edges):
        self.sequence_edges = [[c, l, r, -1, 'd', -1, 0, r - l + 1, 0.0] for (c, l, r) in edges]


def _oracle_rows(ora, intervals, graph_edges, names, min_overlap=500, min_mapq=20):
    """(record, interval, first, last) in the reference's order: interval list order, file order."""
    from oracle.coral_oracle import alignment_run
    g = _G(graph_edges)
    out = []
    for j, (c, s, e) in enumerate(intervals):
        for read in ora.fetch(c, s, e):
            if read.mapq >= min_mapq and read.query_name in names:
                ids = alignment_run(g, [read.reference_name, read.reference_start, read.reference_end], min_overlap)
                if ids:
                    out.append((read._i, j, ids[0], ids[-1]))
    return out


def _windows(t, refs, half=400_000):
    """Per contig with records: a window around the median record start (where the amplicon's reads pile up)."""
    out = {}
    for ti in np.unique(t["tid"][t["tid"] >= 0]):
        mid = int(np.median(t["pos"][t["tid"] == ti]))
        out[refs[int(ti)][0]] = (max(0, mid - half), mid + half)
    return out


def _random_edges(rng, refs, windows, n_cuts, gaps=True):
    """Disjoint sorted edges inside each contig's window: random cut points, some edges < 500 bp, some holes between
    edges (two amplicon intervals)."""
    edges = []
    for c, (lo, hi) in windows.items():
        cuts = np.sort(rng.choice(np.arange(lo, hi), size=n_cuts, replace=False)).tolist()
        # force a few short edges
        for k in range(2, len(cuts) - 1, 7):
            cuts[k] = min(cuts[k], cuts[k - 1] + int(rng.integers(2, 400)))
        cuts = sorted(set(cuts))
        for k in range(len(cuts) - 1):
            if gaps and rng.random() < 0.08:
                continue
            edges.append((c, cuts[k], cuts[k + 1] - 1))
    order = {c: i for i, (c, _) in enumerate(refs)}
    edges.sort(key=lambda e: (order[e[0]], e[1]))
    return edges


def _gpu_rows(ctx, refs, intervals, edges, **name_set):
    tid = {c: i for i, (c, _) in enumerate(refs)}
    rows = ctx.alignment_paths([tid[a[0]] for a in intervals], [a[1] for a in intervals], [a[2] for a in intervals],
                               [tid[e[0]] for e in edges], [e[1] for e in edges], [e[2] for e in edges], **name_set)
    return list(zip(rows["rec"].tolist(), rows["iv"].tolist(), rows["first"].tolist(), rows["last"].tolist()))


@pytest.mark.parametrize("name,kw,n_cuts", [("C1", dict(scale=0.05), 60), ("C3", dict(scale=0.02, n_del=40), 120),
                                            ("C4", dict(scale=0.012), 90)])
def test_alignment_paths_match_oracle(built_lib, dataset, name, kw, n_cuts):
    from coral_b200 import _abi
    from oracle import pysam_shim
    ds = dataset(name, **kw)
    bam = _abi.BamFile(ds["paths"]["bam"], 4)
    ctx = _abi.Context(0)
    ctx.set_references(bam.references)
    ctx.load_records(bam.rec, bam.rec_off)
    ora = pysam_shim.AlignmentFile(ds["paths"]["bam"])
    refs = ds["cfg"].refs
    rng = np.random.default_rng(5)
    t = ctx.record_table()
    keys = ctx.record_keys()
    mapped = np.nonzero(t["tid"] >= 0)[0]
    windows = _windows(t, refs)
    chroms = list(windows)
    n_total = 0
    for trial in range(4):
        edges = _random_edges(rng, refs, windows, n_cuts)
        # intervals: a few large ones, overlapping on purpose, > 32 in the last trial (two chunks of the mask word)
        n_iv = 3 if trial < 3 else 40
        intervals = []
        for _ in range(n_iv):
            c = chroms[int(rng.integers(len(chroms)))]
            a, b = sorted(int(x) for x in rng.integers(windows[c][0], windows[c][1], 2))
            intervals.append((c, a, b + 1))
        pick = rng.choice(mapped, size=max(1, len(mapped) // 2), replace=False)
        names = set(bam.query_names(pick.astype(np.int64)))
        want = _oracle_rows(ora, intervals, edges, names)
        got = _gpu_rows(ctx, refs, intervals, edges, name_rec=pick)
        assert got == want
        # the same set as hashes, and split between the two forms
        got_h = _gpu_rows(ctx, refs, intervals, edges, name_lo=keys["name_lo"][pick], name_hi=keys["name_hi"][pick])
        assert got_h == want
        half = len(pick) // 2
        got_m = _gpu_rows(ctx, refs, intervals, edges, name_rec=pick[:half], name_lo=keys["name_lo"][pick[half:]],
                          name_hi=keys["name_hi"][pick[half:]])
        assert got_m == want
        n_total += len(want)
    assert n_total > 200
    ctx.close(); bam.close()


def test_alignment_paths_thresholds_and_errors(built_lib):
    """Hand-made records against hand-made edges: overlaps of 499 / 500 with the end edges, end edges of 499 / 500 bp,
    runs that shrink to two edges, MAPQ 19 / 20, a name outside the set, a record fetched by two intervals, and the
    records pysam gives no reference_end for."""
    from coral_b200 import _abi
    from oracle import pysam_shim
    from oracle.coral_oracle import alignment_run
    import tempfile, os
    refs = [("chr1", 1_000_000), ("chr2", 1_000_000)]
    # edges on chr1: [0,9999] [10000,10498] (499 bp) [10499,10998] (500 bp) [10999,19999] [20000,29999] [30000,30499]
    # (500 bp) [30500,39999]; a hole; [50000,59999]; chr2: three edges
    edges = [("chr1", 0, 9999), ("chr1", 10000, 10498), ("chr1", 10499, 10998), ("chr1", 10999, 19999),
             ("chr1", 20000, 29999), ("chr1", 30000, 30499), ("chr1", 30500, 39999), ("chr1", 50000, 59999),
             ("chr2", 100, 999), ("chr2", 1000, 1999), ("chr2", 2000, 2999)]
    rng = np.random.default_rng(2)
    bases = lambda n: "".join(rng.choice(list("ACGT"), n))         # noqa: E731
    spans = []
    for s in (9500, 9501, 9502, 9999, 10000, 10400, 10499, 10500, 10998, 10999):
        for e in (20499, 20500, 20501, 29999, 30000, 30499, 30500, 30999, 31000, 31001, 39999, 45000, 50000, 50499, 50500, 50501):
            spans.append((s, e))
    recs, names = [], set()
    for k, (s, e) in enumerate(spans):
        mq = 19 if k % 11 == 0 else 20 + (k % 40)
        nm = "r%d" % k
        if k % 13 != 0:
            names.add(nm)
        recs.append((0, s, util.make_record(nm, 0, s, mq, 0, [("M", e - s)], bases(e - s))))
    recs.append((1, 150, util.make_record("c2", 1, 150, 60, 0, [("M", 600), ("D", 1500), ("M", 400)], bases(1000))))
    names.add("c2")
    recs.sort(key=lambda r: (r[0], r[1]))
    buf, off = util.frame_records([r[2] for r in recs])
    ctx = _abi.Context(0)
    ctx.set_references([r[0] for r in refs])
    ctx.load_records(buf, off)
    t = ctx.record_table()
    g = _G(edges)
    intervals = [("chr1", 0, 40000), ("chr1", 30000, 60001), ("chr2", 0, 3000)]
    qn = [r[2][36:36 + r[2][12] - 1].decode() for r in recs]
    want = []
    for j, (c, s, e) in enumerate(intervals):
        for i in range(len(recs)):
            if refs[int(t["tid"][i])][0] == c and t["pos"][i] < e and t["endpos"][i] > s and t["mapq"][i] >= 20 and qn[i] in names:
                ids = alignment_run(g, [c, int(t["pos"][i]), int(t["endpos"][i])])
                if ids:
                    want.append((i, j, ids[0], ids[-1]))
    name_rec = [i for i in range(len(recs)) if qn[i] in names]
    got = _gpu_rows(ctx, refs, intervals, edges, name_rec=name_rec)
    assert got == want and len(want) > 20
    assert len(set(w[2:] for w in want)) > 5           # several distinct runs
    # empty inputs
    assert _gpu_rows(ctx, refs, [], edges, name_rec=name_rec) == []
    assert _gpu_rows(ctx, refs, intervals, edges) == []
    assert _gpu_rows(ctx, refs, intervals, edges[:2], name_rec=name_rec) == []
    # argument checks: unsorted / overlapping edges, negative fetch start, a record index past the end
    with pytest.raises(_abi.CoralError) as ei:
        _gpu_rows(ctx, refs, intervals, [edges[1], edges[0]] + edges[2:], name_rec=name_rec)
    assert ei.value.code == -2
    with pytest.raises(_abi.CoralError):
        _gpu_rows(ctx, refs, intervals, [("chr1", 0, 10000)] + edges[1:], name_rec=name_rec)
    with pytest.raises(_abi.CoralError):
        _gpu_rows(ctx, refs, [("chr1", -5, 100)], edges, name_rec=name_rec)
    with pytest.raises(_abi.CoralError):
        _gpu_rows(ctx, refs, intervals, edges, name_rec=[len(recs)])
    ctx.close()
    # a CIGAR-less record of the set with MAPQ >= 20 inside an interval: the reference dies with TypeError
    bad = [util.make_record("ok", 0, 100, 60, 0, [("M", 25000)], bases(25000)),
           util.make_record("nocigar", 0, 200, 60, 0, [], bases(20))]
    buf, off = util.frame_records(bad)
    ctx = _abi.Context(0)
    ctx.set_references([r[0] for r in refs])
    ctx.load_records(buf, off)
    assert _gpu_rows(ctx, refs, intervals, edges, name_rec=[0]) == [(0, 0, 0, 4)]
    with pytest.raises(_abi.CoralError) as ei:
        _gpu_rows(ctx, refs, intervals, edges, name_rec=[0, 1])
    assert ei.value.code == -4 and "TypeError" in str(ei.value)
    ctx.close()


@pytest.mark.parametrize("name,kw", [("C2", dict(scale=0.03)), ("C5", dict(scale=0.08, n_seeds=3, chroms=['chr20', 'chr21', 'chr22']))])
def test_cn_segment_index_matches_interval_lookup(built_lib, dataset, name, kw):
    """coral_cn_segment_index (the lookups of hash_alignment_to_seg, ibg:181-210) against the oracle's interval index
    on random non-overlapping CN segments: ends inside a segment, in a hole, before the first / after the last segment,
    exactly on begin / end - 1 / end, and a contig without any segment."""
    from coral_b200 import _abi
    from oracle.intervaltree_shim import IntervalTree
    ds = dataset(name, **kw)
    bam = _abi.BamFile(ds["paths"]["bam"], 4)
    ctx = _abi.Context(0)
    ctx.set_references(bam.references)
    ctx.load_records(bam.rec, bam.rec_off)
    f = ctx.fetch()
    n_aln = int(f["aln_ptr"][-1])
    assert n_aln > 50
    rng = np.random.default_rng(9)
    tids = np.unique(f["tid"])
    skip = int(tids[-1]) if len(tids) > 1 else -1            # this contig gets no CN segment
    ends = np.concatenate([f["rstart"], f["rend"]]).astype(np.int64)
    rows, trees, idx = [], {}, 0
    for t in tids.tolist():
        if t == skip:
            continue
        mine = ends[np.concatenate([f["tid"], f["tid"]]) == t]
        lo, hi = int(mine.min()), int(mine.max())
        cuts = np.unique(np.concatenate([rng.integers(lo - 1000, hi + 1000, 60), rng.choice(mine, 20), rng.choice(mine, 20) + 1]))
        tree = trees.setdefault(t, IntervalTree())
        for k in range(len(cuts) - 1):
            if rng.random() < 0.25:          # a hole
                continue
            b, e = int(cuts[k]), int(cuts[k + 1])
            tree[b:e] = idx
            rows.append((t, b, e, idx))
            idx += 1
    got_lo, got_hi = ctx.cn_segment_index([r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows],
                                          [r[3] for r in rows], n_aln)

    def want(t, p):
        if t not in trees:
            return -2
        hit = trees[t][p]
        assert len(hit) <= 1
        return list(hit)[0].data if hit else -1
    seen = set()
    for k in range(n_aln):
        t, a, b = int(f["tid"][k]), int(f["rstart"][k]), int(f["rend"][k])
        w = (want(t, min(a, b)), want(t, max(a, b)))
        assert (int(got_lo[k]), int(got_hi[k])) == w, (k, t, a, b)
        seen.update(w)
    assert -1 in seen and (skip < 0 or -2 in seen) and len(seen) > 10
    # overlapping or unsorted segments are refused
    with pytest.raises(_abi.CoralError):
        ctx.cn_segment_index([0, 0], [100, 150], [200, 300], [0, 1], n_aln)
    with pytest.raises(_abi.CoralError):
        ctx.cn_segment_index([0, 0], [500, 100], [600, 200], [0, 1], n_aln)
    ctx.close(); bam.close()
