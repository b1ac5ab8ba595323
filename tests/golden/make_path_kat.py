"""Known-answer vectors for the subpath-constraint step, made by EXECUTING THE REFERENCE
(infer_breakpoint_graph.py:1059-1293 compute_path_constraints and path_constraints.py) on seeded random states:
small breakpoint graphs with reads that walk them — one breakpoint, several breakpoints, one / several in-CIGAR
deletions, both kinds on one read, overlapping local alignments, repeated breakpoints, reads across a hole between two
amplicon intervals.  The pipeline fixtures only reach three of the five read classes; these reach all of them.

    python tests/golden/make_path_kat.py        # writes tests/golden/path_kat.json.gz (needs /root/reference)

`build_state` is shared with tests/test_path_kat_cpu.py, which rebuilds the same states on the drop-in classes.
"""
import gzip
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(HERE, "path_kat.json.gz")
N_STATES = 160


def make_state(seed):
    """A neutral description of one state: segments, concordant edges, discordant edges with their read tuples, the
    chimeric / large-indel containers of the reads, and direct calls for the module-level functions."""
    rng = np.random.default_rng(1000 + seed)
    chroms = ["chr1"] if seed % 3 else ["chr1", "chr2"]
    segs, conc = [], []
    for c in chroms:
        pos = int(rng.integers(1_000, 50_000))
        n = int(rng.integers(4, 10))
        for k in range(n):
            size = int(rng.integers(520, 6000)) if rng.random() > 0.15 else int(rng.integers(50, 499))
            segs.append([c, pos, pos + size - 1])
            if k + 1 < n:
                if rng.random() < 0.1:
                    pos += size + int(rng.integers(2_000, 9_000))      # hole: another amplicon interval starts
                else:
                    conc.append([c, pos + size - 1])
                    pos += size
    ends = [(s[0], s[1], '-') for s in segs] + [(s[0], s[2], '+') for s in segs]
    disc = []
    for _ in range(int(rng.integers(1, 5))):
        a, b = (ends[int(x)] for x in rng.choice(len(ends), 2, replace=False))
        if (a, b) not in [(d[0], d[1]) for d in disc] and (b, a) not in [(d[0], d[1]) for d in disc]:
            disc.append([a, b, []])
    seg_of_end = {}
    for k, s in enumerate(segs):
        seg_of_end[(s[0], s[1], '-')] = k
        seg_of_end[(s[0], s[2], '+')] = k

    def arrive(node, length):
        """Local alignment whose walk ENDS at `node`."""
        c, p, o = node
        j = int(rng.integers(-30, 31)) if rng.random() < 0.3 else 0
        return [c, p - length, p + j, '+'] if o == '+' else [c, p + length, p + j, '-']

    def depart(node, length):
        """Local alignment whose walk STARTS at `node`."""
        c, p, o = node
        j = int(rng.integers(-30, 31)) if rng.random() < 0.3 else 0
        return [c, p + j, p + length, '+'] if o == '-' else [c, p + j, p - length, '-']

    chimeric, large_indel = {}, {}
    n_reads = int(rng.integers(3, 12))
    for r in range(n_reads):
        name = "read%d_%d" % (seed, r)
        kind = rng.choice(["one", "many", "del1", "deln", "mixed"], p=[0.25, 0.25, 0.15, 0.15, 0.2])
        if kind in ("one", "many", "mixed"):
            hops = 1 if kind == "one" else int(rng.integers(1, 4))
            rints, tuples, last_exit = [], [], None
            for h in range(hops):
                di = int(rng.integers(len(disc)))
                fwd = rng.random() < 0.5
                n1, n2 = (disc[di][0], disc[di][1]) if fwd else (disc[di][1], disc[di][0])
                for _ in range(8):
                    # a read that entered an edge at one end reaches the next breakpoint at an end of the other kind;
                    # the reference asserts on the rest (path_constraints.py:315) — kept, but rare
                    if h == 0 or n1[2] != last_exit[2] or rng.random() < 0.03:
                        break
                    di = int(rng.integers(len(disc)))
                    fwd = rng.random() < 0.5
                    n1, n2 = (disc[di][0], disc[di][1]) if fwd else (disc[di][1], disc[di][0])
                last_exit = n2
                if h == 0:
                    rints.append(arrive(n1, int(rng.integers(300, 9000))))
                else:       # the previous alignment should reach n1: stretch it (it may not — the reference copes)
                    prev = rints[-1]
                    if prev[0] == n1[0] and rng.random() < 0.8:
                        prev[2] = n1[1]
                rints.append(depart(n2, int(rng.integers(300, 9000))))
                i, j = len(rints) - 2, len(rints) - 1
                tuples.append((di, (i, j) if fwd else (j, i)))
            q, qints = 0, []
            for a in rints:
                ln = abs(a[2] - a[1]) + 1
                back = int(rng.integers(101, 400)) if rng.random() < 0.07 else int(rng.integers(0, 60))
                qs = max(0, q - back)
                qints.append([qs, qs + ln - 1])
                q = qs + ln
            for a in rints:
                if rng.random() < 0.5:
                    a.append(set([int(x) for x in rng.integers(0, 5, 2)]))       # the 5th element hash_alignment_to_seg adds
            chimeric[name] = (qints, rints, [int(x) for x in rng.integers(0, 61, len(rints))], [0.01] * len(rints))
            for di, (i, j) in tuples:
                disc[di][2].append((name, i, j))
        if kind in ("del1", "deln", "mixed"):
            # in-CIGAR deletions: rows [chr, next_start, prev_end, first_start, last_end, mapq]; the breakpoint is
            # (chr, next_start, '-') -> (chr, prev_end, '+') when such nodes exist, else a random discordant edge
            n_del = 1 if kind == "del1" else int(rng.integers(1, 4))
            flip_del = rng.random() < 0.5
            rows = []
            if kind == "mixed" and name in chimeric:
                host = chimeric[name][1][int(rng.integers(len(chimeric[name][1])))]
                lo, hi = min(host[1], host[2]), max(host[1], host[2])
                c = host[0]
            else:
                k0 = int(rng.integers(len(segs)))
                c = segs[k0][0]
                lo = segs[k0][1] - int(rng.integers(0, 3000))
                hi = lo + int(rng.integers(4000, 30000))
            if rng.random() < 0.1:
                lo, hi = hi, lo                                        # the inconsistent rows the reference discards
            cuts = sorted(int(x) for x in rng.integers(min(lo, hi) + 1, max(lo, hi, min(lo, hi) + 3), 2 * n_del))
            for d in range(n_del):
                a, b = cuts[2 * d], cuts[2 * d + 1]
                if rng.random() < 0.75:     # snap to a real pair of nodes so that the path can be valid
                    cand = [(s1, s2) for s1 in segs for s2 in segs if s1[0] == s2[0] == c and s1[2] < s2[1]
                            and min(lo, hi) < s1[2] and s2[1] < max(lo, hi) and (d == 0 or s1[2] > rows[-1][1])]
                    if cand:
                        s1, s2 = cand[int(rng.integers(min(len(cand), 3)))]
                        a, b = s1[2], s2[1]
                rows.append([c, b, a, lo, hi if rng.random() > 0.1 else hi + 7, 60])
            if rng.random() < 0.3:
                rng.shuffle(rows)
            large_indel[name] = rows
            for idx, row in enumerate(rows):
                n1, n2 = (row[0], row[1], '-'), (row[0], row[2], '+')
                di = None
                for k, d in enumerate(disc):
                    if (d[0], d[1]) == (n1, n2):
                        di = k
                if di is None and n1 in seg_of_end and n2 in seg_of_end and rng.random() < 0.8:
                    # the pipeline orients these edges ('-' node, '+' node) (ibg:772); the other orientation is what
                    # lets the several-deletions class get past its first alignment (path_constraints.py:353-362)
                    if kind == "deln" and flip_del:
                        n1, n2 = n2, n1
                    disc.append([n1, n2, []])
                    di = len(disc) - 1
                if di is None:
                    di = int(rng.integers(len(disc)))
                if rng.random() < 0.9:
                    disc[di][2].append((name, idx, idx))
    # module-level calls on the same graph
    calls = []
    span = [min(s[1] for s in segs), max(s[2] for s in segs)]
    for _ in range(12):
        c = chroms[int(rng.integers(len(chroms)))]
        a, b = sorted(int(x) for x in rng.integers(span[0] - 2000, span[1] + 2000, 2))
        calls.append(["alignment_to_path", [c, a, b]])
    for _ in range(12):
        node = ends[int(rng.integers(len(ends)))]
        al = arrive(node, int(rng.integers(100, 12000))) if rng.random() < 0.5 else depart(node, int(rng.integers(100, 12000)))
        calls.append(["chimeric_alignment_to_path_l", al, list(node)])
        calls.append(["chimeric_alignment_to_path_r", al, list(node)])
    for _ in range(8):
        a, b = (ends[int(x)] for x in rng.choice(len(ends), 2, replace=False))
        if a[2] != b[2]:
            calls.append(["traverse_through_sequence_edge", list(a), list(b)])
    return {"seed": seed, "segments": segs, "concordant": conc,
            "discordant": [[list(d[0]), list(d[1]), [list(t) for t in d[2]]] for d in disc],
            "chimeric": {k: [v[0], [[x if not isinstance(x, set) else sorted(x) for x in a] for a in v[1]], v[2], v[3]]
                         for k, v in chimeric.items()},
            "large_indel": large_indel, "calls": calls}


def build_state(state, graph_cls):
    """The state on a graph class (the reference's or the drop-in's): (graph, chimeric_alignments, large_indel)."""
    g = graph_cls()
    for c, l, r in state["segments"]:
        g.add_node((c, l, '-'))
        g.add_node((c, r, '+'))
        g.add_sequence_edge(c, l, r)
    for c, p in state["concordant"]:
        g.add_concordant_edge(c, p, '+', c, p + 1, '-', reads=set())
    for n1, n2, tuples in state["discordant"]:
        reads = set(tuple(t) for t in tuples)
        g.add_discordant_edge(n1[0], n1[1], n1[2], n2[0], n2[1], n2[2], lr_count=len(reads), reads=reads)
    chim = {}
    for k, v in state["chimeric"].items():
        rints = [[set(x) if isinstance(x, list) else x for x in a] for a in v[1]]
        chim[k] = ([list(q) for q in v[0]], rints, list(v[2]), list(v[3]))
    lia = {k: [list(r) for r in rows] for k, rows in state["large_indel"].items()}
    return g, chim, lia


def run_calls(state, g, mod):
    out = []
    for call in state["calls"]:
        fn = getattr(mod, call[0])
        try:
            if call[0] == "alignment_to_path":
                res = fn(g, list(call[1]))
            elif call[0] in ("chimeric_alignment_to_path_l", "chimeric_alignment_to_path_r"):
                res = fn(g, [list(call[1])], 0, tuple(call[2]))
            else:
                res = fn(g, tuple(call[1]), tuple(call[2]))
            out.append([list(x) for x in res])
        except Exception as ex:      # noqa: BLE001 — the exception type is part of the behaviour
            out.append("raise " + type(ex).__name__)
    return out


class _Holder(object):
    """The attributes compute_path_constraints reads (ibg:1059-1323); no intervals -> the record loop is empty."""
    min_bp_match_cutoff_ = 100

    def __init__(self, g, chim, lia):
        self.lr_graph = [g]
        self.path_constraints, self.longest_path_constraints = {}, {}
        self.chimeric_alignments, self.large_indel_alignments = chim, lia
        self.amplicon_intervals, self.ccid2id = [], {}
        self.lr_bamfh = None


def main():
    sys.path.insert(0, ROOT)
    from oracle import ref_runner
    ref_runner.install_shims()
    import logging
    logging.disable(logging.CRITICAL)
    import breakpoint_graph as ref_bg
    import infer_breakpoint_graph as ref_ibg
    import path_constraints as ref_pc
    states = []
    n_paths = n_valid = 0
    for seed in range(N_STATES):
        st = make_state(seed)
        g, chim, lia = build_state(st, ref_bg.BreakpointGraph)
        st["call_results"] = run_calls(st, g, ref_pc)
        h = _Holder(g, chim, lia)
        try:
            ref_ibg.bam_to_breakpoint_nanopore.compute_path_constraints(h)
            st["path_constraints"] = ref_runner.canon_path_constraints(h.path_constraints)
            n_valid += len(st["path_constraints"][0])
        except Exception as ex:      # noqa: BLE001
            st["path_constraints"] = "raise " + type(ex).__name__
        n_paths += sum(len(d[2]) for d in st["discordant"])
        states.append(st)
    with gzip.open(OUT, "wt") as fp:
        json.dump(states, fp, separators=(",", ":"))
    raised = sum(1 for s in states if isinstance(s["path_constraints"], str))
    print("states", len(states), "read tuples", n_paths, "constraints", n_valid, "raising states", raised,
          "size KB", os.path.getsize(OUT) // 1024)


if __name__ == "__main__":
    main()
