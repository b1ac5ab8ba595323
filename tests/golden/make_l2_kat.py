"""Known-answer vectors for the module-level operators of the hot path (SURVEY.md §8c "L2 KATs"), made by EXECUTING THE
REFERENCE's own `cigar_parsing.py` and `breakpoint_utilities.py` (they import as they are; no shim is involved):

  * every `cigar2pos*` on both strands; `alignment_from_satags` incl. the reject path (3-tuple), ties in (qs, qe)
    (first-seen order), a signature outside the nine keys (KeyError) and qe == qs (ZeroDivisionError);
  * `interval_overlap` / `interval_include` / `interval_adjacent` / `interval_overlap_l` / `interval_exclusive`, with
    reversed ('-' strand) intervals; `interval2bp` in all chromosome / coordinate orderings;
  * `alignment2bp_l` (each branch: strand switch, same-strand gap test on both strands, MAPQ gates, the skip pair over a
    low-MAPQ middle alignment, different amplicon intervals) and `alignment2bp`;
  * `cluster_bp_list` incl. the chain case whose result depends on input order, groups below the size threshold, a
    float threshold; `bpc2bp` incl. a unique mode, ties resolved by the median (ceil for '+', floor for '-'; odd and
    even counts), outliers beyond 3 sigma, large coordinates, the rgap slack of `bp_match`, flag-swapped statistics, and
    the remainder; `bp_match` on its own.

    python tests/golden/make_l2_kat.py          # writes tests/golden/l2_kat.json.gz (needs /root/reference)

The vectors pin oracle/coral_oracle.py (tests/test_l2_kat.py, CPU) and the CUDA-backed drop-in modules (same file, GPU).
"""
import gzip
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "l2_kat.json.gz")
REF_SRC = os.environ.get("CORAL_REFERENCE_SRC", "/root/reference/src")
SIGS = ["SM", "MS", "SMS", "SMD", "MDS", "SMDS", "SMI", "MIS", "SMIS"]
CHROMS = ["chr1", "chr2", "chr7", "chr12", "chrX"]


def cigar_of(sig, nums):
    return "".join("%d%s" % (n, c) for n, c in zip(nums, sig))


def cigar_vectors(rng):
    out = []
    for sig in SIGS:
        for strand in "+-":
            for _ in range(6):
                nums = [int(x) for x in rng.integers(1, 40000, len(sig))]
                rl = int(sum(n for n, c in zip(nums, sig) if c in "SMI")) + int(rng.integers(0, 3)) * int(rng.integers(0, 500))
                out.append({"fn": "cigar2pos" + sig, "args": [cigar_of(sig, nums), strand, rl]})
    return out


def sa_entry(rng, rl, sig=None, strand=None, chrom=None):
    sig = sig or SIGS[int(rng.integers(len(SIGS)))]
    nums = [int(x) for x in rng.integers(1, max(2, rl // 2), len(sig))]
    return "%s,%d,%s,%s,%d,%d" % (chrom or CHROMS[int(rng.integers(len(CHROMS)))], int(rng.integers(1, 10 ** 8)),
                                  strand or "+-"[int(rng.integers(2))], cigar_of(sig, nums), int(rng.integers(0, 61)),
                                  int(rng.integers(0, 3000)))


def satag_vectors(rng):
    out = []
    for _ in range(60):
        rl = int(rng.integers(2000, 90000))
        out.append({"fn": "alignment_from_satags", "args": [[sa_entry(rng, rl) for _ in range(int(rng.integers(1, 7)))], rl]})
    rl = 50000
    # ties in (qs, qe): same clip / match numbers on different contigs, first-seen order must survive the sort
    tie = ["chr7,1000,+,100S400M49500S,60,3", "chr2,5000,+,100S400M49500S,30,1", "chr1,9000,+,100S400M49500S,20,0",
           "chr12,77,-,49500S400M100S,11,9", "chrX,5,+,50S450M49500S,60,2"]
    out.append({"fn": "alignment_from_satags", "args": [tie, rl]})
    out.append({"fn": "alignment_from_satags", "args": [tie[::-1], rl]})
    # reject path: an entry without S or without M anywhere in the list -> ([], [], [])
    out.append({"fn": "alignment_from_satags", "args": [[tie[0], "chr1,100,+,500M,60,0"], rl]})
    out.append({"fn": "alignment_from_satags", "args": [["chr1,100,+,200S300I,60,0", tie[1]], rl]})
    out.append({"fn": "alignment_from_satags", "args": [["chr1,100,-,30M20D40M,60,0"], rl]})
    # signature outside the nine keys -> KeyError; H clips are letters too
    out.append({"fn": "alignment_from_satags", "args": [["chr1,100,+,10S20M5D5I15S,60,0"], rl]})
    out.append({"fn": "alignment_from_satags", "args": [[tie[0], "chr1,100,+,10H20M15S,60,0"], rl]})
    # qe == qs -> ZeroDivisionError (cp:268): '-' strand SM with a 1-base match
    out.append({"fn": "alignment_from_satags", "args": [["chr1,100,-,49999S1M,60,0"], rl]})
    out.append({"fn": "alignment_from_satags", "args": [[], rl]})
    return out


def rand_interval(rng, reversed_ok=True, chrom=None, lo=0, hi=200000):
    a, b = sorted(int(x) for x in rng.integers(lo, hi, 2))
    if reversed_ok and rng.random() < 0.3:
        a, b = b, a
    return [chrom or CHROMS[int(rng.integers(len(CHROMS)))], a, b]


def interval_vectors(rng):
    out = []
    for _ in range(80):
        c = CHROMS[int(rng.integers(2))]
        i1, i2 = rand_interval(rng, chrom=c if rng.random() < 0.8 else None), rand_interval(rng, False, chrom=c)
        for fn in ("interval_overlap", "interval_include", "interval_adjacent"):
            out.append({"fn": fn, "args": [i1, i2]})
    for k in range(12):       # adjacency on the nose
        a = int(rng.integers(1000, 5000))
        out.append({"fn": "interval_adjacent", "args": [["chr1", 100, a], ["chr1", a + (k % 3), a + 500]]})
        out.append({"fn": "interval_adjacent", "args": [["chr1", a + (k % 3), a + 500], ["chr1", 100, a]]})
    for _ in range(40):
        c = CHROMS[int(rng.integers(2))]
        lst = [rand_interval(rng, False, chrom=c if rng.random() < 0.8 else None) + [int(rng.integers(5))] for _ in range(int(rng.integers(0, 6)))]
        out.append({"fn": "interval_overlap_l", "args": [rand_interval(rng, chrom=c), lst]})
        out.append({"fn": "interval_include_l", "args": [rand_interval(rng, False, chrom=c), lst]})
        out.append({"fn": "interval_exclusive", "args": [rand_interval(rng, False, chrom=c) + [-1], lst]})
    for _ in range(60):
        r1 = rand_interval(rng) + ["+-"[int(rng.integers(2))]]
        r2 = rand_interval(rng, chrom=r1[0] if rng.random() < 0.6 else None) + ["+-"[int(rng.integers(2))]]
        out.append({"fn": "interval2bp", "args": [r1, r2, ["read%d" % int(rng.integers(99)), int(rng.integers(5)), int(rng.integers(5))],
                                                  int(rng.integers(-200, 3000))]})
    return out


def rand_chimeric(rng, intervals):
    """2-6 local alignments, most inside one of `intervals`, with query intervals, strands and MAPQs that reach every
    branch of alignment2bp_l."""
    n = int(rng.integers(2, 7))
    qint, rint, qual = [], [], []
    q = int(rng.integers(0, 500))
    for k in range(n):
        iv = intervals[int(rng.integers(len(intervals)))] if rng.random() < 0.85 else ["chr2", 10, 10 ** 6]
        ln = int(rng.integers(300, 20000))
        s = int(rng.integers(iv[1] - 2000, iv[2] + 2000 - ln)) if iv[2] - iv[1] > ln else iv[1]
        strand = "+-"[int(rng.integers(2))]
        if k and rng.random() < 0.4:          # follow the previous alignment on the reference: small / large gaps
            p = rint[-1]
            gap = int(rng.choice([0, 50, 99, 100, 101, 150, 900, 5000, -300]))
            strand = p[3]
            if strand == '+':
                s = p[2] + gap
            else:
                s = p[2] - gap - ln
            iv = [p[0], 0, 0]
        rint.append([iv[0], s, s + ln, '+'] if strand == '+' else [iv[0], s + ln, s, '-'])
        qint.append([q, q + ln - 1])
        q += ln + int(rng.choice([0, 1, 30, -50, -99, -100, -101, -400, 600, 2500]))
        qual.append(int(rng.choice([0, 5, 9, 10, 19, 20, 60, 60, 60])))
    return [qint, rint, qual, [0.01] * n]


def candidate_vectors(rng):
    out = []
    for t in range(150):
        c = CHROMS[int(rng.integers(3))]
        ivs = []
        for _ in range(int(rng.integers(1, 4))):
            a = int(rng.integers(10 ** 5, 10 ** 6))
            ivs.append([c if rng.random() < 0.7 else "chr12", a, a + int(rng.integers(20000, 400000)), int(rng.integers(3))])
        ca = rand_chimeric(rng, ivs)
        out.append({"fn": "alignment2bp_l", "args": ["rd%d" % t, ca, 100, 20, 100, ivs]})
        if len(ivs) > 1:
            out.append({"fn": "alignment2bp", "args": ["rd%d" % t, ca, 100, 20, ivs[0], ivs[1]]})
        out.append({"fn": "alignment2bp", "args": ["rd%d" % t, ca, 100, 20, ivs[0], ivs[0]]})
    return out


def nm_vectors(rng):
    """The edit-distance filtered forms (bu:99-126, :189-249; reached only with nm_filter = True, which the reference never
    sets): the same alignments with per-alignment NM rates on both sides of max_nm."""
    out = []
    for t in range(120):
        c = CHROMS[int(rng.integers(3))]
        ivs = []
        for _ in range(int(rng.integers(1, 4))):
            a = int(rng.integers(10 ** 5, 10 ** 6))
            ivs.append([c if rng.random() < 0.7 else "chr12", a, a + int(rng.integers(20000, 400000)), int(rng.integers(3))])
        ca = rand_chimeric(rng, ivs)
        ca[3] = [float(rng.choice([0.0, 0.01, 0.05, 0.0999, 0.1, 0.2])) for _ in ca[1]]
        out.append({"fn": "alignment2bp_nm_l", "args": ["rn%d" % t, ca, 100, 20, 0.1, 100, ivs]})
        out.append({"fn": "alignment2bp_nm", "args": ["rn%d" % t, ca, 100, 20, 0.1, ivs[0], ivs[-1]]})
    return out


def make_bp(k, c1, p1, o1, c2, p2, o2, rgap=0, flag=0, mq=(60, 60)):
    return [c1, int(p1), o1, c2, int(p2), o2, ["r%d" % k, k % 5, k % 5 + 1], int(rgap), int(flag), int(mq[0]), int(mq[1])]


def cluster_vectors(rng):
    out = []
    chain = [100000, 101500, 103000, 100100]          # one cluster; permuted: two
    for order in (chain, [100000, 103000, 101500, 100100], chain[::-1]):
        lst = [make_bp(k, "chr7", p, '+', "chr7", p + 50000, '-') for k, p in enumerate(order)]
        out.append({"fn": "cluster_bp_list", "args": [lst, 3, 2000]})
        out.append({"fn": "cluster_bp_list", "args": [lst, 5, 2000]})       # group below the threshold: one cluster as is
        out.append({"fn": "cluster_bp_list", "args": [lst, 3.5, 2000]})
    for t in range(40):
        n = int(rng.integers(1, 60))
        lst = []
        centers = [(int(rng.integers(10 ** 5, 10 ** 6)), int(rng.integers(10 ** 5, 10 ** 6))) for _ in range(int(rng.integers(1, 5)))]
        for k in range(n):
            c1, c2 = (("chr7", "chr7"), ("chr12", "chr7"), ("chr7", "chr1"))[int(rng.integers(3))]
            a, b = centers[int(rng.integers(len(centers)))]
            spread = int(rng.choice([50, 900, 1990, 2010, 4000]))
            lst.append(make_bp(k, c1, a + int(rng.integers(-spread, spread + 1)), "+-"[int(rng.integers(2))], c2,
                               b + int(rng.integers(-spread, spread + 1)), "+-"[int(rng.integers(2))],
                               rgap=int(rng.integers(-50, 400)), flag=int(rng.integers(2)),
                               mq=(int(rng.integers(20, 61)), int(rng.integers(20, 61)))))
        out.append({"fn": "cluster_bp_list", "args": [lst, float(rng.choice([1, 2, 3, 3.0, 4.2, 7])), int(rng.choice([500, 2000]))]})
    return out


def bpc2bp_vectors(rng):
    out = []

    def cluster(p1s, p2s, o1='+', o2='-', c1="chr7", c2="chr7", rgaps=None, flags=None):
        return [make_bp(k, c1, a, o1, c2, b, o2, rgap=(rgaps[k] if rgaps else 0), flag=(flags[k] if flags else k % 2),
                        mq=(20 + k % 41, 60 - k % 41)) for k, (a, b) in enumerate(zip(p1s, p2s))]
    # unique mode / tie -> median, ceil for '+', floor for '-' ({1000, 1000, 1003, 1003} -> 1002 / 1001)
    for o1, o2 in (('+', '-'), ('-', '+'), ('+', '+'), ('-', '-')):
        out.append({"fn": "bpc2bp", "args": [cluster([1000, 1000, 1003, 1003], [5000, 5001, 5001, 5004], o1, o2), 100]})
        out.append({"fn": "bpc2bp", "args": [cluster([1000, 1000, 1000, 1003], [5000, 5002, 5004, 5006, ][:4], o1, o2), 100]})
        out.append({"fn": "bpc2bp", "args": [cluster([1000, 1002, 1003], [5000, 5003, 5007], o1, o2), 100]})
        out.append({"fn": "bpc2bp", "args": [cluster([1000, 1001, 1004, 1005, 1009, 1010], [77, 77, 78, 78, 79, 80], o1, o2), 100]})
    # outliers beyond 3 sigma (they form the remainder), std floor cutoff / 2.99
    out.append({"fn": "bpc2bp", "args": [cluster([1000] * 12 + [1900], [5000] * 12 + [5001]), 100]})
    out.append({"fn": "bpc2bp", "args": [cluster([1000] * 5 + [1150, 1290], [5000] * 5 + [5100, 4890]), 100]})
    out.append({"fn": "bpc2bp", "args": [cluster([1000] * 5 + [1150, 1290], [5000] * 5 + [5100, 4890]), 50]})
    # identical large coordinates: E2 - E1^2 rounds around zero (the negative-variance fallback)
    for base in (248_000_000, 248_956_421, 155_270_559):
        out.append({"fn": "bpc2bp", "args": [cluster([base] * 7, [base - 1] * 7, c1="chr1", c2="chr1"), 100]})
        out.append({"fn": "bpc2bp", "args": [cluster([base + (k % 3) for k in range(200)], [base - 5 - (k % 2) for k in range(200)],
                                                     c1="chr1", c2="chr1"), 100]})
    # rgap slack of bp_match: members short of / past the called position by more than the cutoff
    for o1, o2 in (('+', '-'), ('-', '+'), ('+', '+'), ('-', '-')):
        p1 = [1000] * 6 + [1000 - 150, 1000 + 150, 1000 - 260, 1000 + 260, 1000 - 99, 1000 + 100]
        p2 = [9000] * 6 + [9000 + 150, 9000 - 150, 9000 + 50, 9000 - 50, 9000 + 100, 9000 - 99]
        rg = [0] * 6 + [200, 200, 200, 200, 0, 0]
        out.append({"fn": "bpc2bp", "args": [cluster(p1, p2, o1, o2, rgaps=rg), 100]})
        rg = [0] * 6 + [100, 130, 400, 400, -20, 90]
        out.append({"fn": "bpc2bp", "args": [cluster(p1, p2, o1, o2, rgaps=rg, flags=[1] * 12), 100]})
    for t in range(60):
        n = int(rng.integers(1, 80))
        a, b = int(rng.integers(10 ** 4, 2 * 10 ** 8)), int(rng.integers(10 ** 4, 2 * 10 ** 8))
        sp = int(rng.choice([2, 30, 300, 1500]))
        p1 = [a + int(x) for x in rng.integers(-sp, sp + 1, n)]
        p2 = [b + int(x) for x in rng.integers(-sp, sp + 1, n)]
        rg = [int(x) for x in rng.integers(-100, 500, n)]
        fl = [int(x) for x in rng.integers(0, 2, n)]
        out.append({"fn": "bpc2bp", "args": [cluster(p1, p2, "+-"[int(rng.integers(2))], "+-"[int(rng.integers(2))], rgaps=rg, flags=fl),
                                             int(rng.choice([100, 100, 50, 200]))]})
    for t in range(80):
        o1, o2 = "+-"[int(rng.integers(2))], "+-"[int(rng.integers(2))]
        b1 = make_bp(0, "chr7", int(rng.integers(900, 1300)), o1, "chr7", int(rng.integers(8800, 9200)), o2, rgap=int(rng.integers(-50, 400)))
        b2 = make_bp(1, "chr7" if rng.random() < 0.9 else "chr1", 1100, o1 if rng.random() < 0.9 else '+', "chr7", 9000, o2)
        out.append({"fn": "bp_match", "args": [b1, b2, b1[7] * 1.2, [100, 100]]})
    # call shapes the reference itself never uses: any rgap (float), one cutoff per coordinate, zero / negative rgap
    for t in range(120):
        o1, o2 = "+-"[int(rng.integers(2))], "+-"[int(rng.integers(2))]
        b1 = make_bp(0, "chr7", int(rng.integers(700, 1500)), o1, "chr7", int(rng.integers(8600, 9400)), o2, rgap=int(rng.integers(-50, 400)))
        b2 = make_bp(1, "chr7" if rng.random() < 0.95 else "chr1", 1100, o1 if rng.random() < 0.95 else '+', "chr7", 9000, o2)
        rgap = float(rng.choice([0.0, -3.5, 17.25, 80.0, 250.5, 1e3])) if t % 3 else float(rng.integers(-10, 600)) * 0.7
        out.append({"fn": "bp_match", "args": [b1, b2, rgap, [int(rng.choice([1, 50, 100, 333])), int(rng.choice([1, 80, 100, 500]))]]})
    return out


def tuplify(fn, args):
    """JSON has no tuples: the read tuple of a breakpoint record / the r argument of interval2bp come back as tuples."""
    def bp(rec):
        return rec[:6] + [tuple(rec[6])] + rec[7:]
    if fn == "interval2bp":
        return [args[0], args[1], tuple(args[2]), args[3]]
    if fn == "cluster_bp_list":
        return [[bp(r) for r in args[0]], args[1], args[2]]
    if fn == "bpc2bp":
        return [[bp(r) for r in args[0]], args[1]]
    if fn == "bp_match":
        return [bp(args[0]), bp(args[1]), args[2], args[3]]
    return args


def run_vector(mods, v):
    """Call v['fn'] of whichever module of `mods` has it -> JSON-able result (exceptions by type name; clusters and
    remainders as positions in the input list, found through the unique read tuple of each record)."""
    fn = next(getattr(m, v["fn"]) for m in mods if hasattr(m, v["fn"]))
    args = tuplify(v["fn"], json.loads(json.dumps(v["args"])))
    try:
        res = fn(*args)
    except Exception as ex:      # noqa: BLE001 — the exception type is part of the behaviour
        return "raise " + type(ex).__name__
    if v["fn"] == "cluster_bp_list":
        where = {r[6]: k for k, r in enumerate(args[0])}
        return [[where[r[6]] for r in cl] for cl in res]
    if v["fn"] == "bpc2bp":
        where = {r[6]: k for k, r in enumerate(args[0])}
        bp, bpr, stats, rest = res
        return {"bp": list(bp[:6]) + [list(bp[6])] + list(bp[7:]), "bpr": [list(t) for t in bpr], "stats": [float(x) for x in stats],
                "rest": [where[r[6]] for r in rest]}
    if v["fn"] == "interval_exclusive":
        return [sorted(res[0]), [list(x) for x in res[1]]]
    return json.loads(json.dumps(res, default=list))


def main():
    sys.path.insert(0, REF_SRC)
    import logging
    logging.disable(logging.CRITICAL)
    import breakpoint_utilities as bu
    import cigar_parsing as cp
    rng = np.random.default_rng(20240215)
    vectors = cigar_vectors(rng) + satag_vectors(rng) + interval_vectors(rng) + candidate_vectors(rng) + \
        cluster_vectors(rng) + bpc2bp_vectors(rng) + nm_vectors(np.random.default_rng(20240216))
    counts = {}
    for v in vectors:
        v["out"] = run_vector((cp, bu), v)
        key = v["fn"] + (" (raises)" if isinstance(v["out"], str) else "")
        counts[key] = counts.get(key, 0) + 1
    with gzip.open(OUT, "wt") as fp:
        json.dump(vectors, fp, separators=(",", ":"))
    print(len(vectors), "vectors,", os.path.getsize(OUT) // 1024, "KB")
    for k in sorted(counts):
        print("  %-36s %d" % (k, counts[k]))
    n_c = sum(1 for v in vectors if v["fn"].startswith("alignment2bp") and "_nm" not in v["fn"] and v["out"])
    print("  alignment2bp_nm* vectors that emit candidates:", sum(1 for v in vectors if "_nm" in v["fn"] and v["out"]))
    print("  alignment2bp* vectors that emit candidates:", n_c)


if __name__ == "__main__":
    main()
