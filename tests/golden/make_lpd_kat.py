"""Known-answer vectors for path_constraints.longest_path_dict (reference path_constraints.py:378-417): inputs
[paths, supports, ...] -> [edge multisets, surviving path indices, supports], produced by the UNMODIFIED reference.
Inputs: the path constraints of the golden pipeline fixtures (c3_small, c4_small) and seeded random walks with
repeated edges, sub-walks, duplicates and empty paths.  Run in the build container (needs /root/reference):

    PYTHONHASHSEED=0 python tests/golden/make_lpd_kat.py
"""
import gzip
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "lpd_kat.json.gz")
REF_SRC = os.environ.get("CORAL_REFERENCE_SRC", "/root/reference/src")


def to_paths(rows):
    """JSON rows -> paths with hashable items (edges and nodes as tuples)."""
    return [[tuple(x) for x in p] for p in rows]


def run_vector(fn, v):
    res = fn([to_paths(v["paths"]), list(v["support"]), list(v["amplicons"])])
    return {"bags": [[[list(e), c] for e, c in bag.items()] for bag in res[0]], "ids": list(res[1]), "support": list(res[2])}


def walk(rng, n_edges):
    """alternating edge / node items; edges drawn from a small alphabet so that walks repeat and contain each other"""
    k = int(rng.integers(0, 9))
    items = []
    for i in range(k):
        items.append(["sc"[i % 2] if rng.random() < 0.8 else "d", int(rng.integers(n_edges))])
        items.append(["chr1", int(rng.integers(1000)), "+-"[int(rng.integers(2))]])
    return items[:-1] if items else []


def main():
    sys.path.insert(0, REF_SRC)
    import path_constraints as pc
    vectors = []
    for case in ("c3_small", "c4_small"):
        with gzip.open(os.path.join(HERE, case + ".json.gz"), "rt") as fp:
            stage = json.load(fp)["stages"]["compute_path_constraints"]["path_constraints"]
        for amp in stage:
            vectors.append({"paths": [r[0] for r in amp], "support": [r[1] for r in amp], "amplicons": [r[2] for r in amp]})
    rng = np.random.default_rng(20240217)
    for t in range(60):
        base = [walk(rng, int(rng.integers(2, 7))) for _ in range(int(rng.integers(0, 12)))]
        paths = []
        for p in base:
            paths.append(p)
            if p and rng.random() < 0.5:            # a sub-walk (edge-aligned slice) and sometimes an exact duplicate
                a = 2 * int(rng.integers(0, (len(p) + 1) // 2))
                b = a + 1 + 2 * int(rng.integers(0, (len(p) - a + 1) // 2))
                paths.append(p[a:b])
            if rng.random() < 0.2:
                paths.append(list(p))
        order = rng.permutation(len(paths))
        paths = [paths[i] for i in order]
        vectors.append({"paths": paths, "support": [int(rng.integers(1, 50)) for _ in paths], "amplicons": [0] * len(paths)})
    for v in vectors:
        v["out"] = run_vector(pc.longest_path_dict, v)
    with gzip.open(OUT, "wt") as fp:
        json.dump(vectors, fp, separators=(",", ":"))
    dropped = sum(len(v["paths"]) - len(v["out"]["ids"]) for v in vectors)
    print(len(vectors), "vectors,", sum(len(v["paths"]) for v in vectors), "paths,", dropped, "dropped as sub-paths,",
          os.path.getsize(OUT) // 1024, "KB")


if __name__ == "__main__":
    main()
