"""oracle/ref_runner.py — TEST INFRASTRUCTURE.  Drives the UNMODIFIED reference
(``/root/reference/src``) through its ``reconstruct`` hot path under ``sys.modules`` shims for the
three dependencies that are not installable in this image (pysam, intervaltree, cvxopt) and dumps a
canonical JSON of the state after every stage.  Only runnable where ``/root/reference`` exists (the
build container); its outputs are committed as fixtures under ``tests/golden/`` and are what pins
``oracle/coral_oracle.py`` and the CUDA path.

Run in a FRESH process per invocation: the reference keeps all containers as class attributes
(``infer_breakpoint_graph.py:22-61``), so a second instance would inherit the first one's data.

    python -m oracle.ref_runner --bam X.bam --cn_seg cn.bed --seed seed.bed --out state.json
"""
import argparse
import io
import json
import os
import sys
import time
import types

def _ref_src():
    """CORAL_REFERENCE_SRC, else /root/reference/src (build container), else the travelling copy oracle/_ref/src
    (oracle/make_ref.py: byte-identical files, git-ignored) — what exists on the GPU box."""
    if os.environ.get("CORAL_REFERENCE_SRC"):
        return os.environ["CORAL_REFERENCE_SRC"]
    here = os.path.dirname(os.path.abspath(__file__))
    for d in ("/root/reference/src", os.path.join(here, "_ref", "src")):
        if os.path.exists(os.path.join(d, "infer_breakpoint_graph.py")):
            return d
    return "/root/reference/src"


REF_SRC = _ref_src()


def install_shims():
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.dirname(here))
    from oracle import pysam_shim, intervaltree_shim
    sys.modules["pysam"] = pysam_shim
    sys.modules["intervaltree"] = intervaltree_shim
    cvx = types.ModuleType("cvxopt")
    cvx.modeling = types.ModuleType("cvxopt.modeling")
    sys.modules["cvxopt"] = cvx
    sys.modules["cvxopt.modeling"] = cvx.modeling
    if REF_SRC not in sys.path:
        sys.path.insert(0, REF_SRC)


def _bp_rows(bps):
    out = []
    for bp in bps:
        row = list(bp[:6]) + [sorted([list(t) for t in bp[-1]])]
        out.append(row)
    return out


def canon_chimeric(ca):
    out = []
    for name, v in ca.items():
        if len(v) == 3:
            out.append([name, [], [], [], []])
        else:
            out.append([name, v[0], [r[:4] for r in v[1]], v[2], v[3]])
    return out


def canon_path(path):
    """A subpath and its reverse are the same constraint (ibg:1282-1287); which orientation is stored depends on the
    iteration order of a set of read tuples (ibg:1070, 1086), so the canonical form is the smaller of the two."""
    fwd = [list(x) for x in path]
    return min(fwd, fwd[::-1], key=repr)


def canon_path_constraints(pc):
    """{amplicon index: [paths, counts, amplicon ids]} -> per amplicon the sorted [canonical path, count, amplicon id]."""
    out = []
    for ai in sorted(pc):
        paths, counts, amps = pc[ai]
        out.append(sorted(([canon_path(p), int(c), int(a)] for p, c, a in zip(paths, counts, amps)), key=repr))
    return out


def graph_text(g, writer):
    import tempfile
    fd, p = tempfile.mkstemp(suffix=".txt")
    os.close(fd)
    try:
        writer(g, p)
        return open(p).read()
    finally:
        os.unlink(p)


def run(bam, cn_seg, seed, min_bp_support=1.0, stop_after=None, log=None):
    install_shims()
    import logging
    logging.basicConfig(stream=io.StringIO() if log is None else open(log, "w"), level=logging.WARNING)
    import global_names
    import infer_breakpoint_graph as ibg
    import breakpoint_graph as bg
    global_names.TSTART = time.time()
    st = {"stages": {}, "timing_s": {}}

    def timed(name, fn, *a):
        t = time.perf_counter()
        fn(*a)
        st["timing_s"][name] = time.perf_counter() - t

    t0 = time.perf_counter()
    b = ibg.bam_to_breakpoint_nanopore(bam, seed)
    st["timing_s"]["open"] = time.perf_counter() - t0
    b.min_bp_cov_factor = min_bp_support
    timed("read_cns", b.read_cns, cn_seg)
    st["stages"]["read_cns"] = {"normal_cov": b.normal_cov, "min_cluster_cutoff": b.min_cluster_cutoff}
    if stop_after == "read_cns":
        return st
    timed("fetch", b.fetch)
    st["stages"]["fetch"] = {
        "n_read_length": len(b.read_length),
        "chimeric": canon_chimeric(b.chimeric_alignments),
        "read_length_chimeric": {r: b.read_length[r] for r in b.chimeric_alignments},
        "nm_stats": list(b.nm_stats),
    }
    if stop_after == "fetch":
        return st
    timed("hash_alignment_to_seg", b.hash_alignment_to_seg)
    timed("find_amplicon_intervals", b.find_amplicon_intervals)
    st["stages"]["find_amplicon_intervals"] = {
        "amplicon_intervals": [list(a) for a in b.amplicon_intervals],
        "connections": sorted([[list(k), sorted(v)] for k, v in b.amplicon_interval_connections.items()]),
        "new_bp_list": _bp_rows(b.new_bp_list),
        "new_bp_ccids": list(b.new_bp_ccids),
    }
    if stop_after == "find_amplicon_intervals":
        return st
    timed("find_smalldel_breakpoints", b.find_smalldel_breakpoints)
    st["stages"]["find_smalldel_breakpoints"] = {
        "large_indel_alignments": [[k, v] for k, v in b.large_indel_alignments.items()],
        "new_bp_list": _bp_rows(b.new_bp_list),
        "new_bp_stats": [list(s) for s in b.new_bp_stats],
        "new_bp_ccids": list(b.new_bp_ccids),
    }
    if stop_after == "find_smalldel_breakpoints":
        return st
    timed("find_breakpoints", b.find_breakpoints)
    st["stages"]["find_breakpoints"] = {
        "new_bp_list": _bp_rows(b.new_bp_list),
        "new_bp_stats": [list(s) for s in b.new_bp_stats],
        "new_bp_ccids": list(b.new_bp_ccids),
        "connections": sorted([[list(k), sorted(v)] for k, v in b.amplicon_interval_connections.items()]),
    }
    if stop_after == "find_breakpoints":
        return st
    timed("build_graph", b.build_graph)
    timed("assign_cov", b.assign_cov)
    graphs = []
    for g in b.lr_graph:
        graphs.append({
            "sequence_edges": [list(se) for se in g.sequence_edges],
            "concordant_edges": [list(ce[:9]) + [sorted(ce[9])] + [ce[10]] for ce in g.concordant_edges],
            "discordant_edges": [list(de[:10]) + [sorted([list(t) for t in de[10]])] + [de[11]] for de in g.discordant_edges],
            "graph_txt": graph_text(g, bg.output_breakpoint_graph_lr),
        })
    st["stages"]["assign_cov"] = {"graphs": graphs}
    if stop_after == "assign_cov":
        return st
    # the step after the path (cycle_decomposition.py:2067): reads -> subpath constraints (ibg:1059-1323)
    timed("compute_path_constraints", b.compute_path_constraints)
    st["stages"]["compute_path_constraints"] = {"path_constraints": canon_path_constraints(b.path_constraints)}
    return st


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bam", required=True)
    ap.add_argument("--cn_seg", required=True)
    ap.add_argument("--seed", required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("--min_bp_support", type=float, default=1.0)
    ap.add_argument("--stop_after", default=None)
    a = ap.parse_args()
    st = run(a.bam, a.cn_seg, a.seed, a.min_bp_support, a.stop_after)
    with open(a.out, "w") as fp:
        json.dump(st, fp)
    print(json.dumps(st["timing_s"]))


if __name__ == "__main__":
    main()
