"""Compact digests of a per-stage pipeline state (the format of oracle/ref_runner.py / tests/util.run_pipeline) for
parity at BASELINE.json's NAMED sizes, where the full state would be hundreds of MB of JSON.

Small things stay as values so that a mismatch can be read (breakpoint coordinates, edge counts, graph.txt);
large things (the chimeric table, read-support sets, large_indel_alignments, name sets, path constraints) become
(count, sha1 of the canonical JSON).  The SAME function digests the reference's state (make_full_digests.py, build
container) and the product's state (tests/test_gpu_fullsize.py, GPU box), so equality of digests = equality of state.

Reference: reconstruct_graph, /root/reference/src/infer_breakpoint_graph.py:1354-1394.
"""
import hashlib
import json

# name: (config factory name, kwargs, overrides, what it is)
FULL_CASES = {
    "full_c1": ("C1", dict(scale=1.0), {}, "C1 at its named size (30x, 2 Mb amplicon, ONT 20 kb)"),
    "full_c2": ("C2", dict(scale=1.0), {}, "C2 at its named size (3 seeds / 6 Mb, CN 40, 50x ONT 30 kb) = the bench workload"),
    "full_c3": ("C3", dict(scale=1.0), {}, "C3 at its named size (60x HiFi 15 kb, 300 in-CIGAR deletions, dense D ops)"),
    "full_c1n": ("C1", dict(scale=1.0, seed=99), {"p_n": 0.001}, "C1 at its named size with 0.1 % N bases (non-ACGT slow path of coverage)"),
    "sub_c4": ("C4", dict(scale=0.25), {}, "C4 sub-sample: all 200 junctions / 4 chromosomes / ONT 50 kb at 25x instead of 100x "
               "(the full 100x file is 4 x the records; the reference-on-shim needs > 60 GB of host RAM for it here)"),
    "sub_c5": ("C5", dict(scale=1.0, n_seeds=6, chroms=['chr19', 'chr20', 'chr21', 'chr22']), {},
               "C5 sub-sample: 4 chromosomes (chr19-22) and 6 seeds at the full 30x with the 0.5 % chimeric noise reads "
               "(the whole genome is 4.6 M records / ~190 GB decoded: does not fit this container)"),
}


def make_cfg(case):
    from sim import configs
    name, kw, over, _ = FULL_CASES[case]
    cfg = configs.CONFIGS[name](**kw)
    for k, v in over.items():
        setattr(cfg, k, v)
    return cfg


def _sha(obj):
    return hashlib.sha1(json.dumps(obj, separators=(",", ":"), sort_keys=True).encode()).hexdigest()


def _big(lst):
    return {"n": len(lst), "sha1": _sha(lst)}


def _bp_rows(rows):
    """ref_runner._bp_rows output: [chr1,pos1,o1,chr2,pos2,o2, sorted support tuples] -> support as (n, sha1)."""
    return [list(r[:6]) + [_big(r[6])] for r in rows]


def _names(x):
    """concordant-edge name set: either a sorted name list or make_golden.digest_names' {n, sha1} already."""
    if isinstance(x, dict):
        return x
    return {"n": len(x), "sha1": hashlib.sha1("\n".join(sorted(x)).encode()).hexdigest()}


def digest_state(stages):
    s = json.loads(json.dumps(stages))
    out = {}
    if "read_cns" in s:
        out["read_cns"] = s["read_cns"]
    if "fetch" in s:
        f = s["fetch"]
        out["fetch"] = {"chimeric": _big(f["chimeric"]), "nm_stats": f["nm_stats"],
                        "n_alignments": sum(len(c[1]) for c in f["chimeric"])}
    if "find_amplicon_intervals" in s:
        f = s["find_amplicon_intervals"]
        out["find_amplicon_intervals"] = {"amplicon_intervals": f["amplicon_intervals"], "connections": f["connections"],
                                          "new_bp_list": _bp_rows(f["new_bp_list"]), "new_bp_ccids": f["new_bp_ccids"]}
    if "find_smalldel_breakpoints" in s:
        f = s["find_smalldel_breakpoints"]
        out["find_smalldel_breakpoints"] = {
            "large_indel_alignments": _big(f["large_indel_alignments"]),
            "n_large_indel_rows": sum(len(v) for _, v in f["large_indel_alignments"]),
            "new_bp_list": _bp_rows(f["new_bp_list"]), "new_bp_stats": f["new_bp_stats"], "new_bp_ccids": f["new_bp_ccids"]}
    if "find_breakpoints" in s:
        f = s["find_breakpoints"]
        out["find_breakpoints"] = {"new_bp_list": _bp_rows(f["new_bp_list"]), "new_bp_stats": f["new_bp_stats"],
                                   "new_bp_ccids": f["new_bp_ccids"], "connections": f["connections"]}
    if "assign_cov" in s:
        gs = []
        for g in s["assign_cov"]["graphs"]:
            gs.append({
                "sequence_edges": g["sequence_edges"],
                "concordant_edges": [ce[:9] + [_names(ce[9])] + [ce[10]] for ce in g["concordant_edges"]],
                "discordant_edges": [de[:10] + [_big(de[10])] + [de[11]] for de in g["discordant_edges"]],
                "graph_txt": g["graph_txt"],
                "graph_txt_sha1": hashlib.sha1(g["graph_txt"].encode()).hexdigest()})
        out["assign_cov"] = {"graphs": gs}
    if "compute_path_constraints" in s:
        pc = s["compute_path_constraints"]["path_constraints"]
        out["compute_path_constraints"] = {"per_amplicon": [_big(a) for a in pc]}
    return out


def compare_digests(got, want):
    """-> list of (stage, key) that differ.  nm_stats (order-dependent f64 sum, unused downstream) within 1e-9."""
    bad = []
    for stage, fields in want.items():
        if stage not in got:
            bad.append((stage, None))
            continue
        for k, v in fields.items():
            a = got[stage].get(k)
            if k == "nm_stats":
                ok = a is not None and a[2] == v[2] and all(abs(x - y) <= 1e-9 * abs(y) for x, y in zip(a[:2], v[:2]))
            else:
                ok = a == v
            if not ok:
                bad.append((stage, k))
    return bad
