"""Shared test helpers: full-pipeline drivers for the oracle and the product, canonical state dumps."""
import sys
import types


def oracle_engine(bam, seed):
    """CPU pipeline = oracle hot path + the (shared, separately pinned) host steps."""
    from coral_b200.host_steps import HostSteps
    from oracle import coral_oracle

    class OracleEngine(HostSteps, coral_oracle.OracleHotPath):
        _ops = coral_oracle

        def __init__(self, bam, seed):
            self._init_state()
            coral_oracle.OracleHotPath.__init__(self, bam)
            self._init_state()
            self._read_seeds(seed)

        def read_cns(self, cns):
            self._parse_cns(cns)
            segs, total = self._median_cn_segments()
            self.normal_coverage(segs, total)

    return OracleEngine(bam, seed)


def run_pipeline(eng, cn_seg, stop_after=None):
    """Drive an engine in the reference's call order; returns canonical per-stage state (ref_runner format)."""
    from oracle.ref_runner import canon_chimeric, _bp_rows
    st = {}
    eng.read_cns(cn_seg)
    st["read_cns"] = {"normal_cov": eng.normal_cov, "min_cluster_cutoff": eng.min_cluster_cutoff}
    eng.fetch()
    st["fetch"] = {"chimeric": canon_chimeric(eng.chimeric_alignments), "nm_stats": list(eng.nm_stats)}
    eng.hash_alignment_to_seg()
    eng.find_amplicon_intervals()
    st["find_amplicon_intervals"] = {
        "amplicon_intervals": [list(a) for a in eng.amplicon_intervals],
        "connections": sorted([[list(k), sorted(v)] for k, v in eng.amplicon_interval_connections.items()]),
        "new_bp_list": _bp_rows(eng.new_bp_list), "new_bp_ccids": list(eng.new_bp_ccids)}
    eng.find_smalldel_breakpoints()
    st["find_smalldel_breakpoints"] = {
        "large_indel_alignments": [[k, v] for k, v in eng.large_indel_alignments.items()],
        "new_bp_list": _bp_rows(eng.new_bp_list), "new_bp_stats": [list(s) for s in eng.new_bp_stats],
        "new_bp_ccids": list(eng.new_bp_ccids)}
    eng.find_breakpoints()
    st["find_breakpoints"] = {
        "new_bp_list": _bp_rows(eng.new_bp_list), "new_bp_stats": [list(s) for s in eng.new_bp_stats],
        "new_bp_ccids": list(eng.new_bp_ccids),
        "connections": sorted([[list(k), sorted(v)] for k, v in eng.amplicon_interval_connections.items()])}
    eng.build_graph()
    eng.assign_cov()
    from coral_b200.breakpoint_graph import output_breakpoint_graph_lr
    import os
    import tempfile
    graphs = []
    for g in eng.lr_graph:
        fd, p = tempfile.mkstemp(suffix=".txt")
        os.close(fd)
        output_breakpoint_graph_lr(g, p)
        txt = open(p).read()
        os.unlink(p)
        graphs.append({
            "sequence_edges": [list(se) for se in g.sequence_edges],
            "concordant_edges": [list(ce[:9]) + [sorted(ce[9])] + [ce[10]] for ce in g.concordant_edges],
            "discordant_edges": [list(de[:10]) + [sorted([list(t) for t in de[10]])] + [de[11]] for de in g.discordant_edges],
            "graph_txt": txt})
    st["assign_cov"] = {"graphs": graphs}
    return st


def jsonify(x):
    import json
    return json.loads(json.dumps(x))
