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


def run_pipeline(eng, cn_seg, stop_after=None, defer_cov=False):
    """Drive an engine in the reference's call order; returns canonical per-stage state (ref_runner format).
    defer_cov: do not look at normal_cov before find_amplicon_intervals needs it — with the asynchronous scan of the
    product, fetch and hash_alignment_to_seg then really run while the record scan is still in flight."""
    from oracle.ref_runner import canon_chimeric, _bp_rows
    st = {}
    eng.read_cns(cn_seg)
    if not defer_cov:
        st["read_cns"] = {"normal_cov": eng.normal_cov, "min_cluster_cutoff": eng.min_cluster_cutoff}
    eng.fetch()
    st["fetch"] = {"chimeric": canon_chimeric(eng.chimeric_alignments), "nm_stats": list(eng.nm_stats)}
    eng.hash_alignment_to_seg()
    if defer_cov:
        st["read_cns"] = {"normal_cov": eng.normal_cov, "min_cluster_cutoff": eng.min_cluster_cutoff}
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
    if stop_after == "assign_cov":
        return st
    from oracle.ref_runner import canon_path_constraints
    eng.compute_path_constraints()
    st["compute_path_constraints"] = {"path_constraints": canon_path_constraints(eng.path_constraints)}
    return st


def spawn_ranks(args, world, timeout=1200):
    """Run tests/run_rank.py as `world` processes (one rank each, all on cuda:0, gloo for the host side) and wait."""
    import os
    import socket
    import subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   PYTHONHASHSEED="0", CUDA_DEVICE_MAX_CONNECTIONS="32", GLOO_SOCKET_IFNAME="lo")
        procs.append(subprocess.Popen([sys.executable, os.path.join(here, "run_rank.py")] + [str(a) for a in args], env=env))
    rcs = []
    try:
        for p in procs:
            rcs.append(p.wait(timeout=timeout))
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    assert rcs == [0] * world, "rank exit codes %s" % rcs


def jsonify(x):
    import json
    return json.loads(json.dumps(x))


# ---- hand-made BAM records for the edge-case tests -------------------------------------------------------------
_OPS = "MIDNSHP=X"
_SEQ = "=ACMGRSVTWYHKDBN"


def make_record(name, tid, pos, mapq, flag, cigar, seq=None, aux=b"", l_seq=None):
    """One BAM alignment record (block_size .. aux) as bytes.  cigar: [(op char, length)]; seq: base string, or None
    for '*' (then l_seq gives the stored length, default 0 — QUAL is then 0xFF bytes of that length with no SEQ only
    when l_seq is 0, as htslib writes it)."""
    import struct
    nm = name.encode() + b"\0"
    cg = b"".join(struct.pack("<I", (ln << 4) | _OPS.index(op)) for op, ln in cigar)
    if seq is None:
        ls = 0 if l_seq is None else l_seq
        sq = bytes((ls + 1) // 2)
        ql = b"\xff" * ls
    else:
        ls = len(seq)
        codes = [_SEQ.index(c) for c in seq] + ([0] if ls & 1 else [])
        sq = bytes((codes[2 * k] << 4) | codes[2 * k + 1] for k in range(len(codes) // 2))
        ql = bytes([30]) * ls
    body = struct.pack("<iiBBHHHIiii", tid, pos, len(nm), mapq, 4680, len(cigar), flag, ls, -1, -1, 0) + nm + cg + sq + ql + aux
    return struct.pack("<i", len(body)) + body


def frame_records(records, align=16):
    """Concatenate records with 16-byte aligned starts -> (uint8 array, uint64 offsets[n+1]); offsets[n] is the end of
    the last record's padded frame, as the host reader produces them."""
    import numpy as np
    buf, off = bytearray(), [0]
    for r in records:
        buf += r
        buf += b"\0" * (-len(buf) % align)
        off.append(len(buf))
    return np.frombuffer(bytes(buf), np.uint8).copy(), np.array(off, np.uint64)


def aux_tag(tag, ty, value):
    import struct
    t = tag.encode() + ty.encode()
    if ty in "ZH":
        return t + value.encode() + b"\0"
    if ty == "A":
        return t + value.encode()
    if ty == "B":
        sub, vals = value
        return t + sub.encode() + struct.pack("<i", len(vals)) + b"".join(struct.pack("<" + {"c": "b", "C": "B", "s": "h", "S": "H", "i": "i", "I": "I", "f": "f"}[sub], v) for v in vals)
    return t + struct.pack("<" + {"c": "b", "C": "B", "s": "h", "S": "H", "i": "i", "I": "I", "f": "f"}[ty], value)
