"""Host-side marshalling between the reference's Python list layouts and the C ABI tables.

No arithmetic of the hot path happens here: lists are flattened to arrays, handed to the CUDA library
through ``_abi.Context`` and the results are re-assembled into the exact Python shapes the reference's
callers expect (lists of lists, tuples ``(name, i, j)``, aliasing of the input records).
"""
import numpy as np

from . import _abi

import threading

_tls = threading.local()     # a coral_ctx is not thread-safe: one default context per thread (and per device)
STRAND = ('+', '-')


def set_default_device(device):
    """Device of this thread's default context (the engine calls this with its own device)."""
    if getattr(_tls, "device", 0) != int(device):
        _tls.device = int(device)
        _tls.ctx = None


def default_context():
    """Per-thread context for the module-level operators (cluster_bp_list, bpc2bp, ...)."""
    ctx = getattr(_tls, "ctx", None)
    if ctx is None:
        ctx = _tls.ctx = _abi.Context(getattr(_tls, "device", 0))
    return ctx


def flatten_bp_list(bp_list):
    """[[chr1,pos1,o1,chr2,pos2,o2,r,rgap,flag,mq1,mq2], ...] -> dict of arrays (chr -> local ids)."""
    ids = {}
    n = len(bp_list)
    a = {k: np.zeros(n, np.int32) for k in ("chr1", "pos1", "chr2", "pos2", "rgap", "mq1", "mq2")}
    a.update({k: np.zeros(n, np.uint8) for k in ("o1", "o2", "flag")})
    for i, bp in enumerate(bp_list):
        a["chr1"][i] = ids.setdefault(bp[0], len(ids))
        a["chr2"][i] = ids.setdefault(bp[3], len(ids))
        a["pos1"][i] = int(bp[1])
        a["pos2"][i] = int(bp[4])
        a["o1"][i] = 0 if bp[2] == '+' else 1
        a["o2"][i] = 0 if bp[5] == '+' else 1
        a["rgap"][i] = bp[7]
        a["flag"][i] = bp[8]
        a["mq1"][i] = bp[9]
        a["mq2"][i] = bp[10]
    return a


def cluster_arrays(ctx, a, min_cluster_size, bp_distance_cutoff, peel_cutoff, match_cutoff, max_peels=0):
    return ctx.cluster_call(a["chr1"], a["pos1"], a["o1"], a["chr2"], a["pos2"], a["o2"], a["rgap"], a["flag"],
                            a["mq1"], a["mq2"], float(min_cluster_size), int(bp_distance_cutoff), float(peel_cutoff),
                            int(match_cutoff), int(max_peels))


def stats_row(row):
    """bp_stats_ of one bpc2bp call; -1.0 in the std slots marks the reference's `except: ... = 0` branch."""
    return [0 if (k in (2, 3) and x == -1.0) else float(x) for k, x in enumerate(row)]


def cluster_bp_list(bp_list, min_cluster_size, bp_distance_cutoff, ctx=None):
    """Drop-in for breakpoint_utilities.cluster_bp_list (reference :252-286); clusters alias the inputs."""
    if len(bp_list) == 0:
        return []
    ctx = ctx or default_context()
    r = cluster_arrays(ctx, flatten_bp_list(bp_list), min_cluster_size, bp_distance_cutoff, float("inf"), 100)
    order, ptr = r["order"], r["cluster_ptr"]
    return [[bp_list[int(i)] for i in order[ptr[c]:ptr[c + 1]]] for c in range(r["n_clusters"])]


def bpc2bp(bp_cluster, bp_distance_cutoff, ctx=None):
    """Drop-in for breakpoint_utilities.bpc2bp (reference :299-388) on one homogeneous cluster."""
    first = bp_cluster[0]
    for b in bp_cluster:
        if (b[0], b[3], b[2], b[5]) != (first[0], first[3], first[2], first[5]):
            raise ValueError("bpc2bp: cluster members must share (chr1, chr2, o1, o2)")
    ctx = ctx or default_context()
    a = flatten_bp_list(bp_cluster)
    r = cluster_arrays(ctx, a, float("inf"), 2000, 0.0, bp_distance_cutoff, max_peels=1)
    assert r["n_clusters"] == 1 and r["n_sub"][0] == 1
    bp = list(first[:-2])
    bp[1] = int(r["row_pos1"][0])
    bp[4] = int(r["row_pos2"][0])
    if r["row_nb"][0] == 0:
        return bp, [], [0, 0, 0, 0, 0, 0], []
    sub = r["sub_of"]
    bpr = [bp_cluster[i][6] for i in range(len(bp_cluster)) if sub[i] == 0]
    rest = [bp_cluster[i] for i in range(len(bp_cluster)) if sub[i] != 0]
    return bp, bpr, stats_row(r["row_stats"][0]), rest
