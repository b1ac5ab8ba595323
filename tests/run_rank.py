"""One RANK of a multi-process test of the peer-memory transport (csrc/xchg.cu): `world` processes share cuda:0 (the
test box has one GPU), map each other's exchange windows through CUDA IPC exactly as one-process-per-GPU ranks do, and
talk host-side over torch.distributed / gloo.  Started by tests/util.py::spawn_ranks with RANK / WORLD_SIZE /
MASTER_ADDR / MASTER_PORT / PYTHONHASHSEED=0 in the environment.

    python tests/run_rank.py xchg <out.json>
    python tests/run_rank.py pipeline <golden case> <out.json>
    python tests/run_rank.py pass <golden case> <out.json>
"""
import faulthandler
import json
import os
import sys
import tempfile

import numpy as np

faulthandler.enable()
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "golden"))
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")


def same(a, b, skip=()):
    """Names of the entries of two result dicts that differ."""
    bad = []
    for k in a:
        if k in skip:
            continue
        x, y = a[k], b.get(k)
        if isinstance(x, np.ndarray):
            if y is None or x.shape != y.shape or not np.array_equal(x, y):
                bad.append(k)
        elif isinstance(x, float):
            if abs(x - y) > 1e-9 * max(1.0, abs(y)):
                bad.append(k)
        elif isinstance(x, (int, np.integer)):
            if int(x) != int(y):
                bad.append(k)
    return bad


def same_clusters(a, b):
    """coral_get_clusters results: the row arrays are only defined for the first n_sub[c] slots of every cluster."""
    bad = same(a, b, skip=("row_pos1", "row_pos2", "row_nb", "row_first", "row_stats"))
    if bad or not b:
        return bad or ["missing"]
    ptr, n_sub = a["cluster_ptr"], a["n_sub"]
    idx = np.array([int(ptr[c]) + k for c in range(len(n_sub)) for k in range(int(n_sub[c]))], np.int64)
    for k in ("row_pos1", "row_pos2", "row_nb", "row_first", "row_stats"):
        if not np.array_equal(np.asarray(a[k])[idx], np.asarray(b[k])[idx]):
            bad.append(k)
    return bad


def xchg_mode(comm, out_path):
    import torch
    from coral_b200 import _abi, sharded
    r, world = comm.rank, comm.world
    dev = torch.device("cuda", 0)
    ctx = _abi.Context(0)
    x = sharded.make_xchg(comm, 0, slot_bytes=(1 << 20, 0, 0, 0, 0, 1 << 16, 0, 0), timeout_s=60)
    bad = []
    for it in range(5):                       # several rounds on one channel: both slot sets, sequence numbers
        def size(q):
            return 16 * ((q * 7 + it * 3) % 11) * (it + 1)         # different sizes per rank and round, some empty
        n = size(r)
        msg = (torch.arange(n, device=dev, dtype=torch.int64) % 251 + r + it).to(torch.uint8)
        torch.cuda.synchronize()
        b, st, sz = ctx.xchg_allgather(x, _abi.CH_SA, msg.data_ptr() if n else 0, n)
        for q in range(world):
            got = ctx.read_device(b + int(st[q]), int(sz[q]))
            want = ((np.arange(size(q), dtype=np.int64) % 251 + q + it) & 255).astype(np.uint8)
            if int(sz[q]) != size(q) or not np.array_equal(got, want):
                bad.append(("allgather", it, q))
        tot = ctx.xchg_allreduce(x, _abi.CH_COV, np.array([r + 1, it, -r], np.int64))
        if tot.tolist() != [world * (world + 1) // 2, it * world, -world * (world - 1) // 2]:
            bad.append(("allreduce", it, tot.tolist()))
    refused = []
    for ptr, n, ch in ((torch.zeros((1 << 20) + 4096, dtype=torch.uint8, device=dev), (1 << 20) + 4096, _abi.CH_SA), (None, 0, _abi.CH_CHIM)):
        try:                                  # beyond the slot size / a channel without slots: refused, nothing launched
            ctx.xchg_allgather(x, ch, ptr.data_ptr() if ptr is not None else 0, n)
            refused.append(False)
        except _abi.CoralError:
            refused.append(True)
    res = {"bad": bad, "refused": refused, "timeouts": x.timeouts()}
    # a peer that never takes part: the wait gives up instead of hanging the GPU
    x.set_timeout(1.0)
    comm.all_gather(0)
    if r == 0:
        try:
            ctx.xchg_allgather(x, _abi.CH_SA, 0, 0)
            res["lonely"] = "returned"
        except _abi.CoralError as ex:
            res["lonely"] = "error: %s" % ex
        res["lonely_timeouts"] = x.timeouts()
    sharded.close_xchg(comm, x)
    ctx.close()
    allres = comm.all_gather(res)
    if r == 0:
        json.dump(allres, open(out_path, "w"))


def pipeline_mode(comm, case, out_path):
    import util
    import test_golden_cpu as G
    from coral_b200 import _abi, sharded
    cfg, data, paths = G.golden_dataset(case, None, outdir=tempfile.mkdtemp(prefix="coral_rank%d_" % comm.rank))
    fh = _abi.BamFile(paths["bam"], 2)
    x = sharded.make_xchg(comm, 0, timeout_s=120)
    eng = sharded.make_sharded_engine(fh, paths["seed"], comm, device=0, xchg=x)
    st = util.jsonify(util.run_pipeline(eng, paths["cn_seg"]))
    tmo = x.timeouts()
    eng.closebam()
    sharded.close_xchg(comm, x)
    import hashlib
    cmp_state = dict(st, fetch=dict(st["fetch"], nm_stats=None))
    dig = hashlib.sha1(json.dumps(cmp_state, sort_keys=True).encode()).hexdigest()
    digs = comm.all_gather((dig, tmo))
    if comm.rank == 0:
        json.dump({"state": st, "digests": [d[0] for d in digs], "timeouts": [d[1] for d in digs]}, open(out_path, "w"))


def pass_mode(comm, case, out_path):
    import bench
    import make_golden
    from sim.generate import simulate
    from coral_b200 import gpu_ops, sharded
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from coral_b200.records import RecordSet
    cfg = make_golden.make_cfg(case)
    data = simulate(cfg)
    recs = RecordSet([c for c, _ in cfg.refs], data.rec, data.rec_off)
    cns, seed = bench.write_inputs(cfg, tempfile.mkdtemp(prefix="coral_sp%d_" % comm.rank))
    eng1 = bam_to_breakpoint_nanopore(None, seed, records=recs, placement="copy")
    plan1 = bench.build_plan(eng1, cns)
    want = bench.device_pass(eng1._ctx, plan1, gpu_ops, pipelined=False)
    eng1.closebam()
    x = sharded.make_xchg(comm, 0, timeout_s=120)
    eng = sharded.make_sharded_engine(recs, seed, comm, device=0, placement="copy", xchg=x)
    plan = bench.build_plan(eng, cns)
    eng.materialise = True
    info = {"replayed": [], "bad": []}
    for it in range(7):
        if it == 4:       # the exchange chain next to the record scan (pre-scan + full scan on its own stream)
            eng._ctx.set_option("pass_overlap_scan", 1)
            eng._ctx.set_option("pass_scan_reserve", 74)
        got = bench.sharded_device_pass(eng, plan, gpu_ops, pipelined=(it > 0))
        info["replayed"].append(bool(got.get("replayed")))
        bad = same(want["fetch"], got["fetch"], skip=("read_first_rec", "nm_sum", "nm_sumsq"))
        bad += ["cand." + k for k in same(want["cand"], got["cand"])]
        if "cl2" in want:
            bad += ["cl2." + k for k in same_clusters(want["cl2"], got.get("cl2", {}))]
        for k in ("normal", "seq"):
            if not (np.array_equal(want[k][0], got[k][0]) and np.array_equal(want[k][1], got[k][1])):
                bad.append(k)
        if not np.array_equal(np.asarray(want["conc"][0]), np.asarray(got["conc"])):
            bad.append("conc")
        merged = eng._smalldel_merge(got["smalldel"][1])
        w = want["smalldel"]
        if [m[2:8] for m in merged] != list(zip(*(w[k].tolist() for k in ("ai", "next_start", "prev_end", "first_start", "last_end", "mapq")))):
            bad.append("smalldel")
        info["bad"].append(bad)
    info.update(timeouts=x.timeouts(), n_reads=int(got["fetch"]["n_reads"]), n_cand=int(got["cand"]["n"]), smalldel_rows=len(merged),
                launches_last_pass=None)
    eng.closebam()
    sharded.close_xchg(comm, x)
    allinfo = comm.all_gather(info)
    if comm.rank == 0:
        json.dump({"ranks": allinfo, "want_reads": int(want["fetch"]["n_reads"]), "want_cand": int(want["cand"]["n"])}, open(out_path, "w"))


def main():
    import torch.distributed as dist
    from coral_b200 import sharded
    dist.init_process_group("gloo")
    comm = sharded.TorchComm()
    mode = sys.argv[1]
    try:
        if mode == "xchg":
            xchg_mode(comm, sys.argv[2])
        elif mode == "pipeline":
            pipeline_mode(comm, sys.argv[2], sys.argv[3])
        else:
            pass_mode(comm, sys.argv[2], sys.argv[3])
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
