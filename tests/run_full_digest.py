"""One full-size case (tests/golden/full_digest.py::FULL_CASES) through the CUDA class -> digest JSON.

Run by tests/test_gpu_fullsize.py in a subprocess with PYTHONHASHSEED=0 (the reference iterates a set of read names,
infer_breakpoint_graph.py:432; the digests were produced with the same hash seed).  The records go to the class as an
in-memory record set (what bench.py's e2e arm does), so no 13 GB BGZF file is written on the GPU box; the BGZF route is
covered at small sizes by tests/test_gpu_pipeline.py and tests/test_gpu_device_ingest.py.

    python tests/run_full_digest.py CASE OUT.json [placement] [--world N]
"""
import json
import os
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "golden"))
import full_digest as FD  # noqa: E402
import util  # noqa: E402


def write_inputs(cfg, outdir):
    cns = os.path.join(outdir, "cn_seg.bed")
    with open(cns, "w") as fp:
        for (c, s, e, cn) in cfg.cn_rows:
            fp.write("%s\t%d\t%d\t%g\n" % (c, s, e, cn))
    seed = os.path.join(outdir, "seed.bed")
    with open(seed, "w") as fp:
        for (c, s, e) in cfg.seeds:
            fp.write("%s\t%d\t%d\n" % (c, s, e))
    return cns, seed


def main():
    case, out = sys.argv[1:3]
    placement = sys.argv[3] if len(sys.argv) > 3 and not sys.argv[3].startswith("--") else "auto"
    world = int(sys.argv[sys.argv.index("--world") + 1]) if "--world" in sys.argv else 1
    import numpy as np
    import torch
    from sim.generate import simulate
    from coral_b200.records import RecordSet
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from coral_b200.host_steps import gc_paused
    cfg = FD.make_cfg(case)
    keep = {}

    def alloc(nbytes):
        keep["t"] = torch.empty(nbytes + 64, dtype=torch.uint8).pin_memory()
        return keep["t"].numpy()[:nbytes]
    t0 = time.time()
    data = simulate(cfg, alloc=alloc)
    t_sim = time.time() - t0
    recs = RecordSet([c for c, _ in cfg.refs], data.rec, data.rec_off, rec_slack=64)
    cns, seed = write_inputs(cfg, tempfile.mkdtemp(prefix="full_" + case))
    t0 = time.time()
    with gc_paused():
        if world == 1:
            eng = bam_to_breakpoint_nanopore(None, seed, records=recs, placement=placement)
            st = util.run_pipeline(eng, cns)
        else:
            st = sharded_states(world, recs, cns, seed)
    t_run = time.time() - t0
    dg = FD.digest_state(st)
    with open(out, "w") as fp:
        json.dump({"case": case, "n_records": int(data.n_records), "decoded_bytes": int(data.rec_off[-1]),
                   "simulate_s": t_sim, "pipeline_s": t_run, "stages": dg}, fp)


def sharded_states(world, recs, cn_seg, seed):
    """`world` contiguous shards of the record set as thread ranks on cuda:0 (ThreadComm = the collective's test double;
    every kernel and every pack / unpack step is the real one).  Rank 0's state is returned."""
    import threading
    from coral_b200 import sharded
    comms = sharded.ThreadComm.make(world)
    outs, err = [None] * world, []

    def body(r):
        try:
            eng = sharded.make_sharded_engine(recs, seed, comms[r], device=0)
            outs[r] = util.run_pipeline(eng, cn_seg)
        except Exception as ex:      # noqa
            import traceback
            traceback.print_exc()
            err.append(ex)
            comms[r]._s.barrier.abort()
    th = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    if err:
        raise err[0]
    return util.jsonify(outs[0])


if __name__ == "__main__":
    main()
