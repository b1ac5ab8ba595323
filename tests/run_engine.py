"""Run one full pipeline (oracle or CUDA product) and dump the canonical per-stage state as JSON.

Executed by the tests in a subprocess with PYTHONHASHSEED=0: the reference iterates over a *set* of read
names in find_interval_i (infer_breakpoint_graph.py:432), so the order of new_bp_list depends on the string
hash seed; the golden fixtures were produced with PYTHONHASHSEED=0 as well.
"""
import faulthandler
import json
import os
import sys

faulthandler.enable()

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import util  # noqa: E402


def sharded_states(world, bam, cn_seg, seed, peer=False):
    """`world` shards of the file as `world` threads on cuda:0, exchanging through ThreadComm (the collective's test
    double; the kernels are the real ones).  Every rank must end in the same state; rank 0's is returned."""
    import threading
    from coral_b200 import _abi, sharded
    fh = _abi.BamFile(bam, 4)
    comms = sharded.ThreadComm.make(world)
    out, err = [None] * world, []

    def body(r):
        try:
            # peer: the exchanges as kernels over "peer" memory (thread ranks share the device: plain addresses)
            x = sharded.make_xchg(comms[r], 0, timeout_s=60) if peer else None
            eng = sharded.make_sharded_engine(fh, seed, comms[r], device=0, xchg=x)
            out[r] = util.jsonify(util.run_pipeline(eng, cn_seg))
            if x is not None:
                sharded.close_xchg(comms[r], x)
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
    for r in range(1, world):
        a, b = dict(out[0]), dict(out[r])
        a["fetch"] = dict(a["fetch"], nm_stats=None); b["fetch"] = dict(b["fetch"], nm_stats=None)
        assert a == b, "rank %d ended in a different state than rank 0" % r
    return out[0]


def main():
    kind, bam, cn_seg, seed, out = sys.argv[1:6]
    if kind.startswith("cuda-sharded"):
        st = sharded_states(int(kind.split(":")[1]), bam, cn_seg, seed, peer=kind.startswith("cuda-sharded-peer:"))
    else:
        if kind == "oracle":
            eng = util.oracle_engine(bam, seed)
        else:
            from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
            eng = bam_to_breakpoint_nanopore(bam, seed, placement={"cuda": "auto", "cuda-async": "mapped_async",
                                                                    "cuda-sync": "mapped", "cuda-lazy": "mapped_lazy",
                                                                    "cuda-copy": "copy", "cuda-ingest": "device_ingest"}[kind])
            assert eng._scan_async == (kind in ("cuda", "cuda-async", "cuda-lazy"))
        st = util.run_pipeline(eng, cn_seg, defer_cov=(kind in ("cuda-async", "cuda-lazy")))
    with open(out, "w") as fp:
        json.dump(st, fp)


if __name__ == "__main__":
    main()
