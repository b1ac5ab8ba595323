"""Run one full pipeline (oracle or CUDA product) and dump the canonical per-stage state as JSON.

Executed by the tests in a subprocess with PYTHONHASHSEED=0: the reference iterates over a *set* of read
names in find_interval_i (infer_breakpoint_graph.py:432), so the order of new_bp_list depends on the string
hash seed; the golden fixtures were produced with PYTHONHASHSEED=0 as well.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import util  # noqa: E402


def main():
    kind, bam, cn_seg, seed, out = sys.argv[1:6]
    if kind == "oracle":
        eng = util.oracle_engine(bam, seed)
    else:
        from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
        eng = bam_to_breakpoint_nanopore(bam, seed)
    st = util.run_pipeline(eng, cn_seg)
    with open(out, "w") as fp:
        json.dump(st, fp)


if __name__ == "__main__":
    main()
