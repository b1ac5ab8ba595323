"""Development aid: cProfile of the host steps of one e2e pass (C2 at --scale) after the scan has been joined.

    python scripts/e2e_profile.py [--scale 1.0]
"""
import argparse
import cProfile
import os
import pstats
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    a = ap.parse_args()
    import torch
    import bench
    from sim.generate import simulate
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from coral_b200.records import RecordSet
    from coral_b200.host_steps import gc_paused
    cfg = bench.workload_config(a.scale)
    pinned = {}

    def alloc(nbytes):
        pinned["t"] = torch.empty(nbytes + 64, dtype=torch.uint8).pin_memory()
        return pinned["t"].numpy()[:nbytes]
    data = simulate(cfg, alloc=alloc)
    recs = RecordSet([c for c, _ in cfg.refs], data.rec, data.rec_off, rec_slack=64)
    cns, seed = bench.write_inputs(cfg, tempfile.mkdtemp(prefix="coral_prof_"))
    for it in range(3):
        with gc_paused():
            eng = bam_to_breakpoint_nanopore(None, seed, records=recs, placement="mapped_async")
            eng.read_cns(cns); eng.fetch(); eng.hash_alignment_to_seg(); eng._ctx.wait_scan()
            pr = cProfile.Profile()
            pr.enable()
            eng.find_amplicon_intervals(); eng.find_smalldel_breakpoints(); eng.find_breakpoints(); eng.build_graph(); eng.assign_cov()
            eng.closebam()
            pr.disable()
            del eng
    pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
    pstats.Stats(pr).sort_stats("tottime").print_stats(30)


if __name__ == "__main__":
    main()
