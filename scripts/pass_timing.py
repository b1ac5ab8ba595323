"""Development aid: host time of every call of one pipelined device pass (enqueue only) against the pass total.

    python scripts/pass_timing.py [--scale 1.0] [--overlap 1]
"""
import argparse
import os as _os
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # streams of a ctx (scan, lane) must not share a hardware queue
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--overlap", type=int, default=1)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--reserve", type=int, default=-1)
    ap.add_argument("--fork", type=int, default=1)
    ap.add_argument("--pinned", type=int, default=1)
    a = ap.parse_args()
    import bench
    from sim.generate import simulate
    from coral_b200 import gpu_ops
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from coral_b200.records import RecordSet
    cfg = bench.workload_config(a.scale)
    data = simulate(cfg)
    recs = RecordSet([c for c, _ in cfg.refs], data.rec, data.rec_off)
    cns, seed = bench.write_inputs(cfg, tempfile.mkdtemp(prefix="coral_pt_"))
    eng = bam_to_breakpoint_nanopore(None, seed, records=recs, placement="copy")
    plan = bench.build_plan(eng, cns)
    ctx = eng._ctx
    ctx.set_option("pass_overlap_scan", a.overlap)
    ctx.set_option("pass_scan_reserve", a.reserve)
    ctx.set_option("pass_fork", a.fork)
    laps = {}

    def one():
        t = [time.perf_counter()]

        def lap(k):
            t.append(time.perf_counter())
            laps.setdefault(k, []).append((t[-1] - t[-2]) * 1e3)
        ctx.pass_begin(pinned_results=bool(a.pinned)); lap("pass_begin")
        ctx.rescan(); lap("rescan")
        ctx.fetch(); lap("fetch")
        cand = ctx.candidates_l(*plan["intervals"]); lap("candidates_l")
        if cand["n"]:
            ctx.cluster_call_candidates(cand["n"], plan["cutoff"], 2000, plan["cutoff"], 100)
        lap("cluster")
        ctx.region_coverage(*plan["normal"]); lap("cov_normal")
        ctx.smalldel(*plan["intervals"]); lap("smalldel")
        ctx.region_coverage(*plan["seq_edges"]); lap("cov_seq")
        ctx.concordant(*plan["conc"]); lap("concordant")
        t_enq = time.perf_counter() - t[0]
        ctx.pass_end(); lap("pass_end")
        laps.setdefault("enqueue_total", []).append(t_enq * 1e3)
        laps.setdefault("total", []).append((time.perf_counter() - t[0]) * 1e3)
    for _ in range(5):
        one()
    laps.clear()
    ctx.reset_stage_times()
    for _ in range(a.reps):
        one()
    out = {k: round(float(np.mean(v)), 4) for k, v in laps.items()}
    out["stage_ms"] = {n: round(ctx.stage_ms(i), 4) for i, n in enumerate(["scan", "coverage", "fetch", "smalldel", "candidates", "cluster", "bpc2bp", "concordant"])}
    out["overlap"] = a.overlap
    out["reserve"] = a.reserve
    out["fork"] = a.fork
    names = ["scan", "coverage", "fetch", "smalldel", "candidates", "cluster", "bpc2bp", "concordant"]
    out["timeline_ms_from_scan_begin"] = {n: [round(ctx.stage_offset_ms(0, 0, i, 0), 4), round(ctx.stage_offset_ms(0, 0, i, 1), 4)]
                                          for i, n in enumerate(names) if ctx.stage_ms(i) >= 0}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
