"""One pass of the rows of SURVEY.md §8f that run on the device, for an ncu launch list:
hash_alignment_to_seg lookups (csrc/segments.cu), compute_path_constraints (csrc/paths.cu) and the HSR breakpoint
search (cluster + peel call), on a C4-shaped sample (200 junctions, so that subpath constraints exist).

    python scripts/next_rows_pass.py [--scale 0.05]      # prints what it did; under ncu: --profile-from-start off
"""
import argparse
import json
import os
import sys
import tempfile
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=0.05)
    a = ap.parse_args()
    import torch
    from sim import configs
    from sim.generate import simulate, write_dataset
    from coral_b200 import _abi, hsr
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    cfg = configs.c4_complex(scale=a.scale)
    data = simulate(cfg)
    d = tempfile.mkdtemp(prefix="coral_next_")
    paths = write_dataset(data, d)
    eng = bam_to_breakpoint_nanopore(paths["bam"], paths["seed"])
    for fn, args in ((eng.read_cns, (paths["cn_seg"],)), (eng.fetch, ())):
        fn(*args)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    t = {}
    t0 = time.perf_counter(); eng.hash_alignment_to_seg(); t["hash_alignment_to_seg_ms"] = (time.perf_counter() - t0) * 1e3
    torch.cuda.cudart().cudaProfilerStop()
    for fn in (eng.find_amplicon_intervals, eng.find_smalldel_breakpoints, eng.find_breakpoints, eng.build_graph, eng.assign_cov):
        fn()
    torch.cuda.cudart().cudaProfilerStart()
    t0 = time.perf_counter(); eng.compute_path_constraints(); t["compute_path_constraints_ms"] = (time.perf_counter() - t0) * 1e3
    t["paths_kernel_ms"] = eng._ctx.stage_ms(8)
    bed = os.path.join(d, "ecdna.bed")
    with open(bed, "w") as fp:
        fp.write("chr1\t100000000\t102000000\nchr12\t60000000\t64000000\n")
    args = types.SimpleNamespace(lr_bam=paths["bam"], cycles=bed, cn_seg=paths["cn_seg"], normal_cov=2.0, bp_match_cutoff=100,
                                 bp_match_cutoff_clustering=2000, output_prefix=os.path.join(d, "out"))
    import contextlib
    import io
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        res = hsr.locate_hsrs(args)
    t["hsr_locate_ms"] = (time.perf_counter() - t0) * 1e3
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    t.update(records=data.n_records, chimeric_reads=len(eng.chimeric_alignments),
             sequence_edges=sum(len(g.sequence_edges) for g in eng.lr_graph),
             subpath_constraints=sum(len(v[0]) for v in eng.path_constraints.values()),
             hsr_breakpoints=len(res["bp_refined"]))
    print(json.dumps(t))


if __name__ == "__main__":
    main()
