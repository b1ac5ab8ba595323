"""Hand-over of the built graph to the reference's own downstream code (build container only: needs /root/reference).

Runs the pipeline (CPU oracle hot path + the shared host steps, the same `build_graph` the CUDA class runs) INSIDE the
reference's module layout — `src/` on sys.path, cvxopt / gurobipy import stubs — and checks that

  * `lr_graph[i]` are instances of the reference's `breakpoint_graph.BreakpointGraph`
    (host_steps.reference_graph_class; reference infer_breakpoint_graph.py:1388-1389 calls `compute_cn_lr` on them);
  * the reference's own writer prints the same graph.txt as the local writer;
  * the reference's accessors run on them (`infer_max_seq_multiplicity`, `nextminus` / `lastplus`, `sort_edges`);
  * the reference's `cycle_decomposition.reconstruct_cycles(args, bb)` (cycle_decomposition.py:2066) accepts the engine
    object: it calls our `compute_path_constraints`, `longest_path_dict`, walks nodes / endnodes / edge lists, and gets as
    far as constructing the Gurobi model (the stub raises there — Gurobi itself is not in this image).

    python tests/run_refgraph.py BAM CN_SEG SEED OUT.json
"""
import json
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import util  # noqa: E402


class ReachedGurobi(Exception):
    pass


def main():
    bam, cn_seg, seed, out = sys.argv[1:5]
    from oracle import ref_runner
    ref_runner.install_shims()                      # reference src/ first on sys.path; pysam / intervaltree / cvxopt stand-ins
    gp = types.ModuleType("gurobipy")

    def _model(*a, **k):
        raise ReachedGurobi()
    gp.Model = _model
    gp.GRB = types.SimpleNamespace(INFEASIBLE=3, BINARY='B', CONTINUOUS='C', INTEGER='I')
    sys.modules["gurobipy"] = gp
    import breakpoint_graph as ref_bg               # the reference's module
    from coral_b200 import host_steps
    assert host_steps.reference_graph_class() is ref_bg.BreakpointGraph
    eng = util.oracle_engine(bam, seed)
    st = util.run_pipeline(eng, cn_seg, stop_after="assign_cov")
    checks = {"n_graphs": len(eng.lr_graph)}
    checks["all_reference_class"] = all(type(g) is ref_bg.BreakpointGraph for g in eng.lr_graph)
    checks["has_cn_methods"] = all(hasattr(g, "compute_cn_lr") and hasattr(g, "infer_max_seq_multiplicity") and
                                   hasattr(g, "merge_edges") and hasattr(g, "del_discordant_edges") for g in eng.lr_graph)
    ref_txt = [ref_runner.graph_text(g, ref_bg.output_breakpoint_graph_lr) for g in eng.lr_graph]
    checks["reference_writer_same_text"] = ref_txt == [g["graph_txt"] for g in st["assign_cov"]["graphs"]]
    acc = []
    for g in eng.lr_graph:
        se = g.sequence_edges[0]
        acc.append([g.infer_max_seq_multiplicity(), g.nextminus(se[0], se[1]), g.lastplus(se[0], se[2]),
                    g.infer_discordant_edge_multiplicities() if len(g.discordant_edges) else []])
        g.sort_edges()                              # idempotent on an already sorted graph
    checks["accessors"] = json.loads(json.dumps(acc))
    checks["sorted_twice_same_text"] = ref_txt == [ref_runner.graph_text(g, ref_bg.output_breakpoint_graph_lr) for g in eng.lr_graph]
    import cycle_decomposition as cd                # reference module, under the gurobipy stub
    args = types.SimpleNamespace(cycle_decomp_alpha=None, postprocess_greedy_sol=False, cycle_decomp_threads=None,
                                 cycle_decomp_time_limit=None, output_prefix=out + ".cd", output_all_path_constraints=False)
    try:
        cd.reconstruct_cycles(args, eng)
        checks["reached_gurobi"] = False
    except ReachedGurobi:
        checks["reached_gurobi"] = True
    checks["path_constraints_amplicons"] = len(eng.path_constraints)
    checks["longest_path_constraints_filled"] = all(len(v) == 3 for v in eng.longest_path_constraints.values())
    with open(out, "w") as fp:
        json.dump({"state": st, "checks": checks}, fp)


if __name__ == "__main__":
    main()
