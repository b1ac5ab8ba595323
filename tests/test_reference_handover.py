"""The graph built by the drop-in is handed to the reference's OWN downstream code (north_star: "CN estimation (cvxopt)
and Gurobi cycle decomposition stay on the reference's code and consume the GPU-built graph unchanged"; reference
infer_breakpoint_graph.py:1388-1389, cycle_decomposition.py:2066).  Build container only: needs /root/reference.
`build_graph` (coral_b200/host_steps.py) is the same code under the CPU oracle engine and the CUDA class, so the
hand-over is tested on the CPU; what fills the graph on the GPU is pinned by tests/test_gpu_pipeline.py."""
import json
import os
import subprocess
import sys

import pytest

import test_golden_cpu as G

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("CORAL_REFERENCE_SRC", "/root/reference/src")

pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree only exists in the build container")


@pytest.mark.parametrize("case", ["c1_small", "c2_small"])
def test_lr_graph_is_the_references_class_and_feeds_cycle_decomposition(case, tmp_path_factory):
    gold = G.load_golden(case)
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    out = str(tmp_path_factory.mktemp("handover") / "handover.json")
    env = dict(os.environ, PYTHONHASHSEED="0")
    subprocess.check_call([sys.executable, os.path.join(HERE, "run_refgraph.py"), paths["bam"], paths["cn_seg"],
                           paths["seed"], out], env=env)
    res = json.load(open(out))
    want = {k: v for k, v in gold["stages"].items() if k != "compute_path_constraints"}
    assert G.compare_states(res["state"], want) == []
    ck = res["checks"]
    assert ck["n_graphs"] >= 1
    assert ck["all_reference_class"] and ck["has_cn_methods"]
    assert ck["reference_writer_same_text"] and ck["sorted_twice_same_text"]
    assert ck["reached_gurobi"], "the reference's reconstruct_cycles did not get as far as building the Gurobi model"
    assert ck["path_constraints_amplicons"] == ck["n_graphs"] and ck["longest_path_constraints_filled"]
    assert all(a[0] >= 2 for a in ck["accessors"])


def test_local_container_is_used_without_the_reference(tmp_path_factory):
    """Without the reference on sys.path (the GPU box) build_graph falls back to the local container, and graph_cls
    overrides both."""
    from coral_b200 import host_steps
    from coral_b200.breakpoint_graph import BreakpointGraph
    assert "breakpoint_graph" not in sys.modules
    assert host_steps.reference_graph_class() is None
    assert host_steps.HostSteps.graph_cls is None
    assert host_steps._LocalBreakpointGraph is BreakpointGraph
