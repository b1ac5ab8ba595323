"""GPU parity, end to end: the CUDA-backed drop-in class driven in the reference's call order reproduces the
state (and graph.txt, CN column 0 because cvxopt is absent) that the UNMODIFIED reference produced."""
import pytest

import test_golden_cpu as G

pytestmark = pytest.mark.gpu

GPU_CASES = ["c1_small", "c1_n", "c2_small", "c2_oneseed", "c3_small", "c4_small", "c5_mini"]


@pytest.mark.parametrize("case", GPU_CASES)
def test_cuda_pipeline_matches_reference(built_lib, case, tmp_path_factory):
    gold = G.load_golden(case)
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    assert data.n_records == gold["n_records"]
    got = G.run_engine("cuda", paths, tmp_path_factory)
    assert G.compare_states(got, gold["stages"]) == []
    for g_got, g_ref in zip(got["assign_cov"]["graphs"], gold["stages"]["assign_cov"]["graphs"]):
        assert g_got["graph_txt"] == g_ref["graph_txt"]          # byte-identical graph.txt


@pytest.mark.parametrize("case,kind", [("c2_small", "cuda-async"), ("c4_small", "cuda-async"), ("c3_small", "cuda-sync"),
                                       ("c5_mini", "cuda-sync"), ("c4_small", "cuda-lazy"), ("c3_small", "cuda-copy")])
def test_cuda_pipeline_placements(built_lib, case, kind, tmp_path_factory):
    """The same parity with every record placement.  "cuda" above = the default: asynchronous scan, with normal_cov read
    right after read_cns as the harness does (which joins the scan there).  "cuda-async" = the same placement with the
    reference's own access pattern: fetch and hash_alignment_to_seg run while the scan is in flight, normal_cov is
    first read when find_amplicon_intervals needs it.  "cuda-sync" = mapped with the scan joined at load, "cuda-lazy" =
    pre-scan at load and the full scan inside the first method that needs it, "cuda-copy" = records copied to HBM."""
    gold = G.load_golden(case)
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    got = G.run_engine(kind, paths, tmp_path_factory)
    assert G.compare_states(got, gold["stages"]) == []
    for g_got, g_ref in zip(got["assign_cov"]["graphs"], gold["stages"]["assign_cov"]["graphs"]):
        assert g_got["graph_txt"] == g_ref["graph_txt"]
