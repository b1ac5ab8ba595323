"""GPU parity of the sharded path: the file cut into 2, 3 and 5 contiguous shards, one engine per
shard on the same device exchanging through the in-process communicator, must reproduce the UNMODIFIED reference's
state after every stage and a byte-identical graph.txt — i.e. exactly the 1-GPU result."""
import pytest

import test_golden_cpu as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case,world", [("c2_small", 2), ("c4_small", 3), ("c5_mini", 2), ("c5_mini", 5), ("c3_small", 2),
                                        ("c1_n", 1)])
def test_sharded_pipeline_matches_reference(built_lib, case, world, tmp_path_factory):
    gold = G.load_golden(case)
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    got = G.run_engine("cuda-sharded:%d" % world, paths, tmp_path_factory)
    assert G.compare_states(got, gold["stages"]) == []
    for g_got, g_ref in zip(got["assign_cov"]["graphs"], gold["stages"]["assign_cov"]["graphs"]):
        assert g_got["graph_txt"] == g_ref["graph_txt"]
