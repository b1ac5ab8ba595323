"""The sharded device pass as ONE stream of kernels per rank (pipelined pass on both contexts, exchanges as kernels over
peer memory, distributed fetch + merge): process ranks on cuda:0 (windows mapped through CUDA IPC) must deliver, in a
recording pass and in replayed passes, exactly the tables of the 1-GPU device pass over the whole file."""
import json

import pytest

import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case,world", [("c2_small", 2), ("c5_mini", 3), ("c3_small", 2), ("c4_small", 4)])
def test_pipelined_sharded_pass_equals_one_gpu_pass(built_lib, case, world, tmp_path):
    out = str(tmp_path / "res.json")
    util.spawn_ranks(["pass", case, out], world)
    res = json.load(open(out))
    for r, info in enumerate(res["ranks"]):
        assert info["bad"] == [[]] * 7, (r, info)
        assert info["timeouts"] == 0
        assert info["n_reads"] == res["want_reads"] and info["n_cand"] == res["want_cand"]
        # pass 0 call by call, pass 1 records, passes 2 and 3 replay (no intermediate synchronisation); passes 4-6 the same
        # with the exchange chain running next to the scan (another call sequence: records once, then replays)
        assert info["replayed"] == [False, False, True, True, False, True, True], (r, info)
    if case == "c3_small":
        assert res["ranks"][0]["smalldel_rows"] > 0
