"""The peer-memory exchange primitives (csrc/xchg.cu) on their own: process ranks on cuda:0 map each other's windows
through CUDA IPC; the kernels (push with system-scope release, bounded wait, sum over slots) are the ones that run
between GPUs (tests/run_rank.py, mode xchg)."""
import json

import pytest

import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [1, 2, 4])
def test_allgather_allreduce_limits_and_timeout(built_lib, world, tmp_path):
    out = str(tmp_path / "res.json")
    util.spawn_ranks(["xchg", out], world, timeout=600)
    res = json.load(open(out))
    assert len(res) == world
    for r, info in enumerate(res):
        assert info["bad"] == [], (r, info)
        assert info["refused"] == [True, True]
        assert info["timeouts"] == 0
    if world > 1:       # rank 0 waited alone for a peer that never came: an error after the timeout, not a hang
        assert res[0]["lonely"].startswith("error") and "did not deliver" in res[0]["lonely"]
        assert res[0]["lonely_timeouts"] >= 1
