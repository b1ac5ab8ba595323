"""The allocation-free entry point (coral_region_coverage_static, include/coral_b200.h: no counterpart in the
reference) against the existing coral_region_coverage, through the C ABI.  Integer sums: the bar is bit-exact.
Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ctx_for(ds):
    from coral_b200 import _abi
    bam = _abi.BamFile(ds["paths"]["bam"], 4)
    ctx = _abi.Context(0)
    ctx.set_references(bam.references)
    ctx.load_records(bam.rec, bam.rec_off)
    return bam, ctx


def _regions(ds, rng, n):
    tids, starts, stops = [], [], []
    for c, s, e in ds["cfg"].background:
        tid = [r[0] for r in ds["cfg"].refs].index(c)
        cuts = np.sort(rng.choice(np.arange(s, e), size=2 * n, replace=False))
        for k in range(n):
            tids.append(tid); starts.append(int(cuts[2 * k])); stops.append(int(cuts[2 * k + 1]))
    o = np.lexsort((starts, tids))
    return np.array(tids, np.int32)[o], np.array(starts, np.int32)[o], np.array(stops, np.int32)[o]


def _n_base_dataset(tmp_path):
    """C1 with 0.1 % N bases: the exact re-walk (the slow list in the static array) runs for most records."""
    from sim import configs
    from sim.generate import simulate, write_dataset
    cfg = configs.CONFIGS["C1"](scale=0.02, seed=99)
    cfg.p_n = 0.001
    data = simulate(cfg)
    return {"data": data, "paths": write_dataset(data, str(tmp_path)), "cfg": cfg}


def test_static_scratch_variant_matches_region_coverage_twice_in_a_row(built_lib, dataset, tmp_path):
    for ds in (dataset("C1", scale=0.05), _n_base_dataset(tmp_path)):
        bam, ctx = _ctx_for(ds)
        rng = np.random.default_rng(11)
        first = _regions(ds, rng, 40)
        second = _regions(ds, rng, 7)          # fewer, different regions: stale sums of call 1 must not leak into call 2
        for tid, start, stop in (first, second):
            want_bs, want_rc = ctx.region_coverage(tid, start, stop)
            got_bs, got_rc = ctx.region_coverage_static(tid, start, stop)
            assert np.array_equal(got_bs, want_bs)
            assert np.array_equal(got_rc, want_rc)
            assert want_bs.sum() > 0 and want_rc.sum() > 0
        ctx.close(); bam.close()


def test_static_scratch_variant_on_a_second_ctx_and_limits(built_lib, dataset):
    from coral_b200 import _abi
    ds = dataset("C3", scale=0.02, n_del=40)
    bam, ctx = _ctx_for(ds)
    rng = np.random.default_rng(5)
    tid, start, stop = _regions(ds, rng, 25)
    want = ctx.region_coverage(tid, start, stop)
    for _ in range(2):
        got = ctx.region_coverage_static(tid, start, stop)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    # beyond the stated maximum: a clear error, nothing run
    n = 4096 + 1
    big = (np.zeros(n, np.int32), np.arange(n, dtype=np.int32) * 10, np.arange(n, dtype=np.int32) * 10 + 5)
    with pytest.raises(_abi.CoralError) as ei:
        ctx.region_coverage_static(*big)
    assert "exceed the static scratch" in str(ei.value)
    # exactly the maximum is accepted and equals the allocating entry point
    ok = tuple(a[:4096] for a in big)
    a, b = ctx.region_coverage_static(*ok), ctx.region_coverage(*ok)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    # zero regions
    e = np.zeros(0, np.int32)
    assert ctx.region_coverage_static(e, e, e)[0].shape == (0,)
    ctx.close(); bam.close()


def test_symbol_is_exported(built_lib):
    L = ctypes.CDLL(built_lib)
    assert hasattr(L, "coral_region_coverage_static")
