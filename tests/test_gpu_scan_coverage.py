"""GPU parity: record-scan table and region coverage vs the oracle (through the C ABI)."""
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


@pytest.mark.parametrize("name,kw", [("C1", dict(scale=0.05)), ("C3", dict(scale=0.02, n_del=40))])
def test_scan_table_matches_oracle_index(built_lib, dataset, name, kw):
    from oracle import pysam_shim
    ds = dataset(name, **kw)
    bam, ctx = _ctx_for(ds)
    ora = pysam_shim.AlignmentFile(ds["paths"]["bam"])
    t = ctx.record_table()
    assert np.array_equal(t["tid"], ora._tid)
    assert np.array_equal(t["pos"], ora._pos)
    assert np.array_equal(t["endpos"], ora._end)
    assert np.array_equal(t["flag"], ora._flag)
    assert np.array_equal(t["mapq"], ora._mapq)
    assert np.array_equal(t["l_seq"], ora._lseq)
    assert np.array_equal(t["n_cigar"], ora._ncig)
    assert np.array_equal(t["infer_len"], ora._infer)
    assert np.array_equal(t["aligned_bases"], ora._mb)
    assert ctx.launch_count >= 1
    ctx.close(); bam.close()


def _regions_for(ds, rng, n=40):
    tids, starts, stops = [], [], []
    for ti, (c, s, e) in enumerate(ds["cfg"].background):
        tid = [r[0] for r in ds["cfg"].refs].index(c)
        cuts = np.sort(rng.choice(np.arange(s, e), size=2 * n, replace=False))
        for k in range(n):
            tids.append(tid); starts.append(int(cuts[2 * k])); stops.append(int(cuts[2 * k + 1]))
    o = np.lexsort((starts, tids))
    return np.array(tids)[o], np.array(starts)[o], np.array(stops)[o]


@pytest.mark.parametrize("name,kw", [("C1", dict(scale=0.05)), ("C3", dict(scale=0.02, n_del=40)),
                                     ("C1", dict(scale=0.03, seed=7))])
def test_region_coverage_matches_oracle(built_lib, dataset, name, kw):
    from oracle import pysam_shim
    ds = dataset(name, **kw)
    bam, ctx = _ctx_for(ds)
    ora = pysam_shim.AlignmentFile(ds["paths"]["bam"])
    rng = np.random.default_rng(3)
    tid, start, stop = _regions_for(ds, rng)
    bs, rc = ctx.region_coverage(tid, start, stop)
    for i in range(len(tid)):
        c = ora.references[tid[i]]
        want = sum(sum(a) for a in ora.count_coverage(c, int(start[i]), int(stop[i]), quality_threshold=0,
                                                      read_callback='nofilter'))
        cnt = len([r for r in ora.fetch(c, int(start[i]), int(stop[i])) if r.infer_read_length()])
        assert bs[i] == want, (i, tid[i], start[i], stop[i])
        assert rc[i] == cnt
    ctx.close(); bam.close()
