"""GPU parity: coral_fetch vs the oracle's fetch (chimeric tuples, read order, nm_stats)."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def gpu_chimeric(ctx, bam, f):
    """Materialise the reference's dict name -> (qint, rint, qual, nm) from the device table."""
    refs = bam.references
    out = {}
    ap = f["aln_ptr"]
    for r in range(f["n_reads"]):
        st = int(f["read_status"][r])
        name = bam.query_name(int(f["read_first_rec"][r]))
        if st == 1:
            continue
        if st == 2:
            out[name] = ([], [], [])
            continue
        assert st == 0, (name, st)
        a, b = int(ap[r]), int(ap[r + 1])
        qint = [[int(f["qs"][k]), int(f["qe"][k])] for k in range(a, b)]
        rint = [[refs[f["tid"][k]], int(f["rstart"][k]), int(f["rend"][k]), '-' if f["strand"][k] else '+'] for k in range(a, b)]
        qual = [int(f["mapq"][k]) for k in range(a, b)]
        nm = [float(f["nm"][k]) / (qint[k - a][1] - qint[k - a][0]) for k in range(a, b)]
        out[name] = (qint, rint, qual, nm)
    return out


@pytest.mark.parametrize("name,kw", [("C1", dict(scale=0.1)), ("C2", dict(scale=0.02)), ("C5", dict(scale=0.02, chroms=['chr1', 'chr2', 'chr3']))])
def test_fetch_matches_oracle(built_lib, dataset, name, kw):
    from coral_b200 import _abi
    from oracle.coral_oracle import OracleHotPath
    ds = dataset(name, **kw)
    bam = _abi.BamFile(ds["paths"]["bam"], 4)
    ctx = _abi.Context(0)
    ctx.set_references(bam.references)
    ctx.load_records(bam.rec, bam.rec_off)
    f = ctx.fetch()
    got = gpu_chimeric(ctx, bam, f)
    ora = OracleHotPath(ds["paths"]["bam"])
    ora.fetch()
    assert list(got.keys()) == list(ora.chimeric_alignments.keys())
    for k in got:
        assert got[k] == ora.chimeric_alignments[k], k
    assert len(got) > 0
    # nm_stats: order-dependent f64 sums; tolerance 1e-9 relative
    n = f["nm_n"]
    mean = f["nm_sum"] / n
    std = math.sqrt(f["nm_sumsq"] / n - mean ** 2)
    assert n == ora.nm_stats[2]
    assert abs(mean - ora.nm_stats[0]) <= 1e-9 * abs(ora.nm_stats[0])
    assert abs(std - ora.nm_stats[1]) <= 1e-9 * abs(ora.nm_stats[1])
    # the 128-bit name sort (what fetch falls back to when two names share the 40 bits it sorts on first): same table
    ctx.set_option("full_name_sort", 1)
    f2 = ctx.fetch()
    assert f2["n_reads"] == f["n_reads"]
    for k in ("read_first_rec", "read_len", "read_status", "aln_ptr", "qs", "qe", "tid", "rstart", "rend", "strand", "mapq", "nm"):
        assert np.array_equal(f2[k], f[k]), k
    with pytest.raises(RuntimeError):
        ctx.set_option("no_such_option", 1)
    ctx.close(); bam.close()
