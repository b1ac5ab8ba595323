"""Coverage track of plot_amplicons.py (reference :376-412; SURVEY.md §8f rank 4, second half): the windows and mean
coverages of coral_b200.plot_amplicons.windowed_coverage against what the UNMODIFIED reference drew
(tests/golden/plot_cov.json.gz, made by tests/golden/make_plot_cov_golden.py).  CPU: the window logic with the oracle's
count_coverage as the operator.  GPU: the same through coral_region_coverage."""
import gzip
import json
import os

import numpy as np
import pytest

import test_golden_cpu as G

HERE = os.path.dirname(os.path.abspath(__file__))


def vectors():
    with gzip.open(os.path.join(HERE, "golden", "plot_cov.json.gz"), "rt") as fp:
        return json.load(fp)


def check(v, rows, max_cov):
    total_len = sum(iv[1] - iv[0] + 1 for ivs in v["intervals_from_graph"].values() for iv in ivs)
    assert len(rows) == len(v["rects"])
    for (chrom, inti, w, ws, cov), (x, width, height) in zip(rows, v["rects"]):
        assert cov == height and ws * 100.0 / total_len == width          # exact: same integer sum, same float expression
    assert [0.0, min(1.25 * max_cov, float("inf"))] == v["ylims"][-2]     # ax.set_ylim(0, min(1.25 * max_cov, cutoff)), :412


def oracle_operator(bam_path):
    from oracle import pysam_shim
    fh = pysam_shim.AlignmentFile(bam_path)

    def region_coverage(tid, start, stop):
        out = [sum(sum(nc) for nc in fh.count_coverage(fh.references[t], s, e, quality_threshold=0, read_callback="nofilter"))
               for t, s, e in zip(tid, start, stop)]
        return np.array(out, np.int64), None
    return fh.references, region_coverage


def test_fixture_shape():
    vs = vectors()
    assert len(vs) == 8 and sum(1 for v in vs if v["plot_bounds"]) == 4
    assert sum(len(v["rects"]) for v in vs) > 10000


@pytest.mark.parametrize("case", ["c1_small", "c4_small"])
def test_window_logic_with_the_oracle_operator(case, tmp_path_factory):
    from coral_b200 import plot_amplicons as pa
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    refs, op = oracle_operator(paths["bam"])
    for v in (v for v in vectors() if v["case"] == case):
        rows, max_cov = pa.windowed_coverage(op, refs, v["intervals_from_graph"], v["sorted_chrs"], v["plot_bounds"])
        check(v, rows, max_cov)
    with pytest.raises(NotImplementedError):
        pa.windowed_coverage(op, refs, v["intervals_from_graph"], v["sorted_chrs"], None, quality_threshold=20)


def test_disjoint_batches():
    from coral_b200 import plot_amplicons as pa
    regs = [(0, 0, 10, "a"), (0, 10, 20, "b"), (0, 15, 25, "c"), (1, 0, 5, "d"), (0, 18, 30, "e"), (0, 25, 40, "f")]
    batches = pa._disjoint_batches(regs)
    assert sorted(r[3] for b in batches for r in b) == list("abcdef")
    for b in batches:
        for p, q in zip(b, b[1:]):
            assert (p[0], p[1]) <= (q[0], q[1]) and (p[0] != q[0] or p[2] <= q[1])


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["c1_small", "c2_small", "c3_small", "c4_small"])
def test_windowed_coverage_on_the_device(built_lib, case, tmp_path_factory):
    from coral_b200 import _abi, plot_amplicons as pa
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    bam = _abi.BamFile(paths["bam"], 2)
    ctx = _abi.Context(0)
    ctx.set_references(bam.references)
    ctx.load_records(bam.rec, bam.rec_off, placement="copy")
    for v in (v for v in vectors() if v["case"] == case):
        rows, max_cov = pa.windowed_coverage(ctx.region_coverage, bam.references, v["intervals_from_graph"], v["sorted_chrs"],
                                             v["plot_bounds"])
        check(v, rows, max_cov)
    ctx.close()
    bam.close()
