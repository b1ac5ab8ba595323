"""HSR mode's breakpoint search (reference src/hsr.py) against what the UNMODIFIED reference
printed and plotted on the same synthetic BAMs (tests/golden/hsr_*.json.gz, made by tests/golden/make_hsr_golden.py).
CPU: the host logic of coral_b200/hsr.py driven by the oracle's operators.  GPU: coral_b200.hsr.locate_hsrs itself
(fetch on the device, clustering + the bpc2bp peel loop as one device call)."""
import contextlib
import gzip
import io
import json
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_hsr_golden as H  # noqa: E402

CASES = list(H.CASES)
_IN = {}


def _inputs(case, tmp_path_factory):
    if case not in _IN:
        _IN[case] = H.hsr_inputs(case, str(tmp_path_factory.mktemp(case)))
    return _IN[case]


def _golden(case):
    with gzip.open(os.path.join(HERE, "golden", case + ".json.gz"), "rt") as fp:
        return json.load(fp)


def _oracle_called(bp_list, t, cut, cut_cl):
    from oracle import coral_oracle as O
    for c in O.cluster_bp_list(bp_list, t, cut_cl):
        if len(c) >= t:
            rest = c
            while len(rest) >= t:
                bp, bpr, st, rest = O.bpc2bp(rest, cut)
                yield bp, bpr, st


@pytest.mark.parametrize("case", CASES)
def test_hsr_host_logic_with_oracle_operators(case, tmp_path_factory):
    from coral_b200 import hsr
    from oracle import coral_oracle
    from oracle.ref_runner import canon_chimeric
    data, args = _inputs(case, tmp_path_factory)
    gold = _golden(case)
    assert data.n_records == gold["n_records"]
    eng = coral_oracle.OracleHotPath(args.lr_bam)
    eng.fetch()
    chim = eng.chimeric_alignments
    if gold["fetch"] is not None:
        assert json.loads(json.dumps(canon_chimeric(chim))) == gold["fetch"]
    ecdna, ext = hsr.read_ecdna_intervals(args.cycles, args.bp_match_cutoff)
    bp_list = hsr.candidate_breakpoints(chim, ecdna)
    t = float(args.normal_cov) * 0.5
    refined, stats = hsr.refine(_oracle_called(bp_list, t, args.bp_match_cutoff, args.bp_match_cutoff_clustering), t,
                                args.bp_match_cutoff)
    pts = hsr.integration_points(refined, ext, hsr.read_cn_segments(args.cn_seg), args.normal_cov)
    lines = ["Fetched %d chimeric alignments." % len(chim),
             "Found %d breakpoints connecting ecDNA and chromosomes." % len(refined)] + [p[0] for p in pts]
    assert lines == gold["lines"]
    assert [[x, y] for _, x, y in pts] == gold["points"]
    assert len(refined) > 0 and len(stats) == len(refined)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_hsr_locate_on_gpu(built_lib, case, tmp_path_factory):
    from coral_b200 import _abi, hsr
    from oracle.ref_runner import canon_chimeric
    data, args = _inputs(case, tmp_path_factory)
    gold = _golden(case)
    if gold["fetch"] is not None:
        fh = _abi.BamFile(args.lr_bam, 4)
        read_length, chim = hsr.fetch(fh)
        assert json.loads(json.dumps(canon_chimeric(chim))) == gold["fetch"]
        assert all(name in read_length for name in chim)
        fh.close()
    for device_ingest in (False, True):        # host reader (zlib) / the .bam bytes inflated and framed on the GPU
        out = io.StringIO()
        with contextlib.redirect_stdout(out):
            res = hsr.locate_hsrs(args, device_ingest=device_ingest)
        lines = [ln for ln in out.getvalue().splitlines() if ln.startswith(("Fetched", "Found", "Breakpoint"))]
        assert lines == gold["lines"]
        assert res["points"] == gold["points"]
        assert len(res["bp_refined"]) == len(res["bp_stats"]) > 0
