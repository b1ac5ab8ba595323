"""Known-answer vectors of the pure operators of the hot path (SURVEY.md §8c "L2"): cigar2pos*, alignment_from_satags,
the interval predicates, interval2bp, alignment2bp(_l), cluster_bp_list, bpc2bp, bp_match — inputs and the outputs the
UNMODIFIED reference produced (tests/golden/make_l2_kat.py, run here where /root/reference exists).

CPU: the oracle restatement reproduces every vector (pins the oracle).  GPU: the drop-in modules — the same callables,
backed by the CUDA library — reproduce every vector."""
import gzip
import importlib.util
import json
import os
import types

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_l2_kat", os.path.join(HERE, "golden", "make_l2_kat.py"))
KAT = importlib.util.module_from_spec(spec)
spec.loader.exec_module(KAT)


def vectors():
    with gzip.open(os.path.join(HERE, "golden", "l2_kat.json.gz"), "rt") as fp:
        return json.load(fp)


def by_fn(vs):
    out = {}
    for v in vs:
        out.setdefault(v["fn"], []).append(v)
    return out


def mismatches(mods, vs):
    bad = []
    for k, v in enumerate(vs):
        got = KAT.run_vector(mods, v)
        if got != v["out"]:
            bad.append((k, v["fn"], v["args"], got, v["out"]))
    return bad


def oracle_namespace():
    """The oracle under the reference's names: one cigar2pos for the nine per-signature functions; interval_include(_l)
    are not on the oracle's path (hsr / the drop-in only) and are skipped."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
    import coral_oracle as O
    ns = types.SimpleNamespace(**{k: getattr(O, k) for k in (
        "alignment_from_satags", "interval_overlap", "interval_overlap_l", "interval_adjacent", "interval_exclusive",
        "interval2bp", "alignment2bp_l", "alignment2bp", "cluster_bp_list", "bp_match", "bpc2bp")})
    for sig in ("SM", "MS", "SMS", "SMD", "MDS", "SMDS", "SMI", "MIS", "SMIS"):
        setattr(ns, "cigar2pos" + sig, lambda cigar, strand, rl, _f=O.cigar2pos: _f(cigar, strand, rl))
    return ns


def test_fixture_shape():
    groups = by_fn(vectors())
    for fn in ("alignment_from_satags", "alignment2bp", "alignment2bp_l", "cluster_bp_list", "bpc2bp", "bp_match",
               "interval2bp", "interval_overlap", "interval_overlap_l", "interval_include", "interval_include_l",
               "interval_adjacent", "interval_exclusive"):
        assert len(groups[fn]) >= 40, fn
    assert sum(1 for fn in groups if fn.startswith("cigar2pos")) == 9
    assert any(isinstance(v["out"], str) for v in groups["alignment_from_satags"])        # the KeyError / ValueError cases
    assert sum(1 for v in groups["alignment2bp"] + groups["alignment2bp_l"] if v["out"]) >= 50


def test_oracle_matches_reference_vectors():
    ns = oracle_namespace()
    vs = [v for v in vectors() if hasattr(ns, v["fn"])]
    assert len(vs) > 1000
    bad = mismatches((ns,), vs)
    assert bad == [], bad[:3]


HOST_ONLY = ("interval_overlap", "interval_overlap_l", "interval_include", "interval_include_l", "interval_adjacent",
             "interval_exclusive", "interval2bp", "alignment2bp_nm", "alignment2bp_nm_l")


def test_host_python_callables_of_the_dropin_match_reference_vectors():
    """The callables of coral_b200.breakpoint_utilities that are plain host Python (predicates, interval2bp and the
    nm-filtered candidate forms the reference never reaches) need no GPU."""
    from coral_b200 import breakpoint_utilities as bu
    vs = [v for v in vectors() if v["fn"] in HOST_ONLY]
    assert len(vs) > 600 and sum(1 for v in vs if "_nm" in v["fn"] and v["out"]) >= 20
    bad = mismatches((bu,), vs)
    assert bad == [], bad[:3]


@pytest.mark.gpu
def test_dropin_modules_match_reference_vectors(built_lib):
    from coral_b200 import breakpoint_utilities as bu, cigar_parsing as cp
    bad = mismatches((cp, bu), vectors())
    assert bad == [], bad[:3]
