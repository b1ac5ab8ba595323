"""CPU test: the drop-in path_constraints module and the host half of compute_path_constraints reproduce what the
UNMODIFIED reference returned on 160 seeded random states (tests/golden/path_kat.json.gz, made by
tests/golden/make_path_kat.py): every read class of infer_breakpoint_graph.py:1086-1279 (one breakpoint, several, one /
several in-CIGAR deletions, both kinds on one read), the discard rules, and the exceptions the reference raises."""
import collections
import gzip
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_path_kat as K  # noqa: E402


def _states():
    with gzip.open(K.OUT, "rt") as fp:
        return json.load(fp)


def _engine(g, chim, lia):
    from coral_b200.host_steps import HostSteps

    class Eng(HostSteps):
        def _concordant_read_runs(self, amplicon_idx, intervals, concordant_reads):
            return []
    e = Eng()
    e._init_state()
    e.lr_graph = [g]
    e.chimeric_alignments, e.large_indel_alignments = chim, lia
    return e


def test_path_functions_match_reference():
    from coral_b200 import path_constraints as pc
    from coral_b200.breakpoint_graph import BreakpointGraph
    seen = collections.Counter()
    for st in _states():
        g, _, _ = K.build_state(st, BreakpointGraph)
        got = K.run_calls(st, g, pc)
        for call, a, b in zip(st["calls"], got, st["call_results"]):
            assert a == b, (st["seed"], call)
            seen[(call[0], "raise" if isinstance(b, str) else "path" if b else "empty")] += 1
    for fn in ("alignment_to_path", "chimeric_alignment_to_path_l", "chimeric_alignment_to_path_r",
               "traverse_through_sequence_edge"):
        assert seen[(fn, "path")] > 50, seen


def test_compute_path_constraints_matches_reference():
    from coral_b200 import path_steps
    from coral_b200.breakpoint_graph import BreakpointGraph
    from oracle.ref_runner import canon_path_constraints
    classes = collections.Counter()
    orig = path_steps.PathConstraintSteps._paths_of_bp_read

    def counted(self, g, rn, split, sdel):
        k = ("one" if len(split) == 1 and not sdel else "many" if len(split) > 1 and not sdel else
             "del1" if not split and len(sdel) == 1 else "deln" if not split else "mixed")
        out = orig(self, g, rn, split, sdel)
        classes[k] += 1
        classes[k + "_kept"] += sum(1 for p in out if len(p) > 5 and path_steps.pc.valid_path(g, p))
        return out
    path_steps.PathConstraintSteps._paths_of_bp_read = counted
    try:
        n_ok = n_raise = n_paths = 0
        for st in _states():
            g, chim, lia = K.build_state(st, BreakpointGraph)
            eng = _engine(g, chim, lia)
            want = st["path_constraints"]
            try:
                eng.compute_path_constraints()
                got = json.loads(json.dumps(canon_path_constraints(eng.path_constraints)))
            except Exception as ex:      # noqa: BLE001
                got = "raise " + type(ex).__name__
            assert got == want, st["seed"]
            if isinstance(want, str):
                n_raise += 1
            else:
                n_ok += 1
                n_paths += len(want[0])
    finally:
        path_steps.PathConstraintSteps._paths_of_bp_read = orig
    assert n_ok > 100 and n_raise > 0 and n_paths > 100
    for k in ("one", "many", "del1", "deln", "mixed"):       # every read class ran and kept at least one path
        assert classes[k] > 20 and classes[k + "_kept"] > 0, classes


def test_longest_path_dict_matches_reference_vectors():
    """path_constraints.longest_path_dict (the hand-over to cycle decomposition) against vectors of the unmodified
    reference: golden pipeline constraints + seeded walks with sub-walks, duplicates, repeated edges, empty paths."""
    import gzip
    import importlib.util
    import json
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    spec = importlib.util.spec_from_file_location("make_lpd_kat", os.path.join(here, "golden", "make_lpd_kat.py"))
    kat = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(kat)
    from coral_b200 import path_constraints as pc
    with gzip.open(os.path.join(here, "golden", "lpd_kat.json.gz"), "rt") as fp:
        vectors = json.load(fp)
    assert len(vectors) >= 60 and sum(len(v["paths"]) - len(v["out"]["ids"]) for v in vectors) > 100
    for k, v in enumerate(vectors):
        assert kat.run_vector(pc.longest_path_dict, v) == v["out"], k
