"""Parity at BASELINE.json's NAMED sizes.

(1) Against the UNMODIFIED reference: tests/golden/{full_c1,full_c2,full_c3,full_c1n,sub_c4,sub_c5}.json.gz hold compact
digests (tests/golden/full_digest.py) of the state the reference itself produced on C1 / C2 / C3 at scale 1.0 and on
stated sub-samples of C4 / C5 (tests/golden/make_full_digests.py: minutes per case in the build container — C2 x1.0 is
75 s of BAM inflate + 74 s of reference Python).  The CUDA class must reproduce every digest: breakpoints, read-support
sets, the chimeric table, normal_cov, per-edge counts, byte-identical graph.txt, path constraints.
(2) Size-independent properties on C2 x1.0 (13 GB of decoded BAM): sums against the generator's own counts, additivity
of coverage, idempotence of the scan, equality of the record placements, invariants of the chimeric table."""
import gzip
import json
import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import full_digest as FD  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2_full():
    from sim import configs
    from sim.generate import simulate
    cfg = configs.CONFIGS["C2"](scale=1.0)
    return cfg, simulate(cfg)


def _load(cfg, data, placement):
    from coral_b200 import _abi
    c = _abi.Context(0)
    c.set_references([r[0] for r in cfg.refs])
    if placement.startswith("mapped"):
        nb = int(data.rec_off[-1])
        buf = np.empty(nb + 64, np.uint8)
        buf[:nb] = data.rec[:nb]
        buf[nb:] = 0
        c.load_records(buf[:nb], data.rec_off, placement=placement, slack=64)
    else:
        c.load_records(data.rec, data.rec_off)
    return c


def _digest_case(case, tmp_path_factory, *extra):
    gold_path = os.path.join(HERE, "golden", case + ".json.gz")
    if not os.path.exists(gold_path):
        pytest.skip("no committed digest for " + case)
    with gzip.open(gold_path, "rt") as fp:
        gold = json.load(fp)
    out = str(tmp_path_factory.mktemp("fd_" + case) / "digest.json")
    env = dict(os.environ, PYTHONHASHSEED="0")
    subprocess.check_call([sys.executable, os.path.join(HERE, "run_full_digest.py"), case, out] + list(extra), env=env)
    got = json.load(open(out))
    assert got["n_records"] == gold["n_records"] and got["decoded_bytes"] == gold["decoded_bytes"]
    assert FD.compare_digests(got["stages"], gold["stages"]) == []
    for a, b in zip(got["stages"]["assign_cov"]["graphs"], gold["stages"]["assign_cov"]["graphs"]):
        assert a["graph_txt"] == b["graph_txt"]                 # byte-identical graph.txt (CN column 0: no cvxopt)
    return got


@pytest.mark.parametrize("case", ["full_c1", "full_c2", "full_c3", "full_c1n", "sub_c4", "sub_c5"])
def test_named_size_matches_the_unmodified_reference(built_lib, case, tmp_path_factory):
    _digest_case(case, tmp_path_factory)


@pytest.mark.parametrize("case,world", [("full_c2", 3), ("sub_c4", 4)])
def test_named_size_sharded_matches_the_unmodified_reference(built_lib, case, world, tmp_path_factory):
    """The same digests from `world` contiguous genomic-bin shards of the file (thread ranks on one GPU): every junction
    of C4 has its two sides on different shards, so the chimeric exchange carries cross-shard reads."""
    _digest_case(case, tmp_path_factory, "auto", "--world", str(world))


def test_full_size_properties(built_lib, c2_full):
    cfg, data = c2_full
    names = [r[0] for r in cfg.refs]
    c = _load(cfg, data, "copy")
    t = c.record_table()
    assert c.n_records == data.n_records == len(t["tid"])
    # against the generator's plan: every kept piece is one record with that placement and reference span
    kept = data.pieces[data.pieces["dropped"] == 0]
    trip = lambda a, b, c_: np.sort((a.astype(np.int64) << 48) | (b.astype(np.int64) << 20) | c_.astype(np.int64))    # noqa: E731
    assert np.array_equal(trip(t["tid"], t["pos"], t["endpos"] - t["pos"]), trip(kept["tid"], kept["ref_start"], kept["ref_len"]))
    # CIGAR sums of a random sample of records, recomputed from the record bytes
    off = np.asarray(data.rec_off, dtype=np.int64)
    for i in np.random.default_rng(1).choice(data.n_records, 300, replace=False).tolist():
        o = int(off[i])
        ncig = int(data.rec[o + 16]) | int(data.rec[o + 17]) << 8
        cg = np.frombuffer(data.rec[o + 36 + int(data.rec[o + 12]):][:4 * ncig].tobytes(), "<u4")
        op, ln = cg & 15, (cg >> 4).astype(np.int64)
        assert int(t["aligned_bases"][i]) == int(ln[(op == 0) | (op == 7) | (op == 8)].sum())
        assert int(t["infer_len"][i]) == int(ln[np.isin(op, (0, 1, 4, 5, 7, 8))].sum())
        assert int(t["endpos"][i] - t["pos"][i]) == int(ln[np.isin(op, (0, 2, 3, 7, 8))].sum())
        assert int(t["n_cigar"][i]) == ncig
    assert int(t["non_acgt"].sum()) == 0
    assert bool((t["endpos"] > t["pos"]).all()) and bool((t["infer_len"] >= t["l_seq"]).all())
    # file order = coordinate order
    key = t["tid"].astype(np.int64) << 32 | t["pos"].astype(np.int64)
    assert bool((np.diff(key) >= 0).all())
    # scan is idempotent
    c.rescan()
    t2 = c.record_table()
    for k in t:
        assert np.array_equal(t[k], t2[k]), k
    # coverage: a region covering a whole contig's records sums their aligned bases; splitting it anywhere keeps the sum
    rng = np.random.default_rng(5)
    for (chrom, lo, hi) in cfg.background:
        tid = names.index(chrom)
        on = t["tid"] == tid
        a, b = int(t["pos"][on].min()), int(t["endpos"][on].max())
        whole, cnt = c.region_coverage([tid], [a], [b])
        assert int(whole[0]) == int(t["aligned_bases"][on].sum()) and int(cnt[0]) == int(on.sum())
        cuts = np.unique(np.concatenate([[a, b], rng.integers(a + 1, b, 40)]))
        parts, pcnt = c.region_coverage([tid] * (len(cuts) - 1), cuts[:-1], cuts[1:])
        assert int(parts.sum()) == int(whole[0])
        assert int(pcnt.sum()) >= int(cnt[0])                   # a record straddling a cut is counted on both sides
        # record counts against the interval rule of fetch(): pos < stop and endpos > start
        for k in range(0, len(cuts) - 1, 7):
            want = int((on & (t["pos"] < cuts[k + 1]) & (t["endpos"] > cuts[k])).sum())
            assert int(pcnt[k]) == want
    # chimeric table: reads = distinct names among SA-bearing records; alignments sorted by (qs, qe) inside a read
    keys = c.record_keys()
    sa = keys["sa_off"] != 0
    pairs = np.stack([keys["name_lo"][sa], keys["name_hi"][sa]], 1)
    f = c.fetch()
    assert f["n_reads"] == len(np.unique(pairs, axis=0))
    assert set(np.unique(f["read_status"]).tolist()) <= {0, 1, 2}
    ap = f["aln_ptr"].astype(np.int64)
    assert int(ap[-1]) == f["n_aln"] == len(f["qs"])
    inside = np.ones(f["n_aln"], bool)
    inside[ap[:-1][ap[:-1] < f["n_aln"]]] = False              # first alignment of each read
    q = f["qs"].astype(np.int64) << 32 | f["qe"].astype(np.int64)
    assert bool((np.diff(q)[inside[1:]] >= 0).all())
    assert bool((f["qe"] > f["qs"]).all())
    # read_first_rec is a record of the read, and no two reads share a name
    fr = f["read_first_rec"].astype(np.int64)
    assert bool((fr < c.n_records).all())
    assert len(np.unique(np.stack([keys["name_lo"][fr], keys["name_hi"][fr]], 1), axis=0)) == f["n_reads"]
    # both placements see the same bytes
    m = _load(cfg, data, "mapped")
    tm = m.record_table()
    for k in t:
        assert np.array_equal(t[k], tm[k]), k
    fm = m.fetch()
    for k in ("read_first_rec", "read_len", "read_status", "aln_ptr", "qs", "qe", "tid", "rstart", "rend", "strand", "mapq", "nm"):
        assert np.array_equal(f[k], fm[k]), k
    assert (fm["nm_sum"], fm["nm_sumsq"], fm["nm_n"]) == (f["nm_sum"], f["nm_sumsq"], f["nm_n"])
    m.close()
    # asynchronous placement: fetch right after the pre-scan, while the full scan is still pulling 5.5 GB over the link
    m = _load(cfg, data, "mapped_async")
    fa = m.fetch()
    for k in ("read_first_rec", "read_len", "read_status", "aln_ptr", "qs", "qe", "tid", "rstart", "rend", "strand", "mapq", "nm"):
        assert np.array_equal(f[k], fa[k]), k
    assert (fa["nm_sum"], fa["nm_sumsq"], fa["nm_n"]) == (f["nm_sum"], f["nm_sumsq"], f["nm_n"])
    ta = m.record_table()                      # joins the scan
    for k in t:
        assert np.array_equal(t[k], ta[k]), k
    ka = m.record_keys()
    for k in keys:
        assert np.array_equal(keys[k], ka[k]), k
    m.close(); c.close()


def test_full_size_sharded_equals_single(built_lib, c2_full, tmp_path_factory):
    """The full C2 file cut into 3 shards (thread ranks on one device, real kernels and messages) ends every stage in
    the state of the 1-GPU run, graph.txt included."""
    import threading
    import bench
    import util
    from coral_b200 import sharded
    from coral_b200.infer_breakpoint_graph import bam_to_breakpoint_nanopore
    from coral_b200.records import RecordSet
    cfg, data = c2_full
    cns, seed = bench.write_inputs(cfg, str(tmp_path_factory.mktemp("c2_full_inputs")))
    rs = RecordSet([r[0] for r in cfg.refs], data.rec, data.rec_off)
    one = util.jsonify(util.run_pipeline(bam_to_breakpoint_nanopore(None, seed, records=rs, placement="copy"), cns))
    assert len(one["find_breakpoints"]["new_bp_list"]) == 12          # C2 is built from 12 junctions
    world = 3
    comms = sharded.ThreadComm.make(world)
    out, err = [None] * world, []

    def body(r):
        try:
            eng = sharded.make_sharded_engine(rs, seed, comms[r], device=0, placement="copy")
            out[r] = util.jsonify(util.run_pipeline(eng, cns))
            eng.closebam()
        except Exception as ex:      # noqa
            import traceback
            traceback.print_exc()
            err.append(ex)
            comms[r]._s.barrier.abort()
    th = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not err, err
    for r in range(world):
        for stage in one:
            a, b = dict(one[stage]), dict(out[r][stage])
            if stage == "fetch":          # f64 partial sums are added in another order: 1e-9 relative
                na, nb = a.pop("nm_stats"), b.pop("nm_stats")
                assert na[2] == nb[2] and all(abs(x - y) <= 1e-9 * abs(x) for x, y in zip(na[:2], nb[:2]))
            assert a == b, (r, stage)
