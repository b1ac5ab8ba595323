"""Pipelined pass (coral_pass_begin / coral_pass_end): the hot path enqueued as one stream of kernels must deliver exactly
what the same calls deliver one by one — on the recording pass, on replayed passes (no intermediate synchronisation), and
when the call sequence or a host input changes between passes."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


_ROWS = ("row_pos1", "row_pos2", "row_nb", "row_first", "row_stats")


def _same(a, b, path=""):
    if isinstance(a, dict) and "cluster_ptr" in a and "n_sub" in a:
        # coral_get_clusters: the row arrays are only defined for the first n_sub[c] slots of every cluster (one per
        # bpc2bp call); the other slots are whatever the buffer held
        assert set(a) == set(b), path
        for k in a:
            if k not in _ROWS:
                _same(a[k], b[k], path + "/" + str(k))
        ptr, n_sub = a["cluster_ptr"], a["n_sub"]
        idx = np.array([int(ptr[c]) + j for c in range(len(n_sub)) for j in range(int(n_sub[c]))], np.int64)
        for k in _ROWS:
            assert np.array_equal(np.asarray(a[k])[idx], np.asarray(b[k])[idx]), path + "/" + k
        return
    if isinstance(a, dict):
        assert set(a) == set(b), path
        for k in a:
            _same(a[k], b[k], path + "/" + str(k))
    elif isinstance(a, (tuple, list)):
        assert len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, "%s[%d]" % (path, i))
    elif isinstance(a, np.ndarray):
        assert np.array_equal(a, b), path
    else:
        assert a == b, path


def _calls(ctx, plan, cutoff, fetch_first=False):
    """fetch_first: the order bench.py uses — what only needs the pre-scan's columns comes first and runs, inside a
    pass, next to the full scan (coral_rescan in a pass = pre-scan + full scan on a second stream)."""
    from coral_b200 import gpu_ops
    out = {}
    ctx.rescan()

    def chimeric():
        out["fetch"] = ctx.fetch()
        cand = ctx.candidates_l(*plan["intervals"])
        out["cand"] = cand
        if cand["n"]:
            out["cl2"] = ctx.cluster_call_candidates(cand["n"], cutoff, 2000, cutoff, 100)
    if fetch_first:
        chimeric()
    out["normal"] = ctx.region_coverage(*plan["normal"])
    out["smalldel"] = ctx.smalldel(*plan["intervals"])
    if len(plan["sd_cand"]["chr1"]):
        out["cl1"] = gpu_ops.cluster_arrays(ctx, plan["sd_cand"], cutoff, 2000, cutoff, 100)
    if not fetch_first:
        chimeric()
    out["seq"] = ctx.region_coverage(*plan["seq"])
    out["conc"] = ctx.concordant(*plan["conc"])
    return out


def _canon(out):
    """The stab rows come out in the order their atomics landed: compare them sorted."""
    cnt, rows = out["conc"]
    o = np.lexsort((rows["mask"], rows["rec"], rows["edge"]))
    out["conc"] = (cnt, {k: v[o] for k, v in rows.items()})
    return out


@pytest.mark.parametrize("cfg_name,scale", [("C2", 0.03), ("C3", 0.01), ("C1", 0.05)])
def test_pipelined_pass_equals_call_by_call(built_lib, dataset, cfg_name, scale):
    from coral_b200 import _abi, gpu_ops
    ds = dataset(cfg_name, scale=scale)
    cfg, data = ds["cfg"], ds["data"]
    names = [r[0] for r in cfg.refs]
    ctx = _abi.Context(0)
    ctx.set_references(names)
    ctx.load_records(data.rec, data.rec_off)
    t = ctx.record_table()
    tid_of = {n: i for i, n in enumerate(names)}
    seeds = sorted((tid_of[c], s, e) for (c, s, e) in cfg.seeds)
    iv = tuple(np.array(x, np.int32) for x in zip(*seeds))
    # regions / edges: the seed intervals cut in three, stab points at the cuts
    reg, pts = [], []
    for (tt, s, e) in seeds:
        a, b = s + (e - s) // 3, s + 2 * (e - s) // 3
        reg += [(tt, s, a), (tt, a + 1, b), (tt, b + 1, e + 1)]
        pts += [(tt, a), (tt, b)]
    normal = [(tid_of[c], lo, lo + 2_000_000) for (c, lo, hi) in cfg.background]
    sd = ctx.smalldel(*iv)
    n_sd = sd["n"]
    sd_cand = {"chr1": sd["ai"].astype(np.int32), "pos1": sd["next_start"], "chr2": sd["ai"].astype(np.int32), "pos2": sd["prev_end"],
               "o1": np.ones(n_sd, np.uint8), "o2": np.zeros(n_sd, np.uint8), "rgap": np.zeros(n_sd, np.int32),
               "flag": np.zeros(n_sd, np.uint8), "mq1": np.full(n_sd, -1, np.int32), "mq2": np.full(n_sd, -1, np.int32)}
    rng = np.random.default_rng(3)
    rb_rec = rng.choice(len(t["tid"]), 40, replace=False).astype(np.uint32)
    rb_edge = rng.integers(0, len(pts), 40).astype(np.uint32)
    plan = {"normal": tuple(np.array(x, np.int32) for x in zip(*sorted(normal))), "intervals": iv, "sd_cand": sd_cand,
            "seq": tuple(np.array(x, np.int32) for x in zip(*reg)),
            "conc": (np.array([p[0] for p in pts], np.int32), np.array([p[1] for p in pts], np.int32), rb_edge, rb_rec)}
    want = _canon(_calls(ctx, plan, 3.0))
    if cfg_name == "C3":
        assert want["smalldel"]["n"] > 0 and "cl1" in want
    assert want["fetch"]["n_reads"] > 0
    if cfg_name == "C1":                    # junctions inside one amplicon interval: alignment2bp_l emits candidates
        assert want["cand"]["n"] > 0 and "cl2" in want
    for k in range(3):                      # recording pass, then two replayed ones
        ctx.pass_begin()
        got = _calls(ctx, plan, 3.0)
        replayed = ctx.pass_end()
        assert replayed == (k > 0)
        _same(_canon(got), want)
    for k in range(3):                      # the overlapped order: another call sequence, so it records first
        ctx.pass_begin()
        got = _calls(ctx, plan, 3.0, fetch_first=True)
        replayed = ctx.pass_end()
        assert replayed == (k > 0)
        _same(_canon(got), want)
    ctx.set_option("pass_overlap_scan", 0)  # the one-kernel scan on the ctx's stream
    for k in range(2):
        ctx.pass_begin()
        got = _calls(ctx, plan, 3.0, fetch_first=True)
        ctx.pass_end()
        _same(_canon(got), want)
    ctx.set_option("pass_overlap_scan", 1)
    # a host input changes (other cutoff -> other cluster calls): the pass re-records from that call on, results stay right
    want2 = _canon(_calls(ctx, plan, 5.0))
    ctx.pass_begin()
    got2 = _calls(ctx, plan, 5.0)
    ctx.pass_end()
    _same(_canon(got2), want2)
    ctx.pass_begin()
    got2 = _calls(ctx, plan, 5.0)
    assert ctx.pass_end() is True
    _same(_canon(got2), want2)
    # an entry point outside the pass protocol inside a pass: it synchronises, everything still arrives
    ctx.pass_begin()
    got3 = _calls(ctx, plan, 5.0)
    tab = ctx.record_table()
    ctx.pass_end()
    _same(_canon(got3), want2)
    assert np.array_equal(tab["endpos"], t["endpos"])
    # result arrays in the ctx's page-locked arena (direct DMA, the chimeric table as one packed copy): same values
    for k in range(2):
        ctx.pass_begin(pinned_results=True)
        got5 = _calls(ctx, plan, 5.0)
        ctx.pass_end()
        _same(_canon(got5), want2)
    # new records invalidate the recorded sizes
    ctx.load_records(data.rec, data.rec_off)
    ctx.pass_begin()
    got4 = _calls(ctx, plan, 5.0)
    assert ctx.pass_end() is False
    _same(_canon(got4), want2)
    ctx.close()
