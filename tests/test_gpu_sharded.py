"""GPU parity of the sharded path: the file cut into 2, 3 and 5 contiguous shards, one engine per
shard on the same device exchanging through the in-process communicator, must reproduce the UNMODIFIED reference's
state after every stage and a byte-identical graph.txt — i.e. exactly the 1-GPU result."""
import pytest

import test_golden_cpu as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case,world", [("c2_small", 2), ("c4_small", 3), ("c5_mini", 2), ("c5_mini", 5), ("c3_small", 2),
                                        ("c1_n", 1)])
def test_sharded_pipeline_matches_reference(built_lib, case, world, tmp_path_factory):
    gold = G.load_golden(case)
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    got = G.run_engine("cuda-sharded:%d" % world, paths, tmp_path_factory)
    assert G.compare_states(got, gold["stages"]) == []
    for g_got, g_ref in zip(got["assign_cov"]["graphs"], gold["stages"]["assign_cov"]["graphs"]):
        assert g_got["graph_txt"] == g_ref["graph_txt"]


@pytest.mark.parametrize("case,world", [("c2_small", 2), ("c4_small", 3), ("c5_mini", 4), ("c3_small", 2)])
def test_sharded_pipeline_over_peer_memory_matches_reference(built_lib, case, world, tmp_path):
    """The same pipeline with the exchanges as kernels over peer memory (csrc/xchg.cu): `world` PROCESSES on cuda:0 that
    map each other's windows through CUDA IPC, as one-process-per-GPU ranks do (tests/run_rank.py)."""
    import json
    import util
    gold = G.load_golden(case)
    out = str(tmp_path / "state.json")
    util.spawn_ranks(["pipeline", case, out], world)
    res = json.load(open(out))
    assert len(set(res["digests"])) == 1, "ranks ended in different states"
    assert res["timeouts"] == [0] * world
    got = res["state"]
    assert G.compare_states(got, gold["stages"]) == []
    for g_got, g_ref in zip(got["assign_cov"]["graphs"], gold["stages"]["assign_cov"]["graphs"]):
        assert g_got["graph_txt"] == g_ref["graph_txt"]


# ---- the packed device-side exchange on its own (csrc/exchange.cu) against the whole file ----------------------------
def _packed_rows(shards, refs):
    """Runs the fetch exchange for `shards` (list of (rec, rec_off)) the way N ranks do, sequentially on one device.
    -> (gathered rows in the layout test_sharded_cpu._fetch_view reads, extra rows per rank, nm partial sums)."""
    import numpy as np
    import torch
    from coral_b200 import _abi, exchange
    dev = torch.device("cuda", 0)

    def cat(msgs):
        sizes = [int(m.shape[0]) for m in msgs]
        return torch.cat(msgs), np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.uint64), sizes
    ctxs = []
    for rec, off in shards:
        c = _abi.Context(0)
        c.set_references(refs)
        c.load_records(rec, off)
        ctxs.append(c)
    buf1, st1, _ = cat([exchange.view(*c.pack_sa_rows(), "|u1", dev).clone() for c in ctxs])
    torch.cuda.synchronize()
    buf2, st2, _ = cat([exchange.view(*c.pack_extra_rows(buf1.data_ptr(), st1), "|u1", dev).clone() for c in ctxs])
    torch.cuda.synchronize()
    extras = [int(buf2[int(s):int(s) + 8].cpu().numpy().view(np.int64)[0]) for s in st2]
    cctx = _abi.Context(0)
    cctx.set_references(refs)
    n, parts = cctx.unpack_rows(buf1.data_ptr(), st1, buf2.data_ptr(), st2)
    assert n == cctx.n_records
    p = cctx.device_view()
    off = exchange.view(p[1], n + 1, "<i8", dev).cpu().numpy()
    text = exchange.view(p[0], int(off[-1]), "|u1", dev).cpu().numpy()
    fm = exchange.view(p[4], n, "<i4", dev).cpu().numpy().astype(np.int64)
    lseq = exchange.view(p[5], n, "<i4", dev).cpu().numpy()
    keys = cctx.record_keys()
    sa = [text[int(off[i]) + int(keys["sa_off"][i]):int(off[i]) + int(keys["sa_off"][i]) + int(keys["sa_len"][i])].tobytes()
          for i in range(n)]
    g = {"names": exchange.row_names(text, off[:-1]), "flag": fm & 0xffff, "l_seq": lseq,
         "text": np.frombuffer(b"".join(sa), np.uint8), "text_off": np.concatenate([[0], np.cumsum([len(x) for x in sa])])}
    for c in ctxs + [cctx]:
        c.close()
    return g, extras, parts


@pytest.mark.parametrize("world", [1, 2, 5])
def test_packed_fetch_exchange_matches_the_file(built_lib, world, tmp_path_factory):
    """SA rows + the first primaries that carry no SA tag themselves (ibg:142-143), gathered from N shards, give fetch
    the same per-read SA lists and read lengths as the whole file gives the reference's loop."""
    import numpy as np
    import test_sharded_cpu as S
    from coral_b200 import _abi, sharded
    from sim import configs
    from sim.bgzf import write_bam
    from sim.generate import simulate
    data = simulate(configs.CONFIGS["C5"](scale=0.004))
    rec, off = data.rec.copy(), np.asarray(data.rec_off, dtype=np.uint64)
    refs = data.cfg.refs
    names = [r[0] for r in refs]
    # make some primaries lose their SA tag (renamed to XY:Z) while the supplementaries keep theirs
    c = _abi.Context(0)
    c.set_references(names)
    c.load_records(rec, off)
    keys = c.record_keys()
    flag = c.record_table()["flag"]
    c.close()
    prim = np.nonzero((keys["sa_off"] != 0) & (flag < 256))[0][::3]
    for i in prim:
        a = int(off[i]) + int(keys["sa_off"][i]) - 3
        assert bytes(rec[a:a + 3]) == b"SAZ"
        rec[a:a + 2] = np.frombuffer(b"XY", np.uint8)
    path = str(tmp_path_factory.mktemp("packed") / "edited.bam")
    write_bam(path, refs, rec, off)
    rows = S._file_rows(path)
    want = S._reference_view(rows)
    cuts = sharded.shard_ranges(off, world)
    shards = [(rec[int(off[a]):int(off[b])], off[a:b + 1] - off[a]) for a, b in cuts]
    g, extras, parts = _packed_rows(shards, names)
    assert S._fetch_view(g) == want
    if world > 1:
        assert sum(extras) > 0          # the case the second message exists for
    # nm_stats partial sums add up to the 1-shard sums
    _, _, whole = _packed_rows([(rec, off)], names)
    assert int(parts[2].sum()) == int(whole[2].sum()) and int(parts[3].sum()) == int(whole[3].sum())
    assert abs(float(parts[0].sum()) - float(whole[0].sum())) <= 1e-9 * max(1.0, abs(float(whole[0].sum())))


def test_packed_exchange_edge_cases(built_lib, tmp_path_factory):
    """An empty shard in the middle, a shard without any SA record, and the argument checks of the entry points."""
    import numpy as np
    import torch
    import test_sharded_cpu as S
    from coral_b200 import _abi, exchange
    from sim import configs
    from sim.bgzf import write_bam
    from sim.generate import simulate
    data = simulate(configs.CONFIGS["C2"](scale=0.01))
    rec, off = data.rec, np.asarray(data.rec_off, dtype=np.uint64)
    refs = data.cfg.refs
    names = [r[0] for r in refs]
    n = len(off) - 1
    path = str(tmp_path_factory.mktemp("packed_edge") / "reads.bam")
    write_bam(path, refs, rec, off)
    want = S._reference_view(S._file_rows(path))
    assert len(want) > 0
    # cuts: [0, a) | empty | [a, b) | [b, n); the first shard is short enough to hold no SA record on most data
    a, b = n // 50, n // 2
    cuts = [(0, a), (a, a), (a, b), (b, n)]
    shards = [(rec[int(off[x]):int(off[y])], off[x:y + 1] - off[x]) for x, y in cuts]
    g, extras, parts = _packed_rows(shards, names)
    assert S._fetch_view(g) == want
    assert int(parts[2][1]) == 0 and float(parts[0][1]) == 0.0          # the empty shard contributes nothing
    # argument checks
    dev = torch.device("cuda", 0)
    c = _abi.Context(0)
    c.set_references(names)
    with pytest.raises(RuntimeError):
        c.pack_sa_rows()                                  # no records loaded
    c.load_records(*shards[2])
    ptr, nb = c.pack_sa_rows()
    assert nb % 16 == 0 and nb >= 64
    msg = exchange.view(ptr, nb, "|u1", dev).clone()
    with pytest.raises(RuntimeError):
        c.pack_extra_rows(msg.data_ptr(), [8])            # misaligned message start
    with pytest.raises(RuntimeError):
        c.pack_extra_rows(msg.data_ptr(), [0] * 17)       # more than 16 ranks
    xptr, xnb = c.pack_extra_rows(msg.data_ptr(), [0])
    x = exchange.view(xptr, xnb, "|u1", dev).clone()
    cc = _abi.Context(0)
    cc.set_references(names)
    with pytest.raises(RuntimeError):
        cc.unpack_rows(x.data_ptr(), [0], msg.data_ptr(), [0])          # message kinds swapped
    nrows, _ = cc.unpack_rows(msg.data_ptr(), [0], x.data_ptr(), [0])
    assert nrows == cc.n_records
    for call in (cc.rescan, lambda: cc.region_coverage([0], [0], [10]), cc.pack_sa_rows, lambda: cc.concordant_rows_only([0], [1000])):
        with pytest.raises(RuntimeError):
            call()                                         # a rows-only ctx holds no records
    # concordant messages: a bad size is refused, an empty message counts nothing
    c.set_params(min_bp_match_cutoff=100)
    c.concordant_rows_only([0], [1000])
    mptr, mnb = c.concordant_pack()
    m = exchange.view(mptr, mnb, "|u1", dev).clone()
    with pytest.raises(RuntimeError):
        c.count_name_msgs(1, m.data_ptr(), [0], [mnb + 8], 0, 0)
    cnt = c.count_name_msgs(1, m.data_ptr(), [0], [mnb], 0, 0)
    assert cnt.shape == (1,) and int(cnt[0]) >= 0
    c.close(); cc.close()
