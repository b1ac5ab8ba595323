"""CPU tests of the multi-GPU exchange logic (coral_b200/sharded.py): world_size-2 `gloo` processes and in-process
thread ranks.  The device-free part of the N > 1 path: what every rank gathers must not depend on how the file was
cut, and must reproduce what the reference's fetch does with the whole file (ibg:139-174 via the oracle shim)."""
import hashlib
import os
import sys
import threading

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from coral_b200 import sharded  # noqa: E402


def _file_rows(bam_path):
    """Every placed record of the file, in file order, as the arrays exchange_fetch_rows wants (name hash = md5)."""
    from oracle import pysam_shim
    fh = pysam_shim.AlignmentFile(bam_path)
    names, sa, flag, mapq, lseq, nm = [], [], [], [], [], []
    for r in fh.fetch():
        names.append(r.query_name)
        sa.append(r.get_tag("SA").encode() if r.has_tag("SA") else None)
        flag.append(r.flag); mapq.append(r.mapping_quality); lseq.append(r.query_length)
        nm.append(r.get_cigar_stats()[0][-1])
    h = [hashlib.md5(n.encode()).digest() for n in names]
    return {"names": names, "sa": sa, "flag": np.array(flag, np.uint32), "mapq": np.array(mapq, np.uint32),
            "l_seq": np.array(lseq, np.uint32), "nm": np.array(nm, np.int32),
            "name_lo": np.array([int.from_bytes(x[:8], "little") for x in h], np.uint64),
            "name_hi": np.array([int.from_bytes(x[8:], "little") for x in h], np.uint64)}


def _local(rows, r0, r1):
    return {"name_lo": rows["name_lo"][r0:r1], "name_hi": rows["name_hi"][r0:r1],
            "has_sa": np.array([s is not None for s in rows["sa"][r0:r1]], bool), "flag": rows["flag"][r0:r1],
            "mapq": rows["mapq"][r0:r1], "l_seq": rows["l_seq"][r0:r1], "nm": rows["nm"][r0:r1],
            "sa_text": lambda i: rows["sa"][r0 + i], "name": lambda i: rows["names"][r0 + i]}


def _fetch_view(g):
    """What fetch derives from gathered rows: per name, in first-SA-row order, the de-duplicated SA entries in file
    order (ibg:145-151) and l_seq of the first row with flag < 256 (ibg:142-143)."""
    sa, rl, order = {}, {}, []
    for k, name in enumerate(g["names"]):
        if g["flag"][k] < 256 and name not in rl:
            rl[name] = int(g["l_seq"][k])
        a, b = int(g["text_off"][k]), int(g["text_off"][k + 1])
        if b > a:
            if name not in sa:
                sa[name] = []
                order.append(name)
            for ent in g["text"][a:b].tobytes().decode()[:-1].split(";"):
                if ent not in sa[name]:
                    sa[name].append(ent)
    return [(n, sa[n], rl.get(n)) for n in order]


def _reference_view(rows):
    """The same thing straight from the whole file, the way the reference's loop does it."""
    sa, rl, order = {}, {}, []
    for k, name in enumerate(rows["names"]):
        if rows["flag"][k] < 256 and name not in rl:
            rl[name] = int(rows["l_seq"][k])
        if rows["sa"][k] is not None:
            if name not in sa:
                sa[name] = []
                order.append(name)
            for ent in rows["sa"][k].decode()[:-1].split(";"):
                if ent not in sa[name]:
                    sa[name].append(ent)
    return [(n, sa[n], rl.get(n)) for n in order]


@pytest.fixture(scope="module")
def c5_rows(tmp_path_factory):
    from sim import configs
    from sim.generate import simulate, write_dataset
    data = simulate(configs.CONFIGS["C5"](scale=0.004))
    paths = write_dataset(data, str(tmp_path_factory.mktemp("shard_c5")))
    return _file_rows(paths["bam"]), paths


def _run_threads(world, fn):
    comms = sharded.ThreadComm.make(world)
    out, err = [None] * world, []

    def body(r):
        try:
            out[r] = fn(comms[r])
        except Exception as ex:      # noqa
            err.append(ex)
            comms[r]._s.barrier.abort()
    th = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not err, err
    return out


@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_fetch_exchange_is_cut_invariant(c5_rows, world):
    rows, _ = c5_rows
    n = len(rows["names"])
    want = _reference_view(rows)
    assert len(want) > 50                                  # chimeric reads exist, many across chromosomes
    cuts = [n * r // world for r in range(world + 1)]
    got = _run_threads(world, lambda c: sharded.exchange_fetch_rows(c, _local(rows, cuts[c.rank], cuts[c.rank + 1])))
    for g in got:
        assert _fetch_view(g) == want
        assert len(g["names"]) == len(g["flag"]) == len(g["text_off"]) - 1


def test_shard_ranges_cover_and_balance():
    off = np.concatenate([[0], np.cumsum(np.random.default_rng(1).integers(1, 400, 1000) * 16)]).astype(np.uint64)
    for world in (1, 2, 3, 8):
        rg = sharded.shard_ranges(off, world)
        assert rg[0][0] == 0 and rg[-1][1] == 1000 and all(rg[k][1] == rg[k + 1][0] for k in range(world - 1))
        sizes = [int(off[b]) - int(off[a]) for a, b in rg]
        assert max(sizes) - min(sizes) <= 2 * 400 * 16
    assert sharded.shard_ranges(np.zeros(1, np.uint64), 4) == [(0, 0)] * 4


def test_member_handles_equal_low_words():
    set_lo = np.array([5, 5, 7], np.uint64); set_hi = np.array([1, 9, 2], np.uint64)
    lo = np.array([5, 5, 5, 7, 8], np.uint64); hi = np.array([9, 1, 3, 2, 0], np.uint64)
    assert sharded._member(lo, hi, set_lo, set_hi).tolist() == [True, True, False, True, False]


def test_smalldel_and_concordant_merges():
    rows = [[("a", 0, 1, 2, 3, 4, 60), ("b", 1, 1, 2, 3, 4, 60)], [("c", 0, 5, 6, 7, 8, 30), ("d", 1, 5, 6, 7, 8, 30)]]
    got = _run_threads(2, lambda c: sharded.merge_smalldel_rows(c, rows[c.rank]))
    assert got[0] == got[1] == [rows[0][0], rows[1][0], rows[0][1], rows[1][1]]
    # read r1 covers p / p+1 on shard 0 and the two flanks on shard 1: only the OR of its records passes the 4-point test
    conc = [[(0, "r1", 3), (0, "r2", 15), (0, "r3", 15)], [(0, "r1", 12), (0, "r4", 4)]]
    got = _run_threads(2, lambda c: sharded.combine_concordant(c, 1, conc[c.rank], [{"r3"}]))
    assert got[0] == got[1] == ([2], [{"r1", "r2", "r3"}])


def _gloo_worker(rank, world, port, bam, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = sharded.TorchComm()
        rows = _file_rows(bam)
        n = len(rows["names"])
        cuts = [n * r // world for r in range(world + 1)]
        g = sharded.exchange_fetch_rows(comm, _local(rows, cuts[rank], cuts[rank + 1]))
        tot = comm.all_reduce_sum(np.array([rank + 1, 10 * (rank + 1)], np.int64))
        merged = sharded.merge_smalldel_rows(comm, [("r%d" % rank, 1 - rank, 0, 0, 0, 0, 0)])
        q.put((rank, _fetch_view(g) == _reference_view(rows), tot.tolist(), [m[0] for m in merged]))
    finally:
        dist.destroy_process_group()


def test_exchange_over_gloo_world2(c5_rows):
    import torch.multiprocessing as mp
    _, paths = c5_rows
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    ps = [ctx.Process(target=_gloo_worker, args=(r, 2, port, paths["bam"], q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=300) for _ in range(2))
    [p.join(60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    for rank, same, tot, merged in res:
        assert same and tot == [3, 30] and merged == ["r1", "r0"]


# ---- transport of the packed device messages (coral_b200/exchange.py) on CPU tensors ---------------------------------
def _msg(rank, nbytes):
    import torch
    return torch.as_tensor(((np.arange(nbytes) * 7 + rank * 13) % 251).astype(np.uint8))


@pytest.mark.parametrize("world", [1, 2, 4])
def test_gather_msgs_thread_ranks(world):
    from coral_b200 import exchange
    sizes = [16 * (3 + 5 * r) for r in range(world)]
    got = _run_threads(world, lambda c: exchange.TensorComm(c).gather_msgs(_msg(c.rank, sizes[c.rank])))
    for buf, st, sz in got:
        assert sz == sizes and all(s % 16 == 0 for s in st)
        for r in range(world):
            assert np.array_equal(buf[st[r]:st[r] + sz[r]].numpy(), _msg(r, sizes[r]).numpy())


def test_smalldel_message_round_trip():
    rows = [("read/%d" % k, k % 3, 1000 + k, 900 + k, 10, 5000, 60, -(2 ** 62) + k, 2 ** 62 - k) for k in range(7)]
    for part in ([], rows[:1], rows):
        msg = sharded.pack_smalldel(part)
        padded = np.concatenate([msg.view(np.uint8), np.zeros(8, np.uint8)])     # transport pads to 16 bytes
        assert sharded.unpack_smalldel(padded) == part


def test_row_names_from_text_slots():
    from coral_b200 import exchange
    slots = [b"readA\0chr1,5,+,10M,60,0;", b"b\0", b"read/long/name\0chr2,7,-,5S5M,3,1;chr3,1,+,5M5S,0,0;"]
    text = np.frombuffer(b"".join(slots), np.uint8)
    off = np.cumsum([0] + [len(x) for x in slots])[:-1]
    assert exchange.row_names(text, off) == ["readA", "b", "read/long/name"]
    assert exchange.row_names(text, off[[2, 0, 1]]) == ["read/long/name", "readA", "b"]       # any slot order
    assert exchange.row_names(text, off[:0]) == []


def _gloo_msg_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from coral_b200 import exchange
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = sharded.TorchComm()
        ok = True
        # eager round only / a message larger than the eager slot (second, exact round) / an empty-payload header
        for eager, sizes in ((4096, [160, 48]), (256, [64, 1024]), (256, [1024, 2048]), (4096, [16, 16])):
            tc = exchange.TensorComm(comm, eager=eager)
            buf, st, sz = tc.gather_msgs(_msg(rank, sizes[rank]))
            ok = ok and sz == sizes and all(s % 16 == 0 for s in st)
            for r in range(world):
                ok = ok and np.array_equal(buf[st[r]:st[r] + sz[r]].numpy(), _msg(r, sizes[r]).numpy())
            ok = ok and tc.n_collectives == (1 if max(sizes) <= eager - 16 else 2)
        # small-deletion rows: host-built messages merged in (interval, rank, block) order
        mine = [("r%d_%d" % (rank, k), (k + rank) % 2, k, k, 0, 9, 60, rank, k) for k in range(3)]
        tc = exchange.TensorComm(comm)
        buf, st, sz = tc.gather_msgs(exchange.host_msg(sharded.pack_smalldel(mine), torch.device("cpu")))
        raw = buf.numpy()
        merged = [r for a, n in zip(st, sz) for r in sharded.unpack_smalldel(raw[a:a + n])]
        merged.sort(key=lambda r: r[1])
        want = [("r%d_%d" % (rk, k), (k + rk) % 2, k, k, 0, 9, 60, rk, k) for rk in range(world) for k in range(3)]
        want.sort(key=lambda r: r[1])
        ok = ok and merged == want
        # concordant-read rows of compute_path_constraints: (interval, file) order restored from per-rank rows
        iv, first, last = _path_rows(rank)
        m = sharded.merge_path_rows(comm, iv, first, last)
        ok = ok and all(np.array_equal(a, b) for a, b in zip(m, _path_rows_whole(world)))
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def _path_rows(rank):
    """Rows of rank `rank`: intervals 0..2, a few records each, edge runs that encode (interval, rank, k)."""
    iv = np.repeat(np.arange(3), [2 + rank, 0 if rank == 1 else 3, 1])
    k = np.arange(len(iv))
    return iv, 100 * iv + 10 * rank, 100 * iv + 10 * rank + k


def _path_rows_whole(world):
    rows = [(int(i), int(f), int(l), r, n) for r in range(world) for n, (i, f, l) in enumerate(zip(*_path_rows(r)))]
    rows.sort(key=lambda t: (t[0], t[3], t[4]))          # interval, then rank (= file) order, then row order
    return tuple(np.array([t[c] for t in rows], np.int64) for c in range(3))


@pytest.mark.parametrize("world", [1, 2, 3])
def test_merge_path_rows_thread_ranks(world):
    comms = sharded.ThreadComm.make(world)
    out = [None] * world

    def body(r):
        out[r] = sharded.merge_path_rows(comms[r], *_path_rows(r))
    th = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    for r in range(world):
        assert all(np.array_equal(a, b) for a, b in zip(out[r], _path_rows_whole(world)))


def test_gather_msgs_over_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    ps = [ctx.Process(target=_gloo_msg_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=300) for _ in range(2))
    [p.join(60) for p in ps]
    assert all(p.exitcode == 0 for p in ps) and all(ok for _, ok in res)


def test_merge_of_packed_smalldel_messages_restores_the_emission_order():
    """The device-packed small-deletion messages (csrc/chimeric_merge.cu: 64-byte header + 48-byte rows) of three ranks,
    one of them empty, merged on the host: interval order first, then rank (= file) order, then the rank's row order —
    the order find_smalldel_breakpoints' loops emit in (ibg:750-762)."""
    import numpy as np
    from coral_b200 import sharded

    def msg(rows):
        hdr = np.zeros(8, np.uint64)
        hdr[0] = len(rows); hdr[6] = 5
        body = np.zeros(len(rows), sharded._SDEL_ROW)
        for k, r in enumerate(rows):
            body[k] = r
        return np.concatenate([hdr.view(np.uint8), body.view(np.uint8)])
    r0 = [(5, 0, 100, 90, 10, 900, 60, 0, 11, 12), (9, 2, 300, 290, 20, 800, 50, 0, 13, 14)]
    r2 = [(1, 0, 120, 95, 15, 950, 40, 0, 21, 22), (2, 1, 220, 200, 25, 850, 30, 0, -23, 24), (2, 1, 260, 240, 25, 850, 30, 0, -23, 24)]
    got = sharded.merge_smalldel_msgs([msg(r0), msg([]), msg(r2)])
    assert [(g[0], g[1], g[2]) for g in got] == [(0, 5, 0), (2, 1, 0), (2, 2, 1), (2, 2, 1), (0, 9, 2)]
    assert got[3][3:8] == (260, 240, 25, 850, 30) and got[1][8:] == (21, 22) and got[2][8:] == (-23, 24)
    assert sharded.merge_smalldel_msgs([msg([])]) == []
