"""Multi-GPU form of the hot path: one process per GPU, each holding a contiguous range of the coordinate-sorted
file (a genomic bin range balanced by decoded bytes), SURVEY.md §8e.

Per-record work (record scan, coverage sums, gap events, stab rows) is local and additive — a record is owned by
exactly one shard — so sequence-edge and normal-coverage sums only need an integer all-reduce.  The one real
exchange is `fetch`: the reference unions the SA strings of all records of a read name in file order and takes
read_length from the first record with flag < 256 anywhere in the file (infer_breakpoint_graph.py:142-151), and
the records of a chimeric read live on several shards.  Every rank exports one small row per relevant record,
the rows are all-gathered in rank (= file) order, and every rank runs the same deterministic fetch on the
gathered rows (coral_load_rows), so candidates and clustering are replicated and identical to the 1-GPU result.

This module holds the communicators, the engine that binds the exchanges to the CUDA contexts (`ShardedEngine`, at
the bottom: packed device messages, csrc/exchange.cu + exchange.py) and — in plain numpy — the statement of what each
exchange must deliver (`exchange_fetch_rows`, `merge_smalldel_rows`, `combine_concordant`): the CPU tests run those
over thread ranks and world_size-2 gloo, and the GPU tests check the device messages against the same file view.
"""
import threading

import numpy as np


# ---- communicators -------------------------------------------------------------------------------------------
class SingleComm(object):
    rank, world = 0, 1

    def all_gather(self, obj):
        return [obj]

    def all_reduce_sum(self, arr):
        return np.array(arr, copy=True)


class TorchComm(object):
    """torch.distributed (NCCL over NVLink on the GPU box, gloo in the CPU tests)."""

    def __init__(self, group=None, device=None):
        import torch.distributed as dist
        self._dist = dist
        self._group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self._device = device

    def all_gather(self, obj):
        out = [None] * self.world
        self._dist.all_gather_object(out, obj, group=self._group)
        return out

    def all_reduce_sum(self, arr):
        import torch
        a = np.ascontiguousarray(arr)
        t = torch.from_numpy(a.copy())
        if self._device is not None:
            t = t.to(self._device)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self._group)
        return t.cpu().numpy()


class ThreadComm(object):
    """N ranks as N threads of one process (shards emulated on one GPU, or on the CPU in tests)."""

    class _Shared(object):
        def __init__(self, world):
            self.world = world
            self.barrier = threading.Barrier(world, timeout=600)
            self.slots = [None] * world

    def __init__(self, shared, rank):
        self._s = shared
        self.rank = rank
        self.world = shared.world

    @classmethod
    def make(cls, world):
        sh = cls._Shared(world)
        return [cls(sh, r) for r in range(world)]

    def all_gather(self, obj):
        self._s.slots[self.rank] = obj
        self._s.barrier.wait()
        out = list(self._s.slots)
        self._s.barrier.wait()
        return out

    def all_reduce_sum(self, arr):
        parts = self.all_gather(np.asarray(arr))
        tot = np.array(parts[0], copy=True)
        for p in parts[1:]:
            tot = tot + p
        return tot


# ---- pure host logic -------------------------------------------------------------------------------------------
def shard_ranges(rec_off, world):
    """Contiguous record ranges [r0, r1) per rank, balanced by decoded bytes (file order = genomic bin order)."""
    rec_off = np.asarray(rec_off, dtype=np.uint64)
    n = len(rec_off) - 1
    total = int(rec_off[-1]) if n > 0 else 0
    cuts = [0]
    for r in range(1, world):
        cuts.append(int(np.searchsorted(rec_off[:-1], np.uint64(total * r // world), side="left")))
    cuts.append(n)
    for k in range(1, len(cuts)):
        cuts[k] = max(cuts[k], cuts[k - 1])
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def _member(lo, hi, set_lo, set_hi):
    """Vectorised membership of (lo, hi) pairs in a set of pairs sorted by (lo, hi)."""
    if len(set_lo) == 0 or len(lo) == 0:
        return np.zeros(len(lo), bool)
    pos = np.searchsorted(set_lo, lo, side="left")
    hit = np.zeros(len(lo), bool)
    live = pos < len(set_lo)
    while live.any():
        idx = np.nonzero(live)[0]
        p = pos[idx]
        same_lo = set_lo[p] == lo[idx]
        ok = same_lo & (set_hi[p] == hi[idx])
        hit[idx[ok]] = True
        nxt = idx[same_lo & ~ok]                   # same low word, other high word: look at the next entry
        pos[nxt] += 1
        live = np.zeros(len(lo), bool)
        live[nxt] = pos[nxt] < len(set_lo)
    return hit


def exchange_fetch_rows(comm, local):
    """The `fetch` exchange.  `local` describes this rank's records in file order:
         name_lo, name_hi (uint64), has_sa (bool), flag, mapq, l_seq (uint32), nm (int32) — arrays of length n;
         sa_text(i) -> bytes (the SA:Z value without its NUL), name(i) -> str.
    Returns the gathered rows in global file order: dict(name_lo, name_hi, flag, mapq, l_seq, nm, text, text_off,
    names[list of str]) — the same on every rank.  A world-N gather carries up to N primaries per name (the first of
    each rank) where a 1-rank export carries one; fetch keeps the first in file order, so its result is the same."""
    lo = np.asarray(local["name_lo"], np.uint64)
    hi = np.asarray(local["name_hi"], np.uint64)
    has_sa = np.asarray(local["has_sa"], bool)
    flag = np.asarray(local["flag"], np.uint32)
    # round 1: names that carry an SA tag anywhere in the file
    sa_idx = np.nonzero(has_sa)[0]
    mine = np.stack([lo[sa_idx], hi[sa_idx]], axis=1) if len(sa_idx) else np.zeros((0, 2), np.uint64)
    allp = np.concatenate(comm.all_gather(mine), axis=0)
    if len(allp):
        o = np.lexsort((allp[:, 1], allp[:, 0]))
        set_lo, set_hi = allp[o, 0], allp[o, 1]
    else:
        set_lo = set_hi = np.zeros(0, np.uint64)
    # round 2: + the first local record with flag < 256 of each such name (read_length comes from the first primary
    # in the file, which need not carry an SA tag itself)
    cand = np.nonzero((flag < 256) & ~has_sa)[0]
    cand = cand[_member(lo[cand], hi[cand], set_lo, set_hi)]
    if len(cand):
        pairs = np.stack([lo[cand], hi[cand]], axis=1)
        _, first = np.unique(pairs, axis=0, return_index=True)
        cand = cand[np.sort(first)]
    rows = np.sort(np.concatenate([sa_idx, cand])).astype(np.int64)
    texts = [local["sa_text"](int(i)) if has_sa[i] else b"" for i in rows]
    off = np.zeros(len(rows) + 1, np.uint64)
    if len(rows):
        off[1:] = np.cumsum([len(t) for t in texts])
    payload = {
        "name_lo": lo[rows], "name_hi": hi[rows], "flag": flag[rows],
        "mapq": np.asarray(local["mapq"], np.uint32)[rows], "l_seq": np.asarray(local["l_seq"], np.uint32)[rows],
        "nm": np.asarray(local["nm"], np.int32)[rows], "text": np.frombuffer(b"".join(texts), np.uint8),
        "text_off": off, "names": [local["name"](int(i)) for i in rows]}
    parts = comm.all_gather(payload)
    out = {k: np.concatenate([p[k] for p in parts]) for k in ("name_lo", "name_hi", "flag", "mapq", "l_seq", "nm", "text")}
    offs, base = [np.zeros(1, np.uint64)], 0
    for p in parts:
        offs.append(p["text_off"][1:] + np.uint64(base))
        base += int(p["text_off"][-1])
    out["text_off"] = np.concatenate(offs)
    out["names"] = [nm for p in parts for nm in p["names"]]
    return out


def merge_smalldel_rows(comm, rows):
    """rows: this rank's (name, interval index, next_start, prev_end, first_start, last_end, mapq) in local emission
    order.  Global emission order = interval list order, then file order, then block order (ibg:750-762); ranks hold
    increasing file ranges, so concatenating in rank order and stable-sorting by interval restores it."""
    allr = [r for part in comm.all_gather(list(rows)) for r in part]
    allr.sort(key=lambda r: r[1])
    return allr


def pack_smalldel(rows):
    """rows: list of (name, ai, next_start, prev_end, first_start, last_end, mapq, name_lo, name_hi) -> one int64
    message: [n, name bytes] + n x 8 columns + the NUL-separated names (padded to 8 bytes)."""
    n = len(rows)
    blob = b"".join(r[0].encode() + b"\0" for r in rows)
    nb = (len(blob) + 7) // 8
    out = np.zeros(2 + 8 * n + nb, np.int64)
    out[0], out[1] = n, len(blob)
    if n:
        out[2:2 + 8 * n] = np.array([r[1:9] for r in rows], np.int64).reshape(-1)
        out[2 + 8 * n:].view(np.uint8)[:len(blob)] = np.frombuffer(blob, np.uint8)
    return out


def unpack_smalldel(msg):
    """Inverse of pack_smalldel on the bytes of one message (uint8 array, possibly padded at the end)."""
    a = np.ascontiguousarray(msg).view(np.uint8)
    head = a[:16].view(np.int64)
    n, nbytes = int(head[0]), int(head[1])
    cols = a[16:16 + 64 * n].view(np.int64).reshape(n, 8)
    names = a[16 + 64 * n:16 + 64 * n + nbytes].tobytes().split(b"\0")[:n]
    return [(names[k].decode(), ) + tuple(int(x) for x in cols[k]) for k in range(n)]


def combine_concordant(comm, n_edges, rows, rbps):
    """rows: this rank's (edge, read name, mask) for every local record covering one of the four points of an edge
    (mask bit0 = p, bit1 = p+1, bit2 = p-d-1, bit3 = p+1+d).  The reference works on NAME sets (ibg:1043-1055), so
    the masks of all records of a name — on any shard — are OR-ed before the 4-point test.
    Returns (ce[8] per edge, ce[9] name set per edge)."""
    acc = [dict() for _ in range(n_edges)]
    for part in comm.all_gather(list(rows)):
        for (e, name, m) in part:
            d = acc[e]
            d[name] = d.get(name, 0) | m
    cnt, names = [], []
    for e in range(n_edges):
        bad = rbps[e]
        cnt.append(sum(1 for nm, m in acc[e].items() if m == 15 and nm not in bad))
        names.append(set(nm for nm, m in acc[e].items() if m & 3))
    return cnt, names


def merge_path_rows(comm, iv, first, last):
    """Concordant-read pass of compute_path_constraints across shards: every rank holds the rows (interval index, first
    sequence edge, last sequence edge) of ITS records in (interval, file) order; ranks hold increasing file ranges, so
    concatenating in rank order and stable-sorting by interval restores the reference's emission order
    (ibg:1303-1310: interval list order, then file order).  -> (iv, first, last), the same on every rank."""
    parts = comm.all_gather((np.asarray(iv, np.int64), np.asarray(first, np.int64), np.asarray(last, np.int64)))
    iv = np.concatenate([p[0] for p in parts])
    o = np.argsort(iv, kind="stable")
    return iv[o], np.concatenate([p[1] for p in parts])[o], np.concatenate([p[2] for p in parts])[o]


# ---- the engine --------------------------------------------------------------------------------------------------
class _ShardView(object):
    """This rank's slice of a record set: the interface the drop-in class needs from its BAM handle."""

    def __init__(self, parent, r0, r1):
        off = np.asarray(parent.rec_off, dtype=np.uint64)
        b0, b1 = int(off[r0]), int(off[r1])
        self.references = tuple(parent.references)
        self.rec = parent.rec[b0:b1]
        self.rec_off = (off[r0:r1 + 1] - np.uint64(b0)).astype(np.uint64)
        self.n_records = r1 - r0
        # bytes readable behind the slice: the next shard's records, or the parent's own slack after the last one
        self.rec_slack = 64 if b1 + 64 <= len(parent.rec) + int(getattr(parent, "rec_slack", 0)) else 0
        self.g0 = r0
        self._parent = parent

    def query_name(self, i):
        return self._parent.query_name(self.g0 + i)

    def query_names(self, idx):
        from .records import query_names
        return query_names(self.rec, self.rec_off, idx)

    def close(self):
        pass


def make_xchg(comm, device, slot_bytes=None, timeout_s=None):
    """This rank's exchange window over peer memory (csrc/xchg.cu), connected to the windows of all ranks of `comm`:
    the IPC handles travel once over the comm (ranks that are threads of one process hand over plain addresses)."""
    import os
    from . import _abi
    x = _abi.Xchg(device, comm.rank, comm.world, _abi.XCHG_DEFAULT_SLOTS if slot_bytes is None else slot_bytes)
    if timeout_s is not None:
        x.set_timeout(timeout_s)
    h, ptr = x.handle()
    allh = comm.all_gather((h, ptr, os.getpid()))
    same = [a[1] if (a[2] == os.getpid() and r != comm.rank) else 0 for r, a in enumerate(allh)]
    x.connect([a[0] for a in allh], same)
    comm.all_gather(0)                 # nobody exchanges before everybody has connected
    return x


def close_xchg(comm, x):
    comm.all_gather(0)                 # a peer may still be writing into this window
    x.close()


class _PeerTransport(object):
    """All-gather of device messages as two kernels over peer memory (NVLink); nothing visits the host."""
    name = "peer"

    def __init__(self, xchg):
        self.x = xchg

    def gather(self, ctx, channel, ptr, nbytes, lane=False):
        return ctx.xchg_allgather(self.x, channel, ptr, nbytes, lane)

    def read(self, ctx, gathered, lane=False):
        b, st, sz = gathered
        return [ctx.read_device(b + int(o), int(n), lane) for o, n in zip(st, sz)]

    def region_counts(self, eng, tid, start, stop):
        from . import _abi
        return eng._ctx.region_coverage_sharded(self.x, _abi.CH_COV, tid, start, stop)


class _NcclTransport(object):
    """The same exchanges through torch.distributed (NCCL on the box, gloo / thread ranks in tests): one collective per
    exchange issued by the host, sizes read back.  The baseline the peer-memory transport is measured against."""
    name = "nccl"

    def __init__(self, comm, device):
        from . import exchange
        self._ex = exchange
        self._comm = comm
        self._tc = exchange.TensorComm(comm)
        self._device = device
        self._keep = []

    def gather(self, ctx, channel, ptr, nbytes, lane=False):
        import torch
        buf, st, sz = self._tc.gather_msgs(self._ex.view(ptr, nbytes, "|u1", torch.device("cuda", self._device)))
        self._keep = (self._keep + [buf])[-8:]          # the gathered buffers must outlive their consumers
        return buf.data_ptr(), np.asarray(st, np.uint64), np.asarray(sz, np.uint64)

    def read(self, ctx, gathered, lane=False):
        b, st, sz = gathered
        for buf in self._keep:
            if buf.data_ptr() == b:
                raw = buf.cpu().numpy()
                return [raw[int(o):int(o) + int(n)] for o, n in zip(st, sz)]
        raise RuntimeError("gathered buffer no longer held")

    def region_counts(self, eng, tid, start, stop):
        bs, rc = eng._ctx.region_coverage(tid, start, stop)
        tot = self._comm.all_reduce_sum(np.concatenate([bs, rc]).astype(np.int64))
        return tot[:len(bs)], tot[len(bs):]


_SDEL_ROW = np.dtype([("rec", "<u4"), ("ai", "<i4"), ("next_start", "<i4"), ("prev_end", "<i4"), ("first_start", "<i4"),
                      ("last_end", "<i4"), ("mapq", "<i4"), ("pad", "<u4"), ("lo", "<i8"), ("hi", "<i8")])


def merge_smalldel_msgs(raws):
    """Gathered coral_smalldel_pack messages (one uint8 array per rank: 64-byte header whose first word is the row count,
    then 48-byte rows, csrc/chimeric_merge.cu SdelRow) -> rows (rank, rec, ai, next_start, prev_end, first_start, last_end,
    mapq, name hash lo, hi) in the reference's emission order (ibg:750-762): interval list order, then rank (= file) order,
    then the rank's own row order (file order, block order)."""
    out = []
    for r, raw in enumerate(raws):
        raw = np.ascontiguousarray(raw).view(np.uint8)
        n = int(np.frombuffer(raw[:8].tobytes(), np.uint64)[0])
        body = np.frombuffer(raw[64:64 + 48 * n].tobytes(), _SDEL_ROW)
        out.extend((r, int(x["rec"]), int(x["ai"]), int(x["next_start"]), int(x["prev_end"]), int(x["first_start"]),
                    int(x["last_end"]), int(x["mapq"]), int(x["lo"]), int(x["hi"])) for x in body)
    out.sort(key=lambda x: x[2])          # stable
    return out


def make_sharded_engine(records, seedfile, comm, device=0, placement="auto", presharded=False, xchg=None):
    """Factory (keeps the CUDA-bound import out of the CPU tests of the exchange logic).
    `records`: the whole file (every rank then takes its range of shard_ranges) or, with presharded=True, this rank's
    own shard — ranks must hold increasing, contiguous ranges of one coordinate-sorted file.
    `xchg`: a connected exchange window (make_xchg): the exchanges then run as kernels over peer memory; without it they
    go through torch.distributed collectives."""
    from . import _abi
    from .infer_breakpoint_graph import bam_to_breakpoint_nanopore

    class ShardedEngine(bam_to_breakpoint_nanopore):
        """bam_to_breakpoint_nanopore over one shard of the file + the exchanges of SURVEY.md §8e.  Same public
        methods, same call order; every rank ends with the same containers as the 1-GPU engine."""

        _auto_mapped = "mapped"      # every shard's scan is joined at load: the exchanges need its columns at once

        def __init__(self):
            self._comm = comm
            if presharded:
                view = records
            else:
                r0, r1 = shard_ranges(records.rec_off, comm.world)[comm.rank]
                view = _ShardView(records, r0, r1)
            bam_to_breakpoint_nanopore.__init__(self, None, seedfile, device=device, records=view, placement=placement)
            self._cctx = _abi.Context(device)          # rows-only ctx: the gathered fetch rows live here
            self._cctx.set_references(view.references)
            # distributed fetch: every rank parses the reads of its residue class of the name hash, the partial tables are
            # gathered and merged; the rows-only ctx runs on the second lane of the records ctx so that, inside a
            # pipelined pass, its kernels are ordered with the packing kernels that feed them
            self._cctx.set_option("fetch_part_world", comm.world)
            self._cctx.set_option("fetch_part_rank", comm.rank)
            self._cctx.set_option("pass_fork", 0)
            self._cctx.adopt_stream(self._ctx.stream_handle(1))
            self._cctx._stream_owner = self._ctx       # the adopted stream must outlive the ctx that runs on it
            self._tp = _PeerTransport(xchg) if xchg is not None else _NcclTransport(comm, device)
            self.materialise = True  # False: the replicated result tables stay on the device (ranks that only compute)
            self._row_names = None
            self._rbt = None
            self._edge_arrays = None
            self._name_hash = {}     # read name -> 128-bit hash as two int64 (for the rbps rows of the concordant combine)
            self.fetch_laps = {}

        @property
        def _chim_ctx(self):
            return self._cctx

        # a12 / a10: additive over records -> integer sum of the shard accumulators
        def _region_counts(self, tid, start, stop):
            return self._tp.region_counts(self, tid, start, stop)

        def _views(self):
            """Tensor views of this shard's name-hash columns (device memory owned by the ctx)."""
            import torch
            from . import exchange
            p = self._ctx.device_view()
            n = self._ctx.n_records
            dev = torch.device("cuda", device)
            return {"name_lo": exchange.view(p[2], n, "<i8", dev), "name_hi": exchange.view(p[3], n, "<i8", dev)}

        # a1: the exchange, on the GPU.  Two packed messages per rank (csrc/exchange.cu) are gathered and unpacked into
        # the rows-only ctx, every rank parses its share of the reads, the partial tables are gathered and merged
        # (csrc/chimeric_merge.cu): nothing visits the host.
        def _run_fetch(self):
            import time
            ctx, cctx, tp = self._ctx, self._cctx, self._tp
            t = [time.perf_counter()]

            def lap(name):      # host wall time of the steps of the exchange (bench.py reports them)
                t.append(time.perf_counter())
                self.fetch_laps[name] = self.fetch_laps.get(name, 0.0) + (t[-1] - t[-2]) * 1e3
            m1 = ctx.pack_sa_rows(); lap("pack_sa_rows")
            b1, st1, _ = tp.gather(ctx, _abi.CH_SA, *m1, lane=True); lap("gather_1")
            m2 = ctx.pack_extra_rows(b1, st1); lap("pack_extra_rows")
            b2, st2, _ = tp.gather(ctx, _abi.CH_EXTRA, *m2, lane=True); lap("gather_2")
            _, (s1, s2, nn, zd) = cctx.unpack_rows(b1, st1, b2, st2); lap("unpack_rows")
            f = cctx.fetch(read_back=False); lap("fetch_own_reads")
            m3 = cctx.pack_chimeric(); lap("pack_chimeric")
            b3, st3, _ = tp.gather(cctx, _abi.CH_CHIM, *m3); lap("gather_3")
            nr, na = cctx.merge_chimeric(b3, st3); lap("merge_chimeric")
            f["n_reads"], f["n_aln"] = nr, na
            if self.materialise:
                cctx.get_chimeric(nr, na, f); lap("read_back")
            # nm_stats (ibg:153-157) run over ALL records: the shard partials are summed in rank order (the f64 sum
            # order differs from the 1-GPU order: within 1e-9 relative, unused downstream)
            def nm_totals():
                f["nm_sum"] = float(s1.sum()); f["nm_sumsq"] = float(s2.sum())
                f["nm_n"] = int(nn.sum()); f["nm_zero_div"] = int(zd.sum())
            nm_totals()
            if cctx._pass_keep is not None:     # inside a pipelined pass fetch() fills its own (rows-only: empty) sums at
                cctx._pass_fin.append(nm_totals)    # pass_end: the totals of the shards go in after that
            self._row_names = None

            def names_of(recs):
                if self._row_names is None:         # one D2H of the gathered text, only when names are materialised
                    import torch
                    from . import exchange
                    dev = torch.device("cuda", device)
                    p = cctx.device_view()
                    n = cctx.n_records
                    off = exchange.view(p[1], n + 1, "<i8", dev).cpu().numpy()
                    text = exchange.view(p[0], int(off[-1]), "|u1", dev).cpu().numpy()
                    self._row_names = exchange.row_names(text, off[:-1])
                    lo = exchange.view(p[2], n, "<i8", dev).cpu().tolist()
                    hi = exchange.view(p[3], n, "<i8", dev).cpu().tolist()
                    self._name_hash.update(zip(self._row_names, zip(lo, hi)))
                return [self._row_names[int(r)] for r in recs]
            return f, names_of

        def find_breakpoints(self):
            self._cctx.set_params(min_del_len=self.min_del_len, min_bp_match_cutoff=self.min_bp_match_cutoff_)
            bam_to_breakpoint_nanopore.find_breakpoints(self)

        # a3: local gap rows (+ the name hash of each), packed on the device, gathered, merged in emission order
        def _smalldel_gather(self, t, s, e):
            """-> (local rows dict, list of per-rank raw messages; filled at pass_end inside a pipelined pass)."""
            rows = self._ctx.smalldel(t, s, e)
            g = self._tp.gather(self._ctx, _abi.CH_SDEL, *self._ctx.smalldel_pack())
            return rows, self._tp.read(self._ctx, g)

        @staticmethod
        def _smalldel_merge(raws):
            return merge_smalldel_msgs(raws)

        def _smalldel_rows(self, t, s, e):
            rows, raws = self._smalldel_gather(t, s, e)
            merged = self._smalldel_merge(raws)
            # names: every rank knows the names of its own records; tens of rows
            mine = sorted(set(x[1] for x in merged if x[0] == self._comm.rank))
            local = dict(zip(mine, self._names_of(np.array(mine, np.int64)))) if mine else {}
            table = {}
            for r, part in enumerate(self._comm.all_gather(local)):
                for rec, nm in part.items():
                    table[(r, rec)] = nm
            out = []
            for x in merged:
                nm = table[(x[0], x[1])]
                self._name_hash[nm] = (x[8], x[9])
                out.append((nm, ) + x[2:8])
            return out

        # a11: name-level set algebra across shards, on name hashes: the local stab rows are packed on the device as
        # (edge, name hash, mask), gathered, and counted by the same kernels on every rank
        def _concordant_counts(self, uniq, rbps):
            import torch
            ctx = self._ctx
            if self._edge_arrays is None or self._edge_arrays[0] is not uniq:
                self._edge_arrays = (uniq, np.array([u[0] for u in uniq], np.int32), np.array([u[1] for u in uniq], np.int32))
            ctx.concordant_rows_only(self._edge_arrays[1], self._edge_arrays[2])
            if self._rbt is None or self._rbt[0] is not rbps:      # (edge, name hash, 16) rows of rbps: built once per graph
                rb = [(k, ) + self._name_hash[nm] + (16, ) for k, names in enumerate(rbps) for nm in names if nm in self._name_hash]
                self._rbt = (rbps, torch.as_tensor(np.array(rb, np.int64).reshape(-1, 4), device=torch.device("cuda", device)))
                torch.cuda.current_stream().synchronize()
            gathered = self._tp.gather(ctx, _abi.CH_CONC, *ctx.concordant_pack())
            rbt = self._rbt[1]
            cnt = ctx.count_name_msgs(len(uniq), gathered[0], gathered[1], gathered[2], rbt.data_ptr(), int(rbt.shape[0]))
            return cnt, gathered

        def _concordant_support(self, uniq, rbps):
            cnt, gathered = self._concordant_counts(uniq, rbps)
            # ce[9] wants the read NAMES covering p or p+1: every rank contributes the names of its own records
            # (its message rows are in the order of its stab rows)
            raws = self._tp.read(self._ctx, gathered)
            rows = self._ctx.concordant_rows()
            mine = np.frombuffer(raws[self._comm.rank][64:].tobytes(), np.int64).reshape(-1, 4)
            sel = np.nonzero((rows["mask"] & 3) != 0)[0]
            urec, first = np.unique(rows["rec"][sel], return_index=True)
            keys = mine[sel[first]]
            table = {}
            for part in self._comm.all_gather((keys[:, 1].tolist(), keys[:, 2].tolist(), self._names_of(urec.astype(np.int64)))):
                table.update(zip(zip(part[0], part[1]), part[2]))
            self._ce_hash = table          # (hash lo, hi) -> name over every ce[9]: compute_path_constraints' name set
            out = [set() for _ in uniq]
            for raw in raws:
                g = np.frombuffer(raw[64:].tobytes(), np.int64).reshape(-1, 4)
                g = g[(g[:, 3] & 3) != 0]
                for e, lo, hi in g[:, :3].tolist():
                    out[e].add(table[(lo, hi)])
            return [int(x) for x in cnt], out

        # §8f-2: the name set spans shards, so it travels as name hashes; every rank runs the pass on its own records,
        # the rows are gathered in rank (= file) order and put back in (interval, file) order
        def _concordant_read_runs(self, amplicon_idx, intervals, concordant_reads):
            if not intervals or not concordant_reads:
                return []
            hash_of = {nm: h for h, nm in getattr(self, "_ce_hash", {}).items()}
            hash_of.update({nm: h for nm, h in self._name_hash.items() if nm not in hash_of})
            keys = np.array([hash_of[nm] for nm in concordant_reads], np.int64).reshape(-1, 2)
            iv, first, last = self._path_rows(self.lr_graph[amplicon_idx], intervals, name_lo=keys[:, 0].view(np.uint64),
                                              name_hi=keys[:, 1].view(np.uint64))
            _, first, last = merge_path_rows(self._comm, iv, first, last)
            return self._distinct_runs(first, last)

        def closebam(self):
            self._cctx.close()
            bam_to_breakpoint_nanopore.closebam(self)

    return ShardedEngine()
