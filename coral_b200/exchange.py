"""Tensor form of the multi-GPU exchanges: the selection of the rows a rank contributes, the
all-gather and the combine all run on torch tensors — CUDA tensors over NCCL/NVLink on the GPU box (the rows never
visit the host), CPU tensors over gloo in the tests.  `sharded.py` holds the plain numpy statement of the same
exchanges; tests check that both gather the same rows.

PyTorch is plumbing here: device memory views, sort/unique/scatter, torch.distributed.  The arithmetic of the path
stays in the CUDA kernels behind the C ABI.
"""
import numpy as np
import torch

_BIG = 1 << 62


class DevPtr(object):
    """A raw device address as a __cuda_array_interface__ object, so torch can view it without a copy."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def view(ptr, n, typestr, device):
    if n == 0:
        return torch.zeros(0, dtype={"|u1": torch.uint8, "<i8": torch.int64, "<i4": torch.int32}[typestr], device=device)
    return torch.as_tensor(DevPtr(ptr, (n,), typestr), device=device)


# ---- variable-length all-gather of tensors -------------------------------------------------------------------
class TensorComm(object):
    """all_gather of variable-length tensors (dim 0) + integer all-reduce over a torch.distributed group, or over
    in-process thread ranks (sharded.ThreadComm), or a single rank."""

    def __init__(self, comm):
        self._c = comm
        self.rank, self.world = comm.rank, comm.world
        self._dist = getattr(comm, "_dist", None)
        self._group = getattr(comm, "_group", None)

    def all_gather(self, t):
        """t: tensor [n, ...] (same trailing shape and dtype on every rank) -> list of the ranks' tensors."""
        if self.world == 1:
            return [t]
        if self._dist is None:                      # thread ranks share the process: hand the tensors over as they are
            return self._c.all_gather(t)
        n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
        sizes = torch.zeros(self.world, dtype=torch.int64, device=t.device)
        self._dist.all_gather_into_tensor(sizes, n, group=self._group)
        sizes = sizes.tolist()
        m = max(sizes)
        pad = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[:t.shape[0]] = t
        out = torch.empty((self.world * m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        self._dist.all_gather_into_tensor(out, pad.contiguous(), group=self._group)
        return [out[r * m:r * m + sizes[r]] for r in range(self.world)]

    def all_reduce_sum(self, t):
        if self.world == 1:
            return t
        if self._dist is None:
            return torch.as_tensor(self._c.all_reduce_sum(t.cpu().numpy())).to(t.device)
        t = t.clone()
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self._group)
        return t


def _member_pos(set_lo, set_hi, lo, hi):
    """Index in the (lo, hi)-sorted set of every query pair, or -1.  Exact also when low words collide."""
    n = set_lo.shape[0]
    out = torch.full_like(lo, -1)
    if n == 0 or lo.shape[0] == 0:
        return out
    pos = torch.searchsorted(set_lo, lo)
    mult = 1
    if n > 1 and bool((set_lo[1:] == set_lo[:-1]).any()):
        mult = int(torch.unique_consecutive(set_lo, return_counts=True)[1].max())
    for k in range(mult):
        p = (pos + k).clamp(max=n - 1)
        ok = (out < 0) & (set_lo[p] == lo) & (set_hi[p] == hi)
        out = torch.where(ok, p, out)
    return out


def _gather_bytes(recs, starts, lens):
    """Concatenation of recs[starts[i] : starts[i] + lens[i]] for all i, as one uint8 tensor."""
    total = int(lens.sum())
    if total == 0:
        return torch.zeros(0, dtype=torch.uint8, device=recs.device)
    offs = torch.cumsum(lens, 0) - lens
    idx = torch.repeat_interleave(starts - offs, lens) + torch.arange(total, device=recs.device)
    return recs[idx]


def fetch_exchange(tc, tv):
    """Tensor form of sharded.exchange_fetch_rows.  tv: this rank's view, tensors on one device —
         recs uint8, off int64[n+1], name_lo/name_hi int64[n] (bit patterns of the uint64 hashes),
         flag_mapq, l_seq, nm, sa_off, sa_len int32[n].
    Returns the gathered rows (same on every rank, global file order): dict of tensors
         name_lo, name_hi int64; flag_mapq, l_seq, nm, sa_off, sa_len int32; text uint8; text_off int64[m+1];
         name_len int64 (bytes of the NUL-terminated name at the start of each text slot)."""
    recs, off = tv["recs"], tv["off"]
    lo, hi = tv["name_lo"], tv["name_hi"]
    flag = tv["flag_mapq"].to(torch.int64) & 0xffff
    has_sa = tv["sa_off"] != 0
    dev = lo.device
    # round 1: names that carry an SA tag anywhere in the file
    sa_idx = torch.nonzero(has_sa).flatten()
    allp = torch.cat(tc.all_gather(torch.stack([lo[sa_idx], hi[sa_idx]], 1)), 0)
    uniq = torch.unique(allp, dim=0) if allp.shape[0] else allp
    set_lo, set_hi = uniq[:, 0].contiguous(), uniq[:, 1].contiguous()
    # round 2: + the first local record with flag < 256 of each such name
    cand = torch.nonzero((flag < 256) & ~has_sa).flatten()
    grp = _member_pos(set_lo, set_hi, lo[cand], hi[cand])
    keep = grp >= 0
    cand, grp = cand[keep], grp[keep]
    if cand.shape[0]:
        first = torch.full((set_lo.shape[0],), _BIG, dtype=torch.int64, device=dev).scatter_reduce(0, grp, cand, "amin")
        prim = first[first < _BIG]
    else:
        prim = cand
    rows = torch.sort(torch.cat([sa_idx, prim]))[0]
    # text slot of a row = NUL-terminated read name + SA text
    base = off[rows]
    name_len = recs[base + 12].to(torch.int64)
    sa_off = tv["sa_off"][rows].to(torch.int64)
    sa_len = tv["sa_len"][rows].to(torch.int64)
    starts = torch.stack([base + 36, base + sa_off], 1).flatten()
    lens = torch.stack([name_len, sa_len], 1).flatten()
    blob = _gather_bytes(recs, starts, lens)
    cols = torch.stack([lo[rows], hi[rows], tv["flag_mapq"][rows].to(torch.int64), tv["l_seq"][rows].to(torch.int64),
                        tv["nm"][rows].to(torch.int64), name_len, sa_len], 1)
    g_cols = torch.cat(tc.all_gather(cols), 0)
    g_text = torch.cat(tc.all_gather(blob), 0)
    slot = g_cols[:, 5] + g_cols[:, 6]
    text_off = torch.zeros(g_cols.shape[0] + 1, dtype=torch.int64, device=dev)
    text_off[1:] = torch.cumsum(slot, 0)
    has = g_cols[:, 6] > 0
    return {"name_lo": g_cols[:, 0].contiguous(), "name_hi": g_cols[:, 1].contiguous(),
            "flag_mapq": g_cols[:, 2].to(torch.int32), "l_seq": g_cols[:, 3].to(torch.int32), "nm": g_cols[:, 4].to(torch.int32),
            "sa_off": torch.where(has, g_cols[:, 5], torch.zeros_like(slot)).to(torch.int32), "sa_len": g_cols[:, 6].to(torch.int32),
            "text": g_text, "text_off": text_off, "name_len": g_cols[:, 5].contiguous()}


def row_names(rows):
    """Host side: read names of the gathered rows (one D2H of the text)."""
    text = rows["text"].cpu().numpy().tobytes()
    off = rows["text_off"].cpu().numpy()
    ln = rows["name_len"].cpu().numpy()
    return [text[int(off[i]):int(off[i]) + int(ln[i]) - 1].decode() for i in range(len(ln))]


def _lexsort_rows(keys):
    """Row order that sorts `keys` [n, c] lexicographically (stable radix passes, last column first)."""
    o = torch.arange(keys.shape[0], device=keys.device)
    for c in range(keys.shape[1] - 1, -1, -1):
        o = o[torch.sort(keys[o, c], stable=True)[1]]
    return o


def _or_reduce(keys, mask):
    """Unique rows of `keys` (sorted) with the bitwise OR (4 bits) of their masks."""
    o = _lexsort_rows(keys)
    k, m = keys[o], mask[o]
    head = torch.ones(k.shape[0], dtype=torch.bool, device=k.device)
    head[1:] = (k[1:] != k[:-1]).any(1)
    seg = torch.cumsum(head.to(torch.int64), 0) - 1
    n = int(seg[-1]) + 1
    out = torch.zeros(n, dtype=torch.int64, device=k.device)
    for b in range(4):
        out |= torch.zeros_like(out).scatter_reduce(0, seg, (m >> b) & 1, "amax") << b
    return k[head], out


def concordant_combine(tc, n_edges, edge, lo, hi, mask, rb_edge, rb_lo, rb_hi):
    """Tensor form of sharded.combine_concordant on name HASHES: local rows (edge, name hash, mask) are OR-reduced
    per (edge, name), all-gathered, reduced again; ce[8] = #names with all four points covered that are not in
    rbps (given as (edge, name hash) rows, identical on every rank).
    Returns (count int64[n_edges], keys int64[m, 3] = (edge, lo, hi), mask int64[m]) — same on every rank."""
    dev = edge.device
    keys = torch.stack([edge, lo, hi], 1)
    if keys.shape[0]:
        keys, mask = _or_reduce(keys, mask)
    g = torch.cat(tc.all_gather(torch.cat([keys, mask.reshape(-1, 1)], 1)), 0)
    if g.shape[0]:
        keys, mask = _or_reduce(g[:, :3].contiguous(), g[:, 3].contiguous())
    else:
        keys, mask = g[:, :3], g[:, 3]
    full = mask == 15
    if rb_edge.shape[0] and keys.shape[0]:
        # membership of every key in the rbps rows: sort both together, a key is hit when an rbps row shares its run
        nk = keys.shape[0]
        both = torch.cat([keys, torch.stack([rb_edge, rb_lo, rb_hi], 1)], 0)
        is_rb = torch.zeros(both.shape[0], dtype=torch.int64, device=dev)
        is_rb[nk:] = 1
        o = _lexsort_rows(both)
        b = both[o]
        head = torch.ones(b.shape[0], dtype=torch.bool, device=dev)
        head[1:] = (b[1:] != b[:-1]).any(1)
        seg = torch.cumsum(head.to(torch.int64), 0) - 1
        seg_hit = torch.zeros(int(seg[-1]) + 1, dtype=torch.int64, device=dev).scatter_reduce(0, seg, is_rb[o], "amax")
        hit = torch.zeros(both.shape[0], dtype=torch.bool, device=dev)
        hit[o] = seg_hit[seg] > 0
        full = full & ~hit[:nk]
    cnt = torch.bincount(keys[:, 0][full], minlength=n_edges) if keys.shape[0] else torch.zeros(n_edges, dtype=torch.int64, device=dev)
    return cnt, keys, mask
