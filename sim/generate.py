"""Synthetic long-read BAM generator for the BASELINE.json configs.

Geometry (where each read's pieces land on the reference) is planned here with numpy; the
per-record bytes (CIGAR from an ONT/HiFi error model, SEQ, QUAL, aux tags incl. minimap2-dialect
``SA:Z``) are produced by ``sim/simgen.cpp``.  Everything is deterministic in ``seed``.

The generator emits *decompressed BAM alignment records* (the form the C-ABI consumes) and can
also wrap them into a coordinate-sorted BGZF ``.bam`` plus the ``cn_seg.bed`` / ``seed.bed`` text
inputs of reference call sites ``infer_breakpoint_graph.py:64-69`` and ``:75-98``.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_simgen.so")

CHR_SIZES = {
    'chr1': 248956422, 'chr2': 242193529, 'chr3': 198295559, 'chr4': 190214555, 'chr5': 181538259,
    'chr6': 170805979, 'chr7': 159345973, 'chr8': 145138636, 'chr9': 138394717, 'chr10': 133797422,
    'chr11': 135086622, 'chr12': 133275309, 'chr13': 114364328, 'chr14': 107043718, 'chr15': 101991189,
    'chr16': 90338345, 'chr17': 83257441, 'chr18': 80373285, 'chr19': 58617616, 'chr20': 64444167,
    'chr21': 46709983, 'chr22': 50818468, 'chrX': 156040895, 'chrY': 57227415,
}
CHR_ORDER = ['chr%d' % i for i in range(1, 23)] + ['chrX', 'chrY']

PIECE_DT = np.dtype([
    ("read_id", "<i4"), ("tid", "<i4"), ("ref_start", "<i4"), ("ref_len", "<i4"),
    ("strand", "u1"), ("mapq", "u1"), ("dropped", "u1"), ("pad", "u1"),
    ("forced_beg", "<i4"), ("forced_end", "<i4"),
])


class _Params(ctypes.Structure):
    _fields_ = [("p_mis", ctypes.c_double), ("p_ins", ctypes.c_double), ("p_del", ctypes.c_double),
                ("indel_mean", ctypes.c_double), ("p_n", ctypes.c_double), ("dense_del_boost", ctypes.c_double),
                ("seed", ctypes.c_uint64), ("hard_clip_supp", ctypes.c_int32), ("record_align", ctypes.c_int32),
                ("n_threads", ctypes.c_int32), ("name_style", ctypes.c_int32)]


def build_simgen(force: bool = False) -> str:
    src = os.path.join(_HERE, "simgen.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-pthread", "-std=c++17", src, "-o", _SO])
    return _SO


_lib = None


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build_simgen())
        _lib.sim_write.restype = ctypes.c_int64
    return _lib


@dataclass
class Cycle:
    segments: list  # [(chr, start, end_exclusive, '+'|'-')]
    cn: float


@dataclass
class SimConfig:
    name: str
    refs: list                      # [(chr, length)] in header order
    cn_rows: list                   # [(chr, start, end_exclusive, cn)]
    seeds: list                     # [(chr, start, end_inclusive)]
    background: list                # [(chr, start, end_exclusive)] covered at `cov`
    cov: float                      # coverage of a CN-2 region
    cycles: list = field(default_factory=list)
    deletions: list = field(default_factory=list)   # [(chr, a, b, carrier_fraction)]
    read_mean: int = 20000
    model: str = "ont"              # 'ont' | 'hifi'
    noise_frac: float = 0.0         # random inter-region chimeric reads (fraction of background reads)
    p_n: float = 0.0
    dense_del_boost: float = 0.0
    low_mapq_frac: float = 0.02
    seed: int = 20240215
    name_style: int = 0
    min_piece: int = 200


@dataclass
class SimData:
    cfg: SimConfig
    rec: np.ndarray        # uint8 record bytes (records start at rec_off[i], 16-byte aligned)
    rec_off: np.ndarray    # uint64 [n+1]
    n_reads: int
    aligned_bases: int     # sum of M ops
    pieces: np.ndarray
    qlen: np.ndarray
    read_len: np.ndarray

    @property
    def n_records(self):
        return len(self.rec_off) - 1


def _read_lengths(rng, n, cfg):
    if cfg.model == "hifi":
        ln = rng.normal(cfg.read_mean, 0.1 * cfg.read_mean, n)
        return np.clip(ln, 2000, 2 * cfg.read_mean).astype(np.int64)
    ln = rng.gamma(4.0, cfg.read_mean / 4.0, n)
    return np.clip(ln, 1000, 4 * cfg.read_mean).astype(np.int64)


def _plan_pieces(cfg: SimConfig, rng):
    tid_of = {c: i for i, (c, _) in enumerate(cfg.refs)}
    cols = {k: [] for k in ("read_id", "k", "tid", "ref_start", "ref_len", "strand")}
    next_read = 0

    def emit(rid, k, tid, rs, rl, st):
        cols["read_id"].append(rid); cols["k"].append(k); cols["tid"].append(tid)
        cols["ref_start"].append(rs); cols["ref_len"].append(rl); cols["strand"].append(st)

    n_bg_total = 0
    for (c, s, e) in cfg.background:
        n = int((e - s) * cfg.cov / cfg.read_mean)
        if n <= 0:
            continue
        ln = _read_lengths(rng, n, cfg)
        st = rng.integers(s, max(s + 1, e - 1000), n)
        ln = np.minimum(ln, e - st)
        rid = np.arange(next_read, next_read + n)
        emit(rid, np.zeros(n, np.int64), np.full(n, tid_of[c]), st, ln, rng.integers(0, 2, n))
        next_read += n
        n_bg_total += n
    for cyc in cfg.cycles:
        segs = cyc.segments
        lens = np.array([e - s for (_, s, e, _) in segs], np.int64)
        cum = np.concatenate([[0], np.cumsum(lens)])
        Lc = int(cum[-1])
        n = int(Lc * cyc.cn * cfg.cov / 2.0 / cfg.read_mean)
        if n <= 0:
            continue
        ln = np.minimum(_read_lengths(rng, n, cfg), Lc - 1)
        u = rng.integers(0, Lc, n)
        rev = rng.integers(0, 2, n)  # read sequenced from the opposite strand of the cycle
        seg = np.searchsorted(cum, u, side="right") - 1
        pos = u.copy(); rem = ln.copy()
        rid = np.arange(next_read, next_read + n)
        seg_s = np.array([s for (_, s, _, _) in segs]); seg_e = np.array([e for (_, _, e, _) in segs])
        seg_t = np.array([tid_of[c] for (c, _, _, _) in segs]); seg_m = np.array([1 if o == '-' else 0 for (_, _, _, o) in segs])
        tmp = []
        for k in range(2 * len(segs) + 2):
            act = rem > 0
            if not act.any():
                break
            take = np.where(act, np.minimum(rem, cum[seg + 1] - pos), 0)
            off = pos - cum[seg]
            rs = np.where(seg_m[seg] == 0, seg_s[seg] + off, seg_e[seg] - off - take)
            tmp.append((rid[act], np.full(act.sum(), k), seg_t[seg][act], rs[act], take[act], seg_m[seg][act]))
            pos = pos + take; rem = rem - take
            adv = act & (pos >= cum[seg + 1])
            seg = np.where(adv, (seg + 1) % len(segs), seg)
            pos = np.where(pos >= Lc, pos - Lc, pos)
        for (r_, k_, t_, rs_, rl_, st_) in tmp:
            # opposite-strand reads traverse the same pieces in reverse order with flipped strands
            rv = rev[r_ - next_read] == 1
            emit(r_, np.where(rv, 1000 - k_, k_), t_, rs_, rl_, np.where(rv, 1 - st_, st_))
        next_read += n
    n_noise = int(n_bg_total * cfg.noise_frac)
    if n_noise > 0 and cfg.background:
        reg = np.array([(tid_of[c], s, e) for (c, s, e) in cfg.background], np.int64)
        w = (reg[:, 2] - reg[:, 1]).astype(float); w /= w.sum()
        rid = np.arange(next_read, next_read + n_noise)
        for k in range(2):
            ri = rng.choice(len(reg), n_noise, p=w)
            ln = np.maximum(_read_lengths(rng, n_noise, cfg) // 2, 500)
            st = rng.integers(reg[ri, 1], np.maximum(reg[ri, 1] + 1, reg[ri, 2] - 1000))
            ln = np.minimum(ln, reg[ri, 2] - st)
            emit(rid, np.full(n_noise, k), reg[ri, 0], st, ln, rng.integers(0, 2, n_noise))
        next_read += n_noise
    arr = {k: np.concatenate([np.asarray(x, np.int64) for x in v]) if v else np.zeros(0, np.int64) for k, v in cols.items()}
    order = np.lexsort((arr["k"], arr["read_id"]))
    for k in arr:
        arr[k] = arr[k][order]
    return arr, next_read


def simulate(cfg: SimConfig, n_threads: int = 0, record_align: int = 16, alloc=None) -> SimData:
    """`alloc(nbytes) -> uint8 ndarray` lets the caller provide the record buffer (e.g. pinned host memory)."""
    lib = _load()
    rng = np.random.Generator(np.random.PCG64(cfg.seed))
    arr, n_reads = _plan_pieces(cfg, rng)
    npc = len(arr["read_id"])
    pieces = np.zeros(npc, PIECE_DT)
    for k in ("read_id", "tid", "ref_start", "ref_len", "strand"):
        pieces[k] = arr[k]
    low = rng.random(npc) < cfg.low_mapq_frac
    pieces["mapq"] = np.where(low, rng.integers(0, 20, npc), 60)
    pieces["dropped"] = (pieces["ref_len"] < cfg.min_piece).astype(np.uint8)
    # forced in-CIGAR deletions
    f_piece, f_off, f_len = [], [], []
    tid_of = {c: i for i, (c, _) in enumerate(cfg.refs)}
    rs = pieces["ref_start"].astype(np.int64); re_ = rs + pieces["ref_len"]
    for (c, a, b, frac) in cfg.deletions:
        hit = (pieces["tid"] == tid_of[c]) & (rs + 300 <= a) & (b + 300 <= re_) & (pieces["dropped"] == 0)
        idx = np.nonzero(hit)[0]
        idx = idx[rng.random(len(idx)) < frac]
        f_piece.append(idx); f_off.append(a - rs[idx]); f_len.append(np.full(len(idx), b - a))
    if f_piece:
        fp = np.concatenate(f_piece); fo = np.concatenate(f_off); fl = np.concatenate(f_len)
        o = np.lexsort((fo, fp)); fp, fo, fl = fp[o], fo[o], fl[o]
        # drop forced deletions overlapping the previous one within the same piece
        keep = np.ones(len(fp), bool)
        prev_end = -1; prev_p = -1
        for i in range(len(fp)):
            if fp[i] == prev_p and fo[i] < prev_end + 50:
                keep[i] = False
                continue
            prev_p = fp[i]; prev_end = fo[i] + fl[i]
        fp, fo, fl = fp[keep], fo[keep], fl[keep]
        beg = np.searchsorted(fp, np.arange(npc), side="left"); end = np.searchsorted(fp, np.arange(npc), side="right")
        pieces["forced_beg"] = beg; pieces["forced_end"] = end
    else:
        fo = np.zeros(1, np.int64); fl = np.zeros(1, np.int64)
    f_off_a = np.ascontiguousarray(fo, dtype=np.int32); f_len_a = np.ascontiguousarray(fl, dtype=np.int32)

    if cfg.model == "hifi":
        p_mis, p_ins, p_del = 0.001, 0.002, 0.002
    else:
        p_mis, p_ins, p_del = 0.02, 0.015, 0.015
    prm = _Params(p_mis, p_ins, p_del, 2.0, cfg.p_n, cfg.dense_del_boost, cfg.seed, 1, record_align,
                  n_threads or (os.cpu_count() or 1), cfg.name_style)
    qlen = np.zeros(npc, np.uint32); nm = np.zeros(npc, np.uint32); nops = np.zeros(npc, np.uint32)
    qs = np.zeros(npc, np.uint32); read_len = np.zeros(max(n_reads, 1), np.uint32)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    lib.sim_plan(P(pieces), ctypes.c_int64(npc), ctypes.c_int64(n_reads), ctypes.byref(prm), P(f_off_a), P(f_len_a),
                 P(qlen), P(nm), P(nops), P(qs), P(read_len))
    live = np.nonzero(pieces["dropped"] == 0)[0]
    order = live[np.lexsort((live, pieces["ref_start"][live], pieces["tid"][live]))].astype(np.int64)
    names = b"".join(c.encode() + b"\x00" for c, _ in cfg.refs)
    name_off = np.zeros(len(cfg.refs), np.int32)
    o = 0
    for i, (c, _) in enumerate(cfg.refs):
        name_off[i] = o; o += len(c) + 1
    names_buf = ctypes.create_string_buffer(names, len(names))
    rec_off = np.zeros(len(order) + 1, np.uint64)
    args = [P(pieces), ctypes.c_int64(npc), ctypes.byref(prm), P(f_off_a), P(f_len_a), P(qlen), P(nm), P(qs),
            P(read_len), P(order), ctypes.c_int64(len(order)), names_buf, P(name_off)]
    total = lib.sim_write(*args, None, P(rec_off))
    rec = alloc(max(int(total), 1) + 64) if alloc is not None else np.zeros(max(int(total), 1), np.uint8)
    lib.sim_write(*args, P(rec), P(rec_off))
    return SimData(cfg, rec[:int(total)], rec_off, n_reads, 0, pieces, qlen, read_len)


def write_dataset(data: SimData, outdir: str, level: int = 1) -> dict:
    """Write <outdir>/reads.bam, cn_seg.bed, seed.bed; return their paths."""
    from .bgzf import write_bam
    os.makedirs(outdir, exist_ok=True)
    bam = os.path.join(outdir, "reads.bam")
    write_bam(bam, data.cfg.refs, data.rec, data.rec_off, level)
    cns = os.path.join(outdir, "cn_seg.bed")
    with open(cns, "w") as fp:
        for (c, s, e, cn) in data.cfg.cn_rows:
            fp.write("%s\t%d\t%d\t%g\n" % (c, s, e, cn))
    seed = os.path.join(outdir, "seed.bed")
    with open(seed, "w") as fp:
        for (c, s, e) in data.cfg.seeds:
            fp.write("%s\t%d\t%d\n" % (c, s, e))
    return {"bam": bam, "cn_seg": cns, "seed": seed}
