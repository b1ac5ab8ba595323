"""oracle/pysam_shim.py — TEST INFRASTRUCTURE (CPU oracle); never imported by the product.

A minimal stand-in for the parts of ``pysam`` that the reference's hot path calls.  ``pysam``
(htslib) is an un-vendored, unpinned dependency of the reference (``pyproject.toml:32``,
``requirements.txt:8``) and is not installable in this image, so its behaviour is restated here
from the SAM/BAM specification and pysam's documented semantics:
**parity unpinned against real pysam**.

Reference call sites served (``infer_breakpoint_graph.py``):
  :65   ``pysam.AlignmentFile(path, 'rb')``            :140  ``fetch()``  (all placed records, file order)
  :131/:1033 ``count_coverage(contig, start, stop, quality_threshold=0, read_callback='nofilter')``
  :141-154 ``query_name, flag, query_length, get_tag('SA:Z:'), mapping_quality, get_cigar_stats()``
  :751/:1031/:1043 ``fetch(contig, start, stop)``      :756  ``get_blocks()``   :1031 ``infer_read_length()``
  :1306-1310 ``mapq, reference_name, reference_start, reference_end``   :1330 ``close()``

Inner loops that are C inside pysam are C here too (``oracle/covlib.c``), so that timing the
reference-on-shim is not unfairly slow.
"""
import ctypes
import os
import struct
import subprocess
import zlib

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_covlib.so")
_lib = None


def build_covlib(force=False):
    src = os.path.join(_HERE, "covlib.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", src, "-o", _SO])
    return _SO


def _cov():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build_covlib())
        _lib.ora_count_coverage.restype = ctypes.c_int64
        _lib.ora_blocks.restype = ctypes.c_int64
    return _lib


def _P(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def inflate_bgzf(path):
    """Concatenate the inflated payloads of every BGZF member."""
    raw = open(path, "rb").read()
    out = []
    p = 0
    n = len(raw)
    while p < n:
        if raw[p:p + 4] != b"\x1f\x8b\x08\x04":
            raise ValueError("not a BGZF member at %d" % p)
        xlen = struct.unpack_from("<H", raw, p + 10)[0]
        q = p + 12
        bsize = None
        while q < p + 12 + xlen:
            si1, si2, slen = raw[q], raw[q + 1], struct.unpack_from("<H", raw, q + 2)[0]
            if si1 == 66 and si2 == 67:
                bsize = struct.unpack_from("<H", raw, q + 4)[0]
            q += 4 + slen
        if bsize is None:
            raise ValueError("BGZF member without BC field")
        payload = raw[p + 12 + xlen: p + bsize + 1 - 8]
        isize = struct.unpack_from("<I", raw, p + bsize + 1 - 4)[0]
        if isize:
            out.append(zlib.decompress(payload, -15))
        p += bsize + 1
    return b"".join(out)


def parse_bam(data):
    """-> (refs [(name, len)], record offsets uint64[n+1] into `data`)."""
    if data[:4] != b"BAM\x01":
        raise ValueError("not BAM")
    l_text = struct.unpack_from("<i", data, 4)[0]
    p = 8 + l_text
    n_ref = struct.unpack_from("<i", data, p)[0]
    p += 4
    refs = []
    for _ in range(n_ref):
        ln = struct.unpack_from("<i", data, p)[0]
        name = data[p + 4:p + 4 + ln - 1].decode()
        rl = struct.unpack_from("<i", data, p + 4 + ln)[0]
        refs.append((name, rl))
        p += 8 + ln
    offs = []
    n = len(data)
    unpack = struct.Struct("<i").unpack_from
    while p < n:
        offs.append(p)
        p += 4 + unpack(data, p)[0]
    offs.append(p)
    return refs, np.asarray(offs, dtype=np.uint64)


_AUX_FIXED = {ord('A'): 1, ord('c'): 1, ord('C'): 1, ord('s'): 2, ord('S'): 2, ord('i'): 4, ord('I'): 4, ord('f'): 4}
_AUX_FMT = {ord('c'): "<b", ord('C'): "<B", ord('s'): "<h", ord('S'): "<H", ord('i'): "<i", ord('I'): "<I", ord('f'): "<f"}


class AlignedSegment(object):
    __slots__ = ("_f", "_i", "_o")

    def __init__(self, f, i):
        self._f = f
        self._i = i
        self._o = int(f._off[i])

    # --- core fields -------------------------------------------------------------------------
    @property
    def query_name(self):
        d = self._f._data
        o = self._o
        return d[o + 36:o + 36 + d[o + 12] - 1].decode()

    @property
    def flag(self):
        return int(self._f._flag[self._i])

    @property
    def mapping_quality(self):
        return int(self._f._mapq[self._i])

    mapq = mapping_quality

    @property
    def query_length(self):
        return int(self._f._lseq[self._i])

    @property
    def reference_id(self):
        return int(self._f._tid[self._i])

    @property
    def reference_name(self):
        t = int(self._f._tid[self._i])
        return self._f.references[t] if t >= 0 else None

    @property
    def reference_start(self):
        return int(self._f._pos[self._i])

    @property
    def reference_end(self):
        # pysam: aend / reference_end is None for unmapped or CIGAR-less records
        if (self.flag & 4) or self._f._ncig[self._i] == 0:
            return None
        return int(self._f._end[self._i])

    # --- CIGAR-derived ------------------------------------------------------------------------
    def get_blocks(self):
        lib = _cov()
        buf = self._f._cbuf
        n = lib.ora_blocks(buf, ctypes.c_uint64(self._o), None)
        out = np.empty(2 * max(n, 1), np.int64)
        lib.ora_blocks(buf, ctypes.c_uint64(self._o), _P(out))
        v = out[:2 * n].tolist()
        return [(v[2 * k], v[2 * k + 1]) for k in range(n)]

    def infer_read_length(self):
        n = int(self._f._infer[self._i])
        return n if n > 0 else None

    def _aux_items(self):
        d = self._f._data
        o = self._o
        bs = struct.unpack_from("<i", d, o)[0]
        l_name = d[o + 12]
        ncig = struct.unpack_from("<H", d, o + 16)[0]
        lseq = struct.unpack_from("<I", d, o + 20)[0]
        p = o + 36 + l_name + 4 * ncig + (lseq + 1) // 2 + lseq
        end = o + 4 + bs
        while p + 3 <= end:
            tag = d[p:p + 2]
            ty = d[p + 2]
            p += 3
            if ty in _AUX_FIXED:
                sz = _AUX_FIXED[ty]
                yield tag, ty, p, sz
                p += sz
            elif ty == ord('Z') or ty == ord('H'):
                q = d.index(b"\x00", p)
                yield tag, ty, p, q - p
                p = q + 1
            elif ty == ord('B'):
                sub = d[p]
                cnt = struct.unpack_from("<i", d, p + 1)[0]
                sz = _AUX_FIXED[sub] * cnt
                yield tag, ty, p, 5 + sz
                p += 5 + sz
            else:
                raise ValueError("bad aux type %r" % ty)

    def get_tag(self, tag):
        """pysam matches on the first two characters of `tag` (bam_aux_get)."""
        want = tag[:2].encode()
        d = self._f._data
        for t, ty, p, sz in self._aux_items():
            if t == want:
                if ty == ord('Z') or ty == ord('H'):
                    return d[p:p + sz].decode()
                if ty == ord('A'):
                    return chr(d[p])
                if ty in _AUX_FMT:
                    return struct.unpack_from(_AUX_FMT[ty], d, p)[0]
                raise KeyError(tag)
        raise KeyError("tag '%s' not present" % tag)

    def has_tag(self, tag):
        try:
            self.get_tag(tag)
            return True
        except KeyError:
            return False

    def get_cigar_stats(self):
        """(base counts per op MIDNSHP=XB + NM tag value, block counts)."""
        d = self._f._data
        o = self._o
        l_name = d[o + 12]
        ncig = struct.unpack_from("<H", d, o + 16)[0]
        cig = np.frombuffer(d, dtype="<u4", count=ncig, offset=o + 36 + l_name)
        ops = (cig & 15).astype(np.int64)
        lens = (cig >> 4).astype(np.int64)
        base = np.bincount(ops, weights=lens, minlength=11)[:11].astype(np.int64).tolist()
        blocks = np.bincount(ops, minlength=11)[:11].astype(np.int64).tolist()
        try:
            nm = self.get_tag("NM")
            base[10] = int(nm)
        except KeyError:
            pass
        return base, blocks


_MEMORY_BAMS = {}


def register_memory_bam(key, refs, data, rec_off):
    """Make ``AlignmentFile("mem:<key>")`` open an already decoded record set instead of a file: `refs` [(name, length)],
    `data` = bytes holding the alignment records, `rec_off` = uint64[n+1] record offsets into it (any framing — the shim
    only ever addresses records through their offsets).  What bench.py's reference legs use at BASELINE.json's full
    size, so that the CPU arm starts where the GPU arm's e2e starts: from decoded records in host memory."""
    if not isinstance(data, bytes):
        data = bytes(data)
    _MEMORY_BAMS[key] = (list(refs), data, np.ascontiguousarray(rec_off, dtype=np.uint64))
    return "mem:" + key


class AlignmentFile(object):
    def __init__(self, path, mode="rb"):
        assert mode == "rb"
        self.filename = path
        if isinstance(path, str) and path.startswith("mem:"):
            self.refs, self._data, self._off = _MEMORY_BAMS[path[4:]]
        else:
            self._data = inflate_bgzf(path)
            self.refs, self._off = parse_bam(self._data)
        self.references = tuple(r[0] for r in self.refs)
        self.lengths = tuple(r[1] for r in self.refs)
        self._tid_of = {r: i for i, r in enumerate(self.references)}
        n = len(self._off) - 1
        self._n = n
        self._cbuf = ctypes.c_char_p(self._data)  # borrowed pointer; self._data keeps it alive
        self._tid = np.zeros(n, np.int32); self._pos = np.zeros(n, np.int32); self._end = np.zeros(n, np.int32)
        self._flag = np.zeros(n, np.int32); self._mapq = np.zeros(n, np.int32); self._lseq = np.zeros(n, np.int64)
        self._ncig = np.zeros(n, np.int32); self._infer = np.zeros(n, np.int64); self._mb = np.zeros(n, np.int64)
        if n:
            _cov().ora_index(self._cbuf, _P(self._off), ctypes.c_int64(n), _P(self._tid), _P(self._pos), _P(self._end),
                             _P(self._flag), _P(self._mapq), _P(self._lseq), _P(self._ncig), _P(self._infer), _P(self._mb))

    # fetch() with no region on an indexed BAM: all records with a reference id, file order.
    # fetch(contig, start, stop): pos < stop and endpos > start (htslib iterator rule).
    def _select(self, contig=None, start=None, stop=None):
        if contig is None:
            return np.nonzero(self._tid >= 0)[0]
        if contig not in self._tid_of:
            raise ValueError("invalid contig `%s`" % contig)
        t = self._tid_of[contig]
        if start is None:
            start = 0
        if stop is None:
            stop = self.lengths[t]
        if start < 0:
            raise ValueError("start out of range (%i)" % start)
        if stop < start:
            raise ValueError("stop out of range")
        m = (self._tid == t) & (self._pos < stop) & (self._end > start)
        return np.nonzero(m)[0]

    def fetch(self, contig=None, start=None, stop=None, region=None, reference=None, end=None, until_eof=False):
        if reference is not None:
            contig = reference
        if end is not None:
            stop = end
        for i in self._select(contig, start, stop):
            yield AlignedSegment(self, int(i))

    def count_coverage(self, contig, start=None, stop=None, region=None, quality_threshold=15,
                       read_callback="all", reference=None, end=None):
        """Returns four arrays like pysam; only their *sums* are used by the reference (:131,:1033),
        so the A-array carries the total and C,G,T are zero-length-safe zeros."""
        if quality_threshold != 0 or read_callback != "nofilter":
            raise NotImplementedError("shim restates only quality_threshold=0, read_callback='nofilter'")
        idx = self._select(contig, start, stop).astype(np.int64)
        total = _cov().ora_count_coverage(self._cbuf, _P(self._off), _P(idx), ctypes.c_int64(len(idx)),
                                          ctypes.c_int64(start), ctypes.c_int64(stop))
        return ([int(total)], [0], [0], [0])

    def close(self):
        pass
