"""BGZF / BAM container writer for the synthetic data sets.

Only used to hand the *same* records to the oracle's reader and to the product's C++ reader as a
real coordinate-sorted ``.bam`` file (the interface of reference call site
``infer_breakpoint_graph.py:65``).  Compression is zlib raw-deflate per <=64 KiB block.
"""
import struct
import zlib

_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")
_MAX_BLOCK = 0xFF00


def bgzf_block(data: bytes, level: int = 1) -> bytes:
    assert len(data) <= 0x10000
    co = zlib.compressobj(level, zlib.DEFLATED, -15)
    payload = co.compress(data) + co.flush()
    bsize = len(payload) + 25  # total block size - 1
    assert bsize < 0x10000
    head = struct.pack("<BBBBIBBHBBHH", 0x1F, 0x8B, 8, 4, 0, 0, 0xFF, 6, 0x42, 0x43, 2, bsize)
    tail = struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data))
    return head + payload + tail


def bam_header_bytes(refs, text: str = None) -> bytes:
    """refs: list of (name, length)."""
    if text is None:
        text = "@HD\tVN:1.6\tSO:coordinate\n" + "".join("@SQ\tSN:%s\tLN:%d\n" % r for r in refs)
    t = text.encode()
    out = [b"BAM\x01", struct.pack("<i", len(t)), t, struct.pack("<i", len(refs))]
    for name, ln in refs:
        nb = name.encode() + b"\x00"
        out += [struct.pack("<i", len(nb)), nb, struct.pack("<i", ln)]
    return b"".join(out)


def write_bam(path, refs, record_bytes, rec_off=None, level: int = 1, threads: int = 0):
    """Write header + records.  ``record_bytes`` is a bytes-like of concatenated records; when
    ``rec_off`` is given (n+1 offsets, possibly with alignment padding between records) the padding
    is removed so the file is a standard BAM stream.  The header goes into its own block(s), the record
    stream is cut every 0xFF00 bytes; blocks are deflated on ``threads`` threads (0: all cores; zlib
    releases the GIL) and written in order."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    import numpy as np

    buf = memoryview(record_bytes)
    if rec_off is None:
        stream = buf
    else:
        off = np.asarray(rec_off, dtype=np.uint64)
        n = len(off) - 1
        raw = np.frombuffer(buf, dtype=np.uint8)
        starts = off[:-1].astype(np.int64)
        if n and np.all(starts % 4 == 0) and len(raw) % 4 == 0:
            lens = raw.view("<i4")[starts // 4].astype(np.int64) + 4
        else:
            lens = np.array([struct.unpack_from("<i", buf, int(s))[0] + 4 for s in starts], np.int64)
        dst = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        out = np.empty(int(dst[-1]), np.uint8)
        for s, d, ln in zip(starts.tolist(), dst[:-1].tolist(), lens.tolist()):
            out[d:d + ln] = raw[s:s + ln]
        stream = memoryview(out)
    nthreads = threads or os.cpu_count() or 1

    def blocks(mv):
        for a in range(0, len(mv), _MAX_BLOCK):
            yield bytes(mv[a:a + _MAX_BLOCK])
    with open(path, "wb") as fp, ThreadPoolExecutor(nthreads) as pool:
        for chunk in blocks(memoryview(bam_header_bytes(refs))):
            fp.write(bgzf_block(chunk, level))
        batch = []
        for chunk in blocks(stream):
            batch.append(chunk)
            if len(batch) == 64 * nthreads:
                for blk in pool.map(lambda c: bgzf_block(c, level), batch):
                    fp.write(blk)
                batch = []
        for blk in pool.map(lambda c: bgzf_block(c, level), batch):
            fp.write(blk)
        fp.write(_EOF)
