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


def write_bam(path, refs, record_bytes, rec_off=None, level: int = 1):
    """Write header + records.  ``record_bytes`` is a bytes-like of concatenated records; when
    ``rec_off`` is given (n+1 offsets, possibly with alignment padding between records) the padding
    is removed so the file is a standard BAM stream."""
    import numpy as np

    buf = memoryview(record_bytes)
    with open(path, "wb") as fp:
        pending = bytearray(bam_header_bytes(refs))

        def flush(force=False):
            nonlocal pending
            while len(pending) >= _MAX_BLOCK or (force and pending):
                chunk = bytes(pending[:_MAX_BLOCK])
                del pending[:_MAX_BLOCK]
                fp.write(bgzf_block(chunk, level))

        flush(force=True)  # header in its own block(s)
        if rec_off is None:
            pending += buf
            flush(force=True)
        else:
            off = np.asarray(rec_off, dtype=np.uint64)
            for i in range(len(off) - 1):
                s = int(off[i])
                bs = struct.unpack_from("<i", buf, s)[0]
                pending += buf[s:s + 4 + bs]
                if len(pending) >= _MAX_BLOCK:
                    flush()
            flush(force=True)
        fp.write(_EOF)
