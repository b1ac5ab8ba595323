"""CPU tests of the device-side BAM ingest's host-verifiable pieces (csrc/inflate.cuh, csrc/ingest.cu — SURVEY.md §8f
rank 3): the DEFLATE core and every per-element framing step are __host__ __device__; coral_ingest_host_check runs
them on the CPU and must reproduce the host reader (zlib + serial framing, csrc/bam_reader.cpp) byte for byte.
The kernels themselves are exercised by tests/test_gpu_device_ingest.py (-m gpu)."""
import os
import struct
import zlib

import numpy as np
import pytest

from sim import bgzf


def _bam_bytes(path):
    return np.fromfile(path, dtype=np.uint8)


def _same_as_host_reader(built, path):
    from coral_b200 import _abi
    bam = _abi.BamFile(path, 2)
    got = _abi.ingest_host_check(_bam_bytes(path))
    assert np.array_equal(got["rec_off"], bam.rec_off)
    assert np.array_equal(got["rec"], bam.rec[:len(got["rec"])])
    n = bam.n_records
    bam.close()
    return got, n


def test_ingest_matches_host_reader_on_simulated_bams(built_lib, dataset):
    for cfg, scale in (("C1", 0.01), ("C3", 0.004)):
        ds = dataset(cfg, scale=scale)
        got, n = _same_as_host_reader(built_lib, ds["paths"]["bam"])
        assert n == ds["data"].n_records
        assert got["used_serial"] == 0                      # the speculative framing found the chain
        assert got["n_candidates"] >= n                     # every true record start is a candidate
        assert got["n_candidates"] <= n + max(8, n // 100)  # ... and chance hits are rare
        from coral_b200 import _abi
        forced = _abi.ingest_host_check(_bam_bytes(ds["paths"]["bam"]), mode=1)
        assert forced["used_serial"] == 1
        assert np.array_equal(forced["rec_off"], got["rec_off"]) and np.array_equal(forced["rec"], got["rec"])


def _raw_member(payload, deflated):
    """A BGZF member around an arbitrary raw-DEFLATE body (to reach block types zlib's levels do not all emit)."""
    head = b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00"
    size = len(head) + 2 + len(deflated) + 8
    return head + struct.pack("<H", size - 1) + deflated + struct.pack("<II", zlib.crc32(payload) & 0xFFFFFFFF, len(payload))


def _record(name, tid=0, pos=100, cigar=((0, 50),), seq_len=50, aux=b"", rng=None):
    nm = name.encode() + b"\0"
    cig = b"".join(struct.pack("<I", (ln << 4) | op) for op, ln in cigar)
    seq = bytes(rng.integers(0, 256, (seq_len + 1) // 2, dtype=np.uint8)) if rng is not None else b"\x11" * ((seq_len + 1) // 2)
    qual = bytes(rng.integers(0, 94, seq_len, dtype=np.uint8)) if rng is not None else b"\x28" * seq_len
    core = struct.pack("<iiBBHHHiiii", tid, pos, len(nm), 60, 4680, len(cigar), 0, seq_len, -1, -1, 0)
    body = core + nm + cig + seq + qual + aux
    return struct.pack("<i", len(body)) + body


def _write(path, members):
    with open(path, "wb") as fp:
        for m in members:
            fp.write(m)
        fp.write(bgzf._EOF)


def _deflate(data, level, strategy=zlib.Z_DEFAULT_STRATEGY):
    co = zlib.compressobj(level, zlib.DEFLATED, -15, 9, strategy)
    return co.compress(data) + co.flush()


def test_every_deflate_block_type_and_member_boundaries(built_lib, tmp_path):
    """Stored, fixed-Huffman and dynamic-Huffman blocks (levels 0 / Z_FIXED / 1 / 6 / 9), long matches and runs
    (distance 1), records straddling member boundaries, members cut at arbitrary byte positions, an empty member."""
    rng = np.random.default_rng(7)
    refs = [("chr1", 100000), ("chr2", 50000)]
    stream = bgzf.bam_header_bytes(refs)
    recs = []
    for k in range(400):
        aux = b"NMC\x03" + (b"SAZ" + b"chr2,500,+,10S40M,60,1;" + b"\0" if k % 7 == 0 else b"")
        recs.append(_record("read%05d" % k if k % 5 else "r" * (1 + k % 200), tid=k % 2, pos=10 * k,
                            cigar=((4, 10), (0, 20 + k), (2, 700), (0, 20)), seq_len=50 + 37 * (k % 40), aux=aux,
                            rng=rng if k % 3 else None))
    stream += b"".join(recs)
    cuts = [0]
    while cuts[-1] < len(stream):
        cuts.append(min(len(stream), cuts[-1] + int(rng.integers(1, 65000))))
    members = []
    for i in range(len(cuts) - 1):
        data = stream[cuts[i]:cuts[i + 1]]
        kind = i % 6
        body = (_deflate(data, 0) if kind == 0 else _deflate(data, 6, zlib.Z_FIXED) if kind == 1 else
                _deflate(data, 1) if kind == 2 else _deflate(data, 9) if kind == 3 else
                _deflate(data, 6, zlib.Z_HUFFMAN_ONLY) if kind == 4 else _deflate(data, 6, zlib.Z_RLE))
        members.append(_raw_member(data, body))
        if i == 3:
            members.append(_raw_member(b"", _deflate(b"", 6)))
    path = str(tmp_path / "mixed.bam")
    _write(path, members)
    got, n = _same_as_host_reader(built_lib, path)
    assert n == 400 and got["used_serial"] == 0


def test_records_the_conditions_reject_fall_back_to_the_serial_chase(built_lib, tmp_path):
    """A record with refID beyond the reference table is framed by the host reader (it only follows block_size); the
    speculative framing does not see it as a candidate, the chain breaks, and the serial chase gives the same frames."""
    refs = [("chr1", 100000)]
    stream = bgzf.bam_header_bytes(refs) + _record("a") + _record("weird", tid=7) + _record("b") + _record("c")
    path = str(tmp_path / "weird.bam")
    _write(path, [_raw_member(stream, _deflate(stream, 6))])
    got, n = _same_as_host_reader(built_lib, path)
    assert n == 4 and got["used_serial"] == 1


def test_chance_hits_inside_a_record_do_not_change_the_framing(built_lib, tmp_path):
    """A complete, valid-looking record embedded in another record's aux bytes is a candidate, but it is not on the
    chain from the first record."""
    refs = [("chr1", 100000)]
    inner = _record("inner")
    outer = _record("outer", aux=b"XXBC" + struct.pack("<I", len(inner)) + inner)
    stream = bgzf.bam_header_bytes(refs) + _record("a") + outer + _record("b")
    path = str(tmp_path / "nested.bam")
    _write(path, [_raw_member(stream, _deflate(stream, 6))])
    got, n = _same_as_host_reader(built_lib, path)
    assert n == 3 and got["used_serial"] == 0 and got["n_candidates"] >= 4


def test_empty_and_malformed_files(built_lib, tmp_path):
    from coral_b200 import _abi
    refs = [("chr1", 100000)]
    head = bgzf.bam_header_bytes(refs)
    path = str(tmp_path / "empty.bam")
    _write(path, [_raw_member(head, _deflate(head, 6))])
    got, n = _same_as_host_reader(built_lib, path)
    assert n == 0 and len(got["rec"]) == 0
    # corrupt DEFLATE body / truncated record / not BGZF
    good = _raw_member(head + _record("a"), _deflate(head + _record("a"), 6))
    bad = bytearray(good)
    bad[30] ^= 0x55
    for blob in (bytes(bad) + bgzf._EOF, b"not a bam file at all", _raw_member(head + _record("a")[:-5], _deflate(head + _record("a")[:-5], 6))):
        with pytest.raises(_abi.CoralError):
            _abi.ingest_host_check(np.frombuffer(blob, dtype=np.uint8))


def test_random_files_against_the_host_reader(built_lib, tmp_path):
    """Seeded random record sets (repetitive / run / random aux payloads, 0-5 CIGAR ops, unmapped records), members cut
    at random positions and compressed with random zlib level and strategy."""
    rng = np.random.default_rng(1)
    total = 0
    for it in range(12):
        refs = [("chr%d" % (i + 1), 1000000) for i in range(int(rng.integers(1, 6)))]
        stream = bgzf.bam_header_bytes(refs)
        base = bytes(rng.integers(65, 70, 5000, dtype=np.uint8))
        recs = []
        for k in range(int(rng.integers(1, 300))):
            kind = int(rng.integers(0, 4))
            aux = (b"XZZ" + base[:int(rng.integers(0, 5000))] + b"\0" if kind == 0 else
                   b"XZZ" + bytes([65 + int(rng.integers(0, 3))]) * int(rng.integers(0, 3000)) + b"\0" if kind == 1 else
                   b"XBBC" + struct.pack("<I", 2000) + bytes(rng.integers(0, 256, 2000, dtype=np.uint8)) if kind == 2 else b"")
            recs.append(_record("q%d" % k, tid=int(rng.integers(-1, len(refs))), pos=int(rng.integers(-1, 900000)),
                                cigar=((0, int(rng.integers(1, 1000))),) * int(rng.integers(0, 5)),
                                seq_len=int(rng.integers(0, 3000)), aux=aux, rng=rng if rng.random() < 0.5 else None))
        stream += b"".join(recs)
        cuts = [0]
        while cuts[-1] < len(stream):
            cuts.append(min(len(stream), cuts[-1] + int(rng.integers(1, 65536))))
        strategies = [zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FILTERED]
        members = []
        for i in range(len(cuts) - 1):
            data = stream[cuts[i]:cuts[i + 1]]
            members.append(_raw_member(data, _deflate(data, int(rng.integers(0, 10)), strategies[int(rng.integers(0, 5))])))
        path = str(tmp_path / ("f%d.bam" % it))
        _write(path, members)
        got, n = _same_as_host_reader(built_lib, path)
        assert n == len(recs)
        total += n
    assert total > 500
