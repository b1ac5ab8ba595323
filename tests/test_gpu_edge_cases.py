"""GPU parity on hand-made records: the corners of the BAM record layout the simulator does not produce (every CIGAR
op, IUPAC and N bases, odd / zero l_seq, unmapped and CIGAR-less records, every aux type, aux blocks longer than a
staged segment, 60 000-op CIGARs, 1- and 254-character names), empty and malformed inputs, and the placements."""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

REFS = [("chr1", 5_000_000), ("chr2", 3_000_000)]


def _edge_records():
    T = util.aux_tag
    rng = np.random.default_rng(11)
    bases = lambda n: "".join(rng.choice(list("ACGT"), n))         # noqa: E731
    recs = []
    # plain record; NM stored as every integer type pysam accepts
    for k, ty in enumerate("cCsSiI"):
        recs.append(util.make_record("plain%d" % k, 0, 1000 + 10 * k, 60, 0, [("M", 100)], bases(100), T("NM", ty, 7 + k)))
    # every op, hard + soft clips at both ends; IUPAC codes and N inside the aligned part; odd l_seq
    cig = [("H", 9), ("S", 5), ("M", 10), ("I", 2), ("D", 3), ("N", 4), ("=", 6), ("X", 7), ("P", 1), ("M", 8), ("S", 3), ("H", 2)]
    q = 5 + 10 + 2 + 6 + 7 + 8 + 3
    seq = list(bases(q))
    for i in (6, 9, 18, 25, 30):
        seq[i] = "NRYKM"[i % 5]
    recs.append(util.make_record("allops", 0, 2000, 37, 16, cig, "".join(seq), T("XA", "A", "x") + T("NM", "i", 12)))
    assert q % 2 == 1
    # SEQ '*' with a CIGAR (secondary alignment), then a mapped record without CIGAR
    recs.append(util.make_record("noseq", 0, 2500, 0, 256, [("M", 50)], None, T("NM", "C", 0)))
    recs.append(util.make_record("nocigar", 0, 2600, 0, 0, [], bases(20)))
    # in-CIGAR gaps around min_del_len (600): 599D, 600D, 601D, 700N, and two short D's adding up to > 600 with an I between
    cig = [("M", 200), ("D", 599), ("M", 100), ("D", 600), ("M", 100), ("D", 601), ("M", 100), ("N", 700), ("M", 100),
           ("D", 400), ("I", 5), ("D", 300), ("M", 50)]
    recs.append(util.make_record("gaps", 0, 10000, 60, 0, cig, bases(200 + 100 * 4 + 5 + 50), T("NM", "S", 3000)))
    # aux: every type, SA in the middle, a B array longer than a staged segment BEFORE the tags the path reads
    big = T("ZB", "B", ("s", list(range(-3000, 3000)))) + T("ZC", "B", ("I", [1, 2, 3])) + T("ZD", "B", ("f", [1.5, -2.0]))
    sa = "chr2,4001,-,40S60M,60,2;chr1,90001,+,70M30S,13,0;"
    aux = T("XB", "c", -5) + T("XC", "f", 3.25) + T("XD", "H", "1AE301") + big + T("SA", "Z", sa) + T("XE", "Z", "tail") + T("NM", "s", 321)
    recs.append(util.make_record("bigaux", 0, 20000, 60, 0, [("M", 60), ("S", 40)], bases(100), aux))
    # second SA tag in the same record is ignored (pysam returns the first); NM absent
    aux = T("SA", "Z", "chr2,5001,+,10S90M,60,1;") + T("SA", "Z", "chr2,7001,+,90M10S,60,1;")
    recs.append(util.make_record("twosa", 0, 20100, 60, 2048, [("H", 10), ("M", 90)], bases(90), aux))
    # 60 000 CIGAR ops (1M1I...), the u16 limit is 65 535
    n = 30000
    recs.append(util.make_record("long", 0, 30000, 60, 0, [("M", 1), ("I", 1)] * n, bases(2 * n), T("NM", "i", n)))
    # names of 1 and 254 characters
    recs.append(util.make_record("z", 0, 100000, 60, 0, [("M", 33)], bases(33), T("NM", "C", 1)))
    recs.append(util.make_record("n" * 254, 0, 100010, 60, 0, [("M", 31)], bases(31), T("NM", "C", 2)))
    # second contig, record ending exactly at a region boundary used below; then unmapped records at the end
    recs.append(util.make_record("c2a", 1, 0, 60, 0, [("M", 1000)], bases(1000), T("NM", "C", 0)))
    recs.append(util.make_record("c2b", 1, 999, 60, 16, [("S", 1), ("M", 1)], "AN", T("NM", "C", 0)))
    recs.append(util.make_record("unmapped_placed", 1, 1500, 0, 4, [], bases(10)))
    recs.append(util.make_record("unmapped", -1, -1, 0, 4, [], bases(12)))
    return recs


@pytest.fixture(scope="module")
def edge(tmp_path_factory):
    from oracle import pysam_shim
    from sim.bgzf import write_bam
    recs = _edge_records()
    rec, off = util.frame_records(recs)
    path = str(tmp_path_factory.mktemp("edge") / "edge.bam")
    write_bam(path, REFS, rec, off)
    return {"rec": rec, "off": off, "path": path, "ora": pysam_shim.AlignmentFile(path), "n": len(recs)}


def _ctx(rec, off, placement="copy"):
    from coral_b200 import _abi
    c = _abi.Context(0)
    c.set_references([r[0] for r in REFS])
    if placement.startswith("mapped"):
        buf = np.zeros(len(rec) + 64, np.uint8)
        buf[:len(rec)] = rec
        c.load_records(buf[:len(rec)], off, placement=placement, slack=64)
        if placement in ("mapped_async", "mapped_lazy"):
            # what the pre-scan wrote (the columns fetch reads while the full scan is in flight) against what the full
            # scan writes into the same columns; fetch must not wait for the scan and must give the same table
            f0 = c.fetch()
            c.wait_scan()
            f1 = c.fetch()
            for k in f0:
                assert np.array_equal(f0[k], f1[k]), k
    else:
        c.load_records(rec, off)
    return c


def _table_matches(t, ora):
    for k, want in (("tid", ora._tid), ("pos", ora._pos), ("endpos", ora._end), ("flag", ora._flag), ("mapq", ora._mapq),
                    ("l_seq", ora._lseq), ("n_cigar", ora._ncig), ("infer_len", ora._infer), ("aligned_bases", ora._mb)):
        assert np.array_equal(t[k], want), k


@pytest.mark.parametrize("placement", ["copy", "mapped", "mapped_async", "mapped_lazy"])
def test_edge_records_scan_table(built_lib, edge, placement):
    c = _ctx(edge["rec"], edge["off"], placement)
    ora = edge["ora"]
    assert c.n_records == edge["n"] == ora._n
    _table_matches(c.record_table(), ora)
    c.rescan()                                   # idempotent
    _table_matches(c.record_table(), ora)
    # SA tag location and NM as the reference reads them (get_tag('SA:Z:'), get_cigar_stats()[0][-1])
    keys = c.record_keys()
    from oracle.pysam_shim import AlignedSegment
    for i in range(edge["n"]):
        seg = AlignedSegment(ora, i)
        o = int(edge["off"][i])
        if seg.has_tag("SA"):
            a = o + int(keys["sa_off"][i])
            assert bytes(edge["rec"][a:a + int(keys["sa_len"][i])]).decode() == seg.get_tag("SA"), i
        else:
            assert keys["sa_off"][i] == 0
        assert int(keys["nm"][i]) == int(seg.get_cigar_stats()[0][-1]), i
    # names: equal names hash equal, different names differ
    h = set(zip(keys["name_lo"].tolist(), keys["name_hi"].tolist()))
    assert len(h) == edge["n"]
    c.close()


def test_edge_records_coverage(built_lib, edge):
    c = _ctx(edge["rec"], edge["off"])
    ora = edge["ora"]
    regions = [(0, 0, 900), (0, 900, 1050), (0, 1050, 2004), (0, 2004, 2010), (0, 2010, 2021), (0, 2021, 2600), (0, 2600, 10150),
               (0, 10150, 10800), (0, 10800, 12000), (0, 12000, 13600), (0, 19990, 20061), (0, 29999, 30001), (0, 30001, 45000),
               (0, 59999, 60000), (0, 60000, 60001), (0, 100000, 100041), (1, 0, 999), (1, 999, 1000), (1, 1000, 1001), (1, 1400, 1600)]
    tid, start, stop = (np.array(x, np.int32) for x in zip(*regions))
    bs, rc = c.region_coverage(tid, start, stop)
    for i, (t, a, b) in enumerate(regions):
        name = REFS[t][0]
        want = sum(sum(x) for x in ora.count_coverage(name, a, b, quality_threshold=0, read_callback='nofilter'))
        cnt = len([r for r in ora.fetch(name, a, b) if r.infer_read_length()])
        assert (int(bs[i]), int(rc[i])) == (want, cnt), (regions[i], int(bs[i]), want, int(rc[i]), cnt)
    # additivity: adjacent regions add up to their union
    u_bs, _ = c.region_coverage([0, 1], [0, 0], [2600, 1001])
    assert int(u_bs[0]) == int(bs[:6].sum()) and int(u_bs[1]) == int(bs[16:19].sum())
    c.close()


def test_edge_records_smalldel_rows(built_lib, edge):
    """Gaps between consecutive aligned blocks larger than min_del_len (ibg:750-762), on the 'gaps' record."""
    c = _ctx(edge["rec"], edge["off"])
    c.set_params(min_del_len=600, min_bp_match_cutoff=100)
    rows = c.smalldel([0], [9000], [14000])
    got = sorted(zip(rows["next_start"].tolist(), rows["prev_end"].tolist()))
    # blocks of 'gaps': D 599 and D 600 are not > 600; D 601, N 700 and the 400D+5I+300D pair (blocks 700 apart) are
    p = 10000
    ends, starts = [], []
    cur = p
    for op, ln in [("M", 200), ("D", 599), ("M", 100), ("D", 600), ("M", 100), ("D", 601), ("M", 100), ("N", 700), ("M", 100),
                   ("D", 400), ("I", 5), ("D", 300), ("M", 50)]:
        if op == "M":
            starts.append(cur); cur += ln; ends.append(cur)
        elif op in "DN":
            cur += ln
    want = sorted((starts[k + 1], ends[k]) for k in range(len(starts) - 1) if starts[k + 1] - ends[k] > 600)
    assert got == want and len(want) == 3
    c.close()


def test_empty_and_malformed_inputs(built_lib, edge):
    from coral_b200 import _abi
    names = [r[0] for r in REFS]
    c = _abi.Context(0)
    c.set_references(names)
    # no records at all
    c.load_records(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert c.n_records == 0
    bs, rc = c.region_coverage([0], [0], [1000])
    assert int(bs[0]) == 0 and int(rc[0]) == 0
    f = c.fetch()
    assert f["n_reads"] == 0 and f["nm_n"] == 0          # the caller raises ZeroDivisionError like ibg:159
    # a record start that is not 16-byte aligned
    rec, off = edge["rec"], edge["off"].copy()
    bad = off.copy(); bad[1] += 4
    with pytest.raises(RuntimeError):
        c.load_records(rec, bad)
    # block_size pointing past the frame: the scan reports a format error
    broken = rec.copy()
    broken[int(off[2]):int(off[2]) + 4] = np.frombuffer(np.array([1 << 20], "<i4").tobytes(), np.uint8)
    with pytest.raises(RuntimeError):
        c.load_records(broken, off)
    # an aux block with an unknown type byte
    broken = rec.copy()
    o = int(off[0])
    bs0 = int(np.frombuffer(rec[o:o + 4].tobytes(), "<i4")[0])
    broken[o + 4 + bs0 - 2] = ord("?")                   # type byte of the NM tag of record 0 (tag NM, type c, 1 byte)
    with pytest.raises(RuntimeError):
        c.load_records(broken, off)
    # unsorted / overlapping regions are refused
    c.load_records(rec, off)
    with pytest.raises(RuntimeError):
        c.region_coverage([0, 0], [100, 50], [200, 150])
    with pytest.raises(RuntimeError):
        c.region_coverage([0, 0], [100, 150], [200, 250])
    c.close()


@pytest.mark.parametrize("placement", ["mapped", "mapped_async", "mapped_lazy"])
def test_unpinned_buffer_ending_at_a_page_boundary(built_lib, edge, placement):
    """A caller's plain (not page-locked) buffer whose last record ends 16 bytes before a page boundary: the 64 bytes of
    slack the ABI promises the kernels lie mostly on the NEXT page, which must be registered with the rest
    (csrc/abi.cu: cudaHostRegister(nbytes + 64)) — else the device over-read faults."""
    import mmap
    from coral_b200 import _abi
    rec, off = edge["rec"], edge["off"]
    nb = int(off[-1])
    page = mmap.PAGESIZE
    npages = (nb + 16 + page - 1) // page
    mm = mmap.mmap(-1, (npages + 1) * page)                        # anonymous, page aligned, not pinned
    buf = np.frombuffer(mm, np.uint8)
    start = npages * page - 16 - nb
    assert start >= 0 and start % 16 == 0 and (start + nb + 16) % page == 0
    buf[start:start + nb] = rec[:nb]
    c = _abi.Context(0)
    c.set_references([r[0] for r in REFS])
    c.load_records(buf[start:start + nb], off, placement=placement, slack=page)
    c.wait_scan()
    _table_matches(c.record_table(), edge["ora"])
    f = c.fetch()
    c2 = _ctx(rec, off)
    f2 = c2.fetch()
    for k in f2:
        assert np.array_equal(f[k], f2[k]), k
    c.close(); c2.close()        # (the anonymous mapping is released with its last view)
