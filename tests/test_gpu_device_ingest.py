"""GPU tests of the device-side BAM ingest (csrc/ingest.cu): BGZF inflate + record framing on
the GPU against the host reader (zlib) — record table, name hashes, SA/NM columns, read names — and the whole pipeline
from placement="device_ingest" against the reference-generated fixtures.  First run on a B200 in round 2 (4 passed,
gpurun visit 1); the decoder core and the framing steps are also verified on the CPU (tests/test_ingest_cpu.py)."""
import os

import numpy as np
import pytest

import test_golden_cpu as G
import test_ingest_cpu as H
from sim import bgzf

pytestmark = pytest.mark.gpu


def _ingest(path):
    from coral_b200 import _abi
    ctx = _abi.Context(0)
    db = _abi.DeviceBam(path, ctx)
    ctx.set_references(db.references)
    db.run()
    return ctx, db


def _host(path):
    from coral_b200 import _abi
    bam = _abi.BamFile(path, 2)
    ctx = _abi.Context(0)
    ctx.set_references(bam.references)
    ctx.load_records(bam.rec, bam.rec_off, placement="copy")
    return ctx, bam


def _same(path, expect_serial=0):
    c1, db = _ingest(path)
    c2, bam = _host(path)
    assert db.references == bam.references and db.lengths == bam.lengths
    assert db.n_records == bam.n_records
    assert db.framed_bytes == int(bam.rec_off[-1])
    assert c1.ingest_used_serial_chase() == expect_serial
    t1, t2 = c1.record_table(), c2.record_table()
    for k in t2:
        assert np.array_equal(t1[k], t2[k]), k
    if bam.n_records:
        idx = np.unique(np.concatenate([np.arange(0, bam.n_records, max(1, bam.n_records // 200)), [bam.n_records - 1]]))
        assert db.query_names(idx) == bam.query_names(idx)
        k1, k2 = c1.record_keys(), c2.record_keys()          # name hashes, SA location, NM: every aux block decoded alike
        for k in k2:
            assert np.array_equal(k1[k], k2[k]), k
    n = bam.n_records
    c1.close(); c2.close(); bam.close()
    return n


def test_device_ingest_matches_host_reader(built_lib, dataset):
    for cfg, scale in (("C1", 0.01), ("C3", 0.004), ("C2", 0.01)):
        ds = dataset(cfg, scale=scale)
        assert _same(ds["paths"]["bam"]) == ds["data"].n_records


def test_device_ingest_block_types_fallback_and_empty(built_lib, tmp_path):
    H.test_every_deflate_block_type_and_member_boundaries(built_lib, tmp_path)      # writes mixed.bam
    assert _same(str(tmp_path / "mixed.bam")) == 400
    H.test_records_the_conditions_reject_fall_back_to_the_serial_chase(built_lib, tmp_path)
    assert _same(str(tmp_path / "weird.bam"), expect_serial=1) == 4
    H.test_empty_and_malformed_files(built_lib, tmp_path)
    assert _same(str(tmp_path / "empty.bam")) == 0
    from coral_b200 import _abi
    head = bgzf.bam_header_bytes([("chr1", 100000)])
    good = H._raw_member(head + H._record("a"), H._deflate(head + H._record("a"), 6))
    bad = bytearray(good)
    bad[30] ^= 0x55                                   # inside the Huffman-table header of the member's DEFLATE body
    p = str(tmp_path / "corrupt.bam")
    with open(p, "wb") as fp:
        fp.write(H._raw_member(head, H._deflate(head, 6)) + bytes(bad) + bgzf._EOF)
    ctx = _abi.Context(0)
    db = _abi.DeviceBam(p, ctx)
    ctx.set_references(db.references)
    with pytest.raises(_abi.CoralError):
        db.run()
    ctx.close()


@pytest.mark.parametrize("case", ["c2_small", "c4_small"])
def test_pipeline_from_device_ingest_matches_reference(built_lib, case, tmp_path_factory):
    gold = G.load_golden(case)
    cfg, data, paths = G.golden_dataset(case, tmp_path_factory)
    got = G.run_engine("cuda-ingest", paths, tmp_path_factory)
    assert G.compare_states(got, gold["stages"]) == []
    for g_got, g_ref in zip(got["assign_cov"]["graphs"], gold["stages"]["assign_cov"]["graphs"]):
        assert g_got["graph_txt"] == g_ref["graph_txt"]
