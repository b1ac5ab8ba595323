"""CPU tests: the C ABI library loads and exports every declared symbol; the C++ BAM reader and the
oracle's Python reader agree byte-for-byte on the framed records."""
import ctypes
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "coral_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(coral_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(built_lib):
    L = ctypes.CDLL(built_lib)
    syms = declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(L, s), "missing symbol %s" % s


def test_create_without_gpu_fails_loudly(built_lib):
    import torch
    if torch.cuda.is_available():
        return
    from coral_b200 import _abi
    import pytest
    with pytest.raises(_abi.CoralError):
        _abi.Context(0)


def test_bam_reader_matches_oracle_reader(built_lib, dataset):
    from coral_b200 import _abi
    from oracle import pysam_shim
    ds = dataset("C1", scale=0.01)
    bam = _abi.BamFile(ds["paths"]["bam"], 4)
    ora = pysam_shim.AlignmentFile(ds["paths"]["bam"])
    assert bam.references == ora.references
    assert bam.lengths == ora.lengths
    assert bam.n_records == ora._n == ds["data"].n_records
    assert (bam.rec_off % 16 == 0).all()
    for i in list(range(0, bam.n_records, 97)) + [bam.n_records - 1]:
        o = int(bam.rec_off[i]); oo = int(ora._off[i])
        ln = 4 + int(np.frombuffer(bam.rec[o:o + 4].tobytes(), "<i4")[0])
        assert bam.rec[o:o + ln].tobytes() == ora._data[oo:oo + ln]
    # the simulator's in-memory framing equals what comes back from disk
    assert np.array_equal(bam.rec_off, ds["data"].rec_off)
    assert np.array_equal(bam.rec, ds["data"].rec)
    bam.close()
