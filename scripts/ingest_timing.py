"""Round-2 first look at the device-side ingest (csrc/ingest.cu): host reader (zlib threads + framing + pinned copy)
against coral_ingest_open + coral_ingest_run on the same .bam, and the record tables compared.  One JSON line.

    python scripts/ingest_timing.py [--scale 0.1] [--threads 0]
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=float, default=0.1)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    from sim import configs
    from sim.generate import simulate, write_dataset
    from coral_b200 import _abi
    cfg = configs.CONFIGS["C2"](scale=a.scale)
    data = simulate(cfg)
    paths = write_dataset(data, tempfile.mkdtemp(prefix="coral_ingest_"))
    out = {"config": "C2", "scale": a.scale, "records": int(data.n_records), "bam_bytes": os.path.getsize(paths["bam"]),
           "decoded_framed_bytes": int(data.rec_off[-1])}
    host_ms, dev_ms, open_ms = [], [], []
    for _ in range(a.reps):
        t = time.perf_counter()
        bam = _abi.BamFile(paths["bam"], a.threads)
        c2 = _abi.Context(0)
        c2.set_references(bam.references)
        c2.load_records(bam.rec, bam.rec_off, placement="copy")
        host_ms.append((time.perf_counter() - t) * 1e3)
        t = time.perf_counter()
        c1 = _abi.Context(0)
        db = _abi.DeviceBam(paths["bam"], c1)         # file read + member table + header (host)
        c1.set_references(db.references)
        open_ms.append((time.perf_counter() - t) * 1e3)
        t = time.perf_counter()
        db.run()                                      # H2D of the compressed bytes, inflate, framing, load-time scan
        dev_ms.append((time.perf_counter() - t) * 1e3)
        t1, t2 = c1.record_table(), c2.record_table()
        same = all(np.array_equal(t1[k], t2[k]) for k in t2) and db.n_records == bam.n_records
        out["serial_chase"] = c1.ingest_used_serial_chase()
        c1.close(); c2.close(); bam.close()
    out.update({"host_reader_plus_copy_load_ms": host_ms, "device_open_ms": open_ms, "device_run_ms": dev_ms, "tables_identical": bool(same)})
    print(json.dumps(out))


if __name__ == "__main__":
    main()
