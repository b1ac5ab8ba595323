#!/bin/bash
# HSR search from the .bam through the device ingest: test + timing beside the host reader.
set -u
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_hsr.py tests/test_gpu_device_ingest.py -q -m gpu ) > gpurun_out/v18_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit18.txt
( cd tests && timeout 600 python - <<'PY'
import contextlib, io, json, os, sys, tempfile, time
sys.path.insert(0, os.path.abspath("..")); sys.path.insert(0, "golden")
import make_hsr_golden as H
from coral_b200 import hsr
res = {}
for case in H.CASES:
    data, args = H.hsr_inputs(case, tempfile.mkdtemp())
    row = {"records": int(data.n_records), "bam_bytes": os.path.getsize(args.lr_bam)}
    for mode in (False, True, False, True):
        t = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()):
            hsr.locate_hsrs(args, device_ingest=mode)
        row.setdefault("device_ingest_ms" if mode else "host_reader_ms", []).append(round((time.perf_counter() - t) * 1e3, 1))
    res[case] = row
print(json.dumps(res))
PY
) > gpurun_out/v18_hsr_timing.json 2> gpurun_out/v18_hsr_timing.err
tail -3 gpurun_out/v18_tests.log; cat gpurun_out/rc_visit18.txt; cat gpurun_out/v18_hsr_timing.json; tail -3 gpurun_out/v18_hsr_timing.err
