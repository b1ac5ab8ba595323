#!/bin/bash
# Round 2, visit 2: whole GPU suite (ingest tests un-gated), ingest timing at C2 x1.0, source-level ncu capture of the scan.
set -u
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/gpu_tests.log 2>&1
echo "gpu tests rc=$?" > gpurun_out/rc_visit2.txt
( timeout 400 python scripts/ingest_timing.py --scale 1.0 --reps 2 ) > gpurun_out/ingest_timing_full.json 2> gpurun_out/ingest_timing_full.err
echo "ingest timing rc=$?" >> gpurun_out/rc_visit2.txt
python bench.py --only scan --steps 3 --warmup 3 > gpurun_out/scan_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_scan_records -s 3 -c 1 -o gpurun_out/r2_scan_base \
    python bench.py --only scan --steps 3 --warmup 3 > gpurun_out/scan_ncu.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc_visit2.txt
tail -3 gpurun_out/gpu_tests.log; cat gpurun_out/rc_visit2.txt; cat gpurun_out/ingest_timing_full.json; cat gpurun_out/scan_plain.log | tail -1
