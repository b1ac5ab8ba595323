#!/bin/bash
set -u
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_pass.py -x -q -m gpu ) > gpurun_out/v5_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit5.txt
python bench.py --only scan --steps 3 --warmup 3 > gpurun_out/v5_scan_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_scan_records -s 3 -c 1 -o gpurun_out/r2_scan_v8 \
    python bench.py --only scan --steps 3 --warmup 3 > gpurun_out/v5_scan_ncu.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc_visit5.txt
tail -4 gpurun_out/v5_tests.log; cat gpurun_out/rc_visit5.txt
