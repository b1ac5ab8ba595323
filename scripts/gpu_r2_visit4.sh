#!/bin/bash
# Round 2, visit 4: the 8-ops-per-lane scan kernel — parity tests, then scan-only timing of the build variants.
set -u
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_scan_coverage.py tests/test_gpu_edge_cases.py tests/test_gpu_pass.py "tests/test_gpu_pipeline.py::test_cuda_pipeline_matches_reference" -x -q -m gpu ) > gpurun_out/v4_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit4.txt
for v in "" _b3 _b3c4 _b4c2; do
  ( CORAL_B200_LIB=$PWD/coral_b200/libcoral_b200$v.so timeout 300 python bench.py --only scan --steps 10 --warmup 3 ) > gpurun_out/v4_scan$v.json 2> gpurun_out/v4_scan$v.err
  echo "scan$v rc=$?" >> gpurun_out/rc_visit4.txt
done
tail -4 gpurun_out/v4_tests.log; cat gpurun_out/rc_visit4.txt; cat gpurun_out/v4_scan*.json
