#!/bin/bash
set -u
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_pass.py tests/test_gpu_scan_coverage.py tests/test_gpu_edge_cases.py -x -q -m gpu ) > gpurun_out/v6_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit6.txt
for v in "" _nm; do
  ( CORAL_B200_LIB=$PWD/coral_b200/libcoral_b200$v.so timeout 300 python bench.py --only scan --steps 10 --warmup 3 ) > gpurun_out/v6_scan$v.json 2> gpurun_out/v6_scan$v.err
  echo "scan$v rc=$?" >> gpurun_out/rc_visit6.txt
done
( CORAL_B200_LIB=$PWD/coral_b200/libcoral_b200_nm.so timeout 600 python -m pytest tests/test_gpu_scan_coverage.py tests/test_gpu_edge_cases.py "tests/test_gpu_pipeline.py::test_cuda_pipeline_matches_reference" -x -q -m gpu ) > gpurun_out/v6_tests_nm.log 2>&1
echo "tests nm rc=$?" >> gpurun_out/rc_visit6.txt
( timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/v6_bench.json 2> gpurun_out/v6_bench.err
echo "bench rc=$?" >> gpurun_out/rc_visit6.txt
( timeout 2400 python -m pytest tests/test_gpu_fullsize.py -q -m gpu ) > gpurun_out/v6_fullsize.log 2>&1
echo "fullsize rc=$?" >> gpurun_out/rc_visit6.txt
tail -4 gpurun_out/v6_tests.log; tail -3 gpurun_out/v6_tests_nm.log; cat gpurun_out/rc_visit6.txt; cat gpurun_out/v6_scan*.json; tail -8 gpurun_out/v6_fullsize.log
