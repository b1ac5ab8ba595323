#!/bin/bash
# Round 2, visit 3: pipelined pass tests, bench with / without the pipelined pass, then the whole GPU suite.
set -u
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_pass.py tests/test_gpu_edge_cases.py -x -q -m gpu ) > gpurun_out/v3_pass_tests.log 2>&1
echo "pass tests rc=$?" > gpurun_out/rc_visit3.txt
( timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/v3_bench.json 2> gpurun_out/v3_bench.err
echo "bench rc=$?" >> gpurun_out/rc_visit3.txt
( timeout 600 python bench.py --steps 20 --warmup 5 --no-pipeline --no-cpu ) > gpurun_out/v3_bench_nopipe.json 2> gpurun_out/v3_bench_nopipe.err
echo "bench nopipe rc=$?" >> gpurun_out/rc_visit3.txt
( timeout 2400 python -m pytest tests -m gpu -q ) > gpurun_out/v3_gpu_tests.log 2>&1
echo "gpu tests rc=$?" >> gpurun_out/rc_visit3.txt
tail -5 gpurun_out/v3_pass_tests.log; cat gpurun_out/rc_visit3.txt; tail -15 gpurun_out/v3_gpu_tests.log
