#!/bin/bash
# One GPU-box visit: parity tests (new rows first), the default bench line, the host profile of the e2e pass.
# Everything lands in gpurun_out/ (merged back by gpurun).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
( time timeout 600 python -m pytest tests/test_gpu_paths.py tests/test_gpu_pipeline.py -q -m gpu -x ) > gpurun_out/pytest_new.log 2>&1
echo "pytest_new rc=$?" >> gpurun_out/rc.txt
( time timeout 900 python -m pytest tests -q -m gpu --deselect tests/test_gpu_paths.py --deselect tests/test_gpu_pipeline.py ) > gpurun_out/pytest_rest.log 2>&1
echo "pytest_rest rc=$?" >> gpurun_out/rc.txt
( time timeout 600 python bench.py ) > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/rc.txt
( time timeout 300 python bench.py --cprofile --only e2e --steps 3 ) > gpurun_out/cprofile_e2e.txt 2>&1
echo "cprofile rc=$?" >> gpurun_out/rc.txt
tail -3 gpurun_out/pytest_new.log; tail -3 gpurun_out/pytest_rest.log; cat gpurun_out/rc.txt; cut -c1-600 gpurun_out/bench.json
