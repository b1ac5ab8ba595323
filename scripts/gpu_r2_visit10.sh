#!/bin/bash
# Peer-memory exchanges + distributed fetch + pipelined sharded pass: thread-rank tests on one GPU.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 600 python -m pytest tests/test_gpu_xchg.py -x -q -m gpu ) > gpurun_out/v10_xchg.log 2>&1
echo "xchg rc=$?" > gpurun_out/rc_visit10.txt
( timeout 1500 python -m pytest tests/test_gpu_sharded.py -x -q -m gpu ) > gpurun_out/v10_sharded.log 2>&1
echo "sharded rc=$?" >> gpurun_out/rc_visit10.txt
( timeout 1500 python -m pytest tests/test_gpu_sharded_pass.py -x -q -m gpu ) > gpurun_out/v10_spass.log 2>&1
echo "sharded pass rc=$?" >> gpurun_out/rc_visit10.txt
( timeout 600 python -m pytest tests/test_gpu_pass.py tests/test_gpu_fetch.py tests/test_gpu_pipeline.py -x -q -m gpu ) > gpurun_out/v10_other.log 2>&1
echo "other rc=$?" >> gpurun_out/rc_visit10.txt
tail -25 gpurun_out/v10_xchg.log; tail -25 gpurun_out/v10_sharded.log; tail -40 gpurun_out/v10_spass.log; tail -5 gpurun_out/v10_other.log; cat gpurun_out/rc_visit10.txt
