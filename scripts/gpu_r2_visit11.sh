#!/bin/bash
# Peer-memory exchanges over process ranks (CUDA IPC on one GPU), distributed fetch, pipelined sharded pass, new coverage walk,
# batched pass checks + side-stream result copies, e2e_from_file.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 900 python -m pytest tests/test_gpu_xchg.py -x -q -m gpu ) > gpurun_out/v11_xchg.log 2>&1
echo "xchg rc=$?" > gpurun_out/rc_visit11.txt
( timeout 1500 python -m pytest tests/test_gpu_sharded.py -q -m gpu ) > gpurun_out/v11_sharded.log 2>&1
echo "sharded rc=$?" >> gpurun_out/rc_visit11.txt
( timeout 1500 python -m pytest tests/test_gpu_sharded_pass.py -q -m gpu ) > gpurun_out/v11_spass.log 2>&1
echo "sharded pass rc=$?" >> gpurun_out/rc_visit11.txt
( timeout 1200 python -m pytest tests/test_gpu_pass.py tests/test_gpu_fetch.py tests/test_gpu_scan_coverage.py tests/test_gpu_pipeline.py tests/test_zz_no_alloc.py tests/test_gpu_edge_cases.py tests/test_gpu_breakpoints.py -x -q -m gpu ) > gpurun_out/v11_other.log 2>&1
echo "other rc=$?" >> gpurun_out/rc_visit11.txt
( timeout 300 python scripts/pass_timing.py --overlap 0 --fork 1 ) > gpurun_out/v11_pass.json 2> gpurun_out/v11_pass.err
( timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/v11_bench.json 2> gpurun_out/v11_bench.err
echo "bench rc=$?" >> gpurun_out/rc_visit11.txt
tail -15 gpurun_out/v11_xchg.log; grep -v "^$" gpurun_out/v11_sharded.log | tail -70; tail -40 gpurun_out/v11_spass.log; tail -8 gpurun_out/v11_other.log; cat gpurun_out/rc_visit11.txt; cat gpurun_out/v11_pass.json
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/v11_bench.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches", "parity_ok", "stage_ms_per_step")})
    print(d["roofline"]["frac"], d["roofline"]["whole_pass"]); print(d["e2e"]["ms_per_step"]); print(d.get("e2e_from_file"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/v11_bench.err").read()[-3000:])
PY
