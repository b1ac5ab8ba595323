#!/bin/bash
# Size-ordered record queue of the scan (options 0 / 1 / 2), trimmed fetch chain: tests, scan-only timing, the bench line.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 1500 python -m pytest tests/test_gpu_fetch.py tests/test_gpu_pass.py tests/test_gpu_scan_coverage.py tests/test_gpu_pipeline.py tests/test_gpu_edge_cases.py tests/test_gpu_breakpoints.py tests/test_gpu_cluster.py tests/test_gpu_sharded_pass.py -x -q -m gpu ) > gpurun_out/v14_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit14.txt
for o in 0 1 2; do
  ( CORAL_SCAN_ORDER=$o timeout 300 python bench.py --only scan --steps 20 --warmup 5 ) > gpurun_out/v14_scan_order$o.json 2> gpurun_out/v14_scan$o.err
done
( timeout 300 python scripts/pass_timing.py --overlap 0 --fork 1 ) > gpurun_out/v14_pass.json 2> gpurun_out/v14_pass.err
( timeout 900 python bench.py --steps 20 --warmup 5 --no-file ) > gpurun_out/v14_bench.json 2> gpurun_out/v14_bench.err
echo "bench rc=$?" >> gpurun_out/rc_visit14.txt
tail -4 gpurun_out/v14_tests.log; cat gpurun_out/rc_visit14.txt; cat gpurun_out/v14_scan_order*.json; cat gpurun_out/v14_pass.json
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/v14_bench.json").read().strip().splitlines()[-1])
    print(d["steps"], d["ms_per_step"], d["value"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"]["kernel_ms"], d["roofline"]["whole_pass"]["frac"], d["e2e"]["ms_per_step"], d["parity_ok"])
    print(d["stage_ms_per_step"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/v14_bench.err").read()[-2000:])
PY
