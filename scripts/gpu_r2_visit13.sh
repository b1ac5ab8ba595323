#!/bin/bash
# After trimming the exchange chain: sharded tests again; N=1 bench line with the collector paused in the timed loop.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 1500 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_sharded_pass.py -q -m gpu ) > gpurun_out/v13_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit13.txt
( timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/v13_bench.json 2> gpurun_out/v13_bench.err
echo "bench rc=$?" >> gpurun_out/rc_visit13.txt
grep -v "^$" gpurun_out/v13_tests.log | grep -v "^    \|^kwargs\|^popenargs" | tail -30 | cut -c1-300; cat gpurun_out/rc_visit13.txt
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/v13_bench.json").read().strip().splitlines()[-1])
    print(d["steps"], d["ms_per_step"], d["value"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"]["whole_pass"]["frac"], d["e2e"]["ms_per_step"], d["parity_ok"], d["cpu_baseline"]["kind"], d["cpu_baseline"]["value"])
    print(d["e2e_from_file"]["device_ingest"], d["e2e_from_file"]["host_reader"]["ms_per_pass"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/v13_bench.err").read()[-2000:])
PY
