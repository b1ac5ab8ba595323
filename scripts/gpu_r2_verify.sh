#!/bin/bash
# Verification of HEAD on one GPU: the whole GPU suite (-x, as the driver runs it), smoke(), the bench line.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( time timeout 1800 python -m pytest tests -x -q -m gpu ) > gpurun_out/verify_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_verify.txt
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/verify_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/rc_verify.txt
( timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/verify_bench.json 2> gpurun_out/verify_bench.err
echo "bench rc=$?" >> gpurun_out/rc_verify.txt
( timeout 600 python scripts/ingest_timing.py --scale 0.25 --reps 3 ) > gpurun_out/verify_ingest.json 2> gpurun_out/verify_ingest.err
echo "ingest rc=$?" >> gpurun_out/rc_verify.txt
cat gpurun_out/verify_ingest.json
tail -4 gpurun_out/verify_tests.log; tail -2 gpurun_out/verify_smoke.log; cat gpurun_out/rc_verify.txt
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/verify_bench.json").read().strip().splitlines()[-1])
    print(d["steps"], d["ms_per_step"], d["value"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"]["whole_pass"]["frac"], d["e2e"]["ms_per_step"], d["e2e"]["stage_ms"], d["parity_ok"], d["cpu_baseline"]["kind"])
    print(d["e2e_from_file"]["device_ingest"]["ms_per_pass"], d["e2e_from_file"]["host_reader"]["ms_per_pass"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/verify_bench.err").read()[-2000:])
PY
