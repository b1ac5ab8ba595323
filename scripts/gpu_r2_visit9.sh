#!/bin/bash
# Round-2 re-entry visit: the no-alloc entry point alone, the whole GPU suite, the N=1 bench line, the launch list of one device pass.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu ) > gpurun_out/v9_zz_alone.log 2>&1
echo "zz alone rc=$?" > gpurun_out/rc_visit9.txt
( time timeout 1500 python -m pytest tests -x -q -m gpu ) > gpurun_out/v9_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/rc_visit9.txt
( timeout 600 python bench.py --steps 20 --warmup 5 ) > gpurun_out/v9_bench.json 2> gpurun_out/v9_bench.err
echo "bench rc=$?" >> gpurun_out/rc_visit9.txt
( timeout 300 python scripts/pass_timing.py --overlap 0 --fork 1 ) > gpurun_out/v9_pass.json 2> gpurun_out/v9_pass.err
( timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/v9_launches.csv python bench.py --profile --no-cpu ) > gpurun_out/v9_ncu.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc_visit9.txt
tail -3 gpurun_out/v9_zz_alone.log; tail -5 gpurun_out/v9_tests.log; cat gpurun_out/rc_visit9.txt
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/v9_bench.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches", "parity_ok", "stage_ms_per_step")})
    print(d["roofline"]); print(d["e2e"]); print(d.get("clocks")); print(d.get("cpu_baseline"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/v9_bench.err").read()[-2000:])
PY
cat gpurun_out/v9_pass.json
