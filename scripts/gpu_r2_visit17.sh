#!/bin/bash
# After fusing push + wait, repacking the messages and trimming the chain: exchange, sharded, pass and fetch tests; bench lines.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 1800 python -m pytest tests/test_gpu_xchg.py tests/test_gpu_sharded.py tests/test_gpu_sharded_pass.py tests/test_gpu_pass.py tests/test_gpu_fetch.py tests/test_gpu_pipeline.py -q -m gpu ) > gpurun_out/v17_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit17.txt
( timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-file ) > gpurun_out/v17_bench.json 2> gpurun_out/v17_bench.err
echo "bench rc=$?" >> gpurun_out/rc_visit17.txt
( timeout 600 python bench.py --steps 20 --warmup 5 --workload C1n --no-cpu --no-file ) > gpurun_out/v17_bench_C1n.json 2> gpurun_out/v17_bench_C1n.err
echo "bench C1n rc=$?" >> gpurun_out/rc_visit17.txt
grep -v "^$" gpurun_out/v17_tests.log | grep -v "^    \|^kwargs\|^popenargs" | tail -25 | cut -c1-400; cat gpurun_out/rc_visit17.txt
python - <<'PY'
import json
for f in ("v17_bench", "v17_bench_C1n"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d["value"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"]["whole_pass"]["frac"], d["e2e"]["ms_per_step"], d["parity_ok"], d["stage_ms_per_step"])
    except Exception as e:
        print(f, "parse failed", e); print(open("gpurun_out/%s.err" % f).read()[-2000:])
PY
