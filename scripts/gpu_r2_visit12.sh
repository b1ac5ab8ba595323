#!/bin/bash
# Sharded tests after the fixes (destroy order, nm totals, cluster compare), pinned result arrays, where the bench loses 0.4 ms.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 1500 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_sharded_pass.py tests/test_gpu_pass.py tests/test_gpu_fetch.py tests/test_gpu_scan_coverage.py -q -m gpu ) > gpurun_out/v12_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit12.txt
for p in 1 0; do ( timeout 300 python scripts/pass_timing.py --overlap 0 --fork 1 --pinned $p ) > gpurun_out/v12_pass_pinned$p.json 2> gpurun_out/v12_pass$p.err; done
( timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-file ) > gpurun_out/v12_bench_a.json 2> gpurun_out/v12_bench_a.err
( CORAL_BENCH_SAMPLER_S=0.5 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-file ) > gpurun_out/v12_bench_b.json 2> gpurun_out/v12_bench_b.err
( timeout 600 python bench.py --steps 200 --warmup 5 --no-cpu --no-file ) > gpurun_out/v12_bench_c.json 2> gpurun_out/v12_bench_c.err
grep -v "^$" gpurun_out/v12_tests.log | grep -v "^    \|^kwargs\|^popenargs" | tail -40 | cut -c1-300; cat gpurun_out/rc_visit12.txt
cat gpurun_out/v12_pass_pinned1.json; echo; cat gpurun_out/v12_pass_pinned0.json; echo
python - <<'PY'
import json
for k in "abc":
    try:
        d = json.loads(open("gpurun_out/v12_bench_%s.json" % k).read().strip().splitlines()[-1])
        print(k, d["steps"], d["ms_per_step"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"]["whole_pass"]["frac"], d["e2e"]["ms_per_step"], d["clocks"])
    except Exception as e:
        print(k, "parse failed", e); print(open("gpurun_out/v12_bench_%s.err" % k).read()[-2000:])
PY
