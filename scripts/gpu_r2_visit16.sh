#!/bin/bash
# Overlapped exchange chain in the sharded pass test (process ranks, one GPU); N=1 lines of the C4 / C5 sub-sample workloads.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( timeout 900 python -m pytest tests/test_gpu_sharded_pass.py -q -m gpu ) > gpurun_out/v16_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit16.txt
for w in C4 C5; do
  ( timeout 900 python bench.py --steps 20 --warmup 5 --workload $w --no-cpu --no-file ) > gpurun_out/v16_bench_$w.json 2> gpurun_out/v16_bench_$w.err
  echo "bench $w rc=$?" >> gpurun_out/rc_visit16.txt
done
grep -v "^$" gpurun_out/v16_tests.log | grep -v "^    \|^kwargs\|^popenargs" | tail -20 | cut -c1-400; cat gpurun_out/rc_visit16.txt
python - <<'PY'
import json
for w in ("C4", "C5"):
    try:
        d = json.loads(open("gpurun_out/v16_bench_%s.json" % w).read().strip().splitlines()[-1])
        print(w, d["config"]["records_per_gpu"], round(d["ms_per_step"], 3), round(d["value"] / 1e6, 1), d["roofline"]["frac"], d["roofline"]["kernel_ms"], d["roofline"]["whole_pass"]["frac"], round(d["e2e"]["ms_per_step"], 1), d["parity_ok"], d["stage_ms_per_step"])
    except Exception as e:
        print(w, "parse failed", e); print(open("gpurun_out/v16_bench_%s.err" % w).read()[-2000:])
PY
