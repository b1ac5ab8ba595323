#!/bin/bash
# GPU-box visit 5: the final tree (placement "auto" = asynchronous scan, bench e2e default mapped_async) — all GPU
# tests, then the default bench line.
set -u
mkdir -p gpurun_out
( time timeout 275 python -m pytest tests -q -m gpu ) > gpurun_out/pytest_all5.log 2>&1
echo "pytest_all rc=$?" > gpurun_out/rc5.txt
( time timeout 55 python bench.py ) > gpurun_out/bench5.json 2> gpurun_out/bench5.err
echo "bench rc=$?" >> gpurun_out/rc5.txt
tail -15 gpurun_out/pytest_all5.log | cut -c1-400; cat gpurun_out/rc5.txt
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench5.json").read().strip().splitlines()[0])
    print("value", d["value"], "ms", d["ms_per_step"], "e2e", json.dumps(d["e2e"]))
except Exception as ex:
    print("bench5 unreadable", ex)
PY
