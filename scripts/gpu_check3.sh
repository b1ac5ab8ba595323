#!/bin/bash
# GPU-box visit 3: asynchronous scan — all GPU tests, then the bench line (e2e with the overlap) and the sync e2e beside it.
set -u
mkdir -p gpurun_out
( time timeout 1000 python -m pytest tests -q -m gpu -x ) > gpurun_out/pytest_all3.log 2>&1
echo "pytest_all rc=$?" > gpurun_out/rc3.txt
( time timeout 600 python bench.py ) > gpurun_out/bench3.json 2> gpurun_out/bench3.err
echo "bench rc=$?" >> gpurun_out/rc3.txt
( time timeout 600 python bench.py --e2e-placement mapped --steps 3 ) > gpurun_out/bench3_sync.json 2> gpurun_out/bench3_sync.err
echo "bench_sync rc=$?" >> gpurun_out/rc3.txt
tail -4 gpurun_out/pytest_all3.log; cat gpurun_out/rc3.txt
python - <<'PY'
import json
for f in ("gpurun_out/bench3.json", "gpurun_out/bench3_sync.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[0])
        print(f, "value", d["value"], "ms", d["ms_per_step"], "e2e", json.dumps(d["e2e"]))
    except Exception as ex:
        print(f, "unreadable", ex)
PY
