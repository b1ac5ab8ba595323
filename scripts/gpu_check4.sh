#!/bin/bash
# GPU-box visit 4: pre-scan v2 (staged aux block + SA side buffer), lazy placement, L2 vectors through the drop-in
# modules — all GPU tests, then the bench line with the default (sync) e2e and the asynchronous e2e beside it.
set -u
mkdir -p gpurun_out
( time timeout 700 python -m pytest tests -q -m gpu ) > gpurun_out/pytest_all4.log 2>&1
echo "pytest_all rc=$?" > gpurun_out/rc4.txt
( time timeout 300 python bench.py ) > gpurun_out/bench4.json 2> gpurun_out/bench4.err
echo "bench rc=$?" >> gpurun_out/rc4.txt
( time timeout 300 python bench.py --e2e-placement mapped_async --steps 3 ) > gpurun_out/bench4_async.json 2> gpurun_out/bench4_async.err
echo "bench_async rc=$?" >> gpurun_out/rc4.txt
tail -15 gpurun_out/pytest_all4.log | cut -c1-400; cat gpurun_out/rc4.txt
python - <<'PY'
import json
for f in ("gpurun_out/bench4.json", "gpurun_out/bench4_async.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[0])
        print(f, "value", d["value"], "ms", d["ms_per_step"], "e2e", json.dumps(d["e2e"]))
    except Exception as ex:
        print(f, "unreadable", ex)
PY
