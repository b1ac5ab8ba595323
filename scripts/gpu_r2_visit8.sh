#!/bin/bash
set -u
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_pass.py tests/test_gpu_cluster.py tests/test_gpu_fetch.py -x -q -m gpu ) > gpurun_out/v8_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_visit8.txt
python - <<'PY' > gpurun_out/v8_all.json 2> gpurun_out/v8_all.err
import subprocess, sys
for ov, fk, r in ((0, 0, -1), (0, 1, -1), (1, 0, 74)):
    p = subprocess.run([sys.executable, "scripts/pass_timing.py", "--overlap", str(ov), "--fork", str(fk), "--reserve", str(r)], capture_output=True, text=True)
    print(p.stdout.strip().splitlines()[-1] if p.stdout.strip() else p.stderr[-800:])
PY
tail -3 gpurun_out/v8_tests.log; cat gpurun_out/rc_visit8.txt; cat gpurun_out/v8_all.json
