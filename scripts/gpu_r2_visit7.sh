#!/bin/bash
set -u
mkdir -p gpurun_out
python - <<'PY' > gpurun_out/v7_all.json 2> gpurun_out/v7_all.err
import subprocess, sys
for r in (-1, 18, 74, 296):
    p = subprocess.run([sys.executable, "scripts/pass_timing.py", "--overlap", "1", "--reserve", str(r)], capture_output=True, text=True)
    print(p.stdout.strip().splitlines()[-1] if p.stdout.strip() else p.stderr[-500:])
PY
cat gpurun_out/v7_all.json
