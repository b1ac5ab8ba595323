#!/bin/bash
# Overlapped scan inside the pass (pre-scan, full scan on its own stream with block slots left free) against the default.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
python - <<'PY' > gpurun_out/v15_all.json 2> gpurun_out/v15_all.err
import subprocess, sys
for ov, r in ((0, -1), (1, -1), (1, 74), (1, 148), (1, 296)):
    p = subprocess.run([sys.executable, "scripts/pass_timing.py", "--overlap", str(ov), "--fork", "1", "--reserve", str(r)], capture_output=True, text=True)
    print(p.stdout.strip().splitlines()[-1] if p.stdout.strip() else p.stderr[-800:])
PY
cat gpurun_out/v15_all.json
