#!/bin/bash
# Final round-2 visit on one GPU: the whole GPU suite, the bench line, then the ncu evidence (launch list of one device
# pass, full capture of the scan kernel and of the inflate kernel).  Each ncu pass runs only after its command exited 0 without ncu.
set -u
mkdir -p gpurun_out
export CUDA_DEVICE_MAX_CONNECTIONS=32
( time timeout 1800 python -m pytest tests -x -q -m gpu ) > gpurun_out/final_tests.log 2>&1
echo "tests rc=$?" > gpurun_out/rc_final.txt
( timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
echo "bench rc=$?" >> gpurun_out/rc_final.txt
( timeout 600 python bench.py --impl reference --steps 1 --warmup 0 ) > gpurun_out/final_reference.json 2> gpurun_out/final_reference.err
echo "reference rc=$?" >> gpurun_out/rc_final.txt
( timeout 300 python bench.py --profile --no-cpu --no-file ) > gpurun_out/final_profile_plain.log 2>&1 && \
( timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/r2_launches_final.csv python bench.py --profile --no-cpu --no-file ) > gpurun_out/final_ncu_launches.log 2>&1
echo "launch list rc=$?" >> gpurun_out/rc_final.txt
( timeout 300 python bench.py --only scan --steps 3 --warmup 3 ) > gpurun_out/final_scan_plain.json 2>&1 && \
( timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_scan_records -s 3 -c 1 -f -o gpurun_out/r2_scan_final \
    python bench.py --only scan --steps 3 --warmup 3 ) > gpurun_out/final_ncu_scan.log 2>&1
echo "scan capture rc=$?" >> gpurun_out/rc_final.txt
( timeout 600 python scripts/ingest_timing.py --scale 0.25 --reps 3 ) > gpurun_out/final_ingest.json 2> gpurun_out/final_ingest.err
echo "ingest rc=$?" >> gpurun_out/rc_final.txt
( timeout 900 ncu --set full --clock-control none -k regex:k_inflate_members -c 1 -f -o gpurun_out/r2_inflate \
    python scripts/ingest_timing.py --scale 0.05 --reps 1 ) > gpurun_out/final_ncu_inflate.log 2>&1
echo "inflate capture rc=$?" >> gpurun_out/rc_final.txt
for w in C3 C1n; do
  ( timeout 600 python bench.py --steps 20 --warmup 5 --workload $w --no-cpu --no-file ) > gpurun_out/final_bench_$w.json 2> gpurun_out/final_bench_$w.err
  echo "bench $w rc=$?" >> gpurun_out/rc_final.txt
done
tail -4 gpurun_out/final_tests.log; cat gpurun_out/rc_final.txt; cat gpurun_out/final_ingest.json; tail -c 600 gpurun_out/final_reference.json
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/final_bench.json").read().strip().splitlines()[-1])
    print(d["steps"], d["ms_per_step"], d["value"], d["gpu_launches"], d["roofline"]["frac"], d["roofline"]["kernel_ms"], d["roofline"]["whole_pass"]["frac"], d["e2e"]["ms_per_step"], d["parity_ok"], d["cpu_baseline"]["kind"], d["cpu_baseline"]["value"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/final_bench.err").read()[-2000:])
PY
