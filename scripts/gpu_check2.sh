#!/bin/bash
# GPU-box visit 2: all GPU tests, bench, the launch list of the 8f rows.
set -u
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -q -m gpu ) > gpurun_out/pytest_all.log 2>&1
echo "pytest_all rc=$?" > gpurun_out/rc2.txt
( time timeout 600 python bench.py ) > gpurun_out/bench2.json 2> gpurun_out/bench2.err
echo "bench rc=$?" >> gpurun_out/rc2.txt
( timeout 300 python scripts/next_rows_pass.py ) > gpurun_out/next_rows.json 2> gpurun_out/next_rows.err
echo "next_rows rc=$?" >> gpurun_out/rc2.txt
( timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 600 --csv \
    --log-file gpurun_out/launches_next_rows.csv python scripts/next_rows_pass.py ) > gpurun_out/ncu_next_rows.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc2.txt
tail -4 gpurun_out/pytest_all.log; cat gpurun_out/rc2.txt; cat gpurun_out/next_rows.json; cut -c1-300 gpurun_out/bench2.json
