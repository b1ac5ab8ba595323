#!/bin/bash
set -u
mkdir -p gpurun_out
( timeout 300 python scripts/pass_timing.py --overlap 1 ) > gpurun_out/v7_pt1.json 2> gpurun_out/v7_pt1.err
( timeout 300 python scripts/pass_timing.py --overlap 0 ) > gpurun_out/v7_pt0.json 2> gpurun_out/v7_pt0.err
cat gpurun_out/v7_pt1.json gpurun_out/v7_pt0.json; tail -3 gpurun_out/v7_pt1.err
