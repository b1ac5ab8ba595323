#!/bin/bash
# Round 2, first GPU visit: supervised first run of the device-side ingest kernels (never run on a B200 so far), each
# step under its own timeout so that a hang costs seconds, then the timing script.
set -u
mkdir -p gpurun_out
( CORAL_TEST_UNVERIFIED=1 timeout 240 python -m pytest tests/test_unverified_device_ingest.py -q -m gpu -x ) > gpurun_out/ingest_tests.log 2>&1
echo "ingest tests rc=$?" > gpurun_out/rc_ingest.txt
( timeout 200 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/ingest_timing.py --scale 0.005 --reps 1 ) > gpurun_out/ingest_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/rc_ingest.txt
( timeout 240 python scripts/ingest_timing.py --scale 0.1 ) > gpurun_out/ingest_timing.json 2> gpurun_out/ingest_timing.err
echo "timing rc=$?" >> gpurun_out/rc_ingest.txt
tail -5 gpurun_out/ingest_tests.log; tail -5 gpurun_out/ingest_memcheck.log; cat gpurun_out/rc_ingest.txt; cat gpurun_out/ingest_timing.json
