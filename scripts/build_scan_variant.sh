#!/bin/bash
# Development aid: build libcoral_b200 with a variant of the record-scan kernel (extra -D flags) under another name.
#   scripts/build_scan_variant.sh NAME -DSC_MINB=3 -DSC_CH=4096   ->  coral_b200/libcoral_b200_NAME.so
set -e
name=$1; shift
cd "$(dirname "$0")/.."
python -m coral_b200.build > /dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC,-O2,-pthread --expt-relaxed-constexpr \
     "$@" -c coral_b200/csrc/scan_records.cu -o coral_b200/build/scan_records_$name.o
objs=$(ls coral_b200/build/*.o | grep -v "scan_records")
nvcc -shared -o coral_b200/libcoral_b200_$name.so $objs coral_b200/build/scan_records_$name.o -lz -Xcompiler -pthread
echo coral_b200/libcoral_b200_$name.so
