#!/bin/bash
# full ncu capture of the fused-verify kernels at the bench shape (after a plain run of the same command)
set -u
mkdir -p gpurun_out
timeout 300 python tools/ncu_probe.py > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"row_scan|seq_sample|verify_stream" -s 6 -c 4 \
    -f -o gpurun_out/prof python tools/ncu_probe.py > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"
