#!/bin/bash
# full ncu capture of the fused-verify kernels at the bench shape (after a plain run of the same command)
set -u
mkdir -p gpurun_out
timeout 300 python tools/ncu_probe.py ${1:-} > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"row_pair|row_ring|row_resident|row_scan|seq_sample" -s 6 -c 4 \
    -f -o gpurun_out/prof python tools/ncu_probe.py ${1:-} > gpurun_out/ncu_full.log 2>&1
echo "full capture exit $?"
ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv 2>/dev/null
ncu -i gpurun_out/prof.ncu-rep --page source --csv --kernel-name regex:"row_pair|row_ring" --launch-skip 0 --launch-count 1 > gpurun_out/prof_src_ring.csv 2>/dev/null
ncu -i gpurun_out/prof.ncu-rep --page source --csv --kernel-name regex:seq_sample --launch-skip 0 --launch-count 1 > gpurun_out/prof_src_sample.csv 2>/dev/null
