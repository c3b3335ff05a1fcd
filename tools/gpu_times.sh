#!/bin/bash
# per-kernel durations (ncu, cold-ish) of the fused verify with and without the nucleus filter
mkdir -p gpurun_out
for p in 0.0 0.9; do
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"row_pair|row_ring|seq_sample" --csv python tools/ncu_probe.py $p 2>/dev/null | grep -E "row_pair|row_ring|seq_sample" | awk -F'","' '{print $5, $(NF-2), $NF}' | tail -8
  echo "--- top_p $p"
done
