#!/bin/bash
# One GPU-box visit: parity tests, the bench line, the ncu launch list of the same command and
# one full capture of the step's kernels.  Outputs land in gpurun_out/ (copied to profiles/ by hand).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_ref.json
if [ "${1:-}" = "ncu" ]; then
  timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches.csv python bench.py --steps 8 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
  echo "launch list exit $?"
  bash tools/gpu_ncu.sh
fi
