#!/bin/bash
# quick iteration visit: parity tests + pass statistics / timing probe
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python tools/stats_probe.py > gpurun_out/stats.log 2>&1; tail -8 gpurun_out/stats.log
