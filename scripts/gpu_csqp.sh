#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "csqp or solver" > gpurun_out/pytest_csqp.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_csqp.log
WORKLOADS="${WORKLOADS:-C1_ur5_csqp C3_solo12_csqp C3b_solo12_bench}" EXTRA="--no-cpu" bash scripts/gpu_sweep.sh
