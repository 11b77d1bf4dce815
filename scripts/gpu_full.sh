#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
WORKLOADS="${WORKLOADS:-C1_ur5_csqp C3_solo12_csqp C3b_solo12_bench C4_lq_nx64 C5_talos_sqp C4_lq_nx32 C4_lq_nx16}" EXTRA="--no-cpu" bash scripts/gpu_sweep.sh
