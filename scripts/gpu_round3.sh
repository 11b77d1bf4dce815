#!/bin/bash
# v8 evidence: sweep of every workload with the CPU arm, then ncu full captures of the two CSQP kernels
set -u
mkdir -p gpurun_out
WORKLOADS="C2_ur5_sqp C1_ur5_sqp C1_ur5_csqp C3_solo12_csqp C3b_solo12_bench C4_lq_nx64 C4_clq_nx64 C4_lq_nx32 C4_clq_nx32 C4_lq_nx16 C4_lq_nx8 C5_talos_sqp C5_talos_csqp" bash scripts/gpu_sweep.sh
W=C3_solo12_csqp K=k_csqp_direction_stream O=ncu_c3_stream B=512 bash scripts/gpu_prof.sh
W=C5_talos_csqp K=k_csqp_direction_stream O=ncu_c5_direct B=148 bash scripts/gpu_prof.sh
ls -la gpurun_out
