#!/bin/bash
# final evidence of the round: full GPU suite, default bench (+ reference arm), launch list, sweep with the CPU arm
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
WORKLOADS="C2_ur5_sqp C1_ur5_sqp C1_ur5_csqp C3_solo12_csqp C3b_solo12_bench C4_lq_nx64 C4_clq_nx64 C4_lq_nx32 C4_clq_nx32 C4_lq_nx16 C4_lq_nx8 C5_talos_sqp C5_talos_csqp" bash scripts/gpu_sweep.sh
