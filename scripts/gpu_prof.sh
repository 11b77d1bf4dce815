#!/bin/bash
# ncu full capture of one kernel of one workload: W=<workload> K=<kernel regex> O=<out name> [B=<batch>]
set -u
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${K} -s ${SKIP:-3} -c 1 -o gpurun_out/${O} -f python bench.py --workload ${W} --steps 3 --warmup 3 --no-cpu --no-e2e --no-extra ${B:+--batch $B} > gpurun_out/${O}.log 2>&1; echo "ncu ${O} rc=$?"
