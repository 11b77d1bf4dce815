#!/bin/bash
# per-phase clock64 timing of the condensed ADMM kernel (CTA 0) at a few batch sizes:  W=<workload> BS="1 148 296 2048"
set -u
mkdir -p gpurun_out
W=${W:-C3_solo12_csqp}
for B in ${BS:-1 148 296 2048}; do
  echo "== $W B=$B"
  MIMGPU_ADMM_TIMING=1 python bench.py --workload $W --batch $B --steps 2 --warmup 1 --no-cpu --no-e2e --no-extra 2>&1 | grep -E "admm timing|\"value\"" | tail -2 | cut -c1-260
done
