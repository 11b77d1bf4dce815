#!/bin/bash
# bench a few workloads on each tuning variant of the library:  VARIANTS="base v1 v2" WS="C3_solo12_csqp C5_talos_csqp"
set -u
mkdir -p gpurun_out
for v in ${VARIANTS:-base}; do
  if [ $v = base ]; then unset MIMGPU_LIB; else export MIMGPU_LIB=$PWD/mim_solvers_b200/csrc/libmimgpu_$v.so; fi
  for w in ${WS:-C3_solo12_csqp}; do
    echo "== $v $w"
    MIMGPU_ADMM_TIMING=${TIMING:-1} python bench.py --workload $w --steps ${STEPS:-2} --warmup 1 --no-cpu --no-e2e --no-extra 2>&1 | grep -E "admm timing|\"value\"|rror" | tail -2 | cut -c1-200
  done
done
