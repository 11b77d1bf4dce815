#!/bin/bash
set -u
mkdir -p gpurun_out; : > gpurun_out/iters.txt
for w in ${WORKLOADS:-C1_ur5_csqp C3_solo12_csqp}; do for it in ${ITERS:-1 2 25 26 50}; do
  r=$(timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-e2e --no-cpu --qp-iters $it 2>>gpurun_out/iters.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.2f ms'%d['ms_per_step'])")
  echo "$w iters=$it $r" | tee -a gpurun_out/iters.txt
done; done
