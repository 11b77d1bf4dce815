#!/bin/bash
# one-warp-per-problem variant of the record kernel (MIMGPU_REC_W1=1): parity, then A/B
set -u
TAG=${TAG:-r02_v9}
O=gpurun_out; mkdir -p $O
L=$O/rec_w1_$TAG.log
: > $L
echo "== parity, MIMGPU_CSQP_REC=1 MIMGPU_REC_W1=1" >> $L
MIMGPU_CSQP_REC=1 MIMGPU_REC_W1=1 timeout 600 python -m pytest tests -m gpu -q -x -k "csqp or CSQP or Csqp" 2>&1 | tail -12 >> $L
run() {
  local label=$1; shift
  for w in ${WS:-C1_ur5_csqp}; do
    echo "== $label $w" >> $L
    env "$@" timeout 300 python bench.py --workload $w --steps ${STEPS:-3} --warmup 1 --no-cpu --no-e2e --no-extra 2>&1 \
      | grep -E "\"value\"|rror" | tail -1 | cut -c1-200 >> $L
  done
}
run stream MIMGPU_CSQP_REC=0
run rec MIMGPU_CSQP_REC=1
run rec_w1 MIMGPU_CSQP_REC=1 MIMGPU_REC_W1=1
run rec_w1_ns2 MIMGPU_CSQP_REC=1 MIMGPU_REC_W1=1 MIMGPU_REC_NS=2
run rec_w1_ns4 MIMGPU_CSQP_REC=1 MIMGPU_REC_W1=1 MIMGPU_REC_NS=4
if [ "${NCU:-0}" = 1 ]; then
  MIMGPU_CSQP_REC=1 MIMGPU_REC_W1=1 W=C1_ur5_csqp K=k_csqp_direction_rec KS=k_csqp_direction_recILi12ELi6ELi4ELb1 O=ncu_${TAG}_csqp_rec_w1_ur5 bash scripts/gpu_prof.sh >> $L 2>&1
fi
cat $L
