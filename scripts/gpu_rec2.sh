#!/bin/bash
# record-kernel knobs on one workload: ring depth (MIMGPU_REC_NS), resident problems per SM (MIMGPU_REC_MAXCTA), CTA size
set -u
TAG=${TAG:-r02_v8}
O=gpurun_out; mkdir -p $O
L=$O/rec_knobs_$TAG.log
: > $L
echo "== parity, MIMGPU_CSQP_REC=1 (both_kernels + fullsize)" >> $L
MIMGPU_CSQP_REC=1 timeout 600 python -m pytest tests -m gpu -q -x -k "both_kernels or fullsize_csqp or csqp_vs_oracle or csqp_direction_host" 2>&1 | tail -5 >> $L
run() {
  local label=$1; shift
  for w in ${WS:-C1_ur5_csqp}; do
    echo "== $label $w" >> $L
    env "$@" timeout 300 python bench.py --workload $w --steps ${STEPS:-3} --warmup 1 --no-cpu --no-e2e --no-extra 2>&1 \
      | grep -E "\"value\"|rror" | tail -1 | cut -c1-200 >> $L
  done
}
run stream MIMGPU_CSQP_REC=0
run rec_ns3 MIMGPU_CSQP_REC=1
run rec_ns2 MIMGPU_CSQP_REC=1 MIMGPU_REC_NS=2
run rec_ns4 MIMGPU_CSQP_REC=1 MIMGPU_REC_NS=4
run rec_ns6 MIMGPU_CSQP_REC=1 MIMGPU_REC_NS=6
run rec_ns8 MIMGPU_CSQP_REC=1 MIMGPU_REC_NS=8
run rec_ns3_cta4 MIMGPU_CSQP_REC=1 MIMGPU_REC_MAXCTA=4
run rec_ns3_cta6 MIMGPU_CSQP_REC=1 MIMGPU_REC_MAXCTA=6
run rec_ns6_cta5 MIMGPU_CSQP_REC=1 MIMGPU_REC_NS=6 MIMGPU_REC_MAXCTA=5
run rec_ns4_nt64 MIMGPU_CSQP_REC=1 MIMGPU_REC_NS=4 MIMGPU_STREAM_NT=64
if [ "${NCU:-0}" = 1 ]; then
  MIMGPU_CSQP_REC=1 W=C1_ur5_csqp K=k_csqp_direction_rec KS=k_csqp_direction_recILi12ELi6E O=ncu_${TAG}_csqp_rec_ur5 bash scripts/gpu_prof.sh >> $L 2>&1
fi
cat $L
