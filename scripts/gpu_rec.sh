#!/bin/bash
# Record-based CSQP kernel of the small shapes (kernels_rec.cuh, MIMGPU_CSQP_REC=1): parity under the switch,
# then A/B against the streaming kernel on the workloads in WS; optional ncu capture (NCU=1).
#   TAG=r02_v7 WS="C1_ur5_csqp" bash scripts/gpu_rec.sh  ->  gpurun_out/rec_<TAG>.log
set -u
TAG=${TAG:-r02_v7}
O=gpurun_out
mkdir -p $O
L=$O/rec_$TAG.log
: > $L
if [ "${TESTS:-1}" = 1 ]; then
  echo "== parity, MIMGPU_CSQP_REC=1" >> $L
  MIMGPU_CSQP_REC=1 timeout 600 python -m pytest tests -m gpu -q -x -k "csqp or CSQP or Csqp" 2>&1 | tail -15 >> $L
fi
run() {
  local label=$1; shift
  for w in ${WS:-C1_ur5_csqp}; do
    echo "== $label $w" >> $L
    env "$@" timeout 300 python bench.py --workload $w --steps ${STEPS:-3} --warmup 1 --no-cpu --no-e2e --no-extra 2>&1 \
      | grep -E "\"value\"|rror" | tail -1 | cut -c1-230 >> $L
  done
}
run stream MIMGPU_CSQP_REC=0
run rec MIMGPU_CSQP_REC=1
run rec_nt32 MIMGPU_CSQP_REC=1 MIMGPU_STREAM_NT=32
run rec_nt64 MIMGPU_CSQP_REC=1 MIMGPU_STREAM_NT=64
if [ "${NCU:-0}" = 1 ]; then
  MIMGPU_CSQP_REC=1 W=C1_ur5_csqp K=k_csqp_direction_rec KS=k_csqp_direction_recILi12ELi6E O=ncu_${TAG}_csqp_rec_ur5 bash scripts/gpu_prof.sh >> $L 2>&1
  MIMGPU_CSQP_REC=0 W=C1_ur5_csqp K=k_csqp_direction_stream KS=k_csqp_direction_streamILi12ELi6E O=ncu_${TAG}_csqp_stream_ur5 bash scripts/gpu_prof.sh >> $L 2>&1
fi
cat $L
