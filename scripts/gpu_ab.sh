#!/bin/bash
# A/B of kernel switches on a few workloads, with parity under the same switches first:
#   TAG=r02_v10 WS="C1_ur5_csqp" TESTS="csqp or CSQP or Csqp" \
#   ARMS="stream:MIMGPU_CSQP_REC=0 rec:MIMGPU_REC_W1=0 w1: w1_ns2:MIMGPU_REC_NS=2" \
#   [TEST_ENV="MIMGPU_REC_W1=0"] [NCU_ENV="..." NCU_K=k_csqp_direction_rec NCU_KS=k_csqp_direction_recILi12ELi6ELi4ELb1] \
#   bash scripts/gpu_ab.sh                      ->  gpurun_out/ab_<TAG>.log (+ the ncu reductions of gpu_prof.sh)
# An arm is label:VAR=val,VAR=val (empty = the defaults); MIMGPU_LIB=<variant .so of build_variant.sh> is one such VAR.
# (This replaces the one-off visit scripts of the record-kernel work: profiles/variants_r02_v6_stream_nt.log,
#  rec_r02_v7.log, rec_knobs_r02_v8.log, rec_w1_r02_v9.log, rec_w1_r02_v10.log were produced this way.)
set -u
TAG=${TAG:-ab}
O=gpurun_out; mkdir -p $O
L=$O/ab_$TAG.log
: > $L
if [ -n "${TESTS:-}" ]; then
  echo "== parity (-k '$TESTS'), env: ${TEST_ENV:-defaults}" >> $L
  env ${TEST_ENV:-X=0} timeout 900 python -m pytest tests -m gpu -q -x -k "$TESTS" 2>&1 | tail -8 >> $L
fi
for arm in ${ARMS:-base:}; do
  label=${arm%%:*}; vars=${arm#*:}
  for w in ${WS:-C1_ur5_csqp}; do
    echo "== $label $w" >> $L
    env $(echo "${vars:-X=0}" | tr ',' ' ') timeout 300 python bench.py --workload $w --steps ${STEPS:-3} --warmup 1 --no-cpu --no-e2e --no-extra 2>&1 \
      | grep -E "\"value\"|rror" | tail -1 | cut -c1-220 >> $L
  done
done
if [ -n "${NCU_K:-}" ]; then
  env ${NCU_ENV:-X=0} W=${NCU_W:-C1_ur5_csqp} K=$NCU_K KS=${NCU_KS:-$NCU_K} O=ncu_${TAG} bash scripts/gpu_prof.sh >> $L 2>&1
fi
cat $L
