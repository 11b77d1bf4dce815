#!/bin/bash
# CTA size of the streaming CSQP kernel on the small shapes (UR5 class): env switch MIMGPU_STREAM_NT and the
# single-/two-warp library variants of scripts/build_variant.sh (w1, w1s2, w2), then parity under the switch.
#   WS="C1_ur5_csqp" [TESTS=1] bash scripts/gpu_stream_nt.sh  ->  gpurun_out/stream_nt_<TAG>.log
set -u
TAG=${TAG:-r02_v6}
O=gpurun_out
mkdir -p $O
L=$O/stream_nt_$TAG.log
: > $L
run() { # label, then env assignments
  local label=$1; shift
  for w in ${WS:-C1_ur5_csqp}; do
    echo "== $label $w" >> $L
    env "$@" python bench.py --workload $w --steps ${STEPS:-3} --warmup 1 --no-cpu --no-e2e --no-extra 2>&1 \
      | grep -E "\"value\"|rror" | tail -1 | cut -c1-230 >> $L
  done
}
C=$PWD/mim_solvers_b200/csrc
run base X=0
run nt32 MIMGPU_STREAM_NT=32
run nt64 MIMGPU_STREAM_NT=64
for v in ${VARIANTS:-w1 w1s2 w2}; do
  [ -f $C/libmimgpu_$v.so ] && run $v MIMGPU_LIB=$C/libmimgpu_$v.so
done
if [ "${TESTS:-1}" = 1 ]; then
  echo "== parity, MIMGPU_STREAM_NT=32" >> $L
  MIMGPU_STREAM_NT=32 python -m pytest tests -m gpu -q -x -k "csqp or CSQP or Csqp" 2>&1 | tail -4 >> $L
  for v in ${TEST_VARIANTS:-w1s2}; do
    echo "== parity, variant $v" >> $L
    MIMGPU_LIB=$C/libmimgpu_$v.so python -m pytest tests -m gpu -q -x -k "csqp or CSQP or Csqp" 2>&1 | tail -4 >> $L
  done
  echo "== full GPU suite, default library" >> $L
  python -m pytest tests -m gpu -q -x 2>&1 | tail -4 >> $L
fi
cat $L
