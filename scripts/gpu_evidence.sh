#!/bin/bash
# One GPU-box visit that produces the round's evidence under gpurun_out/ (copied to profiles/ afterwards):
#   TAG=r02_v1 [TESTS=1] [SWEEP="C2_ur5_sqp ..."] [NCU=1|list|0] bash scripts/gpu_evidence.sh
# Order matters: every command runs WITHOUT ncu first; numbers printed under ncu are never bench values.
set -u
TAG=${TAG:-r02}
O=gpurun_out
mkdir -p $O
if [ "${TESTS:-1}" = 1 ]; then python -m pytest tests -m gpu -q 2>&1 | tail -6 > $O/tests_$TAG.log; fi
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$TAG.log 2>&1
python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err
python bench.py --impl reference > $O/bench_ref_$TAG.json 2>/dev/null
python scripts/peaks.py > $O/peaks_$TAG.json 2>/dev/null
for w in ${SWEEP:-}; do python bench.py --workload $w --steps ${SWEEP_STEPS:-5} --no-extra 2>/dev/null; done > $O/sweep_$TAG.jsonl
if [ "${NCU:-1}" != 0 ]; then
  # launch list of the default bench command (short: 20 steps), then (NCU=1) one full capture per dominant kernel
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_$TAG.csv \
      python bench.py --steps 20 --warmup 3 --no-cpu > $O/ncu_list_$TAG.log 2>&1
fi
if [ "${NCU:-1}" = 1 ]; then
  W=C2_ur5_sqp K=k_sqp_direction_small_gj KS=k_sqp_direction_small_gjILi12ELi6E O=ncu_${TAG}_small_gj_sqp bash scripts/gpu_prof.sh
  W=C3_solo12_csqp K=k_csqp_direction_admm KS=k_csqp_direction_admmILi36ELi12ELi1E O=ncu_${TAG}_csqp_admm_solo12 B=296 bash scripts/gpu_prof.sh
  W=C5_talos_csqp K=k_csqp_direction_admm KS=k_csqp_direction_admmILi76ELi32ELi2E O=ncu_${TAG}_csqp_admm_talos B=148 bash scripts/gpu_prof.sh
fi
