#!/bin/bash
# sweep of the BASELINE workloads (device-resident arm + CPU baseline), one JSON line each
set -u
mkdir -p gpurun_out
: > gpurun_out/sweep.jsonl
for w in ${WORKLOADS:-C1_ur5_csqp C3_solo12_csqp C3b_solo12_bench C4_lq_nx64 C4_clq_nx64 C5_talos_sqp C5_talos_csqp}; do
  timeout 900 python bench.py --workload $w --steps ${STEPS:-5} --warmup 3 --no-e2e ${EXTRA:-} >> gpurun_out/sweep.jsonl 2>> gpurun_out/sweep.err; echo "$w rc=$?"
done
python - <<'PY'
import json
for l in open('gpurun_out/sweep.jsonl'):
    d=json.loads(l); r=d['roofline']; c=d.get('cpu_baseline') or {}
    print(d['config']['workload'], 'QP/s %.0f'%d['value'], 'ms %.2f'%d['ms_per_step'], r['bound'], 'frac %.3f'%r['frac'], 'fp64 TF %.2f'%r['fp64_tflops_achieved'], 'cpu %.0f'%c.get('value',0), 'cores', c.get('cores'))
PY
