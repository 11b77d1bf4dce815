#!/bin/bash
# quick visit: small-kernel parity tests + default bench (no CPU/e2e arms)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "small or sqp_vs_oracle or sqp_golden" > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_quick.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e ${BENCH_ARGS:-} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_quick.json'));print(d['value'],d['ms_per_step'],d['roofline']['frac'])"
