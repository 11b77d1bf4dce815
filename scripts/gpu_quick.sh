#!/bin/bash
# quick visit: SQP parity tests + default bench (no CPU/e2e arms)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_quick.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e ${BENCH_ARGS:-} > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_quick.json'));print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['batch_summary'])"
