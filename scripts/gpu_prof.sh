#!/bin/bash
# ncu full capture of one kernel of one workload: W=<workload> K=<kernel regex> O=<out name> [B=<batch>] [KS=<mangled-name substring for the line join, e.g. k_csqp_direction_admmILi36ELi12ELi1E>] [KEEP_REP=1]
# The report is reduced ON THE BOX to what profiles/ keeps (raw metrics as metric,unit,value; hottest source
# lines) and then deleted: gpurun brings back at most 64 MiB and one report of a CSQP kernel is 35-40 MB.
set -u
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${K} -s ${SKIP:-3} -c 1 -o gpurun_out/${O} -f python bench.py --workload ${W} --steps 3 --warmup 3 --no-cpu --no-e2e --no-extra ${B:+--batch $B} > gpurun_out/${O}.log 2>&1; echo "ncu ${O} rc=$?"
R=gpurun_out/${O}.ncu-rep
if [ -f $R ]; then
  ncu -i $R --page raw --csv > gpurun_out/${O}.raw.csv 2>/dev/null
  python scripts/ncu_transpose.py gpurun_out/${O}.raw.csv > gpurun_out/${O}.full.csv
  python scripts/ncu_keys.py gpurun_out/${O}.raw.csv > gpurun_out/${O}.keys.txt 2>&1
  rm -f gpurun_out/${O}.raw.csv
  python scripts/ncu_lines.py $R ${KS:-$K} --top 60 > gpurun_out/${O}.lines.txt 2>&1
  [ "${KEEP_REP:-0}" = 1 ] || rm -f $R
fi
