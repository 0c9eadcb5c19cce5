#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
for v in "S2T_LAG_JOIN=1" "S2T_LAG_JOIN=0" "S2T_LATE_TRIGGER=0" "S2T_SM_SHARES=90,58,-1" "S2T_SM_SHARES=110,38,-1" "S2T_SM_SHARES=0,0,0" "S2T_SM_SHARES=148,48,-1"; do
  env $v timeout 300 python bench.py --config c2 --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['ms_per_step'],4), 'ms', d['launches_per_step'], 'launches', 'gemm_ms', round(d['roofline']['gemm_ms_per_step'],3))"
done
timeout 300 python tools/timeline.py gpurun_out/timeline.csv > gpurun_out/timeline.txt 2>&1; head -3 gpurun_out/timeline.txt | tail -1
