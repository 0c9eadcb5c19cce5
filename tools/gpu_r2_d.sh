#!/bin/bash
set -u
for v in "S2T_X=1" "S2T_LAG_JOIN=0" "S2T_MULTI_STREAM=0" "S2T_FUSED_ATTN=0"; do
  rm -f gpurun_out/grad_tolerances.jsonl
  env $v timeout 600 python -m pytest tests/test_engine_gpu.py -q -k "full_size and c5" 2>&1 | tail -1
  python - <<PY
import json
for l in open('gpurun_out/grad_tolerances.jsonl'):
    d=json.loads(l)
    for r in d['widened']+d['over_5e-2']:
        if r[0].startswith('xctc_norm15') or r[0].startswith('ctc_norm15.w'):
            print('$v', r[0], round(r[1],4), 'noise', round(r[3],4))
PY
done
