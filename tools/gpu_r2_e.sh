#!/bin/bash
# full GPU suite + smoke + bench lines per config (1 GPU)
set -u
mkdir -p gpurun_out
rm -f gpurun_out/grad_tolerances.jsonl
timeout 1500 python -m pytest tests -m gpu -q -rf > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
for c in c2 c3 c4 c5; do
  timeout 600 python bench.py --config $c --steps 30 --warmup 5 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"
done
timeout 600 python bench.py --config c2 --batch 40 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c2_b40.json 2> gpurun_out/bench_c2_b40.err; echo "bench b40 rc=$?"
timeout 900 python bench.py --config c3 --bucketed --steps 60 --warmup 3 --utterances 2048 > gpurun_out/bench_c3_bucketed_n1.json 2> gpurun_out/bench_c3_bucketed_n1.err; echo "bucketed rc=$?"; tail -3 gpurun_out/bench_c3_bucketed_n1.err
for c in c2 c3 c4 c5 c2_b40 c3_bucketed_n1; do python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_$c.json'))
    r=d.get('roofline') or {}
    print('$c', round(d['ms_per_step'],3),'ms', round(d['value']),'a-s/s e2e',round(d['e2e']['value']), 'launches',d.get('launches_per_step'),'gemm frac',r.get('frac') and round(r['frac'],3), d.get('bucketing'))
except Exception as e: print('$c', 'ERR', e)
PY
done
