#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py -q -k "empty_utterances or full_size" 2>&1 | tail -5
for v in "S2T_X=1" "S2T_FUSED_ATTN=0" "S2T_LAG_JOIN=0" "S2T_SM_SHARES=0,0,0"; do
  env $v timeout 300 python bench.py --config c2 --batch 40 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('b40 $v', round(d['ms_per_step'],4), 'ms', 'gemm_ms', round(d['roofline']['gemm_ms_per_step'],3))"
done
timeout 900 python bench.py --config c3 --bucketed --steps 60 --warmup 3 --utterances 2048 > gpurun_out/bench_c3_bucketed_n1.json 2> gpurun_out/bench_c3_bucketed_n1.err; echo "bucketed rc=$?"; tail -3 gpurun_out/bench_c3_bucketed_n1.err; cut -c1-1500 gpurun_out/bench_c3_bucketed_n1.json
