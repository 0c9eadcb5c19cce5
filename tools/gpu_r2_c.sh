#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gemm_gpu.py -k "fused_bias" -x -q 2>&1 | tail -2
timeout 180 python -m pytest tests/test_kernels_gpu.py -k "attn_fused or reduce_batched" -x -q 2>&1 | tail -2
timeout 120 python tools/attn_trace.py > gpurun_out/attn_trace.txt 2>&1; echo "trace rc=$?"; cat gpurun_out/attn_trace.txt
timeout 300 python tools/kernel_bench.py 2>&1 | head -8
for v in "S2T_FUSED_ATTN=1" "S2T_FUSED_ATTN=0" "S2T_LATE_TRIGGER=0" "S2T_SM_SHARES=110,38,-1" "S2T_SM_SHARES=90,58,-1"; do
  env $v timeout 300 python bench.py --config c2 --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['ms_per_step'],4), 'ms', d['launches_per_step'], 'launches', 'gemm_ms', round(d['roofline']['gemm_ms_per_step'],3))"
done
timeout 300 python tools/timeline.py gpurun_out/timeline.csv > gpurun_out/timeline.txt 2>&1; head -24 gpurun_out/timeline.txt | tail -22
