#!/bin/bash
# Round-2 kernel pass: the new kernels' unit tests first (each under its own timeout: a barrier bug must not hang the
# box), then the whole GPU suite, then the C2 bench and the kernel timeline.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/grad_tolerances.jsonl
timeout 180 python -m pytest tests/test_gemm_gpu.py -k "fused_bias" -x -q > gpurun_out/t_wgrad.log 2>&1; rc1=$?; echo "wgrad-colsum rc=$rc1"; tail -3 gpurun_out/t_wgrad.log
timeout 180 python -m pytest tests/test_kernels_gpu.py -k "attn_fused or reduce_batched or peaked" -x -q > gpurun_out/t_attn.log 2>&1; rc2=$?; echo "attn/ln rc=$rc2"; tail -3 gpurun_out/t_attn.log
if [ $rc1 -ne 0 ] || [ $rc2 -ne 0 ]; then
  grep -n "Error\|assert\|^E " gpurun_out/t_wgrad.log gpurun_out/t_attn.log | head -40
fi
timeout 1200 python -m pytest tests -m gpu -q -rf > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --config c2 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
head -c 400 gpurun_out/bench_c2.json; echo
timeout 300 python tools/timeline.py gpurun_out/timeline.csv > gpurun_out/timeline.txt 2>&1; echo "timeline rc=$?"
timeout 300 python tools/kernel_bench.py > gpurun_out/kernel_bench.txt 2>&1; echo "kbench rc=$?"; cat gpurun_out/kernel_bench.txt
head -30 gpurun_out/timeline.txt
