#!/bin/bash
# One-GPU measurement pass of a round (run under gpurun from the repo root): GPU tests, smoke, both bench arms, the
# per-kernel table, kernel timelines with and without programmatic dependent launch, and — only after the plain runs
# have exited 0 — the ncu launch list (durations + DRAM bytes + tensor-pipe activity) and one full capture of the
# dominant GEMM kernel.  Everything lands in gpurun_out/; the summaries are copied into profiles/ by hand.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python bench.py --no-cpu-baseline --profile-kernels > /dev/null 2> gpurun_out/kernel_table.txt; echo "table rc=$?"
S2T_PDL=0 python tools/timeline.py gpurun_out/timeline_nopdl.csv > gpurun_out/timeline_nopdl.txt 2>&1
python tools/timeline.py gpurun_out/timeline.csv > gpurun_out/timeline.txt 2>&1
python tools/chain_only.py > gpurun_out/chain_only.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
  --clock-control none -c 2400 --csv --log-file gpurun_out/launches_mem.csv \
  python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline > gpurun_out/ncu_mem.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_kernel -s 60 -c 4 -f -o gpurun_out/gemm_full \
  python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -30
