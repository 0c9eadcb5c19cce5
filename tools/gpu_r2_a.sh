#!/bin/bash
# Round-2 first GPU pass: the whole GPU test suite (no -x: every failure is wanted), smoke, bench per config, the recipe
# batch, the bucketed workload, and the reference arm.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
rm -f gpurun_out/grad_tolerances.jsonl
timeout 1500 python -m pytest tests -m gpu -q -rf --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
for c in c2 c3 c4 c5; do
  timeout 600 python bench.py --config $c --steps 30 --warmup 5 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "bench $c rc=$?"
done
timeout 600 python bench.py --config c2 --batch 40 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c2_b40.json 2> gpurun_out/bench_c2_b40.err; echo "bench b40 rc=$?"
timeout 900 python bench.py --config c3 --bucketed --steps 60 --warmup 3 --utterances 2048 > gpurun_out/bench_c3_bucketed_n1.json 2> gpurun_out/bench_c3_bucketed_n1.err; echo "bucketed rc=$?"
timeout 600 python bench.py --impl reference --steps 10 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
head -c 600 gpurun_out/bench_c2.json; echo
ls -la gpurun_out | tail -30
