#!/bin/bash
# Data-parallel pass of round 3 (run under `gpurun --gpus N`): for every world size given on the command line, the fixed
# C2 step (checks the clean teardown: stderr must say "destroy_process_group returned") and the length-bucketed C3
# workload with the collective-free capture warm-up (the N=8 deadlock of round 2).  Every run sits under its own timeout;
# bench.py adds a 180 s NCCL timeout and a watchdog.  usage: bash tools/gpu_r3_dp.sh 1 2   (or 1 2 4 8)
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in "$@"; do
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --config c3 --bucketed --utterances 8192 --steps 60 --warmup 3 --no-cpu-baseline 2> gpurun_out/r3_bench_c3_bucketed_n1.err | grep '^{' > gpurun_out/r3_bench_c3_bucketed_n1.json; echo "c3 bucketed n1 rc=$?"
    continue
  fi
  timeout 400 $TR --nproc-per-node $n --master-port 2961$n bench.py --gpus $n --steps 30 --warmup 5 --no-cpu-baseline 2> gpurun_out/r3_bench_c2_n$n.err | grep '^{' > gpurun_out/r3_bench_c2_n$n.json; echo "c2 n$n rc=$?"
  grep -h "teardown:" gpurun_out/r3_bench_c2_n$n.err | sort | uniq -c
  timeout 700 $TR --nproc-per-node $n --master-port 2962$n bench.py --gpus $n --config c3 --bucketed --utterances 8192 --steps 60 --warmup 3 2> gpurun_out/r3_bench_c3_bucketed_n$n.err | grep '^{' > gpurun_out/r3_bench_c3_bucketed_n$n.json; echo "c3 bucketed n$n rc=$?"
  grep -h "teardown:" gpurun_out/r3_bench_c3_bucketed_n$n.err | sort | uniq -c; tail -2 gpurun_out/r3_bench_c3_bucketed_n$n.err
done
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r3_bench_c*_n*.json")):
    try:
        d = json.load(open(f)); b = d.get("bucketing", {})
        print(f, round(d["ms_per_step"], 3), "ms", round(d["value"]), "a-s/s", d.get("dp_check"), b.get("padding_efficiency"), b.get("rank_time_ms_min"), b.get("rank_time_ms_max"))
    except Exception as e:
        print(f, "ERR", e)
PY
