#!/bin/bash
# 8-GPU pass: C2 weak scaling at N=8 and N=4, the bucketed C3 workload at N=8/4/2, DP kernel timeline of rank 0.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4; do
  timeout 600 $TR --nproc-per-node $n --master-port 2951$n bench.py --gpus $n --steps 30 --warmup 5 2> gpurun_out/bench_c2_n$n.err | grep '^{' > gpurun_out/bench_c2_n$n.json; echo "c2 n$n rc=$?"
done
for n in 8 4 2; do
  timeout 900 $TR --nproc-per-node $n --master-port 2952$n bench.py --gpus $n --config c3 --bucketed --utterances 8192 --steps 60 --warmup 3 2> gpurun_out/bench_c3_bucketed_n$n.err | grep '^{' > gpurun_out/bench_c3_bucketed_n$n.json; echo "c3 bucketed n$n rc=$?"; tail -2 gpurun_out/bench_c3_bucketed_n$n.err
done
timeout 300 $TR --nproc-per-node 8 --master-port 29531 tools/timeline_dp.py gpurun_out/timeline_dp_n8.csv > gpurun_out/timeline_dp_n8.txt 2>&1; echo "timeline rc=$?"
python - <<PY
import json
for f in ["bench_c2_n8","bench_c2_n4","bench_c3_bucketed_n8","bench_c3_bucketed_n4","bench_c3_bucketed_n2"]:
    try:
        d=json.load(open(f"gpurun_out/{f}.json"))
        print(f, round(d["ms_per_step"],3), "ms", round(d["value"]), "a-s/s", d.get("dp_check"), d.get("bucketing",{}).get("padding_efficiency"), d.get("bucketing",{}).get("rank_time_ms_min"), d.get("bucketing",{}).get("rank_time_ms_max"))
    except Exception as e:
        print(f, "ERR", e)
PY
