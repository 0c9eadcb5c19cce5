"""Kernel timeline of one data-parallel step on rank 0 (torchrun, world >= 2): where the bucketed NCCL all-reduces run
relative to the backward kernels.  usage: torchrun --nproc-per-node 2 tools/timeline_dp.py [out.csv]"""
import os
import sys
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
import bench
from s2t_b200 import frontend
from s2t_b200.engine import EncoderConfig, S2TEncoderEngine

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/timeline_dp.csv"
eng = S2TEncoderEngine(EncoderConfig(vocab=bench.VOCAB, dropout=0.1))
dist.broadcast(eng.store.flat, src=0)
eng.enable_data_parallel(n_buckets=int(os.environ.get("S2T_DP_BUCKETS", "4")))
n = 160000
P = eng.plan(8, frontend.num_frames(n), bench.TARGET_LEN, wave_len=n)
w, ns, tg, tl = bench.synth_batch(8, 10.0, rank)
eng.set_inputs(P, wave=w, num_samples=ns, targets=tg, target_lengths=tl)
eng.train_step(P)
torch.cuda.synchronize()
eng.capture(P)
for _ in range(5):
    eng.run(P)
torch.cuda.synchronize()
dist.barrier()
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        eng.run(P)
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()]
    evs.sort(key=lambda e: e.time_range.start)
    marks = [i for i, e in enumerate(evs) if "counter_add" in e.name]
    step = evs[marks[1]:marks[2]] if len(marks) >= 3 else evs
    t0 = step[0].time_range.start
    with open(out, "w") as f:
        f.write("start_us,dur_us,name\n")
        for e in step:
            nm = e.name.split("(")[0].replace("void ", "").replace(", ", ";")[:60]
            f.write(f"{e.time_range.start - t0:.2f},{e.time_range.end - e.time_range.start:.2f},{nm}\n")
    end = max(e.time_range.end for e in step) - t0
    print(f"kernels {len(step)} span {end:.1f} us")
    for e in step:
        if "nccl" in e.name.lower():
            print(f"nccl  start {e.time_range.start - t0:8.1f}  dur {e.time_range.end - e.time_range.start:7.1f}  {e.name[:50]}")
torch.cuda.synchronize()
dist.barrier()
for plan in list(eng.plans.values()):
    plan.graph = None
os._exit(0)
