"""Kernel timeline of one CUDA-graph replay of the bench step (torch.profiler / CUPTI): per-kernel start, duration and
stream, the union busy time and the idle gaps on the device.  usage: python tools/timeline.py [out.csv]"""
import sys
import torch
sys.path.insert(0, ".")
import bench
from s2t_b200 import frontend
from s2t_b200.engine import EncoderConfig, S2TEncoderEngine

out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/timeline.csv"
torch.cuda.set_device(0)
eng = S2TEncoderEngine(EncoderConfig(vocab=bench.VOCAB, dropout=0.1))
n = 160000
P = eng.plan(8, frontend.num_frames(n), bench.TARGET_LEN, wave_len=n)
w, ns, tg, tl = bench.synth_batch(8, 10.0, 0)
eng.set_inputs(P, wave=w, num_samples=ns, targets=tg, target_lengths=tl)
eng.train_step(P); torch.cuda.synchronize()
eng.capture(P)
for _ in range(5):
    eng.run(P)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        eng.run(P)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()]
evs.sort(key=lambda e: e.time_range.start)
def _sid(e):
    try:
        return e.device_resource_id if hasattr(e, "device_resource_id") else -1
    except Exception:
        return -1
# split replays at counter_add
marks = [i for i, e in enumerate(evs) if "counter_add" in e.name]
step = evs[marks[1]:marks[2]] if len(marks) >= 3 else evs
t0 = step[0].time_range.start
sids = {}
for e in step:
    sids.setdefault(_sid(e), len(sids))
rows = [((e.time_range.start - t0), (e.time_range.end - e.time_range.start), e.name.split("(")[0].replace("void ", "").replace(", ", ";")[:60], sids[_sid(e)]) for e in step]
with open(out, "w") as f:
    f.write("start_us,dur_us,name,stream\n")
    for r in rows:
        f.write(f"{r[0]:.2f},{r[1]:.2f},{r[2]},{r[3]}\n")
end = max(r[0] + r[1] for r in rows)
busy, cur_s, cur_e = 0.0, None, None
for s_, d, _, _ in rows:
    e_ = s_ + d
    if cur_e is None or s_ > cur_e:
        if cur_e is not None:
            busy += cur_e - cur_s
        cur_s, cur_e = s_, e_
    else:
        cur_e = max(cur_e, e_)
busy += cur_e - cur_s
print(f"kernels {len(rows)}  span {end:.1f} us  device busy (union) {busy:.1f} us  idle {end - busy:.1f} us  sum of durations {sum(r[1] for r in rows):.1f} us")
import collections
agg = collections.defaultdict(lambda: [0, 0.0])
for _, d, nme, _ in rows:
    agg[nme][0] += 1; agg[nme][1] += d
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"{k:62s} {c:4d} {t:9.1f} us  avg {t / c:6.1f}")
