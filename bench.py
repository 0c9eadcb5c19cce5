#!/usr/bin/env python
"""bench.py — encoder+CTC fwd/bwd audio-seconds/second of the S2T speech-encoder hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # our arm (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path on this box's host cores

Workload (BASELINE.json configs[1], SURVEY.md §8 "C2"): s2t_transformer_s (12 layers, d=256, ffn=2048, 4 heads,
Conv1d x2 subsampler) + CTC head V=10000 + CTC loss, end to end from raw 16 kHz synthetic waveform (Kaldi fbank
+ utterance CMVN on the GPU), batch 8 x 10 s utterances per GPU, bf16 storage / fp32 accumulate, recipe dropout
0.1, forward + backward (optimizer step excluded as in the metric; gradient all-reduce included for N >= 2).
One "step" = one pass of the hot path over one batch.  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "encoder+CTC fwd/bwd audio-sec/sec"
UNIT = "audio-s/s"
SAMPLE_RATE = 16000
VOCAB = 10000
TARGET_LEN = 60           # ~6 tokens / s
MICRO_DEFAULT = 1          # utterance groups run side by side (s2t_b200/experimental/microbatch.py)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=8, help="utterances per GPU per step")
    ap.add_argument("--seconds", type=float, default=10.0, help="utterance duration")
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of CUDA-graph replay")
    ap.add_argument("--micro", type=int, default=MICRO_DEFAULT,
                    help="utterance groups of the batch run side by side on parallel streams (1 = one chain)")
    ap.add_argument("--config", default="c2", choices=["c2", "c3", "c4", "c5"],
                    help="BASELINE.json configs[1..4] at their stated size (s2t_b200/configs.py); c2 is the headline")
    ap.add_argument("--bucketed", action="store_true",
                    help="length-bucketed workload: log-normal utterance durations, batch_by_size(max_tokens), batches "
                         "sharded round-robin over ranks, (B, T) quantised to the plan bucket set")
    ap.add_argument("--max-tokens", type=int, default=40000, help="--bucketed: frames per batch (recipe: 40000)")
    ap.add_argument("--utterances", type=int, default=512, help="--bucketed: size of the synthetic duration set")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-kernels", action="store_true", help="also print a per-kernel time table (stderr)")
    return ap.parse_args()


def workload_name(a) -> str:
    from s2t_b200.configs import CONFIGS, DESCRIPTIONS
    shape = (f"length-bucketed batches (max_tokens {a.max_tokens}, log-normal durations)" if getattr(a, "bucketed", False)
             else f"{a.batch} x {a.seconds:g} s per GPU")
    return (f"{DESCRIPTIONS[a.config]}(V={VOCAB}) from raw 16 kHz waveform, {shape}, "
            f"fbank+CMVN on device, dropout {CONFIGS[a.config]['dropout']:g}, fwd+bwd")


def workload_config(a, world: int) -> dict:
    """The ``config`` object of the fixed-shape bench line.  Built from the command line alone so that the reference arm
    prints the SAME object for the same flags (the driver compares the two arms' configs); figures that need the engine
    (parameter count, activation footprint) are top-level keys of our line instead."""
    from s2t_b200 import frontend
    n_samp = int(a.seconds * SAMPLE_RATE)
    return {"workload": workload_name(a), "global_batch_utts": world * a.batch,
            "audio_s_per_step": world * a.batch * n_samp / SAMPLE_RATE, "name": a.config,
            "frames": frontend.num_frames(n_samp), "vocab": VOCAB, "parallelism": f"dp{world}",
            "cuda_graph": not a.no_graph, "micro_batches": a.micro,
            "l2": "no flush: per-step working set (activations+logits+grads, > 400 MB) exceeds the 126 MB L2"}


def synth_batch(batch: int, seconds: float, seed: int):
    """Synthetic int16-scale waveform: randn * 3000 + a 440 Hz tone; random targets."""
    import torch
    g = torch.Generator().manual_seed(seed)
    n = int(seconds * SAMPLE_RATE)
    t = torch.arange(n, dtype=torch.float32) / SAMPLE_RATE
    wave = torch.randn((batch, n), generator=g) * 3000.0 + 1500.0 * torch.sin(2 * 3.141592653589793 * 440.0 * t)
    num = torch.full((batch,), n, dtype=torch.int32)
    targets = torch.randint(4, VOCAB, (batch, TARGET_LEN), generator=g).to(torch.int32)
    tlen = torch.full((batch,), TARGET_LEN, dtype=torch.int32)
    return wave, num, targets, tlen


# --------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md "clocks DURING the timed region")
# --------------------------------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own PyTorch path on the host cores
# --------------------------------------------------------------------------------------------------
def _reference_encoder(config: str):
    """(module, kind): the reference's OWN encoder class executed from ``baseline/_ref`` when the offline install of the
    reference left its package tree there (oracle/ref_shim.py imports the unmodified files; kind = "reference"), else
    oracle/encoder.py, the op-for-op restatement pinned bit-level against the reference by oracle/make_golden.py
    (kind = "port").  /root/reference itself does not exist on the GPU box."""
    import torch
    from oracle import encoder as OE
    from s2t_b200.configs import encoder_kwargs
    from dataclasses import fields
    kw = encoder_kwargs(config)
    okw = {k: v for k, v in kw.items() if k in {f.name for f in fields(OE.OracleConfig)}}
    ocfg = OE.OracleConfig(**okw)
    pds = kw.get("encoder_type") == "pds"
    ref_root = ROOT / "baseline" / "_ref"
    if (ref_root / "fairseq" / "models" / "speech_to_text" / "s2t_transformer.py").exists() \
            and os.environ.get("S2T_BENCH_REF_KIND", "") != "port":
        try:
            os.environ["S2T_REFERENCE"] = str(ref_root)
            from oracle import make_golden as MG
            enc = MG.build_reference_encoder(ocfg, pds)
            return enc, "reference", kw
        except Exception as e:  # the installed tree is unusable on this box: say so and fall back to the port
            sys.stderr.write(f"[bench] baseline/_ref present but not usable ({type(e).__name__}: {e}); using the port\n")
    enc = (OE.PDSS2TTransformerEncoder if pds else OE.S2TTransformerEncoder)(ocfg)
    return enc, "port", kw


def cpu_reference_run(batch: int, seconds: float, steps: int, warmup: int, config: str = "c2", budget_s: float = 25.0,
                      threads: int = 0):
    """The reference fairseq CPU path for this workload: torchaudio Kaldi fbank + UtteranceCMVN per
    utterance, then the encoder + fp32 log_softmax + nn.CTCLoss forward/backward in fp32 with recipe dropout, on
    ``threads`` host threads (0 = all cores).  Runs ``steps`` timed steps after ``warmup``, but stops once ``budget_s``
    seconds of timed work have accumulated (the cap is reported)."""
    import torch
    from oracle import encoder as OE
    from oracle import fbank as OF
    cores = threads or (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    try:
        import torchaudio.compliance.kaldi as K
        fb = lambda w: K.fbank(w.unsqueeze(0), num_mel_bins=80, sample_frequency=SAMPLE_RATE).numpy()
    except Exception:  # torchaudio missing: the numpy restatement of the same algorithm
        fb = lambda w: OF.fbank(w.numpy())
    torch.manual_seed(0)
    enc, kind, kw = _reference_encoder(config)
    enc.train()
    wave, num, targets, tlen = synth_batch(batch, seconds, 0)
    tokens = torch.cat([targets.long(), torch.full((batch, 1), 2, dtype=torch.long)], dim=1)
    dual = kw.get("tgt_vocab", 0) > 0
    times = []
    capped = False
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        feats = [torch.from_numpy(OF.utterance_cmvn(fb(wave[b, : int(num[b])]))) for b in range(batch)]
        lens = torch.tensor([f.shape[0] for f in feats])
        src = torch.zeros((batch, int(lens.max()), 80))
        for b, f in enumerate(feats):
            src[b, : f.shape[0]] = f
        enc.zero_grad(set_to_none=True)
        out = enc(src, lens)
        if dual:
            loss = OE.ctc_criterion_total(out, tokens, tokens, ctc_weight=kw["ctc_weight"], xctc_weight=kw["xctc_weight"],
                                          inter_ctc_weight=kw["inter_ctc_weight"], inter_xctc_weight=kw["inter_xctc_weight"])
        else:
            loss = OE.ctc_criterion_loss(out["ctc_logit"][0], out["encoder_padding_mask"][0], tokens)
        loss.backward()
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
            if sum(times) > budget_s and it + 1 < warmup + steps:   # bounded sample: ~10-30 s of CPU work
                capped = True
                break
    audio_s = float(num.sum()) / SAMPLE_RATE
    total = sum(times)
    return {"value": audio_s * len(times) / total, "ms_per_step": 1e3 * total / len(times), "cores": cores, "kind": kind,
            "steps_done": len(times), "capped": capped,
            "sample": f"{len(times)} step(s) of the same {batch} x {seconds:g} s batch after {warmup} warm-up, fp32, dropout "
                      f"{kw['dropout']:g}, torch.set_num_threads({cores})"
                      + (f", stopped at the {budget_s:g} s wall-clock cap" if capped else ""),
            "loss": float(loss.item())}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    budget = float(os.environ.get("S2T_REF_BUDGET_S", "120"))
    r = cpu_reference_run(a.batch, a.seconds, max(1, a.steps), max(1, a.warmup), a.config, budget_s=budget)
    r1 = cpu_reference_run(a.batch, a.seconds, 1, 1, a.config, budget_s=30.0, threads=1)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": r["steps_done"],
        "steps_requested": a.steps, "warmup": max(1, a.warmup), "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": ({"workload": workload_name(a)} if getattr(a, "bucketed", False) else workload_config(a, max(1, a.gpus))),
        "config_note": "the config of the arm this line is compared with; the reference itself ran on rank 0's host cores",
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                         "value_1thread": r1["value"], "wall_clock_cap_s": budget, "capped": r["capped"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def gemm_roofline(torch, eng, P, peaks, replays: int = 20):
    """Roofline of the dominant kernel family, gemm_bf16_kernel / gemm_bf16_splitk_kernel (tcgen05/TMEM GEMMs: every
    linear layer, attention contraction, conv-as-GEMM and the vocabulary projection).  Timed live on the device: the
    arguments of every GEMM launch of one training step are recorded, then exactly those launches (same operands, same
    shapes, same epilogues) are captured back to back into one CUDA graph on one stream and replayed ``replays`` times
    between two CUDA events on that stream.  achieved = algorithmic FLOPs (2*M*N*K*batches per launch, summed over the
    step's launches) / (event time / replays); avg launch duration = event time / (replays * launches).  The operands
    of one step (> 126 MB of activations, weights and gradients) exceed L2, as in the real step."""
    from s2t_b200 import ops
    calls = []
    orig = ops.gemm_raw

    main_stream = torch.cuda.current_stream()

    def rec(**kw):
        kw = dict(kw)
        if not kw.get("max_ctas"):
            kw["max_ctas"] = ops._default_max_ctas      # the SM share in force at this point of the step
        calls.append((kw, torch.cuda.current_stream() != main_stream))
        orig(**kw)

    ops.gemm_raw = rec
    dp, hook = eng.dp, eng.grad_ready_hook
    eng.dp, eng.grad_ready_hook = None, None     # rank 0 alone runs this step: no collective may be issued from it
    try:
        e_all0, e_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e_all0.record()
        eng.train_step(P)
        e_all1.record()
    finally:
        ops.gemm_raw = orig
        eng.dp, eng.grad_ready_hook = dp, hook
    torch.cuda.synchronize()
    step_ms = e_all0.elapsed_time(e_all1)
    fl = sum(2.0 * kw["M"] * kw["N"] * kw["K"] * kw.get("batch", (1, 1))[0] * kw.get("batch", (1, 1))[1] for kw, _ in calls)
    n_side = sum(1 for _, s_ in calls if s_)
    side = torch.cuda.Stream()
    side2 = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()

    def issue():
        # chain launches on one stream, the launches the step puts on side streams (weight gradients, dV/dK) on a second
        # one beside them — the configuration their SM shares (100 + 48 CTAs) are sized for
        side2.wait_stream(side)
        for kw, on_side in calls:
            if on_side:
                with torch.cuda.stream(side2):
                    orig(**kw)
            else:
                orig(**kw)
        side.wait_stream(side2)

    with torch.cuda.stream(side):
        issue()
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=side):
            issue()
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(replays):
            g.replay()
        e1.record(side)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / replays
    n = len(calls)
    peak = float(peaks.get("bf16_tflops_sustained") or 1400.0)
    ach = fl / (ms * 1e-3) / 1e12
    # algorithmic operand/result bytes per launch (every operand read once, the result written once) and the DRAM bytes
    # ncu measured for the same launches (profiles/gemm_traffic.json, committed with the launch list it was taken from)
    esz = lambda t: t.element_size() if t is not None else 2
    alg = 0.0
    for kw, _ in calls:
        nb = kw.get("batch", (1, 1))
        z = nb[0] * nb[1]
        alg += z * (kw["M"] * kw["K"] * 2 + kw["N"] * kw["K"] * 2 + kw["M"] * kw["N"] * esz(kw.get("out")))
    traffic = None
    tf = ROOT / "profiles" / "gemm_traffic.json"
    if tf.exists():
        try:
            traffic = float(json.loads(tf.read_text())["dram_bytes_per_launch"])
        except Exception:
            traffic = None
    return {"bound": "tensor", "kernel": "gemm_bf16_kernel + gemm_bf16_splitk_kernel (all GEMM launches of one step)",
            "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
            "traffic_unit": "DRAM bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum, cold cache)",
            "algorithmic_bytes_per_launch": alg / n,
            "launches_per_step": n, "avg_launch_us": 1e3 * ms / n, "gemm_ms_per_step": ms, "eager_step_ms": step_ms,
            "flop_per_step": fl,
            "side_stream_launches": n_side,
            "how": f"the step's {n} GEMM launches captured alone into one CUDA graph ({n - n_side} chain launches on one "
                   f"stream, the {n_side} side-stream launches of the step on a second stream beside them, same SM shares "
                   f"as in the step), {replays} replays between CUDA events",
            "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback 1.4 PFLOP/s sustained"}


def run_b200(a):
    import torch
    import torch.distributed as dist
    from s2t_b200 import ops
    from s2t_b200.engine import EncoderConfig, S2TEncoderEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import faulthandler
    # a wedged collective must not hang the box: dump every thread's stack and exit instead
    faulthandler.dump_traceback_later(float(os.environ.get("S2T_BENCH_WATCHDOG_S", "600")), exit=True)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (impl b200) needs a CUDA device: s2t_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime
        # a mismatched collective must fail within minutes, not hold an 8-GPU box for the default 10-minute NCCL timeout
        dist.init_process_group("nccl", device_id=torch.device("cuda", local),
                                timeout=datetime.timedelta(seconds=int(os.environ.get("S2T_NCCL_TIMEOUT_S", "180"))))
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())

    from s2t_b200.configs import encoder_kwargs
    cfg = EncoderConfig(**encoder_kwargs(a.config))            # BASELINE.json configs[1..4] at their stated size
    dual = cfg.tgt_vocab > 0
    if a.bucketed:
        return run_bucketed(a, torch, dist, cfg, world, rank, local, peaks)
    if a.micro > 1:
        from s2t_b200.experimental.microbatch import MicroBatchEngine
        eng = MicroBatchEngine(cfg, n_micro=a.micro)
    else:
        eng = S2TEncoderEngine(cfg)
    if world > 1:
        dist.broadcast(eng.store.flat, src=0)                  # same weights on every rank (trainer.py:431)
        eng.enable_data_parallel(n_buckets=int(os.environ.get("S2T_DP_BUCKETS", "4")))
    n_samp = int(a.seconds * SAMPLE_RATE)
    from s2t_b200 import frontend
    T = frontend.num_frames(n_samp)
    P = eng.plan(a.batch, T, TARGET_LEN, wave_len=n_samp)
    # a few different pinned host batches, rotated through the e2e loop
    host = []
    for i in range(4):
        w, n, tg, tl = synth_batch(a.batch, a.seconds, 1000 * rank + i)
        host.append(tuple(t.pin_memory() for t in (w, n, tg, tl)))
    h2d = sum(t.numel() * t.element_size() for t in host[0])
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    audio_s_step = a.batch * n_samp / SAMPLE_RATE               # per rank

    def load(i):
        w, n, tg, tl = host[i % len(host)]
        eng.set_inputs(P, wave=w, num_samples=n, targets=tg, target_lengths=tl,
                       xtargets=tg if dual else None, xtarget_lengths=tl if dual else None)

    dp_check = data_parallel_check(torch, dist, eng, P, host[0], dual, world, rank) if world > 1 else None
    load(0)
    l0 = ops.launch_count
    eng.train_step(P)                                           # eager warm-up (also counts launches per step)
    torch.cuda.synchronize()
    launches_per_step = ops.launch_count - l0
    if not a.no_graph:
        eng.capture(P)
    kind = "train"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident throughput ("value"): inputs already in HBM ---------------------------------------
    for _ in range(max(a.warmup, 3)):
        eng.run(P, kind)
    barrier()
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(a.steps):
        eng.run(P, kind)
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clk = clocks.stop() if rank == 0 else None
    loss_val = float(P.loss.item())

    # ---- end to end through the public call with HOST buffers: H2D of the batch + step + D2H of the loss --------
    for i in range(3):
        load(i)
        eng.run(P, kind)
        loss_host.copy_(P.loss, non_blocking=True)
        torch.cuda.synchronize()
    barrier()
    e0.record()
    for i in range(a.steps):
        load(i)
        eng.run(P, kind)
        loss_host.copy_(P.loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()               # the caller reads the loss every step
    e1.record()
    barrier()
    ms_e2e_sync = max_over_ranks(e0.elapsed_time(e1))

    # the same loop as a pipelined input feed (s2t_b200/feeder.py): the H2D copy of batch i+1 overlaps step i and the
    # loss of step i is read while step i+1 runs — every step still copies its batch from pinned host memory and
    # returns its loss to the host inside the timed region
    from s2t_b200.feeder import InputFeeder
    feeder = InputFeeder(eng, P)
    keys = ("wave", "num_samples", "targets", "target_lengths")

    loss_slots = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event(), torch.cuda.Event()]
    losses = []

    def stage(i):
        b = dict(zip(keys, host[i % len(host)]))
        if dual:     # the XCTC head's target text (synthetic: the same token set)
            b.update(xtargets=b["targets"], xtarget_lengths=b["target_lengths"])
        feeder.stage(**b)

    def pipelined(n):
        stage(0)
        for i in range(n):
            feeder.commit()
            if i + 1 < n:
                stage(i + 1)
            eng.run(P, kind)
            loss_slots[i & 1].copy_(P.loss, non_blocking=True)
            loss_ev[i & 1].record()
            if i > 0:
                loss_ev[(i - 1) & 1].synchronize()
                losses.append(float(loss_slots[(i - 1) & 1][0]))
        loss_ev[(n - 1) & 1].synchronize()
        losses.append(float(loss_slots[(n - 1) & 1][0]))

    if feeder is not None:
        pipelined(3)
        barrier()
        e0.record()
        pipelined(a.steps)
        e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    else:
        ms_e2e = ms_e2e_sync

    roof = gemm_roofline(torch, eng, P, peaks) if rank == 0 else None
    if a.profile_kernels and rank == 0:
        kernel_table(torch, eng, P)
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        r = cpu_reference_run(a.batch, a.seconds, 3, 1, a.config, budget_s=20.0)
        cpu = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
    if world > 1:
        dist.barrier()
    if rank == 0:
        act_mb = sum(t.numel() * t.element_size() for t in vars(P).values() if isinstance(t, torch.Tensor)) / 2 ** 20
        line = {
            "metric": METRIC, "value": world * audio_s_step * a.steps / (ms_total * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_total / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(a, world), "params": eng.num_params(),
            "e2e": {"value": world * audio_s_step * a.steps / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e / a.steps,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "how": "pinned host batch -> InputFeeder (H2D of batch i+1 overlaps step i) -> CUDA-graph step -> loss "
                           "copied to pinned host memory every step, read one step later",
                    "sync_ms_per_step": ms_e2e_sync / a.steps,
                    "sync_value": world * audio_s_step * a.steps / (ms_e2e_sync * 1e-3),
                    "sync_how": "no overlap: copy, step, loss read and stream synchronize in sequence every step"},
            "gpu_launches": launches_per_step * a.steps * 3,     # device-resident loop + the two e2e loops, per rank
            "launches_per_step": launches_per_step, "loss": loss_val, "clocks": clk, "roofline": roof,
            "dp_check": dp_check,
            "cpu_baseline": cpu, "plan_tensor_mb_top_level": round(act_mb, 1),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        teardown(torch, dist, eng)


def teardown(torch, dist, eng) -> None:
    """Leave a multi-rank run the way a host program has to be able to: the captured step graphs hold NCCL kernels of the
    process group's communicator, and destroying the communicator under them can block for ever — so drain, meet at a
    barrier, destroy the GRAPHS first (drop every reference, collect), then call destroy_process_group.  A 30 s timer is
    the safety net (os._exit if the clean path has not finished by then; an 8-GPU box must not be held by a teardown);
    stderr says which path was taken."""
    import gc
    import threading
    import time
    torch.cuda.synchronize()
    dist.barrier()
    for plan in list(eng.plans.values()):
        plan.graph = None
    eng.plans.clear()
    gc.collect()
    torch.cuda.synchronize()
    sys.stdout.flush()
    sys.stderr.flush()
    if os.environ.get("S2T_CLEAN_EXIT", "1") != "1":
        os._exit(0)

    def bail():
        sys.stderr.write("teardown: clean exit did not finish within 30 s, leaving through os._exit(0)\n")
        sys.stderr.flush()
        os._exit(0)

    timer = threading.Timer(30.0, bail)
    timer.daemon = True
    timer.start()
    t0 = time.time()
    dist.destroy_process_group()
    sys.stderr.write(f"teardown: graphs dropped, destroy_process_group returned after {time.time() - t0:.2f} s\n")
    sys.stderr.flush()


def synth_durations(n: int, seed: int = 0):
    """SURVEY.md §8d duration set: log-normal, median 5.5 s, sigma 0.6, clipped to 1-30 s (MuST-C-like); samples."""
    import math
    import numpy as np
    rng = np.random.RandomState(seed)
    secs = np.clip(rng.lognormal(mean=math.log(5.5), sigma=0.6, size=n), 1.0, 30.0)
    return (secs * SAMPLE_RATE).astype(np.int64)


def run_bucketed(a, torch, dist, cfg, world, rank, local, peaks):
    """Length-bucketed data-parallel workload (BASELINE.json configs[2]): utterances ordered by length,
    packed by batch_by_size(max_tokens, multiple of 8), batch i -> rank i mod world, short ranks run a zero-weight dummy
    batch; (B, T) quantised to the plan bucket set, one captured graph per bucket in an LRU plan cache.  A step = one
    batch per rank; K steps cycle through the rank's schedule.  value = real audio seconds of all ranks / max time."""
    import numpy as np
    from s2t_b200 import frontend, ops
    from s2t_b200.batching import BucketSet, rank_schedule
    from s2t_b200.engine import S2TEncoderEngine
    dual = cfg.tgt_vocab > 0
    eng = S2TEncoderEngine(cfg)
    if world > 1:
        dist.broadcast(eng.store.flat, src=0)
        eng.enable_data_parallel(n_buckets=int(os.environ.get("S2T_DP_BUCKETS", "4")))
        eng.collective_free_warmup = True       # ranks capture new plans at different steps (engine.capture)
    samples = synth_durations(a.utterances, 0)
    n_frames = np.array([frontend.num_frames(int(n)) for n in samples], dtype=np.int64)
    bs = BucketSet(max_frames=cfg.max_source_positions)
    sched = rank_schedule(n_frames, a.max_tokens, world, rank, bsz_mult=8, buckets=bs, pad_batch=True)
    eng.plan_cache_bytes = int(float(os.environ.get("S2T_PLAN_CACHE_GB", "120")) * 2 ** 30)     # one plan per frame bucket fits
    n_batches_total = len(sched) * world
    g = torch.Generator().manual_seed(1234 + rank)
    noise = (torch.randn(31 * SAMPLE_RATE, generator=g) * 3000.0)
    steps_host = []
    for rb in sched:
        B, Tq = rb.key
        wl = 400 + 160 * Tq                                 # any sample count that yields <= Tq frames fits
        ns = torch.zeros(B, dtype=torch.int32)              # rows beyond the batch: empty utterances (0 samples, 0 targets)
        ns[: len(rb.indices)] = torch.tensor([int(samples[i]) for i in rb.indices], dtype=torch.int32)
        w = torch.zeros((B, int(ns.max())), dtype=torch.float32)
        for j, n in enumerate(ns.tolist()):
            if n > 0:
                off = int(torch.randint(0, noise.numel() - n, (1,), generator=g))
                w[j, :n] = noise[off:off + n]
        tl = torch.where(ns > 0, torch.clamp((ns.float() / SAMPLE_RATE * 6).to(torch.int32), min=1), torch.zeros_like(ns))
        smax = (int(tl.max()) + 31) // 32 * 32
        tg = torch.randint(4, VOCAB, (B, smax), generator=g).to(torch.int32)
        steps_host.append(dict(key=(B, Tq, smax, wl), dummy=rb.dummy, audio_s=0.0 if rb.dummy else float(ns.sum()) / SAMPLE_RATE,
                               frames=int(n_frames[rb.indices].sum()), padded=B * Tq, real_b=len(rb.indices),
                               t=tuple(x.pin_memory() for x in (w, ns, tg, tl))))

    def run_step(i):
        h = steps_host[i % len(steps_host)]
        P = eng.plan(*h["key"])
        w, ns, tg, tl = h["t"]
        eng.set_inputs(P, wave=w, num_samples=ns, targets=tg, target_lengths=tl, xtargets=tg if dual else None,
                       xtarget_lengths=tl if dual else None, weight=0.0 if h["dummy"] else 1.0)
        if P.graph is None and not a.no_graph:
            c0 = ops.launch_count
            eng.capture(P)                                     # one eager warm-up + one captured issue of the step
            P.launches = (ops.launch_count - c0) // 2
        c0 = ops.launch_count
        eng.run(P, "train")
        if a.no_graph:
            P.launches = ops.launch_count - c0
        return h, P

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    l0 = ops.launch_count
    for i in range(max(len(steps_host), a.warmup, 3)):        # warm-up: one pass over the schedule builds / captures the plans
        run_step(i)
    torch.cuda.synchronize()
    barrier()
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    audio = frames = padded = h2d = dummies = 0.0
    launches = 0
    barrier()
    e0.record()
    for i in range(a.steps):
        h, P = run_step(i)                                     # H2D of the batch + step: the public call with HOST buffers
        loss_host.copy_(P.loss, non_blocking=True)
        audio += h["audio_s"]; frames += 0 if h["dummy"] else h["frames"]; padded += h["padded"]
        dummies += 1.0 if h["dummy"] else 0.0
        h2d += sum(t.numel() * t.element_size() for t in h["t"])
        launches += getattr(P, "launches", 0)
    e1.record()
    barrier()
    ms_local = e0.elapsed_time(e1)
    stats = torch.tensor([ms_local, -ms_local, audio, frames, padded, dummies], device="cuda", dtype=torch.float64)
    if world > 1:
        mx = stats.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    else:
        mx = sm = stats
    ms_total, ms_min = float(mx[0]), -float(mx[1])
    clk = clocks.stop() if rank == 0 else None
    if rank == 0:
        val = float(sm[2]) / (ms_total * 1e-3)
        line = {
            "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(len(steps_host), a.warmup, 3),
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(a), "name": a.config, "utterances": a.utterances, "max_tokens": a.max_tokens,
                       "batches_per_epoch": n_batches_total, "steps_per_rank_per_epoch": len(steps_host),
                       "parallelism": f"dp{world}", "cuda_graph": not a.no_graph,
                       "l2": "no flush: per-step working set exceeds the 126 MB L2"},
            "e2e": {"value": val, "unit": UNIT, "ms_per_step": ms_total / a.steps, "h2d_bytes_per_step": h2d / max(a.steps, 1),
                    "d2h_bytes_per_step": 4, "how": "every step copies its pinned host batch H2D (set_inputs) and its loss D2H "
                    "inside the timed region; value and e2e are the same measurement in this mode"},
            "bucketing": {"real_frames": float(sm[3]), "padded_frames": float(sm[4]),
                          "padding_efficiency": float(sm[3]) / max(float(sm[4]), 1.0),
                          "rank_time_ms_min": ms_min, "rank_time_ms_max": ms_total,
                          "dummy_batches_in_timed_steps": int(sm[5]),
                          "rank0_plan_cache": eng.plan_cache_stats(), "bucket_frames": bs.all_frames()[:8] + ["..."]},
            "gpu_launches": launches, "clocks": clk, "roofline": None, "cpu_baseline": None,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        teardown(torch, dist, eng)


def data_parallel_check(torch, dist, eng, P, host0, dual, world, rank):
    """DP correctness on the hardware (reference semantics: divide by world, then sum,
    legacy_distributed_data_parallel.py:76-160).  One dropout-free step: every rank holds the same weights, rank r
    holds batch r.  (1) the all-reduced flat gradient must be BIT-IDENTICAL on every rank (checksum); (2) rank 0's
    result must equal the mean of the per-batch gradients computed one by one on rank 0 alone (rel-L2 <= 1e-2: bf16
    gradients, different summation order).  Aborts the run on a mismatch."""
    dev = eng.device
    was_training = eng.training
    eng.train(False)       # dropout off (masks are per-rank streams by design; the check needs a pure function of the batch)
    # every rank builds ALL ranks' batches (seeded), rank r runs batch r through the data-parallel step
    batches = [synth_batch(P.B, host0[0].shape[1] / SAMPLE_RATE, 7000 + r) for r in range(world)]

    def put(b):
        w, n, tg, tl = b
        eng.set_inputs(P, wave=w, num_samples=n, targets=tg, target_lengths=tl,
                       xtargets=tg if dual else None, xtarget_lengths=tl if dual else None)

    put(batches[rank])
    eng.train_step(P)
    torch.cuda.synchronize()
    g_dp = eng.store.grad.clone()
    # (1) bit-identical across ranks: compare an order-independent integer checksum of the raw bf16 words
    words = g_dp.view(torch.int16).to(torch.int64)
    ck = torch.stack([words.sum(), (words * (torch.arange(words.numel(), device=dev) % 8191 + 1)).sum()])
    lo, hi = ck.clone(), ck.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    identical = bool((lo == hi).all().item())
    # (2) rank 0: mean of the per-batch gradients, one batch at a time, without the collective
    rel = None
    dp, hook = eng.dp, eng.grad_ready_hook
    eng.dp, eng.grad_ready_hook = None, None
    try:
        if rank == 0:
            acc = torch.zeros_like(g_dp, dtype=torch.float32)
            for b in batches:
                put(b)
                eng.train_step(P)
                acc += eng.store.grad.float()
            acc /= world
            torch.cuda.synchronize()
            rel = float((g_dp.float() - acc).norm() / (acc.norm() + 1e-30))
    finally:
        eng.dp, eng.grad_ready_hook = dp, hook
        eng.train(was_training)
    flag = torch.tensor([1 if (identical and (rel is None or rel <= 1e-2)) else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    out = {"world": world, "bit_identical_across_ranks": identical, "rel_l2_vs_mean_of_per_batch_grads": rel,
           "tolerance": 1e-2, "ok": bool(flag.item())}
    if not out["ok"]:
        if rank == 0:
            print(json.dumps({"dp_check": out, "error": "data-parallel gradient check failed"}), flush=True)
        raise SystemExit(3)
    return out


def kernel_table(torch, eng, P):
    """Per-entry-point device time of one eager step (CUDA events around every C-ABI call) -> stderr."""
    from s2t_b200 import ops
    recs = []
    orig = ops._call

    def timed(name, *args):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        orig(name, *args)
        e1.record()
        recs.append((name, e0, e1))

    ops._call = timed
    dp, hook = eng.dp, eng.grad_ready_hook
    eng.dp, eng.grad_ready_hook = None, None     # rank-0-only step: no collectives
    try:
        eng.train_step(P)
    finally:
        ops._call = orig
        eng.dp, eng.grad_ready_hook = dp, hook
    torch.cuda.synchronize()
    agg = {}
    for name, a, b in recs:
        t, n = agg.get(name, (0.0, 0))
        agg[name] = (t + a.elapsed_time(b), n + 1)
    tot = sum(t for t, _ in agg.values())
    for name, (t, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"{name:28s} calls {n:4d}  total {t:8.3f} ms  avg {1e3 * t / n:8.1f} us  {100 * t / tot:5.1f} %", file=sys.stderr)
    print(f"{'sum':28s} calls {len(recs):4d}  total {tot:8.3f} ms", file=sys.stderr)


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
