#!/usr/bin/env python3
"""Benchmark of the hot path (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload train|stft|istft] [--impl reference]

Default workload `train` = BASELINE.json configs[2]: uPIT 2-speaker BLSTM training, WSJ0-2mix shape (8 kHz, 4-8 s
utterances, STFT 512/128 -> F=257), batch 64 per GPU, one full train step per "step" (forward, uPIT loss, backward,
gradient allreduce when N>1, global-norm clip 0.25, Adam) -- metric "uPIT train audio-sec/sec".  The same JSON line
carries the second half of BASELINE.json's metric, "STFT frames/sec vs HBM roofline" (configs[1], 256/64), under
"stft", and the masked iSTFT under "istft".

Everything timed on the GPU is a libssb200 kernel (hand-written sm_100a CUDA, C ABI in include/ssb200.h); torch
allocates memory, supplies the stream and NCCL.  oracle/ is executed only by the cpu_baseline leg and by
`--impl reference`.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "speech-separation_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

SR, N_FFT, HOP, S, B_PER_GPU = 8000, 512, 128, 2, 64
F = N_FFT // 2 + 1
H = 600
N_BATCHES = 4            # distinct synthetic batches rotated through the timed steps


def peaks():
    """(hbm GB/s, bf16 TF/s burst, bf16 TF/s sustained, source)"""
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), \
            "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback (B200_PROFILING.md)"


def train_lengths(seed, B):
    """cfg3 utterance lengths in samples: U{32000..64000} at 8 kHz, sorted descending."""
    rng = np.random.default_rng(seed)
    return sorted((int(v) for v in rng.integers(32000, 64001, size=B)), reverse=True)


def synth_features(seed, lens_samples, num_spk=S):
    """Magnitude-like synthetic features (T_b, F) per utterance: non-negative, sources sum >= mixture like |STFT|.
    (Feature VALUES do not change the work done; tests/ use real STFT features for parity.)"""
    rng = np.random.default_rng(seed)
    samples = []
    for n in lens_samples:
        T = 1 + n // HOP
        srcs = [rng.random((T, F), dtype=np.float32) * 0.5 for _ in range(num_spk)]
        d = {"mix": np.sum(srcs, axis=0, dtype=np.float32) * rng.random((T, F), dtype=np.float32)}
        for i, s in enumerate(srcs):
            d["source%d" % (i + 1)] = s
        samples.append(d)
    return samples


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.thr.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active")
                                                          for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# kernels launched per C-ABI call (memsets not counted); default 1
KERNELS_PER_CALL = {"colsum_sb16": 2, "bn_stats": 2, "bn_bwd": 3, "grad_norm": 2,
                    # host-side knobs / stream ops of the C ABI: no kernel
                    "stream_wait_value32": 0}


class CallMeter:
    """Wraps ssb200._lib.call: counts kernel launches and (when `timing`) brackets every call with CUDA events on
    the launching stream."""

    def __init__(self):
        from ssb200 import _lib
        self._lib = _lib
        self.orig = _lib.call
        self.launches = 0
        self.timing = False
        self.records = {}
        import torch
        self.main_stream = torch.cuda.current_stream().cuda_stream
        mods = [m for n, m in sys.modules.items() if n.startswith("ssb200") and hasattr(m, "call")]
        for m in mods:
            if getattr(m, "call") is self.orig:
                m.call = self
        self.mods = mods

    def __call__(self, name, *a):
        import torch
        self.launches += KERNELS_PER_CALL.get(name, 1)
        if not self.timing or KERNELS_PER_CALL.get(name, 1) == 0:
            return self.orig(name, *a)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = self.orig(name, *a)
        e1.record()
        main = torch.cuda.current_stream().cuda_stream == self.main_stream
        self.records.setdefault(name, []).append((e0, e1, main))
        return r

    def summary(self, steps):
        """per C-ABI entry: time per step over all streams, and the part launched on the main stream (the critical path;
        the rest runs on the side stream under a recurrence)"""
        out = {}
        for k, v in self.records.items():
            ms = [a.elapsed_time(b) for a, b, _ in v]
            ms_main = [a.elapsed_time(b) for a, b, m in v if m]
            out[k] = {"calls_per_step": len(ms) / steps, "ms_per_step": sum(ms) / steps, "ms_per_call": sum(ms) / len(ms),
                      "ms_main_per_step": sum(ms_main) / steps}
        return out


def dist_setup(n_gpus):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, local, world


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernels, written by tools/ncu_traffic.py
    from `ncu --set full` captures of the same workloads (profiles/ncu_traffic.json, committed; the .ncu-rep files are
    not).  {kernel: {"bytes": ..., "source": ...}}"""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        return {}


def traffic_of(kernel):
    t = ncu_traffic().get(kernel)
    return (t["bytes"], t.get("source")) if t else (None, None)


def timed_region(fn, steps, warmup, world, sampler=None):
    """W untimed + K timed calls of fn(i), barrier + synchronize on both sides, CUDA events, max over ranks (ms)."""
    import torch
    import torch.distributed as dist
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(warmup + i)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    return ms, clocks


def train_flops_per_frame(in_dim=F, out_dim=F * S, passes=1):
    """Algorithmic flops per frame per sequence of one train step, split into the GEMM part (input projections,
    head, and their backward GEMMs) and the recurrence part (h W_hh^T and its backward)  (DESIGN.md)."""
    inproj = 2 * 4 * H * in_dim * 2 + 2 * 4 * H * 2 * H * 2          # both directions, layers 0 and 1
    head = 2 * 2 * H * out_dim
    rec = 2 * 4 * H * H * 2 * 2                                     # both directions, two layers
    gemm = (inproj + head) + (inproj + head) + (2 * 4 * H * 2 * H * 2 + head) + rec   # fwd, dW, dX(l1)+dY, dW_hh
    return passes * gemm, passes * 2 * rec


def pad_batch(samples, keys):
    """list of per-utterance dicts (T_b,F) sorted by descending length -> {key: (T,B,F) tensor}, lens (CPU arm)."""
    import torch
    lens = [s["mix"].shape[0] for s in samples]
    T, B = max(lens), len(samples)
    out = {}
    for k in keys:
        x = torch.zeros(T, B, samples[0][k].shape[1])
        for b, s in enumerate(samples):
            x[:lens[b], b] = torch.from_numpy(s[k])
        out[k] = x
    return out, lens


def cpu_train_sample(n_utts, steps, warmup, seed=0):
    """Times oracle/cpu_step.py (the reference's torch CPU calls) on the first n_utts utterances of batch `seed`.
    Returns (audio-sec/sec, sec/step, cores)."""
    import torch
    from oracle import cpu_step
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    lens_s = train_lengths(seed, B_PER_GPU)[:n_utts]
    samples = synth_features(seed, lens_s)
    keys = ["mix"] + ["source%d" % (i + 1) for i in range(S)]
    pb, lens = pad_batch(samples, keys)
    torch.manual_seed(0)
    tr = cpu_step.CpuTrainer(F, S)
    times = []
    for i in range(warmup + steps):
        h0, c0 = torch.randn(4, n_utts, H), torch.randn(4, n_utts, H)
        t0 = time.perf_counter()
        tr.step(pb["mix"], [pb[k] for k in keys[1:]], lens, h0, c0)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return sum(lens_s) / SR / sec, sec, cores


# The train-step workloads: BASELINE.json configs[2] (the headline), configs[3] (3-speaker: RSH and 6-permutation uPIT),
# configs[4] (16 kHz, 20 s).  lens(seed, B) -> utterance lengths in samples, descending.
WORKLOADS = {
    "train": dict(arch="uPIT", S=2, B=B_PER_GPU, sr=8000, lens=lambda seed, B: train_lengths(seed, B),
                  name="configs[2]: uPIT 2-spk BLSTM(2x600 bidir)+BN+Linear train step, WSJ0-2mix shape, 8 kHz, "
                       "STFT 512/128 (F=257), utterances U[4,8] s, batch 64/GPU"),
    "train_b100": dict(arch="uPIT", S=2, B=100, sr=8000, lens=lambda seed, B: train_lengths(seed, B),
                       name="configs[2] shapes at the reference recipe's default batch 100 (run_train.sh:18): B > 64 takes "
                            "the generic backward recurrence"),
    "upit3": dict(arch="uPIT", S=3, B=B_PER_GPU, sr=8000,
                  lens=lambda seed, B: sorted((int(v) for v in np.random.default_rng(seed).integers(24000, 32001, size=B)), reverse=True),
                  name="configs[3]: uPIT 3-spk (6 permutations, 771 outputs) train step, 8 kHz, 3-4 s, 512/128, batch 64/GPU"),
    "rsh": dict(arch="RSH", S=3, B=B_PER_GPU, sr=8000,
                lens=lambda seed, B: sorted((int(v) for v in np.random.default_rng(seed).integers(24000, 32001, size=B)), reverse=True),
                name="configs[3]: RSH recurrent-selective-hearing train step, 3-spk mixtures (3 passes, greedy assignment), "
                     "8 kHz, 3-4 s, 512/128 (input 2F=514), batch 64/GPU"),
    "long": dict(arch="uPIT", S=2, B=8, sr=16000, lens=lambda seed, B: [320000 - 3840 * b for b in range(B)],
                 name="configs[4]: uPIT 2-spk train step on 16 kHz / 20 s utterances (T=2501), 512/128, batch 8/GPU"),
}


def workload_features(spec, seed):
    """Collated synthetic batch of a workload (features (T_b, F) per utterance; RSH: [mix || ones] input)."""
    from ssb200 import RSH, uPIT
    lens_s = spec["lens"](seed, spec["B"])
    samples = synth_features(seed, lens_s, spec["S"])
    if spec["arch"] == "RSH":
        for d in samples:
            m = d.pop("mix")
            d["combo"] = np.concatenate((m, np.ones_like(m)), axis=1)
        return RSH.Collator("combo")(samples), lens_s
    return uPIT.Collator("mix")(samples), lens_s


def _map_batch(batch, fn):
    from torch.nn.utils.rnn import PackedSequence
    if isinstance(batch, dict):
        return {k: (PackedSequence(fn(v.data), v.batch_sizes) if isinstance(v, PackedSequence) else v) for k, v in batch.items()}
    import copy
    out = copy.copy(batch)                                     # RSH MultiSpkBatch: a list of such dicts
    out.sub_batches = [_map_batch(b, fn) for b in batch.sub_batches]
    return out


def _batch_bytes(batch):
    from torch.nn.utils.rnn import PackedSequence
    if isinstance(batch, dict):
        return sum(v.data.numel() * 4 for v in batch.values() if isinstance(v, PackedSequence))
    return sum(_batch_bytes(b) for b in batch.sub_batches)


def run_train(args, rank, local, world, which="train", detail=True):
    """One train-step workload.  value: device-resident batches; e2e: pinned host batches through Trainer.prefetch/step
    with the loss read back every step.  detail: per-kernel pass, rooflines, latency model, DP checks."""
    import torch
    from ssb200 import RSH, train, uPIT
    spec = WORKLOADS[which]
    arch = RSH if spec["arch"] == "RSH" else uPIT
    Sx, Bx, sr = spec["S"], spec["B"], spec["sr"]
    hbm, tf_burst, tf_sus, peak_src = peaks()
    torch.manual_seed(0)
    kwargs = dict(feat_dim=str(F)) if arch is RSH else dict(feat_dim=str(F), num_spk=str(Sx))
    model = arch.SepDNN(local, **kwargs).cuda()
    trainer = train.Trainer(model, arch, exact=getattr(args, "dp_exact", False))
    meter = CallMeter()
    host, dev, audio_s, frames, tmax = [], [], [], [], []
    for i in range(N_BATCHES):
        # every rank knows every rank's lengths (seeded), so the job totals need no collective
        audio_s.append(sum(sum(spec["lens"](1000 * r + i, Bx)) for r in range(world)) / sr)
        b, lens_s = workload_features(spec, 1000 * rank + i)
        frames.append(sum(1 + n // HOP for n in lens_s))
        tmax.append(1 + max(lens_s) // HOP)
        host.append(_map_batch(b, lambda t: t.pin_memory()))
        dev.append(_map_batch(b, lambda t: t.cuda()))
    h2d = int(np.mean([_batch_bytes(b) for b in host]))
    sink = {}

    # strong scaling (--global-batch G): every rank runs G / (B * world) micro-batches per optimisation step with
    # gradient accumulation, one exchange and one update per step
    n_micro = max(1, getattr(args, "global_batch", 0) // (Bx * world)) if getattr(args, "global_batch", 0) else 1

    def step_dev(i):
        if n_micro == 1:
            sink["loss"], _ = trainer.step(dev[i % N_BATCHES])
        else:
            sink["loss"], _ = trainer.step_accumulate([dev[(i * n_micro + m) % N_BATCHES] for m in range(n_micro)])

    def step_e2e(i):
        # the user-facing loop: the copy of batch i+1 (pinned host memory) is started before step i is enqueued, so it
        # overlaps step i's kernels; every step still copies one full batch H2D and reads its loss back
        if n_micro > 1:
            loss, norm = trainer.step_accumulate([trainer.prefetch(host[(i * n_micro + m) % N_BATCHES]) for m in range(n_micro)])
            sink["host_loss"] = (float(loss), float(norm))
            return
        staged = sink.pop("staged", None)
        if staged is None or sink.get("staged_i") != i:
            staged = trainer.prefetch(host[i % N_BATCHES])
        sink["staged"], sink["staged_i"] = trainer.prefetch(host[(i + 1) % N_BATCHES]), i + 1
        loss, norm = trainer.step(staged)
        sink["host_loss"] = (float(loss), float(norm))          # D2H read of the step's result, every step

    K, W = args.steps, args.warmup
    for i in range(N_BATCHES):          # untimed: the workspace grows to the longest batch once (capacity-sized afterwards)
        step_dev(i)
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if (rank == 0 and detail) else None
    wt = model._wt
    meter.launches, replays0 = 0, wt.graph_replays
    ms, clocks = timed_region(step_dev, K, W, world, sampler)
    launches = (meter.launches + (wt.graph_replays - replays0) * wt.launches_per_refresh()) // (K + W) * K
    tot_audio = sum(audio_s[((W + i) * n_micro + m) % N_BATCHES] for i in range(K) for m in range(n_micro))
    value = tot_audio / (ms / 1e3)
    ms_e2e, _ = timed_region(step_e2e, K, W, world)
    e2e = tot_audio / (ms_e2e / 1e3)
    line = {
        "metric": "uPIT train audio-sec/sec" if which == "train" else "%s train audio-sec/sec" % spec["arch"],
        "value": round(value, 1), "unit": "audio-sec/sec", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": round(ms / K, 3), "higher_is_better": True,
        "scaling": "strong" if n_micro > 1 or getattr(args, "global_batch", 0) else "weak",
        "vs_baseline": None, "dtype": "split-bf16 operands (hi+lo), fp32 accumulate/state", "data": "synthetic",
        "config": {"workload": spec["name"], "global_batch": Bx * world * n_micro, "micro_batches_per_step": n_micro, "frames_per_step_per_gpu": int(np.mean(frames)),
                   "parallelism": "dp%d" % world,
                   "dp_semantics": ("mode B: loss normaliser and BatchNorm statistics over all ranks, gradients summed"
                                    if trainer.exact else
                                    "mode A: per-rank loss/BN, gradients averaged") + " (bucketed NCCL allreduce overlapped "
                                   "with backward)",
                   "l2": "%d distinct batches rotated (T_max %s; workspace sized by capacity, no allocation while timing); "
                         "per-step working set ~2 GB >> 126 MB L2" % (N_BATCHES, ",".join(str(t) for t in tmax))},
        "e2e": {"value": round(e2e, 1), "unit": "audio-sec/sec", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8,
                "ms_per_step": round(ms_e2e / K, 3)},
        "gpu_launches": launches,
    }
    if not detail:
        return line
    line["clocks"] = clocks
    # per-kernel device times (CUDA events around every C-ABI call, on the launching stream), separate untimed pass.
    # Streamed GEMM chunks run on the side stream UNDER the recurrences, so the per-kernel sums exceed the step time.
    meter.timing = True
    PROF = 2
    for i in range(PROF):
        step_dev(i)
    torch.cuda.synchronize()
    meter.timing = False
    prof = meter.summary(PROF)
    fr = float(np.mean(frames[:PROF]))
    passes = Sx if arch is RSH else 1
    gemm_fl, rec_fl = train_flops_per_frame(model.in_dim, model.out_dim, passes)
    rl = {}
    for other in ("gemm_sb16_x2", "gemm_sb16_tn"):   # the batched / weight-gradient launches are the same kernel template
        if other in prof:
            g, x = prof.setdefault("gemm_sb16", {"calls_per_step": 0, "ms_per_step": 0.0, "ms_per_call": 0.0, "ms_main_per_step": 0.0}), prof.pop(other)
            g["calls_per_step"] += x["calls_per_step"]
            g["ms_per_step"] += x["ms_per_step"]
            g["ms_main_per_step"] = g.get("ms_main_per_step", 0.0) + x["ms_main_per_step"]
    # the recurrence flops split evenly between the forward and the backward kernel
    for name, fl in (("gemm_sb16", gemm_fl), ("lstm_fwd", rec_fl / 2), ("lstm_bwd", rec_fl / 2)):
        if name in prof:
            a = fl * fr / (prof[name]["ms_per_step"] * 1e-3) / 1e12
            rl[name] = {"bound": "tensor", "achieved": round(a, 2), "peak": tf_sus, "unit": "TFLOP/s",
                        "frac": round(a / tf_sus, 4), "ms_per_step": round(prof[name]["ms_per_step"], 3)}
            if name == "gemm_sb16":      # every product is 3 bf16 MMAs (hi*hi + hi*lo + lo*hi)
                rl[name]["mma_issued_tflops"] = round(3 * a, 1)
                rl[name]["mma_issued_frac"] = round(3 * a / tf_sus, 4)
    for name, byts in (("pit_loss_fwd", (2 * Sx + 1) * 4.0 * F), ("pit_loss_bwd", (2 * Sx + 1) * 4.0 * F + Sx * F * 8.0)):
        if name in prof and arch is uPIT:
            a = byts * fr / (prof[name]["ms_per_step"] * 1e-3) / 1e9
            rl[name] = {"bound": "hbm", "achieved": round(a, 1), "peak": hbm, "unit": "GB/s", "frac": round(a / hbm, 4),
                        "ms_per_step": round(prof[name]["ms_per_step"], 3)}
    # The recurrences are chain-latency bound, not roofline bound: T_max dependent steps per launch, both directions in
    # one kernel.  What matters is microseconds per time step (phase budgets: profiles/).
    steps_per_launch = float(np.mean(tmax[:PROF]))
    lat = {}
    for name in ("lstm_fwd", "lstm_bwd"):
        if name in prof:
            launches_per_step = prof[name]["calls_per_step"]
            lat[name] = {"us_per_time_step": round(prof[name]["ms_per_step"] * 1e3 / (launches_per_step * steps_per_launch), 3),
                         "time_steps_per_launch": steps_per_launch, "launches_per_train_step": launches_per_step,
                         "ctas": 100 if name == "lstm_fwd" else 80,
                         "batch_per_launch": Bx if (name == "lstm_fwd" or Bx <= 64) else (Bx + 1) // 2,
                         "tensor_flops_per_time_step": 2 * 2 * 4 * H * H * Bx}
    divergence = grad_check = None
    if world > 1:
        import torch.distributed as dist
        # replicas must still hold identical parameters after all those steps
        ref_p = trainer.flat_p.clone()
        dist.broadcast(ref_p, src=0)
        d = (trainer.flat_p - ref_p).abs().max().reshape(1)
        dist.all_reduce(d, op=dist.ReduceOp.MAX)
        divergence = float(d)
        grad_check = trainer.grad_exchange_check(dev[0])
    dom = max(prof, key=lambda k: prof[k]["ms_main_per_step"])        # dominant kernel of the critical path (main stream)
    roofline = dict(rl.get(dom, {"bound": "tensor", "achieved": None, "peak": tf_sus, "unit": "TFLOP/s", "frac": None}))
    tb, tsrc = traffic_of(dom)
    roofline.update({"kernel": dom, "traffic": tb, "traffic_source": tsrc, "peak_source": peak_src,
                     "share_of_step": round(prof[dom]["ms_main_per_step"] / (ms / K), 3),
                     "note": "chain-latency bound kernel: see latency_model (us per time step), not the tensor fraction"})
    line.update({"roofline": roofline, "rooflines": rl, "latency_model": lat, "dp_param_divergence": divergence,
                 "dp_grad_check": grad_check,
                 "kernel_ms_per_step": {k: round(v["ms_per_step"], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms_per_step"])},
                 "kernel_ms_main_stream": {k: round(v["ms_main_per_step"], 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms_main_per_step"]) if v["ms_main_per_step"] > 0},
                 "kernel_ms_note": "CUDA events around each C-ABI call on its own stream; GEMM chunks streamed under the "
                                   "recurrences overlap them, so the sum exceeds ms_per_step"})
    return line


# ---------------------------------------------------------------- STFT (configs[1]) and masked iSTFT legs
STFT_NFFT, STFT_HOP = 256, 64
STFT_F = STFT_NFFT // 2 + 1


def stft_corpus_lengths(n_signals, seed=0):
    """configs[1]: wsj_tr-sized corpus = 20 000 mixtures x 3 signals, N ~ U{16000..80000} samples (2-10 s @ 8 kHz)."""
    rng = np.random.default_rng(seed)
    return rng.integers(16000, 80001, size=n_signals).astype(np.int64)


def cpu_stft_sample(budget_s=8.0):
    """oracle STFT (numpy restatement of the librosa call at extract_feats.py:87,89), one utterance at a time,
    single process, for ~budget_s seconds.  Returns (frames/sec, n_utts, cores=1)."""
    from oracle import stft_oracle
    lens = stft_corpus_lengths(4096, seed=1)
    rng = np.random.default_rng(1)
    sigs = [rng.standard_normal(int(n)).astype(np.float32) * 0.1 for n in lens[:64]]
    stft_oracle.stft_mag(sigs[0], STFT_NFFT, STFT_HOP)
    t0 = time.perf_counter()
    frames = n = 0
    while time.perf_counter() - t0 < budget_s:
        y = sigs[n % len(sigs)]
        frames += stft_oracle.stft_mag(y, STFT_NFFT, STFT_HOP).shape[1]
        n += 1
    return frames / (time.perf_counter() - t0), n, 1


def _cpu_stft_worker(args):
    seed, budget_s = args
    from oracle import stft_oracle
    lens = stft_corpus_lengths(64, seed=seed)
    rng = np.random.default_rng(seed)
    sigs = [rng.standard_normal(int(n)).astype(np.float32) * 0.1 for n in lens[:16]]
    stft_oracle.stft_mag(sigs[0], STFT_NFFT, STFT_HOP)
    t0 = time.perf_counter()
    frames = n = 0
    while time.perf_counter() - t0 < budget_s:
        frames += stft_oracle.stft_mag(sigs[n % len(sigs)], STFT_NFFT, STFT_HOP).shape[1]
        n += 1
    return frames, time.perf_counter() - t0


def cpu_stft_pool_sample(budget_s=4.0):
    """The same per-utterance loop in os.cpu_count() worker processes, the way steps/extract_feats.sh:41-48 fans the
    reference out over --nj SGE tasks.  Returns (aggregate frames/sec, workers)."""
    import multiprocessing as mp
    n = os.cpu_count() or 1
    try:
        with mp.get_context("spawn").Pool(n) as pool:
            res = pool.map_async(_cpu_stft_worker, [(100 + i, budget_s) for i in range(n)]).get(timeout=budget_s + 90)
    except Exception as e:                                     # a reported baseline must never take the bench line down
        sys.stderr.write("cpu_stft_pool_sample failed: %r\n" % (e,))
        return None, n
    return sum(f / t for f, t in res), n


def run_stft(args, rank, local, world, n_mix=20000, e2e_mix=2000):
    """value: the WHOLE corpus (n_mix x 3 signals / world per rank) resident in HBM, one launch per step.
    e2e: a shard of e2e_mix mixtures through pinned host buffers (H2D of the audio + D2H of the features inside
    the timed region), because pinning the whole 34 GB corpus would dominate the bench's wall time."""
    import torch
    from ssb200 import ops
    hbm, _, _, peak_src = peaks()
    K, W = args.steps, args.warmup
    per_rank = n_mix * 3 // world
    lens = stft_corpus_lengths(n_mix * 3)[rank * per_rank:(rank + 1) * per_rank]
    rag = ops.Ragged(lens, STFT_NFFT, STFT_HOP, "cuda", align=4 if ops.STFT_DFT256 else 1)
    n = int(rag.sig_off[-1])
    sig = torch.empty(n, dtype=torch.float32, device="cuda").uniform_(-0.5, 0.5)
    out = torch.empty(rag.total_frames * STFT_F, dtype=torch.float32, device="cuda")
    tot_frames = sum(int((1 + stft_corpus_lengths(n_mix * 3)[r * per_rank:(r + 1) * per_rank] // STFT_HOP).sum())
                     for r in range(world))
    meter = CallMeter()
    ev = []

    def step(i):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.stft(sig, rag, out=out)
        e1.record()
        ev.append((e0, e1))

    sampler = ClockSampler(local) if rank == 0 else None
    ms, clocks = timed_region(step, K, W, world, sampler)
    launches = meter.launches // (K + W) * K
    kms = float(np.mean([a.elapsed_time(b) for a, b in ev[W:]]))
    value = tot_frames * K / (ms / 1e3)
    bytes_per_frame = STFT_HOP * 4 + STFT_F * 4
    ach = bytes_per_frame * rag.total_frames / (kms * 1e-3) / 1e9
    # e2e shard
    ne = e2e_mix * 3 // world
    rag_e = ops.Ragged(lens[:ne], STFT_NFFT, STFT_HOP, "cuda", align=4 if ops.STFT_DFT256 else 1)
    n_e = int(rag_e.sig_off[-1])
    h_sig = torch.empty(n_e, dtype=torch.float32).uniform_(-0.5, 0.5).pin_memory()
    h_out = torch.empty(rag_e.total_frames * STFT_F, dtype=torch.float32).pin_memory()

    def step_e2e(i):
        ops.stft_host(h_sig, rag_e, h_out, chunks=8)     # H2D, kernel, D2H of utterance runs overlapped on 3 streams

    ms_e, _ = timed_region(step_e2e, K, W, world)
    e2e = rag_e.total_frames * world * K / (ms_e / 1e3)
    del sig, out
    return {
        "metric": "STFT frames/sec", "value": round(value, 0), "unit": "frames/sec", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": round(ms / K, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "configs[1]: STFT magnitude features, 8 kHz, 256-pt periodic Hann / hop 64 (F=129), "
                               "wsj_tr-sized corpus: %d mixtures x 3 signals, 2-10 s" % n_mix,
                   "signals_per_gpu": int(per_rank), "frames_per_gpu": int(rag.total_frames),
                   "mode": "magnitude, (F,T) file layout; log/CMVN off (reference has none, SURVEY D2)",
                   "l2": "input %.1f GB + output %.1f GB per step >> 126 MB L2" % (n * 4 / 1e9, rag.total_frames * STFT_F * 4 / 1e9)},
        "e2e": {"value": round(e2e, 0), "unit": "frames/sec", "h2d_bytes_per_step": n_e * 4,
                "d2h_bytes_per_step": rag_e.total_frames * STFT_F * 4, "sample": "%d-mixture shard, pinned host memory, ops.stft_host: 8 utterance runs pipelined over copy/compute/copy streams" % e2e_mix},
        "gpu_launches": launches, "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "stft_dft256_kernel" if ops.STFT_DFT256 else "stft_kernel<256>", "achieved": round(ach, 1), "peak": hbm, "unit": "GB/s",
                     "frac": round(ach / hbm, 4), "traffic": traffic_of("stft")[0], "traffic_source": traffic_of("stft")[1],
                     "peak_source": peak_src, "algorithmic_bytes_per_frame": bytes_per_frame, "kernel_ms": round(kms, 3)},
    }


def run_istft(args, rank, local, world, n_utts=3000, n_src=2):
    """Masked iSTFT (reconstruct_sources.py:39-42): wsj_tt-sized set, 512/128, int16 out, HBM-resident."""
    import torch
    from ssb200 import ops
    hbm, _, _, peak_src = peaks()
    K, W = args.steps, args.warmup
    rng = np.random.default_rng(5)
    lens = rng.integers(32000, 64001, size=n_utts).astype(np.int64)
    rag = ops.Ragged(lens, N_FFT, HOP, "cuda")
    sig = torch.empty(int(rag.sig_off[-1]), dtype=torch.float32, device="cuda").uniform_(-0.3, 0.3)
    spec = ops.stft(sig, rag, complex_out=True)
    mask = torch.rand(rag.total_frames * F * n_src, device="cuda")
    out, _ = ops.istft_masked(spec, mask, rag, n_src)
    ev = []

    def step(i):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.istft_masked(spec, mask, rag, n_src, out=out)
        e1.record()
        ev.append((e0, e1))

    ms, _ = timed_region(step, K, W, world)
    kms = float(np.mean([a.elapsed_time(b) for a, b in ev[W:]]))
    fs = rag.total_frames * n_src
    byts = F * 8 + F * 4 + HOP * 2
    ach = byts * fs / (kms * 1e-3) / 1e9
    return {"metric": "masked iSTFT frames/sec (per source)", "value": round(fs * world * K / (ms / 1e3), 0),
            "unit": "frame-sources/sec", "n_gpus": world,
            "config": {"workload": "mask x mixture-phase iSTFT, 512/128, %d utts x %d sources, int16 out" % (n_utts, n_src)},
            "roofline": {"bound": "hbm", "kernel": "istft_kernel<512>", "achieved": round(ach, 1), "peak": hbm,
                         "unit": "GB/s", "frac": round(ach / hbm, 4), "traffic": traffic_of("istft")[0],
                         "traffic_source": traffic_of("istft")[1], "peak_source": peak_src,
                         "algorithmic_bytes_per_frame_source": byts, "kernel_ms": round(kms, 3)}}


def run_separate(args, rank, local, world, which="train"):
    """Evaluation end to end on the device (SURVEY 8 f3, ssb200/pipeline.py:separate): host waveforms -> STFT ->
    BLSTM masks -> mask x mixture-phase iSTFT -> int16 audio on the host; what run_eval.sh does through three
    programs and two sets of .npz files (steps/extract_feats.py test branch, compute_masks, reconstruct_sources.py)."""
    import torch
    from ssb200 import pipeline, uPIT
    K, W = args.steps, args.warmup
    spec = WORKLOADS[which]
    torch.manual_seed(0)
    model = uPIT.SepDNN(local, feat_dim=str(F), num_spk=str(S)).cuda()
    model.eval()
    rng = np.random.default_rng(11 + rank)
    batches = []
    for i in range(N_BATCHES):
        lens = spec["lens"](5000 + 1000 * rank + i, spec["B"])
        batches.append([rng.uniform(-0.3, 0.3, size=n).astype(np.float32) for n in lens])
    audio = [sum(len(w) for w in b) / spec["sr"] for b in batches]
    sink = {}

    def step(i):
        # two slots: batch i+1 is prepared and enqueued while the device works on batch i; every step still ends with
        # one batch's int16 audio on the host
        h = pipeline.submit(model, batches[(i + 1) % N_BATCHES], N_FFT, HOP, slot=(i + 1) & 1)
        torch.cuda.current_stream().wait_event(h["done"])      # the slots run on streams of their own: the timing events
        if "pending" in sink:                                  # of this stream must wait for every submitted batch
            sink["wav"] = pipeline.collect(sink["pending"])
        sink["pending"] = h

    for i in range(N_BATCHES):           # untimed: every batch shape once
        step(i)
    ms, _ = timed_region(step, K, W, world)
    pipeline.collect(sink.pop("pending"))
    tot = sum(audio[(W + i) % N_BATCHES] for i in range(K)) * world
    nbytes = int(np.mean([sum(len(w) for w in b) for b in batches]))
    return {"metric": "uPIT separation audio-sec/sec (waveform in, %d int16 waveforms out)" % S,
            "value": round(tot / (ms / 1e3), 1), "unit": "audio-sec/sec", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms / K, 3),
            "config": {"workload": "eval of the shapes of %s: %d utterances per step, STFT %d/%d, S=%d, h0/c0 random per batch "
                                   "like the reference; pipeline.submit/collect on two slots"
                                   % (spec["name"].split(":")[0], spec["B"], N_FFT, HOP, S)},
            "e2e": {"value": round(tot / (ms / 1e3), 1), "unit": "audio-sec/sec", "h2d_bytes_per_step": nbytes * 4,
                    "d2h_bytes_per_step": nbytes * 2 * S}}


# ---------------------------------------------------------------- reference arm (CPU, rank 0 only)
def run_reference(args, rank, world):
    """`--impl reference`: the reference's CPU implementation of the path, on the host cores.  /root/reference is
    Python that cannot be installed (no setup.py) nor copied, and is absent on the GPU box, so this times
    oracle/cpu_step.py -- the same torch.nn CPU library calls the reference's step makes -- kind "port"."""
    if rank != 0:
        return None
    K, W = args.steps, args.warmup
    if args.workload == "stft":
        fps, n, cores = cpu_stft_sample(budget_s=min(60.0, 4.0 * (K + W)))
        return {"impl": "reference", "metric": "STFT frames/sec", "value": round(fps, 0), "unit": "frames/sec",
                "n_gpus": world, "steps": K, "warmup": W, "higher_is_better": True, "data": "synthetic", "dtype": "f64->c64",
                "config": {"workload": "configs[1] STFT 256/64, one utterance at a time (extract_feats.py:62-90), numpy"},
                "cpu_baseline": {"value": round(fps, 0), "unit": "frames/sec", "cores": cores, "kind": "port",
                                 "sample": "%d utterances of 2-10 s" % n},
                "e2e": {"value": round(fps, 0), "unit": "frames/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    # per-step sample: 4 utterances (~3 s of CPU work per step here), 2 when many steps are asked for, so that the
    # whole run ends within a few minutes whatever K and W are
    n_utts = 4 if K + W <= 30 else 2
    aps, sec, cores = cpu_train_sample(n_utts, K, W)
    return {"impl": "reference", "metric": "uPIT train audio-sec/sec", "value": round(aps, 2), "unit": "audio-sec/sec",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": round(sec * 1e3, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "configs[2] uPIT train step (same shapes as the GPU arm), CPU, torch %d threads" % cores,
                       "cpu_batch": n_utts, "gpu_arm_batch": B_PER_GPU,
                       "note": "a bounded sample: %d of the 64 utterances of a batch per CPU step; larger CPU batches run "
                               "SLOWER per audio-second on this kind of box (16 utterances: 11.5 vs 35.7 audio-s/s at 4), so "
                               "the GPU/CPU ratio computed from this arm is the conservative one" % n_utts},
            "cpu_baseline": {"value": round(aps, 2), "unit": "audio-sec/sec", "cores": cores, "kind": "port",
                             "sample": "%d of the 64 utterances of one batch per step" % n_utts},
            "e2e": {"value": round(aps, 2), "unit": "audio-sec/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="train", choices=["train", "stft", "istft", "separate", "separate_long"] +
                    [k for k in WORKLOADS if k != "train"])
    ap.add_argument("--dp-exact", action="store_true", help="data-parallel mode B (N > 1): exact large-batch arithmetic")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="strong scaling: fixed global batch (e.g. 512) run as micro-batches of 64 per rank with gradient accumulation")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the stft/istft/cpu_baseline legs of the train line")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: everything else anyone prints (the arch modules echo their model
    # parameters like the reference does, NCCL prints its version) goes to stderr, at file-descriptor level
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)
    emit = lambda line: os.write(out_fd, (json.dumps(line) + "\n").encode())
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        line = run_reference(args, rank, world)
        if line is not None:
            sys.stdout.flush()
            emit(line)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    rank, local, world = dist_setup(args.gpus)
    if args.workload == "stft":
        line = run_stft(args, rank, local, world)
    elif args.workload == "istft":
        line = run_istft(args, rank, local, world)
    elif args.workload == "separate":
        line = run_separate(args, rank, local, world)
    elif args.workload == "separate_long":
        line = run_separate(args, rank, local, world, "long")
    elif args.workload != "train":
        line = run_train(args, rank, local, world, args.workload)
    else:
        line = run_train(args, rank, local, world)
        if not args.no_extras:
            sub = argparse.Namespace(steps=max(3, min(args.steps, 5)), warmup=3, dp_exact=False, global_batch=0)
            torch.cuda.empty_cache()
            line["stft"] = run_stft(sub, rank, local, world)
            torch.cuda.empty_cache()
            line["istft"] = run_istft(sub, rank, local, world)
            torch.cuda.empty_cache()
            line["separate"] = run_separate(sub, rank, local, world)
            # the other BASELINE.json configurations, a few steps each (same timing rules; value / e2e / ms_per_step)
            others = {}
            for key in ("rsh", "upit3", "long", "train_b100"):
                torch.cuda.empty_cache()
                others[key] = run_train(sub, rank, local, world, key, detail=False)
            torch.cuda.empty_cache()
            others["separate_long"] = run_separate(sub, rank, local, world, "long")
            line["configs"] = others
    if rank == 0 and world == 1 and not args.no_extras:
        if args.workload == "train":
            # 4 utterances per CPU step like the --impl reference arm (larger CPU batches run SLOWER per audio-second on
            # this box: 16 utterances 11.5 vs 35.7 audio-s/s), 12 steps ~ 11 s of CPU work
            aps, sec, cores = cpu_train_sample(4, 10, 2)
            line["cpu_baseline"] = {"value": round(aps, 2), "unit": "audio-sec/sec", "cores": cores, "kind": "port",
                                    "cpu_batch": 4,
                                    "sample": "10 train steps (+2 warm-up) on 4 of the 64 utterances (%.2f s/step)" % sec}
            try:   # the recorded same-box comparator (tools/time_cudnn_ref.py; informative, not the official baseline)
                line["cudnn_comparator_recorded"] = [json.loads(l) for l in open(os.path.join(ROOT, "profiles", "r02_cudnn_comparator.jsonl"))]
            except Exception:
                pass
            fps, n, c1 = cpu_stft_sample(5.0)
            line["stft"]["cpu_baseline"] = {"value": round(fps, 0), "unit": "frames/sec", "cores": c1, "kind": "port",
                                            "sample": "%d utterances, single process" % n}
            pfps, pw = cpu_stft_pool_sample(3.0)
            if pfps is not None:
                line["stft"]["cpu_baseline_pool"] = {"value": round(pfps, 0), "unit": "frames/sec", "cores": pw, "kind": "port",
                                                 "sample": "%d worker processes x 3 s, one utterance at a time each" % pw}
        elif args.workload == "stft":
            fps, n, c1 = cpu_stft_sample(10.0)
            line["cpu_baseline"] = {"value": round(fps, 0), "unit": "frames/sec", "cores": c1, "kind": "port",
                                    "sample": "%d utterances, single process" % n}
            pfps, pw = cpu_stft_pool_sample(5.0)
            if pfps is not None:
                line["cpu_baseline_pool"] = {"value": round(pfps, 0), "unit": "frames/sec", "cores": pw, "kind": "port",
                                         "sample": "%d worker processes x 5 s, one utterance at a time each" % pw}
    if rank == 0:
        sys.stdout.flush()
        emit(line)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
