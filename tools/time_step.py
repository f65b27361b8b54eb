"""Per-C-ABI-call device timing of one uPIT train step on synthetic features (development aid)."""
import argparse
import collections
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200"))
from ssb200 import _lib, uPIT, engine, ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--T", type=int, default=376)
ap.add_argument("--F", type=int, default=257)
ap.add_argument("--S", type=int, default=2)
ap.add_argument("--iters", type=int, default=5)
args = ap.parse_args()

torch.manual_seed(0)
model = uPIT.SepDNN(0, feat_dim=str(args.F), num_spk=str(args.S)).cuda()
rng = np.random.default_rng(0)
lens = sorted(rng.integers(int(args.T * 0.66), args.T + 1, size=args.B).tolist(), reverse=True)
lens[0] = args.T
samples = []
for L in lens:
    d = {"mix": rng.random((L, args.F), dtype=np.float32)}
    for s in range(args.S):
        d["source%d" % (s + 1)] = rng.random((L, args.F), dtype=np.float32) * 0.5
    samples.append(d)
batch = uPIT.Collator("mix")(samples)

records = collections.OrderedDict()
orig_call = _lib.call
timing = {"on": False}


def timed_call(name, *a):
    if not timing["on"]:
        return orig_call(name, *a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = orig_call(name, *a)
    e1.record()
    key = name
    if name == "gemm_sb16":
        key = "gemm_sb16 M=%d N=%d K=%d" % (a[8], a[9], a[10])
    elif name == "gemm_sb16_tn":
        key = "gemm_sb16_tn M=%d N=%d K=%d x%d" % (a[22], a[23], 64 * a[15], a[18])
    records.setdefault(key, []).append((e0, e1))
    return r


for mod in (_lib, uPIT, engine, ops):
    mod.call = timed_call

for it in range(args.iters + 2):
    timing["on"] = it >= 2
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    loss, norm = uPIT.compute_loss(model, 0, batch)
    loss.backward()
    t1.record()
    torch.cuda.synchronize()
    print("iter", it, "step ms %.3f" % t0.elapsed_time(t1), "loss", float(loss))
tot = 0.0
for k, v in records.items():
    ms = sum(a.elapsed_time(b) for a, b in v) / args.iters
    tot += ms
    extra = ""
    if k.startswith("gemm"):
        M_, N_, K_ = (int(x.split("=")[1]) for x in k.split()[1:4])
        fl = 2.0 * M_ * N_ * K_ * (2 if k.endswith("x2") else 1) * (len(v) // args.iters)
        extra = "  %6.1f TF/s algorithmic, %6.1f issued" % (fl / ms / 1e9, 3 * fl / ms / 1e9)
    print("%-40s calls/step %3d   ms/step %8.3f%s" % (k, len(v) // args.iters, ms, extra))
print("sum of kernels ms/step %.3f ; audio-s/step %.1f" % (tot, sum(lens) * 128 / 8000.0))
