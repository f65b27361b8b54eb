"""Development aid: device timeline of one train step -- for every C-ABI call, the stream it was launched on and its
start / end time relative to the start of the step (CUDA events on the launching stream).  Shows which GEMM chunks ran
under which recurrence.   python tools/timeline.py [--T 501] [--B 64]"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200"))
from ssb200 import _lib, engine, ops, train, uPIT  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--T", type=int, default=501)
ap.add_argument("--F", type=int, default=257)
args = ap.parse_args()
torch.manual_seed(0)
model = uPIT.SepDNN(0, feat_dim=str(args.F), num_spk="2").cuda()
rng = np.random.default_rng(0)
lens = sorted(rng.integers(int(args.T * 0.5), args.T + 1, size=args.B).tolist(), reverse=True)
lens[0] = args.T
samples = [{"mix": rng.random((L, args.F), dtype=np.float32), "source1": rng.random((L, args.F), dtype=np.float32) * 0.5,
            "source2": rng.random((L, args.F), dtype=np.float32) * 0.5} for L in lens]
batch = {k: torch.nn.utils.rnn.PackedSequence(v.data.cuda(), v.batch_sizes) for k, v in uPIT.Collator("mix")(samples).items()}
tr = train.Trainer(model, uPIT)
orig = _lib.call
rec, on = [], {"v": False}
main_id = torch.cuda.current_stream().cuda_stream


def timed(name, *a):
    if not on["v"] or name == "stream_wait_value32":
        return orig(name, *a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = orig(name, *a)
    e1.record()
    tag = name
    if name == "gemm_sb16":
        tag = "gemm M=%d N=%d K=%d beta=%g lim=%d" % (a[8], a[9], a[10], a[13], a[16])
    elif name == "gemm_sb16_tn":
        tag = "gemm_tn M=%d N=%d blk0=%d,%d n=%d x%d lim=%d" % (a[22], a[23], a[13], a[14], a[15], a[18], a[28])
    rec.append((tag, "main" if torch.cuda.current_stream().cuda_stream == main_id else "side", e0, e1))
    return r


for mod in (_lib, uPIT, engine, ops, train):
    if hasattr(mod, "call"):
        mod.call = timed
for i in range(3):
    tr.step(batch)
torch.cuda.synchronize()
on["v"] = True
t0 = torch.cuda.Event(enable_timing=True)
t0.record()
tr.step(batch)
t1 = torch.cuda.Event(enable_timing=True)
t1.record()
torch.cuda.synchronize()
print("step %.3f ms (with per-call events)" % t0.elapsed_time(t1))
rows = sorted(((t0.elapsed_time(a), t0.elapsed_time(b), s, n) for n, s, a, b in rec))
for a, b, s, n in rows:
    print("%8.3f %8.3f  %-4s %s" % (a, b, s, n))
