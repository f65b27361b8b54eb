"""Informative same-box comparator: the torch library calls the reference's
train step makes, on the B200 through stock cuDNN / cuBLAS / ATen --

  archs/uPIT.py:115-119,129-147   nn.LSTM(F, 600, 2 layers, bidirectional) on a PackedSequence -> pad ->
                                  BatchNorm1d(1200) -> Linear -> sigmoid
  archs/uPIT.py:178-206           mask * tiled mixture, squared error per permutation, min over permutations
  steps/train_qsub.py:120-122     backward, clip_grad_norm_(0.25), Adam(lr=1e-3).step()

-- on the SAME four synthetic batches bench.py's train leg rotates (configs[2]: B=64, 4-8 s at 8 kHz, 512/128), timed
the same way (CUDA events, warm-up, inputs resident in HBM).  Not the official baseline (that is the CPU arm); it only
answers "does the hand-written path beat torch.nn.LSTM on the same GPU".  Prints one JSON line.

    python tools/time_cudnn_ref.py [--steps 20] [--warmup 5] [--tf32]
"""
import argparse
import itertools
import json
import os
import sys

import torch
import torch.nn as nn
from torch.nn.utils.rnn import PackedSequence, pad_packed_sequence

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200"))
import bench  # noqa: E402  (batch generators only)


class Net(nn.Module):
    def __init__(self, feat_dim, num_spk):
        super().__init__()
        self.blstm = nn.LSTM(feat_dim, 600, num_layers=2, bidirectional=True)
        self.lin = nn.Linear(1200, feat_dim * num_spk)
        self.bn = nn.BatchNorm1d(1200)

    def forward(self, packed, hidden):
        out, _ = self.blstm(packed, hidden)
        out, lens = pad_packed_sequence(out, batch_first=True)
        out = self.bn(out.permute(0, 2, 1)).permute(0, 2, 1)
        return torch.sigmoid(self.lin(out)), lens


def loss_fn(mask, mix, sources):
    S = len(sources)
    masked = mask * torch.cat([mix] * S, dim=2)
    losses = torch.stack([((masked - torch.cat([sources[i] for i in perm], dim=2)) ** 2).sum((1, 2))
                          for perm in itertools.permutations(range(S))])
    best, _ = torch.min(losses, 0)
    return best.sum()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--tf32", action="store_true", help="allow TF32 in cuDNN/cuBLAS (default: fp32 math like 2018)")
    args = ap.parse_args()
    torch.backends.cudnn.allow_tf32 = args.tf32
    torch.backends.cuda.matmul.allow_tf32 = args.tf32
    F, S, B = bench.F, bench.S, bench.B_PER_GPU
    torch.manual_seed(0)
    net = Net(F, S).cuda()
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    import numpy as np
    from ssb200 import uPIT
    coll = uPIT.Collator("mix")
    batches, audio = [], []
    for i in range(bench.N_BATCHES):
        lens_s = bench.train_lengths(i, B)
        audio.append(sum(lens_s) / bench.SR)
        b = coll(bench.synth_features(i, lens_s))
        batches.append({k: PackedSequence(v.data.cuda(), v.batch_sizes) for k, v in b.items()})

    def step(i):
        b = batches[i % len(batches)]
        net.zero_grad()
        hidden = (torch.randn(4, B, 600, device="cuda"), torch.randn(4, B, 600, device="cuda"))
        mask, lens = net(b["mix"], hidden)
        mix, _ = pad_packed_sequence(b["mix"], batch_first=True)
        srcs = [pad_packed_sequence(b["source%d" % (s + 1)], batch_first=True)[0] for s in range(S)]
        loss = loss_fn(mask, mix, srcs) / S / (float(lens.sum()) * F)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(net.parameters(), 0.25)
        opt.step()
        return loss

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    tot = sum(audio[(args.warmup + i) % len(audio)] for i in range(args.steps))
    print(json.dumps({"comparator": "torch.nn.LSTM (cuDNN %s) + BatchNorm1d + Linear + ATen uPIT loss + clip + Adam on the B200"
                                    % torch.backends.cudnn.version(),
                      "math": "tf32" if args.tf32 else "fp32", "value": round(tot / (ms / 1e3), 1), "unit": "audio-sec/sec",
                      "ms_per_step": round(ms / args.steps, 3), "steps": args.steps, "warmup": args.warmup,
                      "config": "configs[2]: B=64, 4-8 s at 8 kHz, STFT 512/128 (F=257), same four batches as bench.py",
                      "torch": torch.__version__, "gpu": torch.cuda.get_device_name(0)}))


if __name__ == "__main__":
    main()
