"""Small workload for compute-sanitizer (racecheck / synccheck / memcheck): one tape-free train step with the side-stream
overlap on (ss_lstm_fwd, ss_lstm_bwd K-split path, ss_gemm_sb16[_x2], flag-gated weight-gradient GEMMs, clip + Adam), one
autograd step, STFT -> masks -> masked iSTFT.  Short sequences: the tools slow the persistent kernels down 10-100x.

    compute-sanitizer --tool racecheck python tools/sanitize_step.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200"))
from ssb200 import ops, pipeline, synth, train, uPIT  # noqa: E402

T = int(os.environ.get("SAN_T", "12"))
n_fft, hop, S, B = 256, 64, 2, 6
F = n_fft // 2 + 1
torch.manual_seed(0)
model = uPIT.SepDNN(0, feat_dim=str(F), num_spk=str(S)).cuda()
lens = [hop * (T - 1 - b) + 5 for b in range(B)]
waves = [synth.mixture(70 + b, n, S) for b, n in enumerate(lens)]
batch, order = pipeline.train_batch([w[0] for w in waves], [w[1] for w in waves], n_fft, hop)
tr = train.Trainer(model, uPIT)
for i in range(2):
    loss, norm = tr.step(batch)
torch.cuda.synchronize()
print("trainer steps ok, loss", float(loss))
loss, norm = uPIT.compute_loss(model, 0, batch)
loss.backward()
torch.cuda.synchronize()
print("autograd step ok, loss", float(loss))
model.eval()
out = pipeline.separate(model, [w[0] for w in waves], n_fft, hop)
print("separate ok", len(out), out[0][0].dtype, out[0][0].shape)
