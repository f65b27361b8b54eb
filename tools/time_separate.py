"""Host-side phase times of pipeline.separate (ms, averaged): where a call's wall time goes beside the kernels."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, "speech-separation_b200")
from ssb200 import ops, pipeline, uPIT  # noqa: E402
from ssb200.uPIT import run_forward  # noqa: E402

import os  # noqa: E402

LOCAL = int(os.environ.get("LOCAL_RANK", "0"))             # under torchrun: one process per GPU, all timing at once
torch.cuda.set_device(LOCAL)
N_FFT, HOP, S, B = 512, 128, 2, 64
F = N_FFT // 2 + 1
torch.manual_seed(0)
model = uPIT.SepDNN(LOCAL, feat_dim=str(F), num_spk=str(S)).cuda().eval()
rng = np.random.default_rng(0)
wavs = [rng.uniform(-0.3, 0.3, size=int(n)).astype(np.float32) for n in rng.integers(32000, 64001, size=B)]
for _ in range(3):
    pipeline.separate(model, wavs, N_FFT, HOP)
acc = {}


def lap(name, t0):
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t1 - t0) * 1e3
    return t1


R = 10
for _ in range(R):
    t = time.perf_counter()
    order = pipeline._order(wavs, HOP)
    rag = ops.Ragged([len(wavs[i]) for i in order], N_FFT, HOP, "cuda")
    t = lap("order+Ragged", t)
    sig = pipeline._upload(wavs, order, "cuda")
    t = lap("concat+H2D", t)
    spec = ops.stft(sig, rag, complex_out=True, layout_tf=True)
    x = ops.ragged_to_padded(spec, rag)
    t = lap("stft+pad", t)
    batch = pipeline._batch(rag, x, "cuda")
    model.hidden = model.init_hidden(B)
    t = lap("batch+hidden", t)
    with torch.no_grad():
        run_forward(model, batch, model.hidden, save=False)
    t = lap("forward", t)
    ws = batch["ws"]
    mask = ops.masks_to_ragged(ws.z, ws.ldz, rag, batch["T"], S)
    t = lap("masks", t)
    wav, off = ops.istft_masked(spec, mask, rag, S, layout_tf=True)
    t = lap("istft", t)
    host = wav.cpu().numpy()
    t = lap("D2H", t)
    out = [[host[off[u] + s * int(rag.L[u]): off[u] + (s + 1) * int(rag.L[u])] for s in range(S)] for u in range(B)]
    t = lap("slices", t)
print("rank", LOCAL, "threads", torch.get_num_threads(), {k: round(v / R, 3) for k, v in acc.items()}, "total",
      round(sum(acc.values()) / R, 3), flush=True)
