import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from ssb200 import ops, synth, evaluate_oracle
from ssb200._lib import call, stream_ptr
from oracle import stft_oracle
mix, srcs = synth.mixture(300, 9000, 2)
rag = ops.Ragged([9000], 512, 128, "cuda")
spec = ops.stft(torch.from_numpy(mix).cuda(), rag, complex_out=True)
sp = spec.cpu().numpy().reshape(257, -1)
ref = stft_oracle.stft(mix, 512, 128)
print("gpu spec: zeros", (np.abs(sp) == 0).sum(), "min", np.abs(sp).min(), "nan", np.isnan(sp).sum(), "maxdiff", np.abs(sp - ref).max())
mags = [ops.stft(torch.from_numpy(s).cuda(), rag).cpu().numpy().reshape(257, -1) for s in srcs]
print("mag zeros", [(m == 0).sum() for m in mags], "nan", [np.isnan(m).sum() for m in mags])
est, sc = evaluate_oracle.oracle_separate([mix], [srcs], 512, 128, False)
for s in range(2):
    e = est[0][s]
    bad = np.argwhere(~np.isfinite(e)).ravel()
    print("est", s, "nonfinite", len(bad), bad[:10], "score", sc[0, s])
