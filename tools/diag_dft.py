import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from ssb200 import ops, synth
from oracle import stft_oracle as so
lens = [129, 127 * 64, 128 * 64 - 1, 128 * 64, 129 * 64 + 5, 300, 64 * 3 + 1, 40000, 12345]
sigs = [synth.mixture(1100 + i, n, 2)[0] for i, n in enumerate(lens)]
for align in (4, 1):
    rag = ops.Ragged(lens, 256, 64, "cuda", align=align)
    host = np.zeros(int(rag.sig_off[-1]), dtype=np.float32)
    for u, s in enumerate(sigs):
        host[int(rag.sig_off[u]):int(rag.sig_off[u]) + lens[u]] = s
    flat = torch.from_numpy(host).cuda()
    for mode in (0,):
        ops.STFT_DFT256 = True
        out = ops.stft(flat, rag)
        torch.cuda.synchronize()
        errs = []
        for u, v in enumerate(ops.stft_views(out, rag)):
            ref = so.stft_mag(sigs[u], 256, 64)
            errs.append(float(np.abs(v.cpu().numpy() - ref).max() / np.abs(ref).max()))
        print("align", align, "mode", mode, "max rel err per utt", ["%.2e" % e for e in errs])
# timing on a configs[1]-like shard
from ssb200._lib import call, LL, stream_ptr
rng = np.random.default_rng(0)
L = rng.integers(16000, 80001, size=60000).astype(np.int64)
rag = ops.Ragged(L, 256, 64, "cuda", align=4)
x = torch.empty(int(rag.sig_off[-1]), device="cuda").uniform_(-0.5, 0.5)
out = torch.empty(rag.total_frames * 129, device="cuda")
tab, u0, n_tiles, total_rows = rag.dft_tiles()
for name, exp in (("full", 0), ("no MMA", 1), ("no producers", 2), ("no stores", 4), ("MMA only", 6), ("producers only", 5), ("stores only", 3)):
    run = lambda: call("stft_dft256", x, tab, rag.n_utts, u0, n_tiles, LL(total_rows), out, exp << 8, stream_ptr())
    for _ in range(2):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        run()
    e1.record()
    torch.cuda.synchronize()
    print("dft %-16s %.3f ms" % (name, e0.elapsed_time(e1) / 3))
