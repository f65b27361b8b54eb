"""Development aid: clock64 phase trace of CTA 0 of the forward recurrence."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200"))
from ssb200 import _lib, uPIT
B, T, F = 64, 376, 257
torch.manual_seed(0)
model = uPIT.SepDNN(0, feat_dim=str(F), num_spk="2").cuda()
rng = np.random.default_rng(0)
samples = [{"mix": rng.random((T, F), dtype=np.float32), "source1": rng.random((T, F), dtype=np.float32),
            "source2": rng.random((T, F), dtype=np.float32)} for _ in range(B)]
batch = uPIT.Collator("mix")(samples)
for _ in range(2):
    loss, _ = uPIT.compute_loss(model, 0, batch)
dbg = torch.zeros(T * 32, dtype=torch.int64, device="cuda")
_lib.call("lstm_debug", dbg)
with torch.no_grad():
    uPIT.compute_loss(model, 0, batch)
torch.cuda.synchronize()
_lib.call("lstm_debug", None)
d = dbg.cpu().numpy().reshape(T, 32)   # holds layer 1 (last launch)
names = ["poll_start", "poll_done", "tma_issued", "mma_first_full", "mma_committed", "epi_accum", "epi_math_done", "epi_published"]
s0 = 100
base = d[s0, 1]
for s in range(s0, s0 + 4):
    print("step", s, " ".join("%s=%d" % (n, d[s, i] - base) for i, n in enumerate(names)))
per = np.diff(d[50:350, 1]).mean()
print("cycles/step %.0f" % per)
seg = d[50:350]
print("poll_done->tma_issued %.0f" % (seg[:, 2] - seg[:, 1]).mean())
print("poll_done->mma_first_full %.0f" % (seg[:, 3] - seg[:, 1]).mean())
print("mma_first_full->mma_committed %.0f" % (seg[:, 4] - seg[:, 3]).mean())
print("mma_committed->epi_accum %.0f" % (seg[:, 5] - seg[:, 4]).mean())
print("epi_accum->epi_math_done %.0f" % (seg[:, 6] - seg[:, 5]).mean())
print("epi_math_done->published %.0f" % (seg[:, 7] - seg[:, 6]).mean())
print("epi_accum->tmem_ld_done %.0f" % (seg[:, 8] - seg[:, 5]).mean())
print("chunk arrivals rel poll_done", [int((seg[:, 10 + k] - seg[:, 1]).mean()) for k in range(1, 10)])
print("publish: stores %.0f bar %.0f fence.proxy %.0f red.release %.0f" % ((seg[:,20]-seg[:,6]).mean(), (seg[:,21]-seg[:,20]).mean(), (seg[:,22]-seg[:,21]).mean(), (seg[:,23]-seg[:,22]).mean()))
print("published->next poll_done %.0f" % (seg[1:, 1] - seg[:-1, 7]).mean())
