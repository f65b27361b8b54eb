"""Development aid: clock64 phase trace of CTA 0 of the dual-context forward recurrence (lstm_fwd_dual.cu)."""
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
dbg = torch.zeros(T * 64, dtype=torch.int64, device="cuda")
_lib.call("lstm_debug", dbg)
with torch.no_grad():
    uPIT.compute_loss(model, 0, batch)
torch.cuda.synchronize()
_lib.call("lstm_debug", None)
d = dbg.cpu().numpy().reshape(T, 64)   # holds layer 1 (last launch)
names = ["poll", "seen", "loads", "land0", "mma", "epi", "math", "bar", "rel"]
base = d[100, 1]
for s in range(100, 103):
    for c in range(2):
        print("step", s, "ctx", c, " ".join("%s=%d" % (n, d[s, 16 * c + i] - base) for i, n in enumerate(names)))
seg = d[50:350]
print("cycles/step %.0f" % np.diff(seg[:, 1]).mean())
for c in range(2):
    o = 16 * c
    print("ctx", c, "poll->seen %.0f seen->loads %.0f seen->land0 %.0f land0->mma %.0f mma->epi %.0f epi->math %.0f math->bar %.0f bar->rel %.0f rel->next seen %.0f" % (
        (seg[:, o + 1] - seg[:, o]).mean(), (seg[:, o + 2] - seg[:, o + 1]).mean(), (seg[:, o + 3] - seg[:, o + 1]).mean(),
        (seg[:, o + 4] - seg[:, o + 3]).mean(), (seg[:, o + 5] - seg[:, o + 4]).mean(), (seg[:, o + 6] - seg[:, o + 5]).mean(),
        (seg[:, o + 7] - seg[:, o + 6]).mean(), (seg[:, o + 8] - seg[:, o + 7]).mean(), (seg[1:, o + 1] - seg[:-1, o + 8]).mean()))
print("ctx0 seen -> ctx1 seen %.0f" % (seg[:, 17] - seg[:, 1]).mean())
print("ctx0 chunk full seen by MMA thread, rel. ctx0 seen:", [int((seg[:, 32 + k] - seg[:, 1]).mean()) for k in range(10)])
print("ctx0 copies 0..5 issued, rel. ctx0 seen:", [int((seg[:, 42 + k] - seg[:, 1]).mean()) for k in range(6)])
print("ctx1 chunk full seen by MMA thread, rel. ctx0 seen:", [int((seg[:, 48 + k] - seg[:, 1]).mean()) for k in range(10)])
print("ctx1(s-1) epi/math/bar/rel rel. ctx0(s) seen:", [int((seg[:-1, 16 + k] - seg[1:, 1]).mean()) for k in (4, 5, 6, 7, 8)])
print("ctx0 epi/math/bar/rel rel. ctx0 seen:", [int((seg[:, k] - seg[:, 1]).mean()) for k in (4, 5, 6, 7, 8)])
print("ctx1 seen/loads rel. ctx0 seen:", [int((seg[:, 16 + k] - seg[:, 1]).mean()) for k in (0, 1, 2)])
