"""Development aid: clock64 phase trace of CTA 0 of the K-split backward recurrence (layer 0 = last launch)."""
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
    loss.backward()
loss, _ = uPIT.compute_loss(model, 0, batch)
dbg = torch.zeros((T + 2) * 32 + 80 * 16 * 4, dtype=torch.int64, device="cuda")
torch.cuda.synchronize()
_lib.call("lstm_debug", dbg)
loss.backward()
torch.cuda.synchronize()
_lib.call("lstm_debug", None)
allv = dbg.cpu().numpy()
d = allv[:(T + 2) * 32].reshape(T + 2, 32)
g = allv[(T + 2) * 32:].reshape(80, 16, 4)
seg = d[50:350]
print("cycles/step %.0f" % np.diff(d[50:350, 1]).mean())
def m(a, b): return (seg[:, b] - seg[:, a]).mean()
print("poll_start->poll_done %.0f" % m(0, 1))
print("poll_done->all loads issued %.0f" % m(1, 2))
print("poll_done->chunk i observed by MMA thread", [int(m(1, 10 + i)) for i in range(10)])
print("poll_done->accum committed (issue side) %.0f" % m(1, 3))
print("poll_done->epilogue sees accum %.0f" % m(1, 4))
print("accum seen->partials sent %.0f" % m(4, 5))
print("sent->peers' partials landed %.0f" % m(5, 6))
print("landed->math+dG stored %.0f" % m(6, 7))
print("bar %.0f ; fence+release %.0f" % (m(7, 8), m(8, 9)))
print("published->next poll_done %.0f" % (seg[1:, 1] - seg[:-1, 9]).mean())

# all-CTA skew (globaltimer ns) at step 108 of the last launch, relative to the earliest "accum seen"
st = 8
base = g[:, st, 0].min()
print("per CTA (dir,cluster,rank): accum_seen / partials_landed / before_bar / released  [ns rel.]")
for cta in range(80):
    if cta % 4 == 0 or cta in (37, 38, 39, 77, 78, 79):
        print(cta, (cta // 40, (cta % 40) // 4, cta % 4), [int(x - base) for x in g[cta, st]])
print("released: min %d max %d ; step period %.0f ns" % ((g[:, st, 3] - base).min(), (g[:, st, 3] - base).max(), np.diff(g[0, :, 3]).mean()))
for dname, sl in (("dir0", slice(0, 40)), ("dir1", slice(40, 80))):
    r = g[sl, st, 3] - base
    print(dname, "released spread", int(r.min()), int(r.max()), "slowest cta", int(np.argmax(r)) + sl.start, "accum_seen spread", int((g[sl, st, 0] - base).min()), int((g[sl, st, 0] - base).max()))
