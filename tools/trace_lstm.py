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
d = dbg.cpu().numpy()[:T * 32].reshape(T, 32)   # holds layer 1 (last launch)
# stamps (csrc/lstm_fwd.cu): thread 0 = fetch + epilogue thread, M = the MMA-issuing thread
names = {0: "gate_start", 1: "gate_open", 2: "copies_issued", 3: "M_first_full", 4: "M_committed", 5: "epi_accum",
         8: "epi_sums", 6: "epi_math", 7: "epi_published", 9: "epi_stores"}
s0 = 100
base = d[s0, 0]
for s in range(s0, s0 + 3):
    print("step", s, " ".join("%s=%d" % (n, d[s, i] - base) for i, n in sorted(names.items(), key=lambda kv: d[s, kv[0]])))
seg = d[50:350]
print("cycles/step %.0f" % np.diff(d[50:350, 5]).mean())
def dt(a, b, shift=0):
    return (seg[shift:, a] - seg[:len(seg) - shift, b]).mean()
print("gate_open->copies_issued (10 bulk copies)          %.0f" % dt(2, 1))
print("gate_open->M_first_full (chunk 0 landed + checked) %.0f" % dt(3, 1))
print("M_first_full->M_committed                          %.0f" % dt(4, 3))
print("M_committed->epi_accum                             %.0f" % dt(5, 4))
print("epi_accum->epi_sums (tmem ld + lo-row exchange)    %.0f" % dt(8, 5))
print("epi_sums->epi_math                                 %.0f" % dt(6, 8))
print("epi_math->epi_published                            %.0f" % dt(7, 6))
print("epi_published->epi_stores                          %.0f" % dt(9, 7))
print("epi_published->next gate_open                      %.0f" % dt(1, 7, 1))
print("chunk valid seen by M rel gate_open", [int((seg[:, 10 + k] - seg[:, 1]).mean()) for k in range(1, 10)])
print("copy i issued rel gate_open", [int((seg[:, 20 + k] - seg[:, 1]).mean()) for k in range(10)])
