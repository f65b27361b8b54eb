"""Drop-in arch module for the reference's `archs/RSH.py` ("recurrent selective hearing").

Same plugin surface as ssb200.uPIT.  The net sees [mixture || attention mask] (2F inputs) and emits ONE
mask (F outputs) per pass; a batch is grouped into sub-batches of equal speaker count; a sub-batch with S speakers runs
S passes that share weights, carry the LSTM state from pass to pass, assign each pass's output greedily to the
not-yet-used source of least squared error, and subtract the emitted mask from the attention channel
(archs/RSH.py:197-259).  Every pass keeps its own activation workspace; the backward walks the passes in reverse,
accumulating weight gradients and handing (dh0, dc0) and the residual-attention gradient to the previous pass.
All device arithmetic is libssb200 kernels; there is no torch/cuDNN or CPU fallback.
"""
import os

import numpy as np
import torch
import torch.nn as nn
from torch.utils.data import Dataset

from . import engine, uPIT
from ._lib import LL, call, stream_ptr
from .uPIT import _read_scp, run_forward, unpack_batch

HID = engine.H


class MultiSpkBatch:
    """sub_batches[n] = collated samples that have n speakers ({} when none) (archs/RSH.py:72-87)."""

    def __init__(self, max_spk, length):
        self.sub_batches, self.sub_batch_lens = [], []
        self.max_spk, self.length = max_spk, length

    def __len__(self):
        return self.length

    def __getitem__(self, idx):
        return self.sub_batches[idx]

    def append(self, elem):
        self.sub_batches.append(elem)


class Collator:
    """Groups by speaker count ('num_spk' entry, else number of source arrays), then sorts each group by
    descending input length and packs it (archs/RSH.py:22-70)."""

    def __init__(self, sort_key):
        self.key = sort_key
        self.sub = uPIT.Collator(sort_key)

    def __call__(self, batch):
        if "num_spk" in batch[0]:
            counts = [int(d["num_spk"]) for d in batch]
        else:
            counts = [len(d) - 1 for d in batch]
        out = MultiSpkBatch(max(counts) + 1, len(batch))
        for n in range(out.max_spk):
            members = [d for d, c in zip(batch, counts) if c == n]
            out.sub_batch_lens.append(len(members))
            out.append(self.sub(members) if members else {})
        return out


def _combo(mix_mag):
    """[mix || ones]: the attention mask starts at 1 everywhere (archs/RSH.py:107-109)."""
    return np.concatenate((mix_mag, np.ones(mix_mag.shape)), axis=1)


class TrainSet(Dataset):
    """feats_train.scp -> {'combo' (T,2F), 'source1'..} (archs/RSH.py:90-120)."""

    def __init__(self, datadir, location=""):
        filelist = datadir + "/feats_train.scp"
        if location:
            os.system("tools/copy_scp_data_to_dir.sh " + filelist + " " + location + " --bwlimit 10000")
        self.list = _read_scp(filelist, location)
        self.collator = Collator("combo")

    def __len__(self):
        return len(self.list)

    def __getitem__(self, idx):
        feat = np.load(self.list[idx])
        mix = feat["mix"].transpose()
        sample = {"combo": _combo(mix)}
        if len(feat.files) == 1:
            sample["source1"] = mix
        else:
            for s in range(len(feat.files) - 1):
                sample["source%d" % (s + 1)] = feat["s%d" % (s + 1)].transpose()
        return sample


class TestSet(Dataset):
    """feats_test.scp + utt2num_spk -> {'combo', 'name', 'num_spk'} (archs/RSH.py:122-138)."""

    def __init__(self, datadir):
        self.list = _read_scp(datadir + "/feats_test.scp")
        self.num_spks = [int(line.rstrip("\n").split(" ")[1]) for line in open(datadir + "/utt2num_spk")]
        self.collator = Collator("combo")

    def __len__(self):
        return len(self.list)

    def __getitem__(self, idx):
        mix = np.abs(np.load(self.list[idx])["mix"]).transpose()
        return {"combo": _combo(mix), "name": os.path.basename(self.list[idx]), "num_spk": self.num_spks[idx]}


class SepDNN(uPIT.SepDNN):
    """LSTM(2F -> 600, 2 layers, bidirectional) -> BatchNorm1d(1200) -> Linear(1200 -> F) -> sigmoid
    (archs/RSH.py:141-187); parameter names and construction order as the reference."""

    def __init__(self, gpuid, **kwargs):
        nn.Module.__init__(self)
        self.gpuid = gpuid
        self.feat_dim = int(kwargs.get("feat_dim", 257))
        self.num_spk = 1                                  # one mask per pass
        self.num_layers = int(kwargs.get("num_layers", 2))    # reference: 2 (archs/RSH.py:155)
        for key in kwargs:
            print("modelparam:", key, kwargs[key])
        self.blstm = nn.LSTM(self.feat_dim * 2, HID, num_layers=self.num_layers, bidirectional=True)
        self.lin = nn.Linear(HID * 2, self.feat_dim)
        self.bn = nn.BatchNorm1d(HID * 2)
        self.in_dim, self.out_dim = 2 * self.feat_dim, self.feat_dim
        self._wt = None
        self._wt_versions = None
        self._ws = {}
        self.last = None

    def forward(self, x):
        """x: PackedSequence of [mix || attention]; returns the mask (B, T, F) and carries self.hidden forward like
        the reference's forward (archs/RSH.py:169-187)."""
        dev = self.lin.weight.device
        batch = unpack_batch({"combo": x}, ["combo"], dev)
        run_forward(self, batch, self.hidden, save=False, x=batch["combo"])
        ws = batch["ws"]
        self.hidden = (ws.hn.clone(), ws.cn.clone())
        mask = torch.empty(ws.TB, self.out_dim, dtype=torch.float32, device=dev)
        call("sigmoid", ws.z, LL(ws.ldz), ws.TB, self.out_dim, mask, stream_ptr())
        return mask.view(ws.T, ws.B, self.out_dim).transpose(0, 1)


def _unpack_sub(model, sub, n_spk, dev):
    batch = unpack_batch(sub, ["combo"], dev, ["source%d" % (i + 1) for i in range(n_spk)])
    F = model.feat_dim
    batch["mix"] = batch["combo"][:, :F].contiguous()          # the mixture half never changes (relu(mix - 0) = mix)
    return batch


def _forward_passes(model, batch_sample, save, hidden_fn=None):
    """All sub-batches and passes of archs/RSH.py:205-257.  Returns (subs, stats, last_n): stats = the device accumulator
    {sum of min losses / S, norm, loss, 1/(S_last * norm)}, subs = what the backward pass needs."""
    dev = model.lin.weight.device
    F = model.feat_dim
    stats = torch.zeros(4, dtype=torch.float32, device=dev)
    subs, first, last_n, hidden = [], True, 1, None
    for n in range(batch_sample.max_spk):
        B = batch_sample.sub_batch_lens[n]
        if B == 0:
            continue
        hidden = model.init_hidden(B) if hidden_fn is None else hidden_fn(B)   # drawn even when n == 0, like the reference (H1)
        if n == 0:
            continue
        batch = _unpack_sub(model, batch_sample[n], n, dev)
        used = torch.zeros(B, dtype=torch.int32, device=dev)
        x, passes = batch["combo"], []
        for k in range(n):
            run_forward(model, batch, hidden, save, slot=(n, k), x=x)
            ws = batch["ws"]
            engine.pit_loss(ws, batch["mix"], batch["src"], batch["lens_dev"], F, 1, n, mode=1,
                            accumulate=not first, used=used, stats=stats)
            first = False
            nxt = torch.empty_like(x)
            call("rsh_residual", x, ws.z, LL(ws.ldz), batch["lens_dev"], ws.T, ws.B, F, 1, nxt, stream_ptr())
            passes.append({"ws": ws, "x": x, "h0": hidden[0].contiguous(), "c0": hidden[1].contiguous(),
                           "losses": ws.losses[:n], "idx": ws.idx, "T": ws.T, "B": ws.B})
            hidden = (ws.hn, ws.cn)                       # state carried into the next pass (H5)
            x = nxt
        subs.append({"n": n, "batch": batch, "passes": passes})
        last_n = n
    model.hidden = hidden if subs else None
    model.last = {"subs": subs, "stats": stats}
    return subs, stats, last_n


def _backward_passes(model, subs, stats, last_n, grads, gloss, hook=None):
    """Backward through every pass in reverse, accumulating the weight gradients into `grads` (written, not added to, by
    the first pass processed).  gloss: None (= 1) or a 1-element device tensor; hook: bucket callback of the last pass."""
    F = model.feat_dim
    p = {k: v.detach() for k, v in model.param_dict().items()}
    acc = False
    todo = sum(s["n"] for s in subs)
    for sub in reversed(subs):
        n, batch = sub["n"], sub["batch"]
        wt = batch["wt"]
        # stats[3] = 1/(S_last * norm): rescale to this sub-batch's 1/(n * norm)
        scale = float(last_n) / n
        if gloss is None:
            gout = None if scale == 1.0 else torch.full((1,), scale, dtype=torch.float32, device=stats.device)
        else:
            gout = (gloss.detach().to(torch.float32) * scale).reshape(1).contiguous()
        carry = dmask = dhn = dcn = None
        for k in reversed(range(n)):
            ps = sub["passes"][k]
            more = k > 0
            todo -= 1
            ps["ws"].bind(ps["T"], ps["B"])
            dx, dh0, dc0 = engine.backward(ps["ws"], wt, p, grads, batch["mix"], batch["src"], batch["lens_dev"],
                                           ps["h0"], ps["c0"], F, 1, gout, dmask_extra=dmask, dhn=dhn, dcn=dcn,
                                           want_dx=more, want_d0=more, training=model.training, accumulate=acc,
                                           stats=stats, hook=hook if todo == 0 else None)
            acc = True
            if more:
                TB = ps["ws"].TB
                new_carry = torch.empty(TB, F, dtype=torch.float32, device=dx.device)
                dmask = torch.empty(TB, F, dtype=torch.float32, device=dx.device)
                call("rsh_residual_bwd", dx, carry, ps["x"], LL(TB), F, new_carry, dmask, stream_ptr())
                carry, dhn, dcn = new_carry, dh0, dc0


class _RSHLoss(torch.autograd.Function):
    """loss over all sub-batches and passes; backward = the hand-written kernels, passes in reverse."""

    @staticmethod
    def forward(ctx, model, batch_sample, save, *params):
        subs, stats, last_n = _forward_passes(model, batch_sample, save)
        ctx.model, ctx.subs, ctx.saved, ctx.stats, ctx.last_n = model, subs, save, stats, last_n
        ctx.names = [k for k, _ in model.named_parameters()]
        loss, norm = stats[2].clone(), stats[1].clone()
        ctx.mark_non_differentiable(norm)
        return loss, norm

    @staticmethod
    def backward(ctx, gloss, gnorm):
        if not ctx.saved:
            raise RuntimeError("RSH loss was computed without saving activations (no_grad / eval)")
        model = ctx.model
        p = model.param_dict()
        grads = {k: torch.zeros_like(p[k]) for k in ctx.names}
        _backward_passes(model, ctx.subs, ctx.stats, ctx.last_n, grads, gloss)
        return (None, None, None) + tuple(grads[k] for k in ctx.names)


def compute_loss(model, epoch, batch_sample, plotdir=""):
    """(loss, norm) as archs/RSH.py:197-259; loss.backward() runs the CUDA backward through all passes."""
    model.zero_grad()
    params = [v for _, v in model.named_parameters()]
    save = torch.is_grad_enabled() and any(q.requires_grad for q in params)
    loss, norm = _RSHLoss.apply(model, batch_sample, save, *params)
    return loss, norm.detach()


def loss_and_grads(model, batch_sample, grads, hook=None, hidden=None):
    """compute_loss + loss.backward() without the autograd tape (same contract as ssb200.uPIT.loss_and_grads): every
    parameter gradient is written straight into grads[name] (e.g. views of ssb200.train.Trainer's flat bucket);
    `hook(tag)` fires for each gradient bucket once the LAST pass has enqueued it.  Returns device scalars (loss, norm)."""
    with torch.no_grad():
        subs, stats, last_n = _forward_passes(model, batch_sample, True, hidden_fn=hidden)
        if not subs:
            raise ValueError("RSH.loss_and_grads: the batch holds no utterance with at least one speaker")
        _backward_passes(model, subs, stats, last_n, grads, None, hook=hook)
    return stats[2], stats[1]


def compute_cv_loss(model, epoch, batch_sample, plotdir=""):
    return compute_loss(model, epoch, batch_sample, plotdir)


def compute_masks(model, batch_sample, out_dir):
    """Writes out_dir/<name>.npz with keys s1..sN (one mask per pass), each (F, T_b) float32
    (archs/RSH.py:262-287; no relu on the residual at test time, :280)."""
    model.zero_grad()
    dev = model.lin.weight.device
    F = model.feat_dim
    for n in range(batch_sample.max_spk):
        B = batch_sample.sub_batch_lens[n]
        if B == 0:
            continue
        sub = batch_sample[n]
        hidden = model.init_hidden(B)
        if n == 0:
            continue
        names = sub["name"]
        batch = unpack_batch(sub, ["combo"], dev)
        x = batch["combo"]
        dicts = [dict() for _ in range(B)]
        with torch.no_grad():
            for k in range(n):
                run_forward(model, batch, hidden, save=False, slot=(n, k % 2), x=x)
                ws = batch["ws"]
                mask = torch.empty(ws.TB, F, dtype=torch.float32, device=dev)
                call("sigmoid", ws.z, LL(ws.ldz), ws.TB, F, mask, stream_ptr())
                nxt = torch.empty_like(x)
                call("rsh_residual", x, ws.z, LL(ws.ldz), batch["lens_dev"], ws.T, ws.B, F, 0, nxt, stream_ptr())
                hidden, x = (ws.hn, ws.cn), nxt
                m = mask.view(ws.T, ws.B, F).cpu().numpy()
                for i in range(B):
                    dicts[i]["s%d" % (k + 1)] = m[:batch["lens"][i], i, :].transpose()
        for i in range(B):
            np.savez_compressed(out_dir + "/" + names[i], **dicts[i])
