"""Drop-in arch module for the reference's `archs/uPIT.py` (utterance-level PIT BLSTM mask estimator).

Exposes exactly the plugin surface the reference trainers touch (steps/train_qsub.py:66,80-95,117-133,
steps/eval_qsub.py:43-72): TrainSet / TestSet (+ .collator), SepDNN, compute_loss, compute_cv_loss, compute_masks.
state_dict keys and parameter initialisation are those of the reference (archs/uPIT.py:115-119) so `.mdl` files
interchange.  All device arithmetic goes through libssb200 (hand-written sm_100a kernels); there is no torch/cuDNN
or CPU fallback: without the CUDA library these functions raise.
"""
import os

import numpy as np
import torch
import torch.nn as nn
from torch.nn.utils.rnn import PackedSequence, pack_sequence
from torch.utils.data import Dataset

from . import engine
from ._lib import LL, call, stream_ptr

HID = engine.H


class Collator:
    """Sorts a batch by descending length of `sort_key` and packs every array (archs/uPIT.py:23-48)."""

    def __init__(self, sort_key):
        self.key = sort_key

    def __call__(self, batch):
        order = np.argsort(np.array([len(d[self.key]) for d in batch]))[::-1]
        out = {}
        for k in batch[0]:
            vals = [batch[i][k] for i in order]
            if isinstance(vals[0], np.ndarray):
                out[k] = pack_sequence([torch.from_numpy(np.ascontiguousarray(v)).float() for v in vals])
            else:
                out[k] = vals
        return out


def _read_scp(path, location=""):
    names = [line.rstrip("\n").split(" ")[1] for line in open(path)]
    return [location + "/" + n for n in names] if location else names


class TrainSet(Dataset):
    """feats_train.scp -> {'mix','source1'..} arrays (T,F) (archs/uPIT.py:51-79)."""

    def __init__(self, datadir, location=""):
        filelist = datadir + "/feats_train.scp"
        if location:
            os.system("tools/copy_scp_data_to_dir.sh " + filelist + " " + location + " --bwlimit 10000")
        self.list = _read_scp(filelist, location)
        self.collator = Collator("mix")

    def __len__(self):
        return len(self.list)

    def __getitem__(self, idx):
        feat = np.load(self.list[idx])
        mix = feat["mix"].transpose()
        sample = {"mix": mix}
        if len(feat.files) == 1:
            sample["source1"] = mix
        else:
            for s in range(len(feat.files) - 1):
                sample["source%d" % (s + 1)] = feat["s%d" % (s + 1)].transpose()
        return sample


class TestSet(Dataset):
    """feats_test.scp (complex mixture STFT) -> {'mix': |X| (T,F), 'name'} (archs/uPIT.py:81-94)."""

    def __init__(self, datadir):
        self.list = _read_scp(datadir + "/feats_test.scp")
        self.collator = Collator("mix")

    def __len__(self):
        return len(self.list)

    def __getitem__(self, idx):
        return {"mix": np.abs(np.load(self.list[idx])["mix"]).transpose(), "name": os.path.basename(self.list[idx])}


class SepDNN(nn.Module):
    """Parameter container with the reference's module names; the math runs in ssb200.engine."""

    def __init__(self, gpuid, **kwargs):
        super().__init__()
        self.gpuid = gpuid
        self.feat_dim = int(kwargs.get("feat_dim", 257))
        self.num_spk = int(kwargs.get("num_spk", 2))
        # the reference hard-codes two layers (archs/uPIT.py:115); `num_layers` is an extension (BASELINE.json
        # configs[0] names a 3-layer net) that keeps the state_dict naming of nn.LSTM (weight_ih_l2, ...)
        self.num_layers = int(kwargs.get("num_layers", 2))
        for key in kwargs:
            print("modelparam:", key, kwargs[key])
        self.blstm = nn.LSTM(self.feat_dim, HID, num_layers=self.num_layers, bidirectional=True)
        self.lin = nn.Linear(HID * 2, self.feat_dim * self.num_spk)
        self.bn = nn.BatchNorm1d(HID * 2)
        self.in_dim = self.feat_dim
        self.out_dim = self.feat_dim * self.num_spk
        self._wt = None
        self._wt_versions = None
        self._ws = {}
        self.last = None          # buffers of the most recent step (losses, indices, ...) for inspection

    def init_hidden(self, batch_size):
        # drawn on the CPU generator and moved, exactly like archs/uPIT.py:121-127 (H1: random on every batch)
        h = torch.randn(2 * self.num_layers, batch_size, HID)
        c = torch.randn(2 * self.num_layers, batch_size, HID)
        dev = self.lin.weight.device
        if dev.type != "cuda":
            return h.to(dev), c.to(dev)
        # through pinned memory: a pageable copy would wait for everything already queued on the stream
        return h.pin_memory().to(dev, non_blocking=True), c.pin_memory().to(dev, non_blocking=True)

    # ---- engine plumbing ----
    def param_dict(self):
        p = {k: v for k, v in self.named_parameters()}
        p["bn.running_mean"] = self.bn.running_mean
        p["bn.running_var"] = self.bn.running_var
        return p

    def weights(self, p, need_grad_copies=True):
        dev = p["lin.weight"].device
        if self._wt is None or self._wt.lin_folded.buf.device != dev:
            self._wt = engine.Weights(self.in_dim, self.out_dim, self.num_layers, dev)
            self._wt_versions = None
        ver = tuple(v._version for v in p.values()) + (need_grad_copies,)
        if ver != self._wt_versions:
            self._wt.refresh({k: v.detach() for k, v in p.items()}, need_grad_copies)
            self._wt_versions = ver
        return self._wt

    def workspace(self, T, B, training, slot=0):
        """The activation buffers of (training, slot), bound to this batch's (T, B).  Buffers are sized by capacity and
        grow (with headroom) only when a batch does not fit, so a shuffled loader's ever-changing T_max allocates
        nothing in steady state; other slots are never touched (a second submit/collect slot may be in flight)."""
        key = (bool(training), slot)
        ws = self._ws.get(key)
        if ws is None or not ws.fits(T, B):
            cap_rows = T * B if ws is None else max(T * B, ws.cap_rows)
            cap_B = B if ws is None else max(B, ws.cap_B)
            if ws is not None:                     # growing: leave headroom so the next longer batch fits too
                cap_rows = cap_rows + cap_rows // 8
                self._ws.pop(key)
                del ws
            ws = engine.Workspace(cap_rows, cap_B, self.in_dim, self.out_dim, self.num_spk, self.num_layers,
                                  self.lin.weight.device, training)
            self._ws[key] = ws
        return ws.bind(T, B)

    def reserve(self, max_frames, batch_size, training=True, slot=0):
        """Size the workspace of (training, slot) once for the longest batch expected (frames x batch)."""
        self.workspace(int(max_frames), int(batch_size), training, slot)

    def forward(self, x):
        """x: PackedSequence of mixture features.  Returns the masks (B, T, feat_dim*num_spk) like the reference
        (archs/uPIT.py:129-147); uses self.hidden as initial state."""
        batch = unpack_batch({"mix": x}, ["mix"], self.lin.weight.device)
        z = run_forward(self, batch, self.hidden, save=False)
        ws = batch["ws"]
        mask = torch.empty(ws.TB, self.out_dim, dtype=torch.float32, device=z.device)
        call("sigmoid", ws.z, LL(ws.ldz), ws.TB, self.out_dim, mask, stream_ptr())
        return mask.view(ws.T, ws.B, self.out_dim).transpose(0, 1)


def _to_device_async(arr, dtype, device):
    """Small host array -> device without a blocking pageable copy (which would stall the launch queue)."""
    t = torch.from_numpy(np.ascontiguousarray(arr, dtype=dtype))
    return t.pin_memory().to(device, non_blocking=True)


def unpack_batch(batch_sample, keys, device, src_keys=()):
    """PackedSequences -> time-major padded device tensors (rows t*B+b) + lengths.  `src_keys` are unpacked side by
    side into one (len(src_keys), T*B, F) tensor returned under 'src'."""
    if isinstance(batch_sample, dict) and batch_sample.get("device_batch"):
        return batch_sample               # already padded on the device (ssb200/pipeline.py: no PackedSequence hop)
    mix = batch_sample[keys[0]]
    assert isinstance(mix, PackedSequence)
    bs = mix.batch_sizes.numpy().astype(np.int64)
    T, B = len(bs), int(bs[0])
    lens = (bs[None, :] > np.arange(B)[:, None]).sum(1).astype(np.int32)
    toff = _to_device_async(np.concatenate([[0], np.cumsum(bs)]), np.int64, device)
    out = {"T": T, "B": B, "lens": lens, "lens_dev": _to_device_async(lens, np.int32, device), "toff": toff}

    def unpack_into(k, dst):
        data = batch_sample[k].data.to(device=device, dtype=torch.float32, non_blocking=True).contiguous()
        C = data.shape[1]
        call("unpack", data, toff, T, B, C, C, 0, dst, stream_ptr())
        return C

    for k in keys:
        C = batch_sample[k].data.shape[1]
        pad = torch.empty(T * B, C, dtype=torch.float32, device=device)
        unpack_into(k, pad)
        out[k] = pad
    if src_keys:
        C = batch_sample[src_keys[0]].data.shape[1]
        src = torch.empty(len(src_keys), T * B, C, dtype=torch.float32, device=device)
        for i, k in enumerate(src_keys):
            unpack_into(k, src[i])
        out["src"] = src
    return out


def run_forward(model, batch, hidden, save, slot=0, x=None, dp=None):
    p = model.param_dict()
    wt = model.weights(p, need_grad_copies=save)
    ws = model.workspace(batch["T"], batch["B"], save, slot)
    batch["ws"], batch["wt"] = ws, wt
    h0, c0 = hidden
    pd = {k: v.detach() for k, v in p.items()}
    return engine.forward(ws, wt, pd, batch["mix"] if x is None else x, batch["lens_dev"], h0.contiguous(), c0.contiguous(),
                          bn_train=model.training, save=save, dp=dp)


class _UPITLoss(torch.autograd.Function):
    """loss = uPIT(model(mix), sources); backward = the hand-written backward kernels."""

    @staticmethod
    def forward(ctx, model, batch, hidden, save, *params):
        run_forward(model, batch, hidden, save)
        ws = batch["ws"]
        S, F = model.num_spk, model.feat_dim
        engine.pit_loss(ws, batch["mix"], batch["src"], batch["lens_dev"], F, S, S, mode=0)
        ctx.model, ctx.batch, ctx.hidden, ctx.saved = model, batch, hidden, save
        ctx.names = [k for k, _ in model.named_parameters()]
        loss, norm = ws.stats[2].clone(), ws.stats[1].clone()
        ctx.mark_non_differentiable(norm)
        return loss, norm

    @staticmethod
    def backward(ctx, gloss, gnorm):
        if not ctx.saved:
            raise RuntimeError("uPIT loss was computed without saving activations (no_grad / eval)")
        model, batch = ctx.model, ctx.batch
        ws, wt = batch["ws"], batch["wt"]
        p = {k: v.detach() for k, v in model.param_dict().items()}
        grads = {k: torch.empty_like(p[k]) for k in ctx.names}
        gout = gloss.detach().to(torch.float32).reshape(1).contiguous()
        engine.backward(ws, wt, p, grads, batch["mix"], batch["src"], batch["lens_dev"], ctx.hidden[0].contiguous(),
                        ctx.hidden[1].contiguous(), model.feat_dim, model.num_spk, gout, training=model.training)
        return (None, None, None, None) + tuple(grads[k] for k in ctx.names)


def compute_loss(model, epoch, batch_sample, plotdir=""):
    """(loss, norm) as archs/uPIT.py:157-206; loss.backward() runs the CUDA backward."""
    dev = model.lin.weight.device
    batch = unpack_batch(batch_sample, ["mix"], dev, ["source%d" % (i + 1) for i in range(model.num_spk)])
    model.zero_grad()
    model.hidden = model.init_hidden(batch["B"])
    params = [v for _, v in model.named_parameters()]
    save = torch.is_grad_enabled() and any(q.requires_grad for q in params)
    loss, norm = _UPITLoss.apply(model, batch, model.hidden, save, *params)
    ws = batch["ws"]
    P = {1: 1, 2: 2, 3: 6, 4: 24}[model.num_spk]
    model.last = {"losses": ws.losses[:P], "indices": ws.idx, "stats": ws.stats, "batch": batch}
    return loss, norm.detach()


def loss_and_grads(model, batch_sample, grads, hook=None, hidden=None, dp=None, accumulate=False):
    """compute_loss + loss.backward() without the autograd tape: every parameter gradient is written (not
    accumulated) straight into grads[name] -- e.g. views of the trainer's flat bucket (ssb200/train.py).  `hook(tag)`
    is called when a gradient bucket ('head', 'l1', 'l0') has been enqueued; `hidden(B)` (optional) supplies (h0, c0)
    instead of model.init_hidden; `dp` (ssb200.train.ExactDP, data-parallel mode B) makes the loss normaliser and the
    BatchNorm statistics those of all ranks' utterances together, so that the SUM of the ranks' gradients is the
    reference's gradient on the concatenated batch; `accumulate` adds to grads instead of writing them (gradient
    accumulation over micro-batches).  Returns device scalars (loss, norm)."""
    dev = model.lin.weight.device
    batch = unpack_batch(batch_sample, ["mix"], dev, ["source%d" % (i + 1) for i in range(model.num_spk)])
    model.hidden = model.init_hidden(batch["B"]) if hidden is None else hidden(batch["B"])
    with torch.no_grad():
        run_forward(model, batch, model.hidden, save=True, dp=dp)
        ws, wt = batch["ws"], batch["wt"]
        S, F = model.num_spk, model.feat_dim
        engine.pit_loss(ws, batch["mix"], batch["src"], batch["lens_dev"], F, S, S, mode=0, dp=dp)
        p = {k: v.detach() for k, v in model.param_dict().items()}
        engine.backward(ws, wt, p, grads, batch["mix"], batch["src"], batch["lens_dev"], model.hidden[0].contiguous(),
                        model.hidden[1].contiguous(), F, S, None, training=model.training, hook=hook, dp=dp,
                        accumulate=accumulate)
    model.last = {"losses": ws.losses, "indices": ws.idx, "stats": ws.stats, "batch": batch}
    return ws.stats[2], ws.stats[1]


def compute_cv_loss(model, epoch, batch_sample, plotdir=""):
    return compute_loss(model, epoch, batch_sample, plotdir)


def compute_masks(model, batch_sample, out_dir):
    """Writes out_dir/<name>.npz with keys s1..sS, each (F, T_b) float32 (archs/uPIT.py:209-225)."""
    name = batch_sample["name"]
    dev = model.lin.weight.device
    batch = unpack_batch(batch_sample, ["mix"], dev)
    model.zero_grad()
    model.hidden = model.init_hidden(batch["B"])
    with torch.no_grad():
        run_forward(model, batch, model.hidden, save=False)
        ws = batch["ws"]
        mask = torch.empty(ws.TB, model.out_dim, dtype=torch.float32, device=dev)
        call("sigmoid", ws.z, LL(ws.ldz), ws.TB, model.out_dim, mask, stream_ptr())
    mask = mask.view(ws.T, ws.B, model.out_dim).cpu().numpy()
    F = model.feat_dim
    for i in range(len(name)):
        m = mask[:batch["lens"][i], i, :].transpose()
        np.savez_compressed(out_dir + "/" + name[i], **{"s%d" % (s + 1): m[s * F:(s + 1) * F] for s in range(model.num_spk)})
