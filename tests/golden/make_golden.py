#!/usr/bin/env python3
"""Generate tests/golden/ref_*.npz by running the UNMODIFIED reference arch modules.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference targets torch 0.4 / python 3.6; four shims are installed *around* it
 -- none of its source is edited or copied:
  (1) sys.modules['plot'] stub (matplotlib absent; archs/uPIT.py:18-19)
  (2) torch.utils.data.dataloader.numpy_type_map (removed symbol; archs/uPIT.py:12)
  (3) collections.Mapping alias (archs/uPIT.py:39)
  (4) identity .cuda() on Tensor / PackedSequence / Module (CPU-only container)

Inputs are regenerated from seeds by the tests (ssb200.synth + oracle STFT +
torch.manual_seed), so the fixtures only hold reference OUTPUTS plus checksums of
the regenerated inputs/weights.
"""
import collections
import collections.abc
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "speech-separation_b200"))

from oracle import stft_oracle  # noqa: E402
from ssb200 import synth  # noqa: E402

CASES = {
    # name: arch, S, n_fft, hop, lengths (samples), seed
    "upit_s2_f257": dict(arch="uPIT", S=2, n_fft=512, hop=128, lens=synth.batch_lengths_cfg1(8, 32000, 384), seed=0),
    "upit_s2_f129": dict(arch="uPIT", S=2, n_fft=256, hop=64, lens=synth.batch_lengths_cfg1(4, 8000, 448), seed=1),
    "upit_s3_f129": dict(arch="uPIT", S=3, n_fft=256, hop=64, lens=synth.batch_lengths_cfg1(5, 6400, 320), seed=2),
    "rsh_s3_f129": dict(arch="RSH", S=3, n_fft=256, hop=64, lens=synth.batch_lengths_cfg1(4, 6400, 320), seed=3),
    "rsh_s2_f257": dict(arch="RSH", S=2, n_fft=512, hop=128, lens=synth.batch_lengths_cfg1(4, 16000, 640), seed=4),
    # the BASELINE.json configurations at their full sizes (SURVEY.md 8 table of shapes)
    # configs[2]: the benched train shape -- B=64, 4-8 s at 8 kHz, 512/128 (T<=501, F=257)
    "cfg3_upit_b64": dict(arch="uPIT", S=2, n_fft=512, hop=128, lens=synth.uniform_lengths(5, 64, 32000, 64000), seed=5),
    # configs[3]: 3-speaker mixtures, B=64, 4 s -- RSH (input 2F=514, 3 passes) and uPIT with 6 permutations (771 outputs)
    "cfg4_rsh_b64": dict(arch="RSH", S=3, n_fft=512, hop=128, lens=synth.uniform_lengths(6, 64, 24000, 32000), seed=6),
    "cfg4_upit_s3_b64": dict(arch="uPIT", S=3, n_fft=512, hop=128, lens=synth.uniform_lengths(7, 64, 24000, 32000), seed=7),
    # configs[4]: 16 kHz, 20 s utterances (T=2501), B=8
    "cfg5_upit_t2501": dict(arch="uPIT", S=2, n_fft=512, hop=128, lens=synth.batch_lengths_cfg1(8, 320000, 3840), seed=8,
                            sr=16000),
}


def install_shims():
    plot = types.ModuleType("plot")
    plot.plot_spec = lambda *a, **k: None
    plot.plot_loss = lambda *a, **k: None
    sys.modules["plot"] = plot
    import torch.utils.data.dataloader as dl
    if not hasattr(dl, "numpy_type_map"):
        dl.numpy_type_map = {}
    if not hasattr(collections, "Mapping"):
        collections.Mapping = collections.abc.Mapping
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, *a, **k: self
        torch.nn.utils.rnn.PackedSequence.cuda = lambda self, *a, **k: self
        torch.nn.Module.cuda = lambda self, *a, **k: self


def make_samples(case):
    """Per-utterance feature dicts exactly as TrainSet.__getitem__ yields them
    (uPIT.py:66-79 / RSH.py:105-120): arrays (T_b, F)."""
    samples = []
    for b, n in enumerate(case["lens"]):
        mix, srcs = synth.mixture(1000 * case["seed"] + b, n, case["S"], case.get("sr", 8000))
        feats = {"mix": stft_oracle.stft_mag(mix, case["n_fft"], case["hop"]).T}
        for i, s in enumerate(srcs):
            feats["source%d" % (i + 1)] = stft_oracle.stft_mag(s, case["n_fft"], case["hop"]).T
        samples.append(feats)
    return samples


def grad_summary(name, g):
    g = g.detach().double().numpy().ravel()
    rng = np.random.default_rng(sum(map(ord, name)))
    r = rng.standard_normal(g.size)
    idx = np.linspace(0, g.size - 1, num=min(256, g.size)).astype(np.int64)
    return np.array([np.linalg.norm(g), float(g @ r)]), g[idx].astype(np.float32)


def run_case(name, case):
    install_shims()
    sys.path.insert(0, "/root/reference/archs")
    m = __import__(case["arch"])
    F = case["n_fft"] // 2 + 1
    samples = make_samples(case)
    if case["arch"] == "RSH":
        for s in samples:
            mixf = s.pop("mix")
            s_new = {"combo": np.concatenate((mixf, np.ones(mixf.shape)), axis=1)}
            s_new.update(s)
            s.clear()
            s.update(s_new)
        collator = m.Collator("combo")
        kwargs = dict(feat_dim=str(F))
    else:
        collator = m.Collator("mix")
        kwargs = dict(feat_dim=str(F), num_spk=str(case["S"]))
    batch = collator(samples)
    torch.manual_seed(case["seed"])
    model = m.SepDNN(-1, **kwargs)
    out = {}
    wsum = {k: np.array([float(v.double().norm()), float(v.double().sum())]) for k, v in model.state_dict().items()}
    for k, v in wsum.items():
        out["w/" + k] = v
    torch.manual_seed(100 + case["seed"])   # fixes init_hidden's randn (H1)
    captured = {}
    orig_forward = model.forward

    def fwd(x):
        y = orig_forward(x)
        captured.setdefault("masks", []).append(y.detach().clone())
        return y
    model.forward = fwd
    captured_min = {}
    orig_min = torch.min

    def spy_min(*a, **k):
        r = orig_min(*a, **k)
        if len(a) == 2 and isinstance(a[1], int) and a[1] == 0 and a[0].dim() == 2:
            captured_min.setdefault("losses", []).append(a[0].detach().clone())
            captured_min.setdefault("indices", []).append(r[1].detach().clone())
        return r
    torch.min = spy_min
    try:
        loss, norm = m.compute_loss(model, 0, batch)
    finally:
        torch.min = orig_min
    loss.backward()
    out["loss"] = np.array(float(loss.detach()), dtype=np.float64)
    out["loss_f32"] = loss.detach().numpy()
    out["norm"] = np.array(float(norm), dtype=np.float64)
    out["losses"] = torch.stack(captured_min["losses"]).numpy()      # (passes, P or S, B)
    out["indices"] = torch.stack(captured_min["indices"]).numpy()    # (passes, B)
    for pi, mk in enumerate(captured["masks"]):
        flat = mk.numpy().ravel()
        idx = np.linspace(0, flat.size - 1, num=2048).astype(np.int64)
        out["mask%d/sample" % pi] = flat[idx]
        out["mask%d/shape" % pi] = np.array(mk.shape)
        out["mask%d/sum" % pi] = np.array(float(mk.double().sum()))
    for k, p in model.named_parameters():
        s, v = grad_summary(k, p.grad)
        out["g/" + k + "/stat"] = s
        out["g/" + k + "/sample"] = v
    out["bn.running_mean/sample"] = model.bn.running_mean.numpy()[::8].copy()
    out["bn.running_var/sample"] = model.bn.running_var.numpy()[::8].copy()
    total = torch.sqrt(sum((p.grad.double() ** 2).sum() for p in model.parameters()))
    out["grad_total_norm"] = np.array(float(total))
    # eval-mode forward (compute_cv_loss semantics: BN running stats, no grad)
    model.eval()
    torch.manual_seed(200 + case["seed"])
    with torch.no_grad():
        cv_loss, cv_norm = m.compute_cv_loss(model, 0, batch)
    out["cv_loss"] = np.array(float(cv_loss), dtype=np.float64)
    out["case"] = np.array(repr(case))
    np.savez_compressed(os.path.join(HERE, "ref_%s.npz" % name), **out)
    print(name, "loss", float(loss), "norm", float(norm), "cv_loss", float(cv_loss),
          "indices", out["indices"].tolist(), "gradnorm", float(total))
    sys.path.pop(0)
    for mod in ("uPIT", "RSH"):
        sys.modules.pop(mod, None)


def main():
    names = sys.argv[1:] or list(CASES)
    for n in names:
        run_case(n, CASES[n])


if __name__ == "__main__":
    main()
