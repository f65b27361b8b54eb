"""The drop-in trainer (ssb200/train_qsub.py) driven the way run_train.sh drives steps/train_qsub.py: feature files on
disk, DataLoader with a worker process, epochs, CV pass every 5th epoch, loss files, checkpoints, resume -- and `.mdl`
interchange with modules built from plain torch.nn under the reference's names (archs/uPIT.py:115-119)."""
import os

import numpy as np
import pytest
import torch

from oracle import cpu_step, stft_oracle
from ssb200 import synth, train_qsub, uPIT

pytestmark = pytest.mark.gpu
N_FFT, HOP, F, S = 256, 64, 129, 2


def _make_data(root, n_utts, seed):
    feat = root / "feats"
    feat.mkdir(parents=True)
    lines = []
    for i in range(n_utts):
        mix, srcs = synth.mixture(seed + i, 4000 + 333 * i, S)
        arrs = {"mix": stft_oracle.stft_mag(mix, N_FFT, HOP)}
        for k, s in enumerate(srcs):
            arrs["s%d" % (k + 1)] = stft_oracle.stft_mag(s, N_FFT, HOP)
        path = feat / ("utt%02d.npz" % i)
        np.savez_compressed(str(path), **arrs)                 # extract_feats.py:90 layout: (F, T) float32 per key
        lines.append("utt%02d %s" % (i, path))
    (root / "feats_train.scp").write_text("\n".join(lines) + "\n")
    return str(root)


def test_train_qsub_epochs_checkpoints_resume_and_mdl_interchange(tmp_path):
    data = _make_data(tmp_path / "tr", 10, 500)
    cv = _make_data(tmp_path / "cv", 4, 600)
    out = tmp_path / "exp"
    (out / "intermediate_models").mkdir(parents=True)
    (out / "train_stats" / "plots").mkdir(parents=True)
    cfg = tmp_path / "model.conf"
    cfg.write_text("feat_dim=%d\nnum_spk=%d\n" % (F, S))
    common = ["uPIT", "0", data, str(out), "--cv-data-dir", cv, "--model-config", str(cfg), "--batch-size", "4"]
    torch.manual_seed(0)
    train_qsub.main(common + ["--num-epochs", "5"])
    im = out / "intermediate_models"
    assert (im / "init.mdl").exists() and (im / "005.mdl").exists() and (out / "final.mdl").exists()
    tl = [l.split() for l in (out / "train_stats" / "train_loss.txt").read_text().splitlines()]
    cl = [l.split() for l in (out / "train_stats" / "cv_loss.txt").read_text().splitlines()]
    assert [l[0] for l in tl] == ["001", "002", "003", "004", "005"] and [l[0] for l in cl] == ["005"]
    losses = [float(l[1]) for l in tl]
    assert all(np.isfinite(losses)) and losses[-1] < losses[0]                 # it learns
    # resume: one more epoch from 005.mdl appends to the same files
    train_qsub.main(common + ["--start-epoch", "5", "--num-epochs", "6"])
    tl = [l.split()[0] for l in (out / "train_stats" / "train_loss.txt").read_text().splitlines()]
    assert tl == ["001", "002", "003", "004", "005", "006"]
    # ---- .mdl interchange: ours -> plain torch.nn modules with the reference's names (strict) ...
    sd = torch.load(str(im / "005.mdl"), map_location="cpu")
    net = cpu_step.CpuNet(F, S)
    net.load_state_dict(sd, strict=True)
    # ... and a state_dict saved from such modules -> ours, giving the same masks
    torch.manual_seed(3)
    net2 = cpu_step.CpuNet(F, S)
    net2.bn.running_mean.normal_()
    net2.bn.running_var.uniform_(0.5, 2.0)
    path = tmp_path / "ref_made.mdl"
    torch.save(net2.state_dict(), str(path))
    model = uPIT.SepDNN(0, feat_dim=str(F), num_spk=str(S)).cuda()
    model.load_state_dict(torch.load(str(path), map_location="cpu"), strict=True)
    model.eval()
    net2.eval()
    ds = uPIT.TrainSet(data)
    batch = ds.collator([ds[i] for i in range(3)])
    h0, c0 = torch.randn(4, 3, 600), torch.randn(4, 3, 600)
    model.hidden = (h0.cuda(), c0.cuda())
    with torch.no_grad():
        got = model(batch["mix"]).cpu()                                          # (B, T, F*S)
        from torch.nn.utils.rnn import pad_packed_sequence
        mix, lens = pad_packed_sequence(batch["mix"])                            # (T, B, F)
        want = net2(mix, lens.tolist(), h0, c0).permute(1, 0, 2)
    for b, L in enumerate(lens.tolist()):
        assert float((got[b, :L] - want[b, :L]).abs().max()) <= 1e-4, b
