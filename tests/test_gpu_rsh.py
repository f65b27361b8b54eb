"""RSH train step through the drop-in arch module (CUDA kernels via the C ABI) vs the golden outputs of the UNMODIFIED
reference archs/RSH.py (tests/golden/ref_rsh_*.npz) and the fp64 oracle.  Tolerances (north_star): loss and every
gradient tensor 1e-3 relative; the greedy source assignment of every pass bit-exact."""
import numpy as np
import pytest
import torch

import common
from oracle import upit_oracle as uo
from ssb200 import RSH

pytestmark = pytest.mark.gpu
TOL = 1e-3


def _batch(case):
    mixes, srcs, _ = common.case_features(case)
    samples = []
    for b in range(len(mixes)):
        d = {"combo": np.concatenate((mixes[b], np.ones(mixes[b].shape)), axis=1)}
        for i in range(case["S"]):
            d["source%d" % (i + 1)] = srcs[i][b]
        samples.append(d)
    return RSH.Collator("combo")(samples), mixes, srcs


def _model(case):
    F = case["n_fft"] // 2 + 1
    torch.manual_seed(case["seed"])
    return RSH.SepDNN(0, feat_dim=str(F)).cuda()


@pytest.mark.parametrize("name", ["rsh_s3_f129", "rsh_s2_f257"])
def test_rsh_step_matches_reference_and_oracle(name):
    z, case = common.load_golden(name)
    S = case["S"]
    batch, mixes, srcs = _batch(case)
    assert batch.max_spk == S + 1 and batch.sub_batch_lens[S] == len(mixes)
    model = _model(case)
    for k, v in model.state_dict().items():
        if v.is_floating_point():
            assert abs(float(v.double().norm()) - z["w/" + k][0]) <= 1e-6 * max(1.0, z["w/" + k][0]), k
    torch.manual_seed(100 + case["seed"])
    loss, norm = RSH.compute_loss(model, 0, batch)
    loss.backward()
    torch.cuda.synchronize()
    assert float(norm) == float(z["norm"])
    assert abs(float(loss.detach()) - float(z["loss"])) <= TOL * float(z["loss"])
    passes = model.last["subs"][0]["passes"]
    for p in range(S):
        assert np.array_equal(passes[p]["idx"].cpu().numpy(), z["indices"][p]), p
        a, b = passes[p]["losses"].cpu().numpy(), z["losses"][p]
        fin = np.isfinite(b)
        assert np.array_equal(np.isfinite(a), fin)
        assert common.rel(a[fin], b[fin]) <= 1e-4
    for k, p in model.named_parameters():
        stat, samp = common.grad_stat(k, p.grad)
        rs, rsamp = z["g/" + k + "/stat"], z["g/" + k + "/sample"]
        assert abs(stat[0] - rs[0]) <= TOL * rs[0], (k, stat, rs)
        assert common.rel(samp, rsamp) <= 5 * TOL, k
    assert np.abs(model.bn.running_mean.cpu().numpy()[::8] - z["bn.running_mean/sample"]).max() <= 1e-5
    assert np.abs(model.bn.running_var.cpu().numpy()[::8] - z["bn.running_var/sample"]).max() <= 1e-5
    # full gradient tensors against the fp64 oracle
    mix, lens = common.pad_time_major(mixes)
    sources = [common.pad_time_major(s)[0] for s in srcs]
    sd = common.case_state_dict(case)
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd.update(params)
    h0, c0 = common.case_hidden(case)
    oloss, onorm, _, oidx, _ = uo.rsh_compute_loss(sd, mix, sources, lens, h0, c0)
    (oloss / onorm).backward()
    worst = 0.0
    for k, p in model.named_parameters():
        r = common.rel(p.grad.cpu().numpy(), params[k].grad.numpy())
        worst = max(worst, r)
        assert r <= TOL, (k, r)
    print(name, "worst grad rel err", worst)
    # eval mode (compute_cv_loss): running stats, no grad
    model.eval()
    torch.manual_seed(200 + case["seed"])
    with torch.no_grad():
        cv_loss, _ = RSH.compute_cv_loss(model, 0, batch)
    assert abs(float(cv_loss) - float(z["cv_loss"])) <= TOL * float(z["cv_loss"])


def test_rsh_masks_mixed_speaker_counts(tmp_path):
    """compute_masks on a test batch mixing 2- and 3-speaker utterances: one mask per pass, no relu on the residual."""
    z, case = common.load_golden("rsh_s3_f129")
    _, mixes, _ = _batch(case)
    model = _model(case)
    model.eval()
    F = case["n_fft"] // 2 + 1
    counts = [3, 2, 3, 2][:len(mixes)]
    samples = [{"combo": np.concatenate((m, np.ones(m.shape)), axis=1), "name": "u%d.npz" % i, "num_spk": counts[i]}
               for i, m in enumerate(mixes)]
    batch = RSH.Collator("combo")(samples)
    assert batch.sub_batch_lens == [0, 0, counts.count(2), counts.count(3)]
    torch.manual_seed(5)
    RSH.compute_masks(model, batch, str(tmp_path))
    # oracle: same hidden draws in the same order (sub-batch 2 first, then 3)
    sd = common.case_state_dict(case)          # (reseeds the generator itself: build it before seeding)
    torch.manual_seed(5)
    for n in (2, 3):
        members = [i for i, c in enumerate(counts) if c == n]
        members.sort(key=lambda i: -mixes[i].shape[0])
        h = torch.randn(4, len(members), 600).double()
        c = torch.randn(4, len(members), 600).double()
        mix, lens = common.pad_time_major([mixes[i] for i in members])
        T, B, _ = mix.shape
        valid = torch.zeros(T, B, 1, dtype=torch.float64)
        for b, L in enumerate(lens):
            valid[:L, b] = 1
        combo = torch.cat([mix, valid.expand(T, B, F)], 2)
        with torch.no_grad():
            for k in range(n):
                mask, (h, c), _ = uo.sepdnn_forward(combo, lens, sd, h, c, training=False)
                combo = (combo - torch.cat([torch.zeros_like(mask), mask], 2)) * valid
                for b, i in enumerate(members):
                    got = np.load(str(tmp_path / ("u%d.npz" % i)))
                    assert got.files == ["s%d" % (j + 1) for j in range(n)]
                    ref = mask[:lens[b], b].numpy().T
                    assert got["s%d" % (k + 1)].shape == ref.shape
                    assert np.abs(got["s%d" % (k + 1)] - ref).max() <= 1e-4, (n, k, i)
