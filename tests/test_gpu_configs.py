"""The five BASELINE.json configurations at their FULL sizes, against golden outputs of the UNMODIFIED reference
(tests/golden/ref_cfg*.npz, produced in the build container by tests/golden/make_golden.py from /root/reference/archs):

  configs[0]  uPIT fwd + PIT loss (+ backward), 2-spk, 8 kHz, 4 s, batch 8                    -> ref_upit_s2_f257
  configs[2]  uPIT training shape: batch 64, 4-8 s, 512/128 (T <= 501, F = 257) -- the BENCHED shape -> ref_cfg3_upit_b64
  configs[3]  3-speaker mixtures, batch 64: RSH (3 passes) and uPIT with 6 permutations       -> ref_cfg4_rsh_b64, ref_cfg4_upit_s3_b64
  configs[4]  16 kHz, 20 s utterances (T = 2501), batch 8: train step                         -> ref_cfg5_upit_t2501
  (configs[1], the STFT corpus, and the masked-iSTFT half of configs[4] live in tests/test_gpu_stft.py.)

Every case runs through BOTH routes a user has -- compute_loss -> loss.backward() (autograd) and the tape-free
loss_and_grads that ssb200.train.Trainer.step uses -- with the GEMM streaming under the recurrences ON (the default) and
OFF, and all four must agree with the reference: loss and every gradient within 1e-3 relative, permutation / greedy
assignment indices bit-exact.  The goldens hold summaries (per-tensor gradient norm, a random projection, 256 samples),
which is what lets the fixtures stay small at these sizes.
"""
import numpy as np
import pytest
import torch

import common
from ssb200 import RSH, engine, uPIT

pytestmark = pytest.mark.gpu
TOL = 1e-3

CASES = {
    "configs0_upit_2spk_b8": "upit_s2_f257",
    "configs2_upit_train_b64_benched_shape": "cfg3_upit_b64",
    "configs3_upit_3spk_6perm_b64": "cfg4_upit_s3_b64",
    "configs3_rsh_3spk_b64": "cfg4_rsh_b64",
    "configs4_upit_16k_20s_t2501": "cfg5_upit_t2501",
}


def _samples(case, arch):
    mixes, srcs, _ = common.case_features(case)
    samples = []
    for b in range(len(mixes)):
        if arch is RSH:
            d = {"combo": np.concatenate((mixes[b], np.ones(mixes[b].shape)), axis=1)}
        else:
            d = {"mix": mixes[b]}
        for i in range(case["S"]):
            d["source%d" % (i + 1)] = srcs[i][b]
        samples.append(d)
    return samples


def _check_grads(z, named_grads):
    worst = 0.0
    for k, g in named_grads:
        stat, samp = common.grad_stat(k, g)
        rs, rsamp = z["g/" + k + "/stat"], z["g/" + k + "/sample"]
        assert abs(stat[0] - rs[0]) <= TOL * rs[0], (k, "norm", stat, rs)
        # projection on a fixed N(0,1) vector: |(g - g_ref) . r| ~ ||g - g_ref||, allow 4 sigma
        assert abs(stat[1] - rs[1]) <= 4 * TOL * rs[0], (k, "projection", stat, rs)
        r = common.rel(samp, rsamp)
        worst = max(worst, r)
        assert r <= 5 * TOL, (k, "samples", r)
    return worst


def _check_selection(z, case, model, arch):
    if arch is RSH:
        passes = model.last["subs"][0]["passes"]
        for p in range(case["S"]):
            assert np.array_equal(passes[p]["idx"].cpu().numpy(), z["indices"][p]), ("greedy assignment of pass", p)
            a, b = passes[p]["losses"].cpu().numpy(), z["losses"][p]
            fin = np.isfinite(b)
            assert np.array_equal(np.isfinite(a), fin)
            assert common.rel(a[fin], b[fin]) <= 1e-4
    else:
        P = z["losses"].shape[1]
        assert np.array_equal(model.last["indices"].cpu().numpy(), z["indices"][0]), "permutation indices"
        assert common.rel(model.last["losses"][:P].cpu().numpy(), z["losses"][0]) <= 1e-4


@pytest.mark.parametrize("cfg", list(CASES))
def test_config_matches_unmodified_reference(cfg):
    z, case = common.load_golden(CASES[cfg])
    arch = RSH if case["arch"] == "RSH" else uPIT
    F = case["n_fft"] // 2 + 1
    batch = arch.Collator("combo" if arch is RSH else "mix")(_samples(case, arch))
    kwargs = dict(feat_dim=str(F)) if arch is RSH else dict(feat_dim=str(F), num_spk=str(case["S"]))
    torch.manual_seed(case["seed"])
    model = arch.SepDNN(0, **kwargs).cuda()
    init = {k: v.detach().clone() for k, v in model.state_dict().items()}
    for k, v in init.items():
        if v.is_floating_point():
            assert abs(float(v.double().norm()) - z["w/" + k][0]) <= 1e-6 * max(1.0, z["w/" + k][0]), k
    results = {}
    try:
        for stream in (True, False):
            engine.STREAM = stream
            # ---- route 1: compute_loss -> backward (what steps/train_qsub.py:117-120 calls)
            model.load_state_dict(init)
            torch.manual_seed(100 + case["seed"])
            loss, norm = arch.compute_loss(model, 0, batch)
            loss.backward()
            torch.cuda.synchronize()
            assert float(norm) == float(z["norm"])
            assert abs(float(loss.detach()) - float(z["loss"])) <= TOL * float(z["loss"])
            _check_selection(z, case, model, arch)
            worst = _check_grads(z, [(k, p.grad) for k, p in model.named_parameters()])
            assert np.abs(model.bn.running_mean.cpu().numpy()[::8] - z["bn.running_mean/sample"]).max() <= 1e-5
            assert np.abs(model.bn.running_var.cpu().numpy()[::8] - z["bn.running_var/sample"]).max() <= 1e-5
            results[("autograd", stream)] = (float(loss.detach()), {k: p.grad.clone() for k, p in model.named_parameters()})
            # ---- route 2: the tape-free path of ssb200.train.Trainer.step
            model.load_state_dict(init)
            grads = {k: torch.full_like(p, float("nan")) for k, p in model.named_parameters()}   # must be WRITTEN, not added to
            torch.manual_seed(100 + case["seed"])
            loss2, norm2 = arch.loss_and_grads(model, batch, grads)
            torch.cuda.synchronize()
            assert float(norm2) == float(z["norm"])
            assert abs(float(loss2) - float(z["loss"])) <= TOL * float(z["loss"])
            _check_selection(z, case, model, arch)
            worst = max(worst, _check_grads(z, list(grads.items())))
            results[("tape-free", stream)] = (float(loss2), grads)
            print(cfg, "stream" if stream else "no-stream", "loss", float(loss2), "ref", float(z["loss"]), "worst sample rel", worst)
    finally:
        engine.STREAM = True
    # the four runs did the same arithmetic on the same inputs: streaming only changes WHEN a GEMM chunk runs (and the
    # order in which the two direction halves are added), so they agree far inside the parity tolerance
    l0, g0 = results[("autograd", True)]
    for key, (l, g) in results.items():
        assert abs(l - l0) <= 1e-6 * abs(l0), key
        for k in g0:
            assert common.rel(g[k].cpu().numpy(), g0[k].cpu().numpy()) <= 1e-4, (key, k)
    # eval mode (compute_cv_loss, archs/uPIT.py:149-154): BN running stats after ONE train step, no grad
    model.load_state_dict(init)
    torch.manual_seed(100 + case["seed"])
    loss, _ = arch.compute_loss(model, 0, batch)
    model.eval()
    torch.manual_seed(200 + case["seed"])
    with torch.no_grad():
        cv_loss, _ = arch.compute_cv_loss(model, 0, batch)
    assert abs(float(cv_loss) - float(z["cv_loss"])) <= TOL * float(z["cv_loss"])
