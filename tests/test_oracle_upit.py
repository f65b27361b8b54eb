"""Pin oracle/upit_oracle.py to the outputs of the unmodified reference (tests/golden/ref_*.npz)."""
import numpy as np
import pytest
import torch

import common
from oracle import upit_oracle as uo

UPIT = ["upit_s2_f129", "upit_s3_f129", "upit_s2_f257"]
RSH = ["rsh_s3_f129", "rsh_s2_f257"]


def _check_weights(z, sd):
    for k, v in sd.items():
        if not v.is_floating_point():
            continue
        ref = z["w/" + k]
        assert abs(float(v.norm()) - ref[0]) <= 1e-6 * max(1.0, ref[0]), "weight regeneration differs: " + k


def _check_grads(z, named_grads, tol):
    for k, g in named_grads.items():
        stat, samp = common.grad_stat(k, g)
        ref_stat, ref_samp = z["g/" + k + "/stat"], z["g/" + k + "/sample"]
        assert abs(stat[0] - ref_stat[0]) <= tol * ref_stat[0], (k, stat, ref_stat)
        assert abs(stat[1] - ref_stat[1]) <= 10 * tol * ref_stat[0] * np.sqrt(g.numel()), (k, stat, ref_stat)
        assert common.rel(samp, ref_samp) <= 10 * tol, k


@pytest.mark.parametrize("name", UPIT)
def test_upit_oracle_matches_reference(name):
    z, case = common.load_golden(name)
    mixes, srcs, _ = common.case_features(case)
    mix, lens = common.pad_time_major(mixes)
    sources = [common.pad_time_major(s)[0] for s in srcs]
    sd = common.case_state_dict(case)
    _check_weights(z, sd)
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd.update(params)
    h0, c0 = common.case_hidden(case)
    loss, norm, losses, indices, mask, stats = uo.upit_compute_loss(sd, mix, sources, lens, h0, c0)
    loss.backward()
    assert float(norm) == float(z["norm"])
    assert abs(float(loss) - float(z["loss"])) <= 2e-6 * float(z["loss"])
    assert np.array_equal(indices.numpy(), z["indices"][0])
    assert common.rel(losses.detach().numpy(), z["losses"][0]) <= 1e-5
    # mask: reference is batch-first (B,T,FS); ours time-major
    mk = mask.detach().permute(1, 0, 2).contiguous().numpy().ravel()
    idx = np.linspace(0, mk.size - 1, num=2048).astype(np.int64)
    assert np.abs(mk[idx] - z["mask0/sample"]).max() <= 2e-5
    _check_grads(z, {k: p.grad for k, p in params.items()}, 2e-4)
    n = mix.shape[0] * mix.shape[1]
    rm, rv = uo.bn_running_update(torch.zeros(1200, dtype=torch.float64), torch.ones(1200, dtype=torch.float64),
                                  stats[0].detach(), stats[1].detach(), n)
    assert np.abs(rm.numpy()[::8] - z["bn.running_mean/sample"]).max() <= 1e-6
    assert np.abs(rv.numpy()[::8] - z["bn.running_var/sample"]).max() <= 1e-6
    # eval-mode (compute_cv_loss) with the updated running stats
    sd_eval = {k: v.detach() for k, v in sd.items()}
    sd_eval["bn.running_mean"], sd_eval["bn.running_var"] = rm, rv
    h0, c0 = common.case_hidden(case, 200)
    with torch.no_grad():
        cv = uo.upit_compute_loss(sd_eval, mix, sources, lens, h0, c0, training=False)[0]
    assert abs(float(cv) - float(z["cv_loss"])) <= 5e-6 * float(z["cv_loss"])


@pytest.mark.parametrize("name", RSH)
def test_rsh_oracle_matches_reference(name):
    z, case = common.load_golden(name)
    mixes, srcs, _ = common.case_features(case)
    mix, lens = common.pad_time_major(mixes)
    sources = [common.pad_time_major(s)[0] for s in srcs]
    sd = common.case_state_dict(case)
    _check_weights(z, sd)
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd.update(params)
    h0, c0 = common.case_hidden(case)
    loss, norm, all_losses, all_idx, masks = uo.rsh_compute_loss(sd, mix, sources, lens, h0, c0)
    loss = loss / norm
    loss.backward()
    assert float(norm) == float(z["norm"])
    assert abs(float(loss) - float(z["loss"])) <= 2e-6 * float(z["loss"])
    for p in range(case["S"]):
        assert np.array_equal(all_idx[p].numpy(), z["indices"][p]), p
        a, b = all_losses[p].detach().numpy(), z["losses"][p]
        fin = np.isfinite(b)
        assert np.array_equal(np.isfinite(a), fin)
        assert common.rel(a[fin], b[fin]) <= 1e-5
        mk = masks[p].detach().permute(1, 0, 2).contiguous().numpy().ravel()
        idx = np.linspace(0, mk.size - 1, num=2048).astype(np.int64)
        assert np.abs(mk[idx] - z["mask%d/sample" % p]).max() <= 2e-5
    _check_grads(z, {k: p.grad for k, p in params.items()}, 2e-4)
