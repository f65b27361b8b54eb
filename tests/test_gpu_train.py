"""Trainer.step (tape-free backward into flat buckets + fused clip/Adam kernels) vs the reference's sequence
compute_loss -> backward -> torch clip_grad_norm_(0.25) -> torch Adam(lr=1e-3).step() (steps/train_qsub.py:117-122)
run through the drop-in module's autograd path, whose gradients test_gpu_upit.py pins to the reference goldens.
Tolerance: parameters after 3 steps within 1e-5 absolute of the torch optimiser's (fp32 arithmetic, different
summation order in the norm)."""
import copy

import numpy as np
import pytest
import torch

import common
from ssb200 import train, uPIT
from test_gpu_upit import _batch, _model

pytestmark = pytest.mark.gpu


def test_trainer_matches_torch_clip_adam():
    z, case = common.load_golden("upit_s2_f129")
    batch, _, _ = _batch(case)
    model_a = _model(case)
    model_b = copy.deepcopy(model_a)
    opt = torch.optim.Adam(model_a.parameters(), lr=1e-3)
    tr = train.Trainer(model_b, uPIT, host_rng=True)     # same CPU-generator (h0, c0) draws on both sides
    for it in range(3):
        torch.manual_seed(40 + it)
        loss_a, norm_a = uPIT.compute_loss(model_a, 0, batch)
        loss_a.backward()
        gn = torch.nn.utils.clip_grad_norm_(model_a.parameters(), 0.25)
        opt.step()
        torch.manual_seed(40 + it)
        loss_b, norm_b = tr.step(batch)
        torch.cuda.synchronize()
        assert abs(float(loss_a) - float(loss_b)) <= 1e-5 * abs(float(loss_a)), it
        assert float(norm_a) == float(norm_b)
        assert abs(float(gn) - float(tr.norm)) <= 1e-5 * float(gn)
        for (k, pa), (_, pb) in zip(model_a.named_parameters(), model_b.named_parameters()):
            d = float((pa.detach() - pb.detach()).abs().max())
            assert d <= 1e-5, (it, k, d)
    for k in ("running_mean", "running_var"):
        assert torch.allclose(getattr(model_a.bn, k), getattr(model_b.bn, k), atol=1e-6)


def test_prefetched_batch_gives_the_same_step():
    """Trainer.prefetch (H2D on a copy stream, overlapping the previous step) must not change the arithmetic."""
    from torch.nn.utils.rnn import PackedSequence
    z, case = common.load_golden("upit_s2_f129")
    batch, _, _ = _batch(case)
    pinned = {k: PackedSequence(v.data.pin_memory(), v.batch_sizes) for k, v in batch.items()}
    model_a = _model(case)
    model_b = copy.deepcopy(model_a)
    tr_a = train.Trainer(model_a, uPIT, host_rng=True)
    tr_b = train.Trainer(model_b, uPIT, host_rng=True)
    staged = tr_b.prefetch(pinned)
    for it in range(3):
        torch.manual_seed(7 + it)
        la, _ = tr_a.step(batch)
        nxt = tr_b.prefetch(pinned)                       # copy of the next batch flies during this step
        torch.manual_seed(7 + it)
        lb, _ = tr_b.step(staged)
        staged = nxt
        assert float(la) == float(lb), it
    for (k, pa), (_, pb) in zip(model_a.named_parameters(), model_b.named_parameters()):
        assert torch.equal(pa.detach(), pb.detach()), k


def test_clip_adam_kernels_unaligned_sizes():
    from ssb200._lib import LL, call, lib, stream_ptr
    import ctypes
    for n in (1, 7, 8192, 8193, 100003):
        g = torch.randn(n, device="cuda") * 3
        p = torch.randn(n, device="cuda")
        ref_p = p.clone().requires_grad_(True)
        ref_p.grad = g.clone()
        opt = torch.optim.Adam([ref_p], lr=1e-3)
        m, v = torch.zeros_like(p), torch.zeros_like(p)
        partial = torch.empty(max(1, lib().ss_optim_parts(ctypes.c_longlong(n))), device="cuda")
        norm = torch.zeros(1, device="cuda")
        for step in (1, 2):
            torch.nn.utils.clip_grad_norm_([ref_p], 0.25)
            opt.step()
            ref_p.grad = g.clone()
            call("grad_norm", g, LL(n), 1.0, partial, norm, stream_ptr())
            call("clip_adam", p, g, m, v, LL(n), norm, 1.0, 0.25, 1e-3, 0.9, 0.999, 1e-8, step, stream_ptr())
        assert abs(float(norm) - float(g.norm())) <= 1e-5 * float(g.norm())
        assert float((p - ref_p.detach()).abs().max()) <= 2e-6, n


def test_step_accumulate_equals_mean_of_micro_batch_gradients():
    """Trainer.step_accumulate (strong scaling: a global batch run as micro-batches) = clip + Adam on the MEAN of the
    micro-batch gradients, each micro-batch with its own loss normaliser and BatchNorm statistics."""
    z, case = common.load_golden("upit_s2_f129")
    batch, _, _ = _batch(case)
    z2, case2 = common.load_golden("upit_s2_f129")
    case2 = dict(case2, seed=case2["seed"] + 7, lens=[n - 100 for n in case2["lens"]])
    batch2, _, _ = _batch(case2)
    model_a = _model(case)
    model_b = copy.deepcopy(model_a)
    opt = torch.optim.Adam(model_a.parameters(), lr=1e-3)
    tr = train.Trainer(model_b, uPIT, host_rng=True)
    torch.manual_seed(50)
    grads = []
    for b in (batch, batch2):
        loss, _ = uPIT.compute_loss(model_a, 0, b)
        loss.backward()
        grads.append([p.grad.clone() for p in model_a.parameters()])
    for p, g1, g2 in zip(model_a.parameters(), *grads):
        p.grad = 0.5 * (g1 + g2)
    torch.nn.utils.clip_grad_norm_(model_a.parameters(), 0.25)
    opt.step()
    torch.manual_seed(50)
    tr.step_accumulate([batch, batch2])
    torch.cuda.synchronize()
    for (k, pa), (_, pb) in zip(model_a.named_parameters(), model_b.named_parameters()):
        assert float((pa.detach() - pb.detach()).abs().max()) <= 1e-5, k
    for k in ("running_mean", "running_var"):
        assert torch.allclose(getattr(model_a.bn, k), getattr(model_b.bn, k), atol=1e-6)
