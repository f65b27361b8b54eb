"""Data-parallel semantics on ONE device: the ranks are threads that take turns on the GPU, the
collectives are a lock-step stand-in for NCCL, the arithmetic is the product's.

Mode B (`exact`): loss normaliser and BatchNorm statistics over all ranks' utterances, gradients summed.  Two ranks with
four utterances each must reproduce the UNMODIFIED reference on the concatenated batch of eight (the golden
ref_upit_s2_f257 = BASELINE configs[0]): same loss, same permutations, summed gradients within 1e-3, same BatchNorm
running statistics.
Mode A (default): every rank computes the reference loss on its own shard; the exchanged gradient is the mean.
"""
import threading

import numpy as np
import pytest
import torch

import common
from ssb200 import uPIT

pytestmark = pytest.mark.gpu


class LockStepGroup:
    """`world` threads; exactly one runs (enqueues GPU work) at a time; a collective hands the turn over and completes
    when every rank has arrived.  Keeps two recurrence kernels from ever being in flight together on the one device."""

    def __init__(self, world):
        self.world, self.cond = world, threading.Condition()
        self.turn = threading.Lock()
        self.slots, self.arrived, self.gen = {}, 0, 0
        self.errors = []

    def _collective(self, rank, t, op):
        torch.cuda.synchronize()
        with self.cond:
            self.slots[rank] = t
            self.arrived += 1
            if self.arrived == self.world:
                ts = [self.slots[r] for r in range(self.world)]
                red = torch.stack(ts).sum(0) if op == "sum" else torch.stack(ts).max(0).values
                for x in ts:
                    x.copy_(red)
                torch.cuda.synchronize()
                self.slots, self.arrived = {}, 0
                self.gen += 1
                self.cond.notify_all()
                return
            gen = self.gen
            self.turn.release()                                # let the next rank run up to the same collective
            while self.gen == gen and not self.errors:
                self.cond.wait(timeout=1.0)
        self.turn.acquire()

    def dp(self, rank):
        g = self

        class _DP:
            def counts(self, B, T, device):
                return torch.tensor([float(B), float(T)], dtype=torch.float32, device=device)

            def allreduce_sum(self, t):
                g._collective(rank, t, "sum")

            def allreduce_max(self, t):
                g._collective(rank, t, "max")
        return _DP()

    def run(self, fn):
        out = [None] * self.world

        def body(r):
            self.turn.acquire()
            try:
                out[r] = fn(r)
                torch.cuda.synchronize()
            except BaseException as e:                         # noqa: BLE001  (re-raised in the main thread)
                self.errors.append(e)
                with self.cond:
                    self.cond.notify_all()
            finally:
                if self.turn.locked():
                    try:
                        self.turn.release()
                    except RuntimeError:
                        pass
        ths = [threading.Thread(target=body, args=(r,)) for r in range(self.world)]
        for t in ths:
            t.start()
        for t in ths:
            t.join(timeout=300)
        if self.errors:
            raise self.errors[0]
        return out


def _shards(case, world):
    mixes, srcs, _ = common.case_features(case)
    samples = []
    for b in range(len(mixes)):
        d = {"mix": mixes[b]}
        for i in range(case["S"]):
            d["source%d" % (i + 1)] = srcs[i][b]
        samples.append(d)
    # the collator's order of the whole batch (descending length), dealt round robin
    order = [int(i) for i in np.argsort(np.array([len(d["mix"]) for d in samples]))[::-1]]
    return [[samples[i] for i in order[r::world]] for r in range(world)], [order[r::world] for r in range(world)], order


def _model(case):
    F = case["n_fft"] // 2 + 1
    torch.manual_seed(case["seed"])
    return uPIT.SepDNN(0, feat_dim=str(F), num_spk=str(case["S"])).cuda()


def test_mode_b_two_ranks_equal_the_reference_on_the_concatenated_batch():
    z, case = common.load_golden("upit_s2_f257")
    world = 2
    shards, members, order = _shards(case, world)
    B = len(order)
    torch.manual_seed(100 + case["seed"])                      # the reference's init_hidden draw for the batch of 8
    h0, c0 = torch.randn(4, B, 600), torch.randn(4, B, 600)
    pos = {u: i for i, u in enumerate(order)}                  # batch column of utterance u in the reference run
    group = LockStepGroup(world)
    models = [_model(case) for _ in range(world)]

    def rank_fn(r):
        cols = [pos[u] for u in members[r]]
        hid = lambda n: (h0[:, cols].contiguous().cuda(), c0[:, cols].contiguous().cuda())
        grads = {k: torch.zeros_like(p) for k, p in models[r].named_parameters()}
        batch = uPIT.Collator("mix")(shards[r])
        loss, norm = uPIT.loss_and_grads(models[r], batch, grads, hidden=hid, dp=group.dp(r))
        return float(loss), float(norm), grads, models[r].last["indices"].cpu().numpy()

    res = group.run(rank_fn)
    for r in range(world):
        assert res[r][1] == float(z["norm"])                                   # the GLOBAL normaliser on every rank
        assert abs(res[r][0] - float(z["loss"])) <= 1e-3 * float(z["loss"])     # the GLOBAL loss on every rank
        assert np.array_equal(res[r][3], z["indices"][0][[pos[u] for u in members[r]]])
        assert np.abs(models[r].bn.running_mean.cpu().numpy()[::8] - z["bn.running_mean/sample"]).max() <= 1e-5
        assert np.abs(models[r].bn.running_var.cpu().numpy()[::8] - z["bn.running_var/sample"]).max() <= 1e-5
    for k in res[0][2]:
        g = res[0][2][k] + res[1][2][k]                                         # what the sum-allreduce delivers
        stat, samp = common.grad_stat(k, g)
        rs, rsamp = z["g/" + k + "/stat"], z["g/" + k + "/sample"]
        assert abs(stat[0] - rs[0]) <= 1e-3 * rs[0], (k, stat, rs)
        assert common.rel(samp, rsamp) <= 5e-3, k


def test_mode_a_rank_gradients_are_the_shard_gradients():
    """Mode A: a rank's pre-exchange gradient is the single-device gradient of its shard (no collective touches the
    math), so the averaged gradient is the mean of the per-shard reference gradients."""
    z, case = common.load_golden("upit_s2_f129")
    shards, _, _ = _shards(case, 2)
    ref = []
    for r in range(2):
        m = _model(case)
        torch.manual_seed(5 + r)
        loss, _ = uPIT.compute_loss(m, 0, uPIT.Collator("mix")(shards[r]))
        loss.backward()
        ref.append({k: p.grad.clone() for k, p in m.named_parameters()})
    for r in range(2):
        m = _model(case)
        grads = {k: torch.zeros_like(p) for k, p in m.named_parameters()}
        torch.manual_seed(5 + r)
        uPIT.loss_and_grads(m, uPIT.Collator("mix")(shards[r]), grads)
        for k in grads:
            assert common.rel(grads[k].cpu().numpy(), ref[r][k].cpu().numpy()) <= 1e-5, (r, k)
