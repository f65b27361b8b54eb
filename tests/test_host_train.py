"""Host-side logic of the train step that needs no GPU: flat bucket layout, utterance sharding, the bucketed
allreduce over gloo with world_size 2, and the CPU baseline step (oracle/cpu_step.py) against the fp64 oracle."""
import os

import numpy as np
import torch
import torch.multiprocessing as mp

import common
from oracle import cpu_step
from oracle import upit_oracle as uo
from ssb200 import train


def _names():
    out = []
    for l, I in ((0, 257), (1, 1200)):
        for suf in ("", "_reverse"):
            out += [("blstm.weight_ih_l%d%s" % (l, suf), 2400 * I), ("blstm.weight_hh_l%d%s" % (l, suf), 2400 * 600),
                    ("blstm.bias_ih_l%d%s" % (l, suf), 2400), ("blstm.bias_hh_l%d%s" % (l, suf), 2400)]
    return out + [("lin.weight", 514 * 1200), ("lin.bias", 514), ("bn.weight", 1200), ("bn.bias", 1200)]


def test_flat_layout_buckets_are_contiguous_and_aligned():
    offs, buckets, total = train.flat_layout(_names())
    assert sum(n for _, n in _names()) == 13392514
    pos = 0
    for bk in train.BUCKET_ORDER:
        o, n = buckets[bk]
        assert o == pos
        pos += n
        for name, (po, pn) in offs.items():
            if train.bucket_of(name) == bk:
                assert o <= po and po + pn <= o + n and po % 64 == 0
    assert pos == total
    spans = sorted(offs.values())
    assert all(a[0] + a[1] <= b[0] for a, b in zip(spans, spans[1:]))


def test_shard_utterances_partitions_and_balances():
    rng = np.random.default_rng(0)
    lens = [int(v) for v in rng.integers(250, 501, size=512)]
    shards = [train.shard_utterances(lens, 8, r) for r in range(8)]
    assert sorted(i for s in shards for i in s) == list(range(512))
    tot = [sum(lens[i] for i in s) for s in shards]
    tmax = [max(lens[i] for i in s) for s in shards]
    assert max(tot) - min(tot) <= 0.01 * max(tot)
    assert max(tmax) - min(tmax) <= 0.01 * max(tmax)
    assert train.shard_utterances([], 2, 1) == []


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    offs, buckets, total = train.flat_layout([("blstm.weight_ih_l0", 1000), ("blstm.bias_hh_l1", 77), ("lin.bias", 5)])
    g = torch.Generator().manual_seed(rank)
    flat = torch.randn(total, generator=g)
    works = train.allreduce_buckets(flat, buckets)
    for w in works:
        w.wait()
    flat /= world
    q.put((rank, flat.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_allreduce_gloo_world2():
    world, port = 2, 29500 + os.getpid() % 2000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    offs, buckets, total = train.flat_layout([("blstm.weight_ih_l0", 1000), ("blstm.bias_hh_l1", 77), ("lin.bias", 5)])
    expect = sum(torch.randn(total, generator=torch.Generator().manual_seed(r)) for r in range(world)) / world
    for r in range(world):
        assert np.allclose(res[r], expect.numpy(), atol=1e-6)


def test_cpu_baseline_step_matches_oracle():
    z, case = common.load_golden("upit_s2_f129")
    mixes, srcs, _ = common.case_features(case)
    mix, lens = common.pad_time_major(mixes)
    sources = [common.pad_time_major(s)[0] for s in srcs]
    sd = common.case_state_dict(case)
    h0, c0 = common.case_hidden(case)
    F = case["n_fft"] // 2 + 1
    tr = cpu_step.CpuTrainer(F, case["S"], state_dict={k: v.clone() for k, v in sd.items()}, dtype=torch.float64)
    before = {k: v.detach().clone() for k, v in tr.net.named_parameters()}
    loss, norm, idx = tr.step(mix, sources, lens, h0, c0)
    assert abs(loss - float(z["loss"])) <= 2e-6 * float(z["loss"])
    assert norm == float(z["norm"])
    assert np.array_equal(idx.numpy(), z["indices"][0])
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd2 = dict(sd)
    sd2.update(params)
    oloss, *_ = uo.upit_compute_loss(sd2, mix, sources, lens, h0, c0)
    oloss.backward()
    for k, p in tr.net.named_parameters():
        assert common.rel(p.grad.numpy() * 1.0, params[k].grad.numpy() * min(1.0, 0.25 / (float(z["grad_total_norm"]) + 1e-6))) <= 1e-6, k
        assert not torch.equal(before[k], p.detach())
