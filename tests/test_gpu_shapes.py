"""Shapes beyond the golden cases, through the same drop-in module:
 * batch 100 (the reference recipe's default, run_train.sh:17) -> the B > 64 kernels, checked against the fp64 oracle;
 * config 5 lengths (16 kHz, 20 s -> T = 2501): size-independent properties -- STFT -> iSTFT with an all-ones mask
   returns the signal, and eval-mode masks of a batch equal those of its utterances run one at a time.  (The config 5
   TRAIN step is checked against the unmodified reference's golden in tests/test_gpu_configs.py.)"""
import os

import numpy as np
import pytest
import torch

import common
from oracle import upit_oracle as uo
from ssb200 import ops, synth, uPIT

pytestmark = pytest.mark.gpu


def _rand_batch(seed, lens, F, S):
    rng = np.random.default_rng(seed)
    samples = []
    for L in lens:
        d = {"mix": rng.random((L, F), dtype=np.float32)}
        for s in range(S):
            d["source%d" % (s + 1)] = rng.random((L, F), dtype=np.float32) * 0.5
        samples.append(d)
    return samples


def test_batch_100_matches_oracle():
    F, S, B = 129, 2, 100
    lens = sorted(np.random.default_rng(1).integers(6, 25, size=B).tolist(), reverse=True)
    samples = _rand_batch(2, lens, F, S)
    batch = uPIT.Collator("mix")(samples)
    torch.manual_seed(9)
    model = uPIT.SepDNN(0, feat_dim=str(F), num_spk=str(S)).cuda()
    sd = {k: (v.detach().double().cpu().clone() if v.is_floating_point() else v.cpu().clone())
          for k, v in model.state_dict().items()}
    torch.manual_seed(21)
    loss, norm = uPIT.compute_loss(model, 0, batch)
    loss.backward()
    torch.cuda.synchronize()
    torch.manual_seed(21)
    h0, c0 = torch.randn(4, B, 600).double(), torch.randn(4, B, 600).double()
    order = np.argsort(np.array(lens))[::-1]
    srt = [samples[i] for i in order]
    mix, tl = common.pad_time_major([s["mix"] for s in srt])
    sources = [common.pad_time_major([s["source%d" % (i + 1)] for s in srt])[0] for i in range(S)]
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd.update(params)
    oloss, onorm, _, oidx, _, _ = uo.upit_compute_loss(sd, mix, sources, tl, h0, c0)
    oloss.backward()
    assert float(norm) == float(onorm)
    assert abs(float(loss.detach()) - float(oloss)) <= 1e-3 * float(oloss)
    assert np.array_equal(model.last["indices"].cpu().numpy(), oidx.numpy())
    for k, p in model.named_parameters():
        assert common.rel(p.grad.cpu().numpy(), params[k].grad.numpy()) <= 1e-3, k


def test_cfg5_stft_istft_roundtrip():
    n, n_fft, hop = 320000, 512, 128
    lens = [n, n - 777, 200000]
    sigs = [synth.mixture(90 + i, L, 2, sr=16000)[0] for i, L in enumerate(lens)]
    rag = ops.Ragged(lens, n_fft, hop, "cuda")
    assert int(rag.T[0]) == 2501
    flat = torch.from_numpy(np.concatenate(sigs)).cuda()
    spec = ops.stft(flat, rag, complex_out=True)
    ones = torch.ones(rag.total_frames * rag.F, device="cuda")
    wav, off = ops.istft_masked(spec, ones, rag, 1, out_f32=True)
    wav = wav.cpu().numpy()
    for u, s in enumerate(sigs):
        L = int(rag.L[u])
        got = wav[off[u]:off[u] + L]
        assert np.abs(got - s[:L]).max() <= 1e-4 * np.abs(s).max(), u


def test_cfg5_masks_batch_invariance(tmp_path):
    F, B, T = 257, 8, 2501
    lens = [T - 37 * b for b in range(B)]
    samples = [{"mix": s["mix"], "name": "u%d.npz" % i} for i, s in enumerate(_rand_batch(4, lens, F, 0))]
    torch.manual_seed(2)
    model = uPIT.SepDNN(0, feat_dim=str(F), num_spk="2").cuda()
    model.eval()
    coll = uPIT.Collator("mix")
    h = torch.randn(4, B, 600)
    c = torch.randn(4, B, 600)
    model.init_hidden = lambda n: (h.cuda(), c.cuda())
    (tmp_path / "all").mkdir()
    uPIT.compute_masks(model, coll(samples), str(tmp_path / "all"))
    (tmp_path / "one").mkdir()
    for i in (0, 3, 7):
        model.init_hidden = lambda n, i=i: (h[:, i:i + 1].contiguous().cuda(), c[:, i:i + 1].contiguous().cuda())
        uPIT.compute_masks(model, coll([samples[i]]), str(tmp_path / "one"))
        a = np.load(str(tmp_path / "all" / ("u%d.npz" % i)))
        b = np.load(str(tmp_path / "one" / ("u%d.npz" % i)))
        for k in ("s1", "s2"):
            assert a[k].shape == (F, lens[i])
            assert np.abs(a[k] - b[k]).max() <= 2e-5, (i, k)


def test_three_layer_net_matches_oracle():
    """`num_layers=3` (BASELINE configs[0] names a 3-layer BLSTM; the reference hard-codes 2, archs/uPIT.py:115): the
    engine's layer loops, the ping-pong of the streamed dX buffers and the Trainer's buckets, against the fp64 oracle."""
    from ssb200 import train
    F, S, B = 129, 2, 5
    lens = [70, 66, 61, 55, 49]                              # T >= 48: the GEMMs stream under the recurrences
    samples = _rand_batch(12, lens, F, S)
    batch = uPIT.Collator("mix")(samples)
    torch.manual_seed(4)
    model = uPIT.SepDNN(0, feat_dim=str(F), num_spk=str(S), num_layers="3").cuda()
    assert "blstm.weight_hh_l2_reverse" in model.state_dict()
    sd = {k: (v.detach().double().cpu().clone() if v.is_floating_point() else v.cpu().clone())
          for k, v in model.state_dict().items()}
    torch.manual_seed(33)
    loss, norm = uPIT.compute_loss(model, 0, batch)
    loss.backward()
    torch.cuda.synchronize()
    torch.manual_seed(33)
    h0, c0 = torch.randn(6, B, 600).double(), torch.randn(6, B, 600).double()
    mix, tl = common.pad_time_major([s["mix"] for s in samples])
    sources = [common.pad_time_major([s["source%d" % (i + 1)] for s in samples])[0] for i in range(S)]
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd.update(params)
    oloss, onorm, _, oidx, _, _ = uo.upit_compute_loss(sd, mix, sources, tl, h0, c0, num_layers=3)
    oloss.backward()
    assert abs(float(loss.detach()) - float(oloss)) <= 1e-3 * float(oloss)
    assert np.array_equal(model.last["indices"].cpu().numpy(), oidx.numpy())
    for k, p in model.named_parameters():
        assert common.rel(p.grad.cpu().numpy(), params[k].grad.numpy()) <= 1e-3, k
    tr = train.Trainer(model, uPIT)                          # flat buckets with the third layer, one fused step
    before = tr.flat_p.clone()
    tr.step(batch)
    torch.cuda.synchronize()
    assert set(tr.buckets) == {"l0", "l1", "head"} and float((tr.flat_p - before).abs().max()) > 0


@pytest.mark.parametrize("B", [24, 40])
def test_padded_row_variants_match_oracle(B):
    """Batches of 17..32 utterances run the 32-row variants of the tensor-memory recurrence kernels, 33..64 the 64-row ones
    (B <= 16: the 16-row ones, covered by every small golden): loss, permutations and gradients against the fp64 oracle."""
    F, S = 129, 2
    lens = sorted(np.random.default_rng(B).integers(5, 40, size=B).tolist(), reverse=True)
    samples = _rand_batch(B + 1, lens, F, S)
    batch = uPIT.Collator("mix")(samples)
    torch.manual_seed(3)
    model = uPIT.SepDNN(0, feat_dim=str(F), num_spk=str(S)).cuda()
    sd = {k: (v.detach().double().cpu().clone() if v.is_floating_point() else v.cpu().clone())
          for k, v in model.state_dict().items()}
    torch.manual_seed(22)
    loss, norm = uPIT.compute_loss(model, 0, batch)
    loss.backward()
    torch.cuda.synchronize()
    torch.manual_seed(22)
    h0, c0 = torch.randn(4, B, 600).double(), torch.randn(4, B, 600).double()
    order = np.argsort(np.array(lens))[::-1]
    srt = [samples[i] for i in order]
    mix, tl = common.pad_time_major([s["mix"] for s in srt])
    sources = [common.pad_time_major([s["source%d" % (i + 1)] for s in srt])[0] for i in range(S)]
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd.update(params)
    oloss, onorm, _, oidx, _, _ = uo.upit_compute_loss(sd, mix, sources, tl, h0, c0)
    oloss.backward()
    assert abs(float(loss.detach()) - float(oloss)) <= 1e-3 * float(oloss)
    assert np.array_equal(model.last["indices"].cpu().numpy(), oidx.numpy())
    for k, p in model.named_parameters():
        assert common.rel(p.grad.cpu().numpy(), params[k].grad.numpy()) <= 1e-3, k


def test_forward_sub_batch_view_equals_one_launch():
    """ss_lstm_fwd with batch_pitch: two half-batch launches (rows 0..49 and 50..99 of a 100-row padded batch, tensor-memory
    kernel) give the same masks as one pass of the generic 128-row kernel (batch rows are independent chains)."""
    from ssb200 import engine
    F, B = 129, 100
    lens = sorted(np.random.default_rng(5).integers(6, 30, size=B).tolist(), reverse=True)
    samples = [{"mix": s["mix"]} for s in _rand_batch(6, lens, F, 0)]
    batch = uPIT.Collator("mix")(samples)
    torch.manual_seed(4)
    model = uPIT.SepDNN(0, feat_dim=str(F), num_spk="2").cuda().eval()
    outs = []
    try:
        for split in (False, True):
            engine.SPLIT_WIDE_BATCH_FWD = split
            torch.manual_seed(30)
            model.hidden = (torch.randn(4, B, 600, device="cuda"), torch.randn(4, B, 600, device="cuda"))
            with torch.no_grad():
                outs.append(model(batch["mix"]).float().cpu().numpy())
    finally:
        engine.SPLIT_WIDE_BATCH_FWD = False
    assert outs[0].shape == outs[1].shape
    assert np.abs(outs[0] - outs[1]).max() <= 2e-5


@pytest.mark.parametrize("lens", [[2], [2, 1, 1], [3, 3], [5, 1]])   # (one frame in total: torch's BatchNorm1d refuses it)
def test_tiny_sequences_match_oracle(lens):
    """One- to five-frame utterances, batch 1..3: the recurrence kernels' first / only steps (no streaming below 48 frames,
    reverse rows that start at the very first step, the extra d(h0) step of the backward kernel)."""
    F, S, B = 129, 2, len(lens)
    samples = _rand_batch(77 + sum(lens), lens, F, S)
    batch = uPIT.Collator("mix")(samples)
    torch.manual_seed(5)
    model = uPIT.SepDNN(0, feat_dim=str(F), num_spk=str(S)).cuda()
    sd = {k: (v.detach().double().cpu().clone() if v.is_floating_point() else v.cpu().clone())
          for k, v in model.state_dict().items()}
    torch.manual_seed(23)
    loss, norm = uPIT.compute_loss(model, 0, batch)
    loss.backward()
    torch.cuda.synchronize()
    torch.manual_seed(23)
    h0, c0 = torch.randn(4, B, 600).double(), torch.randn(4, B, 600).double()
    order = np.argsort(np.array(lens), kind="stable")[::-1]
    srt = [samples[i] for i in order]
    mix, tl = common.pad_time_major([s["mix"] for s in srt])
    sources = [common.pad_time_major([s["source%d" % (i + 1)] for s in srt])[0] for i in range(S)]
    params = {k: v.requires_grad_(True) for k, v in sd.items() if v.is_floating_point() and "running" not in k}
    sd.update(params)
    oloss, onorm, _, oidx, _, _ = uo.upit_compute_loss(sd, mix, sources, tl, h0, c0)
    oloss.backward()
    assert abs(float(loss.detach()) - float(oloss)) <= 1e-3 * float(oloss)
    assert np.array_equal(model.last["indices"].cpu().numpy(), oidx.numpy())
    for k, p in model.named_parameters():
        ref = params[k].grad.numpy()
        assert np.abs(p.grad.cpu().numpy() - ref).max() <= 1e-3 * max(np.abs(ref).max(), 1e-12), k
