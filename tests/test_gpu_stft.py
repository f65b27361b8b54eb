"""Parity of the CUDA STFT / masked iSTFT (through the C ABI) against oracle/stft_oracle.py.
Tolerances: features 1e-4 relative (north_star); wavs +-1 LSB of int16."""
import numpy as np
import pytest
import torch

from oracle import stft_oracle as so
from ssb200 import ops, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _corpus(seed, lens, S=2):
    sigs = []
    for i, n in enumerate(lens):
        mix, srcs = synth.mixture(seed * 100 + i, n, S)
        sigs.append(mix)
    return sigs


def _rel(a, b):
    a = np.asarray(a, dtype=np.complex128 if np.iscomplexobj(a) else np.float64)
    return float(np.abs(a - b).max() / np.abs(b).max()), float(np.linalg.norm(a - b) / np.linalg.norm(b))


@pytest.mark.parametrize("n_fft,hop", [(512, 128), (256, 64), (512, 256), (256, 96), (512, 160)])
@pytest.mark.parametrize("complex_out", [False, True])
@pytest.mark.parametrize("layout_tf", [False, True])
def test_stft_matches_oracle(n_fft, hop, complex_out, layout_tf):
    # ragged lengths incl. the minimum legal one (n_fft//2+1), non multiples of hop, one exact tile boundary
    lens = [32000, n_fft // 2 + 1, 31 * hop, 32 * hop, 33 * hop + 7, 4000, 12345]
    sigs = _corpus(3, lens)
    rag = ops.Ragged(lens, n_fft, hop, DEV)
    flat = torch.from_numpy(np.concatenate(sigs)).to(DEV)
    out = ops.stft(flat, rag, complex_out=complex_out, layout_tf=layout_tf)
    torch.cuda.synchronize()
    for u, v in enumerate(ops.stft_views(out, rag, layout_tf)):
        ref = so.stft(sigs[u], n_fft, hop) if complex_out else so.stft_mag(sigs[u], n_fft, hop)
        got = v.cpu().numpy()
        if layout_tf:
            got = got.T
        assert got.shape == ref.shape == (n_fft // 2 + 1, 1 + lens[u] // hop)
        mx, l2 = _rel(got, ref)
        assert mx <= 1e-4 and l2 <= 1e-4, (u, mx, l2)
        dft = n_fft == 256 and hop == 64 and not complex_out and not layout_tf and ops.STFT_DFT256
        # what the arithmetic actually achieves: fp32 FFT 2e-6; split-bf16 tensor-core DFT (three products) 3e-5
        assert mx <= (3e-5 if dft else 2e-6), (u, mx)


@pytest.mark.parametrize("align", [1, 4])
def test_stft_dft256_matches_oracle_and_fft_kernel(align, monkeypatch):
    # tensor-core path (csrc/stft_mma.cu): tile boundaries (127/128/129 frames), the shortest legal utterance, lengths
    # that are not multiples of the hop, 16-byte aligned (fast loads) and back-to-back (scalar loads) layouts
    lens = [129, 127 * 64, 128 * 64 - 1, 128 * 64, 129 * 64 + 5, 300, 64 * 3 + 1, 40000, 12345, 257 * 64 + 63]
    monkeypatch.setattr(ops, "STFT_DFT256", True)
    sigs = _corpus(11, lens)
    rag = ops.Ragged(lens, 256, 64, DEV, align=align)
    host = np.full(int(rag.sig_off[-1]), np.nan, dtype=np.float32)      # gaps must never be read
    for u, s in enumerate(sigs):
        host[int(rag.sig_off[u]):int(rag.sig_off[u]) + lens[u]] = s
    flat = torch.from_numpy(host).to(DEV)
    out = ops.stft(flat, rag)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(out).all())
    fft = None
    if align == 1:
        monkeypatch.setattr(ops, "STFT_DFT256", False)
        fft = ops.stft_views(ops.stft(flat, rag), rag)
    for u, v in enumerate(ops.stft_views(out, rag)):
        ref = so.stft_mag(sigs[u], 256, 64)
        got = v.cpu().numpy()
        assert got.shape == ref.shape
        mx, l2 = _rel(got, ref)
        assert mx <= 1e-4 and l2 <= 1e-4, (u, mx, l2)      # north_star feature tolerance
        assert mx <= 3e-5 and l2 <= 3e-5, (u, mx, l2)      # what three split-bf16 products achieve
        if fft is not None:
            mx2, _ = _rel(got, fft[u].cpu().numpy().astype(np.float64))
            assert mx2 <= 3e-5, (u, mx2)


@pytest.mark.parametrize("n_fft,hop,dft", [(256, 64, True), (256, 64, False), (512, 128, False)])
def test_stft_log_flag(n_fft, hop, dft, monkeypatch):
    """SS_STFT_LOG (an extension; the reference stores plain magnitudes): log(max(|X|, 1e-8)) of the same features, on the
    FFT kernel and on the tensor-core kernel."""
    monkeypatch.setattr(ops, "STFT_DFT256", dft)
    lens = [5000, 12345, n_fft // 2 + 1]
    sigs = _corpus(5, lens)
    rag = ops.Ragged(lens, n_fft, hop, DEV)
    flat = torch.from_numpy(np.concatenate(sigs)).to(DEV)
    mag = ops.stft(flat, rag)
    lg = ops.stft(flat, rag, log=True)
    want = torch.log(torch.clamp(mag, min=1e-8))
    assert float((lg - want).abs().max()) <= 2e-5
    for u, v in enumerate(ops.stft_views(lg, rag)):
        mag_ref = so.stft_mag(sigs[u], n_fft, hop)
        ref = np.log(np.maximum(mag_ref, 1e-8))
        big = mag_ref > 0.05 * mag_ref.max()            # (an absolute error of 1e-5 * peak is a large RELATIVE error of a tiny bin)
        assert np.abs(v.cpu().numpy() - ref)[big].max() <= 1e-3


def test_stft_dft256_full_size_properties(monkeypatch):
    monkeypatch.setattr(ops, "STFT_DFT256", True)
    # BASELINE-sized shard: homogeneity |STFT(a x)| == a |STFT(x)|, agreement with the FFT kernel, and Parseval per
    # frame against the windowed samples of interior frames
    rng = np.random.default_rng(2)
    lens = rng.integers(16000, 80001, size=256)
    rag = ops.Ragged(lens, 256, 64, DEV, align=4)
    n = int(rag.sig_off[-1])
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.rand(n, device=DEV, generator=g) - 0.5
    X = ops.stft(x, rag)
    Y = ops.stft(-2.5 * x, rag)
    assert float((Y - 2.5 * X).abs().max() / Y.abs().max()) <= 3e-5
    u = 17
    s0, N, T = int(rag.sig_off[u]), int(lens[u]), int(rag.T[u])
    w = torch.hann_window(256, periodic=True, device=DEV, dtype=torch.float64)
    fr = x[s0:s0 + N].double().unfold(0, 256, 64) * w                   # interior frames t = 2 .. (no padding)
    Xu = ops.stft_views(X, rag)[u].double()[:, 2:2 + fr.shape[0]]
    par = (Xu[0] ** 2 + Xu[128] ** 2 + 2 * (Xu[1:128] ** 2).sum(0)) / 256
    assert float(((par - (fr ** 2).sum(1)).abs() / (fr ** 2).sum(1)).max()) <= 1e-4


def test_stft_linearity_full_size():
    # size-independent property at a BASELINE-sized shard: STFT(a*x + b*y) == a*STFT(x) + b*STFT(y)
    rng = np.random.default_rng(0)
    lens = rng.integers(16000, 80001, size=64)
    rag = ops.Ragged(lens, 256, 64, DEV)
    n = int(rag.sig_off[-1])
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.rand(n, device=DEV, generator=g) - 0.5
    y = torch.rand(n, device=DEV, generator=g) - 0.5
    X = ops.stft(x, rag, complex_out=True)
    Y = ops.stft(y, rag, complex_out=True)
    Z = ops.stft(0.25 * x - 1.5 * y, rag, complex_out=True)
    err = (Z - (0.25 * X - 1.5 * Y)).abs().max() / Z.abs().max()
    assert float(err) <= 1e-5


@pytest.mark.parametrize("n_fft,hop", [(512, 128), (256, 64), (512, 256), (256, 96), (512, 160)])
@pytest.mark.parametrize("layout_tf", [False, True])
def test_masked_istft_matches_oracle(n_fft, hop, layout_tf):
    lens = [16000, 29 * hop, 28 * hop + 1, 5000, n_fft + 3, n_fft // 2 + hop]
    S = 2
    F = n_fft // 2 + 1
    rng = np.random.default_rng(5)
    sigs = _corpus(7, lens)
    rag = ops.Ragged(lens, n_fft, hop, DEV)
    specs = [so.stft(s, n_fft, hop) for s in sigs]
    masks = [rng.uniform(0, 1, size=(S, F, sp.shape[1])).astype(np.float32) for sp in specs]
    if layout_tf:
        spec_flat = np.concatenate([sp.T.ravel() for sp in specs])
        mask_flat = np.concatenate([m.transpose(0, 2, 1).ravel() for m in masks])
    else:
        spec_flat = np.concatenate([sp.ravel() for sp in specs])
        mask_flat = np.concatenate([m.ravel() for m in masks])
    d_spec, d_mask = torch.from_numpy(spec_flat).to(DEV), torch.from_numpy(mask_flat).to(DEV)
    wav, off = ops.istft_masked(d_spec, d_mask, rag, S, layout_tf=layout_tf)
    wavf, _ = ops.istft_masked(d_spec, d_mask, rag, S, layout_tf=layout_tf, out_f32=True)
    wav, wavf = wav.cpu().numpy(), wavf.cpu().numpy()
    for u in range(len(lens)):
        L = hop * (specs[u].shape[1] - 1)
        for s in range(S):
            ref = so.masked_istft_int16(specs[u], masks[u][s], hop)
            reff = so.istft(specs[u] * masks[u][s], hop)
            got = wav[off[u] + s * L: off[u] + (s + 1) * L]
            gotf = wavf[off[u] + s * L: off[u] + (s + 1) * L]
            assert got.shape == ref.shape
            assert np.abs(gotf - reff).max() <= 2e-6 * max(1e-3, np.abs(reff).max()), (u, s)
            assert np.abs(got.astype(np.int32) - ref.astype(np.int32)).max() <= 1, (u, s)
            assert (got == ref).mean() >= 0.99


def test_stft_istft_roundtrip_full_size():
    # unmasked round trip (mask == 1) reproduces the interior of the signal; cfg5-sized utterances (16 kHz, 20 s)
    lens = [320000] * 4 + [319871, 160000]
    rag = ops.Ragged(lens, 512, 128, DEV)
    g = torch.Generator(device=DEV).manual_seed(2)
    x = (torch.rand(int(rag.sig_off[-1]), device=DEV, generator=g) - 0.5)
    X = ops.stft(x, rag, complex_out=True)
    ones = torch.ones(rag.total_frames * rag.F, device=DEV)
    y, off = ops.istft_masked(X, ones, rag, 1, out_f32=True)
    for u, n in enumerate(lens):
        L = int(rag.L[u])
        a = x[int(rag.sig_off[u]):int(rag.sig_off[u]) + L]
        b = y[int(off[u]):int(off[u]) + L]
        assert float((a - b).abs().max()) <= 5e-6


@pytest.mark.parametrize("chunks", [1, 3, 8])
def test_stft_host_streaming_equals_single_launch(chunks):
    # host -> host streaming over three streams gives bit-identical features to one launch over the whole corpus
    rng = np.random.default_rng(8)
    lens = rng.integers(3000, 20000, size=23)
    rag = ops.Ragged(lens, 256, 64, DEV)
    h_sig = (torch.rand(int(rag.sig_off[-1])) - 0.5).pin_memory()
    h_out = torch.empty(rag.total_frames * rag.F, dtype=torch.float32).pin_memory()
    for _ in range(2):                                       # second call reuses the cached plan and buffers
        h_out.zero_()
        ops.stft_host(h_sig, rag, h_out, chunks=chunks)
        ref = ops.stft(h_sig.to(DEV), rag).cpu()
        assert torch.equal(h_out, ref)


def test_masked_istft_int16_wraps_like_astype():
    """`(s * 32767).astype('int16')` (steps/reconstruct_sources.py:41-42) does not clip: a mask that pushes the waveform
    past full scale wraps modulo 2^16.  The kernel must wrap the same way (compared modulo 2^16, +-1 LSB)."""
    n_fft, hop = 512, 128
    lens = [9000, 6000]
    sigs = [0.9 * s / np.abs(s).max() for s in _corpus(21, lens)]
    rag = ops.Ragged(lens, n_fft, hop, DEV)
    flat = torch.from_numpy(np.concatenate(sigs).astype(np.float32)).to(DEV)
    spec = ops.stft(flat, rag, complex_out=True)
    mask = torch.full((rag.total_frames * rag.F,), 2.5, device=DEV)           # 2.25 x full scale
    wav, off = ops.istft_masked(spec, mask, rag, 1)
    wav = wav.cpu().numpy()
    wrapped = 0
    for u, sgl in enumerate(sigs):
        spec_u = so.stft(sgl.astype(np.float32), n_fft, hop)
        fl = so.masked_istft_float(spec_u, np.full(spec_u.shape, 2.5, dtype=np.float32), hop)
        ref = np.asarray(fl).astype(np.int64)                                  # trunc toward zero, then wrap to int16
        ref = ((ref + 32768) % 65536 - 32768)
        got = wav[off[u]:off[u] + len(ref)].astype(np.int64)
        d = (got - ref + 32768) % 65536 - 32768
        assert np.abs(d).max() <= 1, u
        wrapped += int((np.abs(fl) >= 32768).sum())
    assert wrapped > 100                                                       # the case really exercises the wrap


@pytest.mark.parametrize("n_fft,hop", [(400, 160), (320, 80), (512, 96), (256, 33), (128, 32), (1024, 256)])
def test_generic_sizes_stft_and_istft(n_fft, hop):
    """Any even n_fft / any hop (the reference hands --fft-dim and --step-size to librosa unchecked): the direct-summation
    kernels behind the same entry points, features 1e-4 and wavs +-1 LSB like the fast sizes."""
    lens = [4000, n_fft // 2 + 1, 7 * hop + 3, 3001]
    sigs = _corpus(31, lens)
    rag = ops.Ragged(lens, n_fft, hop, DEV)
    flat = torch.from_numpy(np.concatenate(sigs)).to(DEV)
    for complex_out, layout_tf in ((False, False), (True, False), (True, True)):
        out = ops.stft(flat, rag, complex_out=complex_out, layout_tf=layout_tf)
        for u, v in enumerate(ops.stft_views(out, rag, layout_tf)):
            ref = so.stft(sigs[u], n_fft, hop) if complex_out else so.stft_mag(sigs[u], n_fft, hop)
            got = v.cpu().numpy().T if layout_tf else v.cpu().numpy()
            assert got.shape == ref.shape
            mx, l2 = _rel(got, ref)
            assert mx <= 1e-4 and l2 <= 1e-4, (u, mx, l2)
    spec = ops.stft(flat, rag, complex_out=True)
    rng = np.random.default_rng(4)
    masks = [rng.random((2, rag.F, int(rag.T[u])), dtype=np.float32) for u in range(len(lens))]
    mask_flat = torch.from_numpy(np.concatenate([m.ravel() for m in masks])).to(DEV)
    wav, off = ops.istft_masked(spec, mask_flat, rag, 2)
    wav = wav.cpu().numpy()
    for u in range(len(lens)):
        spec_u = so.stft(sigs[u], n_fft, hop)
        for s in range(2):
            ref = so.masked_istft_int16(spec_u, masks[u][s], hop)
            got = wav[off[u] + s * len(ref): off[u] + (s + 1) * len(ref)]
            assert np.abs(got.astype(np.int64) - ref.astype(np.int64)).max() <= 1, (u, s)


@pytest.mark.parametrize("layout_tf", [False, True])
def test_log_cmvn_epilogue(layout_tf):
    """The optional K1 epilogue of north_star: log-magnitude, then per-utterance / per-bin mean-variance normalisation."""
    lens = [5000, 12345, 300]
    sigs = _corpus(6, lens)
    rag = ops.Ragged(lens, 512, 128, DEV)
    flat = torch.from_numpy(np.concatenate(sigs)).to(DEV)
    feat = ops.cmvn(ops.stft(flat, rag, log=True, layout_tf=layout_tf), rag, layout_tf=layout_tf)
    for u, v in enumerate(ops.stft_views(feat, rag, layout_tf)):
        got = v.cpu().numpy().T if layout_tf else v.cpu().numpy()
        lg = np.log(np.maximum(so.stft_mag(sigs[u], 512, 128), 1e-8)).astype(np.float64)
        ref = (lg - lg.mean(1, keepdims=True)) / np.sqrt(lg.var(1, keepdims=True) + 1e-8)
        assert np.abs(got.mean(1)).max() <= 1e-4 and np.abs(got.std(1) - 1).max() <= 1e-3
        big = so.stft_mag(sigs[u], 512, 128).min(1) > 1e-4          # rows whose log is well conditioned everywhere
        assert np.abs(got - ref)[big].max() <= 5e-3 if big.any() else True
