"""Cross-check oracle/stft_oracle.py against an independent statement (torch.stft / torch.istft).

The reference's STFT lives in librosa (absent, unpinned): parity is UNPINNED at this
boundary; this test only shows two independent restatements of the librosa-0.6
algorithm agree, and that the shapes are the reference's."""
import numpy as np
import pytest
import torch

from oracle import stft_oracle as so
from ssb200 import synth


@pytest.mark.parametrize("n_fft,hop,n", [(512, 128, 32000), (256, 64, 32000), (512, 128, 12345), (256, 64, 257), (512, 128, 320000)])
def test_stft_matches_torch(n_fft, hop, n):
    y, _ = synth.mixture(7, n, 2)
    S = so.stft(y, n_fft, hop)
    assert S.dtype == np.complex64 and S.shape == (n_fft // 2 + 1, 1 + n // hop)
    St = torch.stft(torch.from_numpy(y).double(), n_fft, hop, window=torch.hann_window(n_fft, periodic=True, dtype=torch.float64),
                    center=True, pad_mode="reflect", return_complex=True).numpy()
    assert np.abs(S - St).max() / np.abs(St).max() < 5e-7
    assert so.stft_mag(y, n_fft, hop).dtype == np.float32


def test_shapes_of_configs():
    assert so.num_frames(32000, 128) == 251 and so.num_frames(32000, 64) == 501
    assert so.num_frames(320000, 128) == 2501


@pytest.mark.parametrize("n_fft,hop,n", [(512, 128, 32000), (256, 64, 8000), (512, 128, 5000)])
def test_istft_matches_torch_and_roundtrip(n_fft, hop, n):
    y, _ = synth.mixture(3, n, 2)
    S = so.stft(y, n_fft, hop)
    rec = so.istft(S, hop)
    T = S.shape[1]
    assert rec.dtype == np.float32 and rec.shape == (hop * (T - 1),)
    m = min(len(rec), n)
    # interior round trip (edges where the window sum-square is small are still exact for hann/4x overlap)
    assert np.abs(rec[:m] - y[:m]).max() < 5e-6
    rt = torch.istft(torch.from_numpy(S), n_fft, hop, window=torch.hann_window(n_fft, periodic=True), center=True).numpy()
    assert np.abs(rec[:len(rt)] - rt).max() < 5e-6


def test_masked_int16_truncates_toward_zero():
    y, _ = synth.mixture(5, 4000, 2)
    S = so.stft(y, 256, 64)
    mask = np.full(S.shape, 0.5, dtype=np.float32)
    w = so.masked_istft_int16(S, mask, 64)
    f = so.masked_istft_float(S, mask, 64)
    assert w.dtype == np.int16
    assert np.array_equal(w, np.trunc(f).astype(np.int16))
