"""CPU oracle for the STFT / iSTFT ends of the hot path.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl
reference` legs may import this module.  The product path
(`speech-separation_b200/`) never does; it fails loudly without its CUDA library.

PARITY UNPINNED at this boundary: the arithmetic of the reference lives in
`librosa` (not vendored under /root/reference, not version-pinned; the code
dates to mid-2018 => librosa 0.6.x) and the reference holds no golden vectors
for it.  What follows restates the published librosa 0.6
algorithm for the exact call sites the reference makes:

  steps/extract_feats.py:87,89   np.abs(librosa.core.stft(audio, n_fft, hop_length))
  steps/extract_feats.py:98,105  librosa.core.stft(audio, n_fft, hop_length)        (complex64)
  steps/reconstruct_sources.py:39-42
        S = mix_spec * mask ; s = librosa.core.istft(S, hop_length) ; (s*32767).astype('int16')

librosa 0.6 defaults relied on: win_length=n_fft, window='hann' (periodic,
scipy.signal.get_window(..., fftbins=True)), center=True, pad_mode='reflect',
dtype=complex64 (stft) / float32 (istft), window-sum-square normalisation with
threshold util.tiny(float32), trim n_fft//2 at both ends.

A second, independent statement (torch.stft / torch.istft) is used in
tests/test_oracle_stft.py as a cross-check.
"""
import numpy as np


def hann_periodic(n_fft):
    """scipy.signal.get_window('hann', n_fft, fftbins=True), float64."""
    n = np.arange(n_fft, dtype=np.float64)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * n / n_fft)


def num_frames(n_samples, hop):
    """librosa frame count with center=True: 1 + (N + 2*(n_fft//2) - n_fft)//hop."""
    return 1 + n_samples // hop


def stft(y, n_fft=512, hop=128):
    """librosa.core.stft(y, n_fft=n_fft, hop_length=hop) of librosa 0.6.

    y: (N,) float32.  Returns complex64 (1+n_fft//2, 1+N//hop).
    Windowed frames are formed in float64 (float64 window * float32 signal) and
    transformed in double precision, then cast to complex64 -- as
    scipy.fftpack.fft on a float64 array did.
    (extract_feats.py:76-89,98,105)
    """
    y = np.asarray(y, dtype=np.float32)
    assert y.ndim == 1 and n_fft % 2 == 0
    if len(y) <= n_fft // 2:
        raise ValueError("reflect padding needs len(y) > n_fft//2")
    w = hann_periodic(n_fft)
    yp = np.pad(y, n_fft // 2, mode="reflect")
    T = 1 + (len(yp) - n_fft) // hop
    idx = np.arange(n_fft)[:, None] + hop * np.arange(T)[None, :]
    frames = yp[idx].astype(np.float64) * w[:, None]
    spec = np.fft.rfft(frames, axis=0)
    return spec.astype(np.complex64)


def stft_mag(y, n_fft=512, hop=128):
    """np.abs(librosa.core.stft(...)) -> float32 (F, T)  (extract_feats.py:87,89)."""
    return np.abs(stft(y, n_fft, hop))


def window_sumsquare(T, n_fft, hop):
    """librosa.filters.window_sumsquare(window='hann', n_frames=T, dtype=float32)."""
    n = n_fft + hop * (T - 1)
    x = np.zeros(n, dtype=np.float32)
    win_sq = hann_periodic(n_fft) ** 2
    for i in range(T):
        s = i * hop
        x[s:min(n, s + n_fft)] += win_sq[:max(0, min(n_fft, n - s))]
    return x


def istft(S, hop=128):
    """librosa.core.istft(S, hop_length=hop) of librosa 0.6 (reconstruct_sources.py:40).

    S: complex (F, T).  Returns float32 (hop*(T-1),).
    Per frame: hermitian-extend, inverse FFT, real part, multiply by the
    periodic Hann synthesis window, overlap-add into a float32 buffer; divide by
    the window sum-square where it exceeds tiny(float32); trim n_fft//2 each side.
    """
    S = np.asarray(S)
    F, T = S.shape
    n_fft = 2 * (F - 1)
    w = hann_periodic(n_fft)
    n = n_fft + hop * (T - 1)
    y = np.zeros(n, dtype=np.float32)
    # irfft ignores the imaginary parts of the DC and Nyquist bins, exactly like
    # taking .real of the ifft of [spec, conj(spec[-2:0:-1])].
    frames = np.fft.irfft(S.astype(np.complex128), n=n_fft, axis=0) * w[:, None]
    for i in range(T):
        s = i * hop
        y[s:s + n_fft] = y[s:s + n_fft] + frames[:, i]
    wss = window_sumsquare(T, n_fft, hop)
    nz = wss > np.finfo(np.float32).tiny
    y[nz] /= wss[nz]
    return y[n_fft // 2: n - n_fft // 2]


def masked_istft_int16(mix_spec, mask, hop=128):
    """reconstruct_sources.py:39-42: (mix_spec*mask) -> istft -> *32767 -> astype(int16).

    astype('int16') truncates toward zero and does not clip (values beyond the
    int16 range wrap; synthetic tests stay in range).
    """
    S = np.multiply(mix_spec, mask)
    s = istft(S, hop)
    wav = s * 32767
    return wav.astype("int16")


def masked_istft_float(mix_spec, mask, hop=128):
    """Float value (s*32767) just before the int16 cast; used to reason about +-1 LSB."""
    return istft(np.multiply(mix_spec, mask), hop) * 32767
