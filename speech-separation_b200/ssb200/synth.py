"""Deterministic synthetic corpus: speech-like band-limited noise.

Shared by tests, bench.py and tests/golden/make_golden.py so that every side
sees the same bytes.  Pure numpy; no reference or oracle code involved.
"""
import numpy as np


def source(rng, n, sr=8000):
    """One source: low-passed Gaussian noise times a slow (2-8 Hz) envelope, peak in [0.05,0.5]."""
    x = rng.standard_normal(n + 64)
    k = np.hanning(33)
    k /= k.sum()
    x = np.convolve(x, k, mode="same")[32:32 + n]
    t = np.arange(n) / float(sr)
    fm = rng.uniform(2.0, 8.0)
    ph = rng.uniform(0, 2 * np.pi)
    x = x * (0.5 + 0.5 * np.sin(2 * np.pi * fm * t + ph))
    peak = rng.uniform(0.05, 0.5)
    x = x * (peak / (np.abs(x).max() + 1e-12))
    return x.astype(np.float32)


def mixture(seed, n, num_spk=2, sr=8000):
    """Returns (mix, [s1..sS]) float32, mix = sum of sources (|mix| <= 1)."""
    rng = np.random.default_rng(seed)
    srcs = [source(rng, n, sr) for _ in range(num_spk)]
    mix = np.sum(srcs, axis=0).astype(np.float32)
    return mix, srcs


def batch_lengths_cfg1(B=8, n=32000, step=384):
    """Distinct descending lengths so the collator order is unambiguous (H4)."""
    return [n - step * b for b in range(B)]


def uniform_lengths(seed, B, lo, hi):
    rng = np.random.default_rng(seed)
    return sorted((int(v) for v in rng.integers(lo, hi + 1, size=B)), reverse=True)
