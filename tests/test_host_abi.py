"""CPU-side checks of the C-ABI boundary: libssb200.so loads, exports every entry point include/ssb200.h declares,
and the host-side tables handed across it (ragged offsets, tensor-core STFT tile table, collator order) are right.
No compute call is made here (no GPU)."""
import os
import re
import subprocess

import numpy as np
import pytest

from ssb200 import _lib, ops, pipeline

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        subprocess.run(["make", "-C", os.path.join(ROOT, "speech-separation_b200", "csrc"), "-j", "8"], check=True,
                       stdout=subprocess.DEVNULL)
    return _lib.lib()


def test_every_declared_entry_point_is_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "ssb200.h")).read()
    names = sorted(set(re.findall(r"\b(ss_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 40
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.ss_version() >= 1
    assert lib.ss_stft_tile_frames() == 32 and lib.ss_stft_dft256_tile_frames() == 128


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libssb200.so")
    with pytest.raises(_lib.SSBError, match="no CPU fallback"):
        _lib.lib()


def test_ragged_offsets_and_dft_tile_table(lib):
    lens = np.array([129, 127 * 64, 128 * 64, 129 * 64 + 5, 300])
    for align in (1, 4):
        rag = ops.Ragged(lens, 256, 64, "cpu", align=align)
        assert np.array_equal(rag.T, 1 + lens // 64)
        assert all(int(o) % align == 0 for o in rag.sig_off)
        assert all(int(rag.sig_off[u + 1] - rag.sig_off[u]) >= lens[u] for u in range(len(lens)))
        tab = rag.dft_tiles().numpy()
        # every frame of every utterance is covered exactly once by tiles of <= 128 frames of one utterance
        seen = {u: np.zeros(int(rag.T[u]), dtype=int) for u in range(len(lens))}
        by_s0 = {int(rag.sig_off[u]): u for u in range(len(lens))}
        for s0, obase, nt, t0 in tab:
            u = by_s0[int(s0)]
            assert int(nt) & 0xffffffff == lens[u] and int(nt) >> 32 == rag.T[u]
            assert int(obase) == int(rag.frame_off[u]) * 129 and t0 % 128 == 0
            seen[u][t0:t0 + 128] += 1
        assert all((v == 1).all() for v in seen.values())
    with pytest.raises(ValueError):
        ops.Ragged([128], 256, 64, "cpu")                  # reflect padding needs N > n_fft/2


def test_pipeline_order_is_the_collators():
    wavs = [np.zeros(n, dtype=np.float32) for n in (9000, 12000, 7777, 12000, 5000)]
    order = pipeline._order(wavs, 64)
    frames = [1 + len(w) // 64 for w in wavs]
    assert order == [int(i) for i in np.argsort(np.array(frames))[::-1]]
    assert [frames[i] for i in order] == sorted(frames, reverse=True)
