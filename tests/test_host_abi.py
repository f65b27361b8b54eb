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
        tab, u0, n_tiles, total = rag.dft_tiles()
        tab, u0 = tab.numpy(), u0.numpy()
        # the row stream: utterance u owns T_u + 3 rows; every tile starts inside the utterance the table names
        assert total == int((rag.T + 3).sum()) and n_tiles == -(-total // 128) and len(u0) == n_tiles
        row_off = np.concatenate([[0], np.cumsum(rag.T + 3)])
        assert np.array_equal(tab[:, 3], row_off[:-1])
        for i in range(n_tiles):
            assert row_off[u0[i]] <= i * 128 < row_off[u0[i] + 1]
        for u in range(len(lens)):
            assert int(tab[u, 0]) == int(rag.sig_off[u]) and int(tab[u, 1]) == int(rag.frame_off[u]) * 129
            assert int(tab[u, 2]) & 0xffffffff == lens[u] and int(tab[u, 2]) >> 32 == rag.T[u]
    with pytest.raises(ValueError):
        ops.Ragged([128], 256, 64, "cpu")                  # reflect padding needs N > n_fft/2


def test_pipeline_order_is_the_collators():
    wavs = [np.zeros(n, dtype=np.float32) for n in (9000, 12000, 7777, 12000, 5000)]
    order = pipeline._order(wavs, 64)
    frames = [1 + len(w) // 64 for w in wavs]
    assert order == [int(i) for i in np.argsort(np.array(frames))[::-1]]
    assert [frames[i] for i in order] == sorted(frames, reverse=True)
