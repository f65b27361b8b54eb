"""f4: oracle-mask evaluation on the device (ssb200/evaluate_oracle.py: ss_stft -> ss_oracle_masks -> ss_istft_masked ->
ss_sdr_sums) against the numpy restatement of steps/evaluate_oracle.py:120-143 (oracle/eval_oracle.py).  The score is
SI-SDR; the reference's mir_eval BSS-eval is not reproducible here (parity unpinned, see the module docstrings)."""
import os

import numpy as np
import pytest

from oracle import eval_oracle
from ssb200 import evaluate_oracle, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_fft,hop,S,hard", [(512, 128, 2, False), (512, 128, 2, True), (256, 64, 3, False), (256, 64, 3, True)])
def test_oracle_masks_estimates_and_si_sdr(n_fft, hop, S, hard):
    lens = [9000, 12345, 8000, 16000]
    waves = [synth.mixture(300 + i, n, S) for i, n in enumerate(lens)]
    est, scores = evaluate_oracle.oracle_separate([w[0] for w in waves], [w[1] for w in waves], n_fft, hop, hard)
    assert scores.shape == (len(lens), S)
    for i, (mix, srcs) in enumerate(waves):
        ref_est = eval_oracle.oracle_estimates(mix, srcs, n_fft, hop, hard)
        for s in range(S):
            L = hop * (len(mix) // hop)
            assert est[i][s].shape == (L,)
            peak = np.abs(ref_est[s]).max()
            if not hard:        # (a hard mask flips on near-ties of the two magnitudes: compare scores only)
                assert np.abs(est[i][s] - ref_est[s][:L]).max() <= 1e-4 * peak, (i, s)
            want = eval_oracle.si_sdr(srcs[s], ref_est[s])
            assert abs(scores[i, s] - want) <= (0.25 if hard else 0.01), (i, s, scores[i, s], want)
            assert scores[i, s] > eval_oracle.si_sdr(srcs[s], mix)     # the oracle mask improves on the raw mixture


def test_evaluate_oracle_cli_files(tmp_path):
    from scipy.io import wavfile
    sr = 8000
    for d in ("mix", "s1", "s2"):
        (tmp_path / "wav" / d).mkdir(parents=True)
    lines = []
    for i, n in enumerate((8000, 11000, 9500)):
        mix, srcs = synth.mixture(400 + i, n, 2)
        for d, x in (("mix", mix), ("s1", srcs[0]), ("s2", srcs[1])):
            wavfile.write(str(tmp_path / "wav" / d / ("u%d.wav" % i)), sr, (x * 32767).astype(np.int16))
        lines.append("u%d %s" % (i, tmp_path / "wav" / "mix" / ("u%d.wav" % i)))
    (tmp_path / "data").mkdir()
    (tmp_path / "data" / "wav.scp").write_text("\n".join(lines) + "\n")
    os.environ.pop("SGE_TASK_ID", None)
    evaluate_oracle.main([str(tmp_path / "data"), "--batch", "2", "--bss-filter", "32"])
    out = tmp_path / "data" / "oracle_soft_mask_eval"
    for k in ("SDR", "SIR", "SAR"):                                    # the reference's six files, its line format
        ses = [l.split() for l in (out / ("session_%ss.txt" % k)).read_text().splitlines()]
        sou = [l.split() for l in (out / ("source_%ss.txt" % k)).read_text().splitlines()]
        assert [x[0] for x in ses] == ["u0", "u1", "u2"] and all(len(x) == 3 for x in sou)
        for a, b in zip(ses, sou):
            assert abs(float(a[1]) - (float(b[1]) + float(b[2])) / 2) < 1e-9 and np.isfinite(float(a[1]))
    sess = [l.split() for l in (out / "session_SISDRs.txt").read_text().splitlines()]
    src = [l.split() for l in (out / "source_SISDRs.txt").read_text().splitlines()]
    assert [s[0] for s in sess] == ["u0", "u1", "u2"] and all(len(s) == 3 for s in src)
    for a, b in zip(sess, src):
        assert abs(float(a[1]) - (float(b[1]) + float(b[2])) / 2) < 1e-9 and np.isfinite(float(a[1]))


@pytest.mark.parametrize("flen", [16, 64])
def test_bss_eval_on_device_matches_restatement(flen):
    """SDR / SIR / SAR with the correlations and projection energies on the device (ss_xcorr, ss_bss_proj_sums) against
    the numpy restatement of BSS Eval (oracle/eval_oracle.py; mir_eval itself is absent: parity unpinned)."""
    lens = [5000, 7000, 4100]
    waves = [synth.mixture(800 + i, n, 2) for i, n in enumerate(lens)]
    est, si, (sdr, sir, sar) = evaluate_oracle.oracle_separate([w[0] for w in waves], [w[1] for w in waves], 256, 64, False,
                                                               bss_flen=flen)
    for i, (mix, srcs) in enumerate(waves):
        full = np.zeros((2, len(mix)))
        for s in range(2):
            full[s, :len(est[i][s])] = est[i][s]                 # zero beyond hop*(T-1), evaluate_oracle.py:142-143
        o_sdr, o_sir, o_sar = eval_oracle.bss_eval_sources(np.stack(srcs), full, flen=flen)
        for s in range(2):
            assert abs(sdr[i, s] - o_sdr[s]) <= 0.02, (i, s, sdr[i, s], o_sdr[s])
            assert abs(sir[i, s] - o_sir[s]) <= 0.05, (i, s, sir[i, s], o_sir[s])
            assert abs(sar[i, s] - o_sar[s]) <= 0.02, (i, s, sar[i, s], o_sar[s])
