"""CPU checks of the evaluator's host arithmetic: SI-SDR from the five device sums equals the direct definition, and
the oracle masks behave like steps/evaluate_oracle.py:133-139 (np.divide / np.argmax semantics)."""
import numpy as np

from oracle import eval_oracle
from ssb200 import synth
from ssb200.evaluate_oracle import si_sdr_from_sums


def test_si_sdr_from_sums_matches_definition():
    rng = np.random.default_rng(0)
    for n in (100, 8000):
        r = rng.standard_normal(n) + 0.3
        e = 0.7 * r + 0.1 * rng.standard_normal(n) - 0.2
        sums = np.array([r.sum(), e.sum(), (r * r).sum(), (e * e).sum(), (r * e).sum(), n], dtype=np.float64)
        assert abs(si_sdr_from_sums(sums) - eval_oracle.si_sdr(r, e)) < 1e-9
        assert abs(eval_oracle.si_sdr(r, 3.0 * e + 5.0) - eval_oracle.si_sdr(r, e)) < 1e-9      # scale / offset invariant


def test_oracle_masks_semantics():
    mix, srcs = synth.mixture(7, 4000, 2)
    spec, soft = eval_oracle.oracle_masks(mix, srcs, 256, 64, hard=False)
    _, hard = eval_oracle.oracle_masks(mix, srcs, 256, 64, hard=True)
    assert soft.shape == (2,) + spec.shape and hard.shape == soft.shape
    assert np.array_equal(hard.sum(0), np.ones(spec.shape))          # exactly one source owns every bin
    est = eval_oracle.oracle_estimates(mix, srcs, 256, 64)
    assert est.shape == (2, 4000) and np.all(est[:, 64 * (4000 // 64):] == 0)
    for i in range(2):                                                # the oracle mask improves on the raw mixture
        assert eval_oracle.si_sdr(srcs[i], est[i]) > eval_oracle.si_sdr(srcs[i], mix)


def test_bss_eval_known_answers():
    """The BSS Eval restatement on cases whose decomposition is known in closed form (white sources, flen >= the filter):
    an estimate that is a short FIR of its own source plus a fraction of the other one has SIR = filter energy /
    leak energy and a very large SAR; white noise added to a clean source gives SAR = SDR = 1 / noise power."""
    rng = np.random.default_rng(0)
    N = 6000
    s = rng.standard_normal((2, N))
    h = np.array([0.8, 0.3, -0.1])
    e0 = np.convolve(s[0], h)[:N] + 0.1 * s[1]
    e1 = s[1] + 0.05 * rng.standard_normal(N)
    sdr, sir, sar = eval_oracle.bss_eval_sources(s, [e0, e1], flen=32)
    assert abs(sir[0] - 10 * np.log10((h ** 2).sum() / 0.01)) < 0.3 and sar[0] > 45 and abs(sdr[0] - sir[0]) < 0.1
    assert abs(sar[1] - 10 * np.log10(1 / 0.0025)) < 0.3 and sir[1] > 35 and abs(sdr[1] - sar[1]) < 0.2
    # a pure delay + gain of the right source is NOT a distortion for BSS Eval (time-invariant filter allowed)
    e = np.concatenate([np.zeros(5), 0.5 * s[0][:-5]])
    sdr, _, _ = eval_oracle.bss_eval_sources(s, [e, s[1]], flen=32)
    assert sdr[0] > 35 and sdr[1] > 100          # (the 5 samples shifted out of the window are the only "distortion")
