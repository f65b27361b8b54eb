"""CPU restatement of the oracle-mask evaluator.  TEST INFRASTRUCTURE ONLY (tests/ compare the CUDA path with it).

Follows steps/evaluate_oracle.py:120-143 of the reference: mixture STFT, source magnitude STFTs, ideal ratio mask
|S_i| / |X| (np.divide) or ideal binary mask (np.argmax over sources), S = X * mask, iSTFT, estimate zero-padded to the
source length.  STFT / iSTFT are oracle/stft_oracle.py (the librosa 0.6 restatement).

Scores.  The reference scores with mir_eval.separation.bss_eval_sources (evaluate_oracle.py:144), a third-party
dependency that is absent from /root/reference and from this image -- PARITY UNPINNED.  The score
restated here is the published scale-invariant SDR (Le Roux et al., "SDR - half-baked or well done?", ICASSP 2019):
    alpha = <e, r> / <r, r>;  SI-SDR = 10 log10( |alpha r|^2 / |e - alpha r|^2 ),  both signals zero-mean.
"""
import numpy as np

from oracle import stft_oracle


def oracle_masks(mix, sources, n_fft, hop, hard=False):
    """mix (N,), sources [S x (N,)] float32 -> (mix_spec (F,T) c64, masks (S,F,T) float)."""
    mix_spec = stft_oracle.stft(mix, n_fft, hop)
    mags = np.stack([np.abs(stft_oracle.stft(s, n_fft, hop)) for s in sources])
    if hard:
        best = np.argmax(mags, axis=0)
        masks = np.stack([(best == i).astype(np.float32) for i in range(len(sources))])
    else:
        with np.errstate(divide="ignore", invalid="ignore"):
            masks = np.stack([np.divide(mags[i], np.abs(mix_spec)) for i in range(len(sources))])
        masks[:, np.abs(mix_spec) == 0] = 0.0        # (exactly-zero mixture bins contribute nothing; see csrc/eval.cu)
    return mix_spec, masks


def oracle_estimates(mix, sources, n_fft, hop, hard=False):
    """Estimated sources (S, N) float64: masked iSTFT, zero beyond hop*(T-1) samples like evaluate_oracle.py:142-143."""
    mix_spec, masks = oracle_masks(mix, sources, n_fft, hop, hard)
    est = np.zeros((len(sources), len(mix)))
    for i in range(len(sources)):
        s = stft_oracle.istft(np.multiply(mix_spec, masks[i]), hop)
        est[i, :len(s)] = s
    return est


def si_sdr(ref, est):
    ref = np.asarray(ref, dtype=np.float64)
    est = np.asarray(est, dtype=np.float64)
    ref = ref - ref.mean()
    est = est - est.mean()
    alpha = np.dot(est, ref) / np.dot(ref, ref)
    target = alpha * ref
    return 10.0 * np.log10(np.dot(target, target) / np.dot(est - target, est - target))
