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


# ---------------------------------------------------------------------------------------------------------------
# BSS Eval (SDR / SIR / SAR with a time-invariant distortion filter of `flen` taps), the score the reference computes
# with mir_eval.separation.bss_eval_sources(..., compute_permutation=False) (steps/evaluate_oracle.py:144,
# steps/evaluate_sources.py:57).  mir_eval is a third-party dependency absent from /root/reference and from this image
# (its version is not pinned by the reference either) -- PARITY UNPINNED.  Restated from the published algorithm
# (Vincent, Gribonval, Fevotte, "Performance measurement in blind audio source separation", IEEE TASLP 2006; the
# bss_eval_sources / _bss_decomp_mtifilt / _project / _bss_source_crit decomposition of BSS Eval v3):
#   s_target + e_spat = projection of the estimate on span{ s_j(. - l), 0 <= l < flen }
#   e_interf          = projection on span{ s_i(. - l), all i }  minus that
#   e_artif           = the rest
#   SDR = |s_t + e_spat|^2 / |e_interf + e_artif|^2,  SIR = |s_t + e_spat|^2 / |e_interf|^2,
#   SAR = |s_t + e_spat + e_interf|^2 / |e_artif|^2       (10 log10 of each)
def _xcorr(a, b, lags):
    """c[l] = sum_m a[m + l] b[m] for l in lags (linear, zero outside the signals), float64."""
    full = np.correlate(np.asarray(a, np.float64), np.asarray(b, np.float64), mode="full")    # index k: lag k - (len(b)-1)
    z = len(b) - 1
    return np.array([full[l + z] if 0 <= l + z < len(full) else 0.0 for l in lags])


def _project(refs, est, flen):
    """Least-squares projection of est (padded to N+flen-1) on the delayed references; returns the projection."""
    refs = np.asarray(refs, np.float64)
    S, N = refs.shape
    est = np.concatenate([np.asarray(est, np.float64), np.zeros(N + flen - 1 - len(est))])
    G = np.zeros((S * flen, S * flen))
    idx = np.arange(flen)
    for i in range(S):
        for j in range(i, S):
            c = _xcorr(refs[i], refs[j], range(-(flen - 1), flen))          # c[tau + flen - 1] = sum_m s_i[m+tau] s_j[m]
            blk = c[(idx[None, :] - idx[:, None]) + flen - 1]               # blk[a][b] = c[b - a]
            G[i * flen:(i + 1) * flen, j * flen:(j + 1) * flen] = blk
            G[j * flen:(j + 1) * flen, i * flen:(i + 1) * flen] = blk.T
    D = np.concatenate([_xcorr(est, refs[i], range(flen)) for i in range(S)])   # D_i[l] = <est, s_i(. - l)>
    try:
        C = np.linalg.solve(G, D).reshape(S, flen)
    except np.linalg.LinAlgError:
        C = np.linalg.lstsq(G, D, rcond=None)[0].reshape(S, flen)
    proj = np.zeros(N + flen - 1)
    for i in range(S):
        proj += np.convolve(C[i], refs[i])[:N + flen - 1]
    return proj


def bss_eval_sources(reference_sources, estimated_sources, flen=512):
    """(sdr, sir, sar) arrays, estimate j scored against reference j (no permutation search)."""
    refs = np.asarray(reference_sources, np.float64)
    S, N = refs.shape
    out = np.zeros((3, S))
    for j in range(S):
        se = np.concatenate([np.asarray(estimated_sources[j], np.float64), np.zeros(flen - 1)])
        p1 = _project(refs[j:j + 1], estimated_sources[j], flen)            # s_target + e_spat
        p2 = _project(refs, estimated_sources[j], flen)
        e_interf, e_artif = p2 - p1, se - p2
        db = lambda num, den: 10.0 * np.log10(num / den) if den > 0 else np.inf
        out[0, j] = db(np.sum(p1 ** 2), np.sum((e_interf + e_artif) ** 2))
        out[1, j] = db(np.sum(p1 ** 2), np.sum(e_interf ** 2))
        out[2, j] = db(np.sum((p1 + e_interf) ** 2), np.sum(e_artif ** 2))
    return out[0], out[1], out[2]
