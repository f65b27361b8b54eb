#!/usr/bin/env python3
"""Drop-in for the reference's `steps/evaluate_oracle.py`: oracle (ideal ratio / ideal binary mask) separation scores of
a data directory, computed on the device.

    python -m ssb200.evaluate_oracle <data-dir> [--hard-mask] [--fft-dim 512] [--step-size 128] [--sample-rate 8000]

Per recording (steps/evaluate_oracle.py:120-144): mixture STFT, magnitude STFTs of the sources, mask_i = |S_i| / |X| or
the arg-max indicator, S_i = X * mask_i, iSTFT, score against the true sources.  Here a batch of recordings goes through
ss_stft (K1) -> ss_oracle_masks -> ss_istft_masked (K7, the very kernels of the separation path) -> ss_sdr_sums, with one
upload of the waveforms and one read-back of 5 sums per (recording, source).

Scores.  The reference calls mir_eval.separation.bss_eval_sources (SDR / SIR / SAR with a 512-tap distortion filter,
no permutation search); mir_eval is a third-party dependency that is neither under /root/reference nor in this image
(PARITY UNPINNED).  Its published algorithm (BSS Eval v3) is restated in oracle/eval_oracle.py and
computed here with the O(N * flen) parts on the device (ss_xcorr, ss_bss_proj_sums) and the least-squares solves on the
host; the six files of the reference (`session_{SDR,SIR,SAR}s.txt`: `<ID> <mean>`; `source_*.txt`: `<ID> <v1> <v2> ..`;
`.k` suffix under $SGE_TASK_ID) are written in its line format, plus the scale-invariant SDR of every estimated source
(session_SISDRs.txt / source_SISDRs.txt).  Both are checked against the numpy oracle in tests/.
"""
import argparse
import glob
import os

import numpy as np
import torch

from . import ops
from ._lib import call, stream_ptr


def get_args(argv=None):
    parser = argparse.ArgumentParser(description="""Computes oracle SI-SDR for a source separation data directory""")
    parser.add_argument("data_dir", metavar="data-dir", type=str, help="Data directory")
    parser.add_argument("--hard-mask", action="store_true", help="Use hard mask", default=False)
    parser.add_argument("--fft-dim", type=int, help="Dimension of FFT", default=512)
    parser.add_argument("--step-size", type=int, help="STFT step size", default=128)
    parser.add_argument("--sample-rate", type=int, help="Audio sample rate", default=8000)
    parser.add_argument("--batch", type=int, help="recordings per device batch (extension)", default=64)
    parser.add_argument("--bss-filter", type=int, default=512,
                        help="distortion-filter length of the BSS Eval scores (mir_eval default 512; 0 = SI-SDR only)")
    return parser.parse_args(argv)


def si_sdr_from_sums(sums):
    """sums (..., 5) fp64 = {sum r, sum e, sum r^2, sum e^2, sum r*e} over n samples (n in sums[..., 5] if present) ->
    zero-mean SI-SDR in dB."""
    sr, se, srr, see, sre, n = (sums[..., k] for k in range(6))
    crr = srr - sr * sr / n
    cee = see - se * se / n
    cre = sre - sr * se / n
    target = cre * cre / crr                      # |alpha r|^2 with alpha = <e,r>/<r,r>
    return 10.0 * np.log10(target / (cee - target))


def bss_eval(ref_all, ref_off, est, est_off, n_ref, n_est, S, flen=512):
    """BSS Eval SDR / SIR / SAR of n recordings x S estimates (estimate j against reference j, no permutation search; the
    decomposition of mir_eval.separation.bss_eval_sources, restated in oracle/eval_oracle.py -- parity unpinned).
    Device: every inner product between delayed signals (ss_xcorr) and the projections' energies (ss_bss_proj_sums);
    host: the (S*flen)^2 and flen^2 least-squares solves, fp64.  Items are ordered (recording, source) like ss_sdr_sums.
    Returns (sdr, sir, sar), each (n, S)."""
    dev = ref_all.device
    n = len(n_ref) // S
    max_ref = int(max(n_ref))
    dv = lambda a, dt: (ops._i64 if dt == np.int64 else ops._i32)(np.asarray(a, dtype=dt), dev)
    # reference x reference, lags -(flen-1) .. flen-1, pairs i <= j of every recording
    pairs = [(r, i, j) for r in range(n) for i in range(S) for j in range(i, S)]
    a_off = dv([ref_off[r * S + i] for r, i, j in pairs], np.int64)
    b_off = dv([ref_off[r * S + j] for r, i, j in pairs], np.int64)
    lens = dv([n_ref[r * S] for r, i, j in pairs], np.int32)
    nl = 2 * flen - 1
    rr = torch.empty(len(pairs), nl, dtype=torch.float64, device=dev)
    call("xcorr", ref_all, a_off, lens, ref_all, b_off, lens, len(pairs), max_ref, -(flen - 1), nl, rr, stream_ptr())
    # estimate x reference, lags 0 .. flen-1: D[(r, e), i][l] = <est_e, s_i delayed by l>
    trip = [(r, e, i) for r in range(n) for e in range(S) for i in range(S)]
    ea_off = dv([est_off[r * S + e] for r, e, i in trip], np.int64)
    ea_len = dv([n_est[r * S + e] for r, e, i in trip], np.int32)
    eb_off = dv([ref_off[r * S + i] for r, e, i in trip], np.int64)
    eb_len = dv([n_ref[r * S + i] for r, e, i in trip], np.int32)
    er = torch.empty(len(trip), flen, dtype=torch.float64, device=dev)
    call("xcorr", est, ea_off, ea_len, ref_all, eb_off, eb_len, len(trip), max_ref, 0, flen, er, stream_ptr())
    rr_h, er_h = rr.cpu().numpy(), er.cpu().numpy().reshape(n, S, S, flen)
    idx = np.arange(flen)
    toe = (idx[None, :] - idx[:, None]) + flen - 1                 # blk[a][b] = c[b - a]
    c1 = np.zeros((n, S, flen))
    c2 = np.zeros((n, S, S, flen))
    pi = 0
    for r in range(n):
        G = np.zeros((S * flen, S * flen))
        for i in range(S):
            for j in range(i, S):
                blk = rr_h[pi][toe]
                pi += 1
                G[i * flen:(i + 1) * flen, j * flen:(j + 1) * flen] = blk
                G[j * flen:(j + 1) * flen, i * flen:(i + 1) * flen] = blk.T
        for e in range(S):
            D = er_h[r, e].reshape(S * flen)
            try:
                c2[r, e] = np.linalg.solve(G, D).reshape(S, flen)
                c1[r, e] = np.linalg.solve(G[e * flen:(e + 1) * flen, e * flen:(e + 1) * flen], er_h[r, e, e])
            except np.linalg.LinAlgError:
                c2[r, e] = np.linalg.lstsq(G, D, rcond=None)[0].reshape(S, flen)
                c1[r, e] = np.linalg.lstsq(G[e * flen:(e + 1) * flen, e * flen:(e + 1) * flen], er_h[r, e, e], rcond=None)[0]
    sums = torch.empty(n * S, 5, dtype=torch.float64, device=dev)
    call("bss_proj_sums", ref_all, dv(ref_off, np.int64), est, dv(est_off, np.int64), dv(n_ref, np.int32), dv(n_est, np.int32),
         torch.from_numpy(c1).to(dev), torch.from_numpy(c2).to(dev), n * S, S, flen, max_ref, sums, stream_ptr())
    q = sums.cpu().numpy().reshape(n, S, 5)
    with np.errstate(divide="ignore"):
        db = lambda num, den: 10.0 * np.log10(num / den)
        return db(q[..., 0], q[..., 1]), db(q[..., 0], q[..., 2]), db(q[..., 3], q[..., 4])


def oracle_separate(mixes, sources, n_fft=512, hop=128, hard=False, device="cuda", bss_flen=0):
    """mixes: list of float32 waveforms; sources[i]: list of that recording's source waveforms (same length as the mix,
    the same number S for the whole batch).  Returns (estimates, si_sdr): estimates[i][s] float32 of hop*(T-1) samples,
    si_sdr (n, S) in dB against the true sources over the full recording (estimate zero beyond its end).  With
    bss_flen > 0 a third value: (sdr, sir, sar) of BSS Eval with that filter length (mir_eval's default: 512)."""
    n, S = len(mixes), len(sources[0])
    lens = [len(m) for m in mixes]
    rag = ops.Ragged(lens, n_fft, hop, device)
    F = rag.F
    up = lambda ws: torch.from_numpy(np.concatenate([np.asarray(w, dtype=np.float32) for w in ws])).to(device)
    spec = ops.stft(up(mixes), rag, complex_out=True)
    refs = [up([sources[i][s] for i in range(n)]) for s in range(S)]
    mags = torch.empty(S, rag.total_frames * F, dtype=torch.float32, device=device)
    for s in range(S):
        ops.stft(refs[s], rag, out=mags[s])
    masks = torch.empty(rag.total_frames * F * S, dtype=torch.float32, device=device)
    call("oracle_masks", spec, mags, ops.LL(mags.stride(0)), rag.d_frame_off, n, int(rag.T.max()), F, S, int(bool(hard)),
         masks, stream_ptr())
    est, est_off = ops.istft_masked(spec, masks, rag, S, out_f32=True)
    # items ordered (recording, source)
    ref_all = torch.cat(refs)
    N_tot = int(rag.sig_off[-1])
    ref_off = np.array([s * N_tot + int(rag.sig_off[i]) for i in range(n) for s in range(S)], dtype=np.int64)
    e_off = np.array([int(est_off[i]) + s * int(rag.L[i]) for i in range(n) for s in range(S)], dtype=np.int64)
    n_ref = np.repeat(np.asarray(lens, dtype=np.int32), S)
    n_est = np.repeat(rag.L.astype(np.int32), S)
    sums = torch.empty(n * S, 5, dtype=torch.float64, device=device)
    call("sdr_sums", ref_all, ops._i64(ref_off, device), est, ops._i64(e_off, device), ops._i32(n_ref, device),
         ops._i32(n_est, device), n * S, int(max(lens)), sums, stream_ptr())
    sums = np.concatenate([sums.cpu().numpy(), n_ref[:, None].astype(np.float64)], axis=1).reshape(n, S, 6)
    est_h = est.cpu().numpy()
    outs = [[est_h[e_off[i * S + s]: e_off[i * S + s] + int(rag.L[i])] for s in range(S)] for i in range(n)]
    if bss_flen:
        return outs, si_sdr_from_sums(sums), bss_eval(ref_all, ref_off, est, e_off, n_ref, n_est, S, bss_flen)
    return outs, si_sdr_from_sums(sums)


def _load(path, sr, offset=None, duration=None):
    """int16 / float wav -> float32 in [-1, 1] (what librosa.core.load yields for a file already at `sr`)."""
    from scipy.io import wavfile
    fs, x = wavfile.read(path)
    if fs != sr:
        raise ValueError("%s: sample rate %d != --sample-rate %d (resampling is not part of the hot path)" % (path, fs, sr))
    if x.ndim > 1:
        x = x.mean(axis=1)
    if x.dtype == np.int16:
        x = x.astype(np.float32) / 32768.0
    elif x.dtype == np.int32:
        x = x.astype(np.float32) / 2147483648.0
    else:
        x = x.astype(np.float32)
    if offset is not None:
        a = int(round(offset * sr))
        x = x[a:a + int(round(duration * sr))]
    return x


def main(argv=None):
    args = get_args(argv)
    job_suffix = ""
    if os.environ.get("SGE_TASK_ID", "undefined") != "undefined":
        job_suffix = "." + os.environ["SGE_TASK_ID"]
    dir_out = args.data_dir + ("/oracle_hard_mask_eval/" if args.hard_mask else "/oracle_soft_mask_eval/")
    os.makedirs(dir_out, exist_ok=True)
    seg_dict = None
    if os.path.isfile(args.data_dir + "/segments" + job_suffix):
        seg_dict = {}
        for line in open(args.data_dir + "/segments" + job_suffix):
            seg = line.rstrip().split()
            seg_dict.setdefault(seg[1], []).append((seg[0], float(seg[2]), float(seg[3])))
    # work list: (ID, [mix file, source files...], offset, duration)
    items = []
    with open(args.data_dir + "/wav.scp" + job_suffix) as listF:
        for line in listF:
            reco_id, filename = line.rstrip().split(" ")
            wav_files = sorted(glob.glob(filename.replace("/mix/", "/*/")))
            if seg_dict is None:
                items.append((reco_id, wav_files, None, None))
            else:
                for seg_id, t0, t1 in seg_dict[reco_id]:             # KeyError for an unknown recording, like the reference
                    items.append((seg_id, wav_files, t0, t1 - t0))
    sessionF = open(dir_out + "/session_SISDRs.txt" + job_suffix, "w")
    sourceF = open(dir_out + "/source_SISDRs.txt" + job_suffix, "w")
    # the reference's six files (steps/evaluate_oracle.py:80-87), same line format
    bssF = [(open(dir_out + "/session_%ss.txt" % k + job_suffix, "w"), open(dir_out + "/source_%ss.txt" % k + job_suffix, "w"))
            for k in ("SDR", "SIR", "SAR")] if args.bss_filter else []
    # batches of recordings with the same number of sources
    pos = 0
    while pos < len(items):
        S = len(items[pos][1]) - 1
        batch = []
        while pos < len(items) and len(batch) < args.batch and len(items[pos][1]) - 1 == S:
            batch.append(items[pos])
            pos += 1
        audio = [[_load(f, args.sample_rate, off, dur) for f in files] for _, files, off, dur in batch]
        res = oracle_separate([a[0] for a in audio], [a[1:] for a in audio], args.fft_dim, args.step_size, args.hard_mask,
                              bss_flen=args.bss_filter)
        for b, (ID, _, _, _) in enumerate(batch):
            sc = res[1][b]
            sessionF.write(ID + " " + str(sum(sc) / S) + "\n")
            sourceF.write(ID + "".join(" " + str(v) for v in sc) + "\n")
            for (sesF, srcF), vals in zip(bssF, res[2] if bssF else ()):
                sesF.write(ID + " " + str(sum(vals[b]) / S) + "\n")
                srcF.write(ID + "".join(" " + str(v) for v in vals[b]) + "\n")
    for f in [sessionF, sourceF] + [g for pair in bssF for g in pair]:
        f.close()


if __name__ == "__main__":
    main()
