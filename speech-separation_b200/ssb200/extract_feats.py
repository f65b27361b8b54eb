#!/usr/bin/env python3
"""Drop-in for the reference's steps/extract_feats.py (same command line, same files written), with the STFT of a
whole shard done by ONE batched CUDA launch (libssb200 ss_stft) instead of one librosa call per file.

    python -m ssb200.extract_feats <data-dir> <train|test> <feat-dir> [--fft-dim 512] [--step-size 128]
                                   [--sample-rate 8000] [--max-batch-samples N]

Contract mirrored (steps/extract_feats.py:36-111): reads <data-dir>/wav.scp[.$SGE_TASK_ID] ("<id> <path>"), optional
segments[.k] ("<seg> <reco> <t0> <t1>"); sources of a mixture are found by globbing path.replace("/mix/","/*/")
sorted (mixture first); train -> <feat-dir>/<id>.npz with float32 |STFT| arrays mix,s1.. shaped (F,T); test ->
complex64 `mix` only; writes feats_<type>.scp[.k] and utt2num_spk[.k].  Audio decode: PCM/float wav via
scipy.io.wavfile, int16 scaled by 1/32768 like librosa.load; files whose rate differs from --sample-rate are
rejected (the reference would resample with resampy, which is file IO outside the accelerated path).
"""
import argparse
import glob
import os

import numpy as np


def get_args(argv=None):
    ap = argparse.ArgumentParser(description="STFT feature extraction on the GPU (reference-compatible files)")
    ap.add_argument("data_dir", metavar="data-dir")
    ap.add_argument("data_type", metavar="data-type", choices=["train", "test"])
    ap.add_argument("feat_dir", metavar="feat-dir")
    ap.add_argument("--fft-dim", type=int, default=512)
    ap.add_argument("--step-size", type=int, default=128)
    ap.add_argument("--sample-rate", type=int, default=8000)
    ap.add_argument("--max-batch-samples", type=int, default=1 << 28, help="samples per GPU launch (memory bound)")
    return ap.parse_args(argv)


def load_wav(path, sr, offset=None, duration=None):
    """float32 mono in [-1,1) like librosa.core.load(path, sr=sr[, offset, duration])."""
    from scipy.io import wavfile
    fs, x = wavfile.read(path)
    if fs != sr:
        raise ValueError("%s: sample rate %d != --sample-rate %d (resampling is not part of this tool)" % (path, fs, sr))
    if x.dtype == np.int16:
        x = x.astype(np.float32) / 32768.0
    elif x.dtype == np.int32:
        x = x.astype(np.float32) / 2147483648.0
    elif x.dtype == np.uint8:
        x = (x.astype(np.float32) - 128.0) / 128.0
    else:
        x = x.astype(np.float32)
    if x.ndim > 1:
        x = x.mean(axis=1)
    if offset is not None:
        a = int(round(offset * sr))
        x = x[a:a + int(round(duration * sr))]
    return np.ascontiguousarray(x, dtype=np.float32)


def job_suffix():
    tid = os.environ.get("SGE_TASK_ID", "undefined")
    return "" if tid == "undefined" else "." + tid


def plan(args, suffix):
    """-> list of (out_id, num_spk, [(key, path, offset, duration)])"""
    segs = {}
    seg_path = os.path.join(args.data_dir, "segments" + suffix)
    use_segs = os.path.isfile(seg_path)                        # steps/extract_feats.py:51-60
    if use_segs:
        for line in open(seg_path):
            s = line.split()
            segs.setdefault(s[1], []).append((s[0], float(s[2]), float(s[3])))
    items = []
    for line in open(os.path.join(args.data_dir, "wav.scp" + suffix)):
        reco, filename = line.rstrip().split(" ")
        files = sorted(glob.glob(filename.replace("/mix/", "/*/")))
        num_spk = max(1, len(files) - 1)
        if args.data_type == "train":
            keyed = [("mix" if i == 0 else "s%d" % i, f) for i, f in enumerate(files)]
        else:
            keyed = [("mix", filename)]
        if use_segs:
            for sid, t0, t1 in segs[reco]:                     # KeyError for a recording without segments, like :71
                items.append((sid, num_spk, [(k, f, t0, t1 - t0) for k, f in keyed]))
        else:
            items.append((reco, num_spk, [(k, f, None, None) for k, f in keyed]))
    return items


def run_batch(batch, args, featF, spkF):
    """One GPU launch for all signals of `batch` (list of plan items with loaded audio)."""
    import torch
    from . import ops
    sigs = [a for _, _, entries in batch for _, a in entries]
    rag = ops.Ragged([len(a) for a in sigs], args.fft_dim, args.step_size, "cuda")
    flat = torch.from_numpy(np.concatenate(sigs)).pin_memory().cuda(non_blocking=True)
    out = ops.stft(flat, rag, complex_out=(args.data_type == "test"))
    host = out.cpu().numpy()
    F, u = rag.F, 0
    for out_id, num_spk, entries in batch:
        arrays = {}
        for key, _ in entries:
            a, b = int(rag.frame_off[u]) * F, int(rag.frame_off[u + 1]) * F
            arrays[key] = host[a:b].reshape(F, int(rag.T[u]))
            u += 1
        path = os.path.join(args.feat_dir, out_id)
        np.savez_compressed(path, **arrays)
        featF.write(out_id + " " + path + ".npz\n")
        spkF.write(out_id + " " + str(num_spk) + "\n")


def main(argv=None):
    args = get_args(argv)
    suffix = job_suffix()
    os.makedirs(args.feat_dir, exist_ok=True)
    items = plan(args, suffix)
    with open(os.path.join(args.data_dir, "feats_" + args.data_type + ".scp" + suffix), "w") as featF, \
            open(os.path.join(args.data_dir, "utt2num_spk" + suffix), "w") as spkF:
        batch, n = [], 0
        for out_id, num_spk, entries in items:
            loaded = [(k, load_wav(f, args.sample_rate, off, dur)) for k, f, off, dur in entries]
            batch.append((out_id, num_spk, loaded))
            n += sum(len(a) for _, a in loaded)
            if n >= args.max_batch_samples:
                run_batch(batch, args, featF, spkF)
                batch, n = [], 0
        if batch:
            run_batch(batch, args, featF, spkF)


if __name__ == "__main__":
    main()
