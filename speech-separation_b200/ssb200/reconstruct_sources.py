#!/usr/bin/env python3
"""Drop-in for the reference's steps/reconstruct_sources.py (same command line and output files), with
mask x mixture-STFT -> iSTFT -> *32767 -> int16 done by ONE batched CUDA launch per shard (libssb200 ss_istft_masked).

    python -m ssb200.reconstruct_sources <data-dir> <exp-dir> [--step-size 128] [--sample-rate 8000]

Contract mirrored (steps/reconstruct_sources.py:29-42): for every "<id> <npz>" line of <data-dir>/feats_test.scp the
complex64 `mix` (F,T) is multiplied by every array of <exp-dir>/masks/<id>.npz (keys s1..sS) and written to
<exp-dir>/wav/<key>/<id>.wav as int16 (truncating cast, no clipping) at --sample-rate.
"""
import argparse
import os

import numpy as np


def get_args(argv=None):
    ap = argparse.ArgumentParser(description="masked iSTFT reconstruction on the GPU (reference-compatible files)")
    ap.add_argument("data_dir", metavar="data-dir")
    ap.add_argument("exp_dir", metavar="exp-dir")
    ap.add_argument("--step-size", type=int, default=128)
    ap.add_argument("--sample-rate", type=int, default=8000)
    ap.add_argument("--max-batch-frames", type=int, default=1 << 21, help="STFT frames per GPU launch (memory bound)")
    return ap.parse_args(argv)


def run_batch(batch, args):
    """batch: list of (id, spec (F,T) complex64, [(key, mask (F,T) float32)]) with the same number of masks each."""
    import torch
    from scipy.io import wavfile
    from . import ops
    n_src = len(batch[0][2])
    F = batch[0][1].shape[0]
    n_fft, hop = 2 * (F - 1), args.step_size
    # Ragged wants sample counts; any N with 1 + N//hop == T gives the same frame bookkeeping
    rag = ops.Ragged([max(hop * (s.shape[1] - 1), n_fft // 2 + 1) for _, s, _ in batch], n_fft, hop, "cuda")
    assert all(int(rag.T[u]) == b[1].shape[1] for u, b in enumerate(batch))
    spec = torch.from_numpy(np.concatenate([np.ascontiguousarray(s, dtype=np.complex64).ravel() for _, s, _ in batch]))
    mask = torch.from_numpy(np.concatenate([np.ascontiguousarray(m, dtype=np.float32).ravel()
                                            for _, _, ms in batch for _, m in ms]))
    wav, off = ops.istft_masked(spec.pin_memory().cuda(non_blocking=True), mask.pin_memory().cuda(non_blocking=True),
                                rag, n_src)
    wav = wav.cpu().numpy()
    for u, (uid, _, ms) in enumerate(batch):
        L = int(rag.L[u])
        for s, (key, _) in enumerate(ms):
            out = os.path.join(args.exp_dir, "wav", key, uid + ".wav")
            wavfile.write(out, args.sample_rate, wav[off[u] + s * L: off[u] + (s + 1) * L])


def main(argv=None):
    args = get_args(argv)
    batch, frames = [], 0
    for line in open(os.path.join(args.data_dir, "feats_test.scp")):
        uid, path = line.rstrip().split(" ")[:2]
        spec = np.load(path)["mix"]
        masks = np.load(os.path.join(args.exp_dir, "masks", uid + ".npz"))
        ms = [(k, masks[k]) for k in masks.files]
        if batch and (len(ms) != len(batch[0][2]) or spec.shape[0] != batch[0][1].shape[0]
                      or frames >= args.max_batch_frames):
            run_batch(batch, args)
            batch, frames = [], 0
        batch.append((uid, spec, ms))
        frames += spec.shape[1] * len(ms)
    if batch:
        run_batch(batch, args)


if __name__ == "__main__":
    main()
