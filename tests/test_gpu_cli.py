"""Drop-in CLIs (ssb200.extract_feats / ssb200.reconstruct_sources) against the oracle restatement of what the
reference's steps/extract_feats.py:36-111 and steps/reconstruct_sources.py:29-42 write.  Features 1e-4 relative,
wavs +-1 LSB."""
import os

import numpy as np
import pytest
from scipy.io import wavfile

from oracle import stft_oracle as so
from ssb200 import extract_feats, reconstruct_sources, synth

pytestmark = pytest.mark.gpu


def _corpus(root, n_utts=3, S=2):
    ids = []
    for u in range(n_utts):
        mix, srcs = synth.mixture(700 + u, 9000 + 777 * u, S)
        for name, x in [("mix", mix)] + [("s%d" % (i + 1), s) for i, s in enumerate(srcs)]:
            d = os.path.join(root, "wav8k", name)
            os.makedirs(d, exist_ok=True)
            wavfile.write(os.path.join(d, "utt%d.wav" % u), 8000, np.round(x * 32767).astype(np.int16))
        ids.append("utt%d" % u)
    data = os.path.join(root, "data")
    os.makedirs(data, exist_ok=True)
    with open(os.path.join(data, "wav.scp"), "w") as f:
        for i in ids:
            f.write("%s %s\n" % (i, os.path.join(root, "wav8k", "mix", i + ".wav")))
    return data, ids


def _audio(root, name, uid):
    fs, x = wavfile.read(os.path.join(root, "wav8k", name, uid + ".wav"))
    return x.astype(np.float32) / 32768.0


@pytest.mark.parametrize("n_fft,hop", [(512, 128), (256, 64)])
def test_extract_feats_and_reconstruct(tmp_path, monkeypatch, n_fft, hop):
    monkeypatch.delenv("SGE_TASK_ID", raising=False)
    root = str(tmp_path)
    data, ids = _corpus(root)
    feat_tr, feat_tt = os.path.join(root, "feats_tr"), os.path.join(root, "feats_tt")
    extract_feats.main([data, "train", feat_tr, "--fft-dim", str(n_fft), "--step-size", str(hop)])
    lines = open(os.path.join(data, "feats_train.scp")).read().splitlines()
    assert lines == ["%s %s.npz" % (i, os.path.join(feat_tr, i)) for i in ids]
    assert open(os.path.join(data, "utt2num_spk")).read().splitlines() == ["%s 2" % i for i in ids]
    for i in ids:
        z = np.load(os.path.join(feat_tr, i + ".npz"))
        assert sorted(z.files) == ["mix", "s1", "s2"]
        for key in z.files:
            ref = so.stft_mag(_audio(root, key, i), n_fft, hop)
            assert z[key].dtype == np.float32 and z[key].shape == ref.shape
            assert np.abs(z[key] - ref).max() <= 1e-4 * np.abs(ref).max()
    extract_feats.main([data, "test", feat_tt, "--fft-dim", str(n_fft), "--step-size", str(hop)])
    exp = os.path.join(root, "exp")
    os.makedirs(os.path.join(exp, "masks"))
    rng = np.random.default_rng(3)
    masks = {}
    for i in ids:
        z = np.load(os.path.join(feat_tt, i + ".npz"))
        assert z.files == ["mix"] and z["mix"].dtype == np.complex64
        ref = so.stft(_audio(root, "mix", i), n_fft, hop)
        assert np.abs(z["mix"] - ref).max() <= 1e-4 * np.abs(ref).max()
        masks[i] = {"s1": rng.random(ref.shape, dtype=np.float32), "s2": rng.random(ref.shape, dtype=np.float32)}
        np.savez_compressed(os.path.join(exp, "masks", i), **masks[i])
    for k in ("s1", "s2"):
        os.makedirs(os.path.join(exp, "wav", k))             # run_eval.sh:72-76 pre-creates these
    reconstruct_sources.main([data, exp, "--step-size", str(hop)])
    for i in ids:
        spec = np.load(os.path.join(feat_tt, i + ".npz"))["mix"]
        for k in ("s1", "s2"):
            fs, got = wavfile.read(os.path.join(exp, "wav", k, i + ".wav"))
            ref = so.masked_istft_int16(spec, masks[i][k], hop)
            assert fs == 8000 and got.dtype == np.int16 and got.shape == ref.shape
            assert np.abs(got.astype(np.int64) - ref.astype(np.int64)).max() <= 1


def test_extract_feats_segments_and_sge_task_suffix(tmp_path, monkeypatch):
    """The `segments[.k]` branch (steps/extract_feats.py:51-60,70-81) under an SGE array task: inputs and outputs carry the
    `.k` suffix, every segment becomes its own feature file cut at [t0, t1) seconds, and a recording that has no
    segment raises KeyError exactly like the reference's `seg_dict[reco_id]`."""
    root = str(tmp_path)
    data, ids = _corpus(root, n_utts=2)
    os.rename(os.path.join(data, "wav.scp"), os.path.join(data, "wav.scp.3"))
    segs = [("utt0-a", "utt0", 0.1, 0.6), ("utt0-b", "utt0", 0.5, 1.0), ("utt1-a", "utt1", 0.0, 0.75)]
    with open(os.path.join(data, "segments.3"), "w") as f:
        for sid, reco, t0, t1 in segs:
            f.write("%s %s %.2f %.2f\n" % (sid, reco, t0, t1))
    monkeypatch.setenv("SGE_TASK_ID", "3")
    feat = os.path.join(root, "feats")
    extract_feats.main([data, "train", feat])
    lines = open(os.path.join(data, "feats_train.scp.3")).read().splitlines()
    assert [l.split()[0] for l in lines] == [s[0] for s in segs]
    assert open(os.path.join(data, "utt2num_spk.3")).read().splitlines() == ["%s 2" % s[0] for s in segs]
    for sid, reco, t0, t1 in segs:
        z = np.load(os.path.join(feat, sid + ".npz"))
        a = int(round(t0 * 8000))
        n = int(round((t1 - t0) * 8000))
        for key in ("mix", "s1", "s2"):
            ref = so.stft_mag(_audio(root, key, reco)[a:a + n], 512, 128)
            assert z[key].shape == ref.shape and np.abs(z[key] - ref).max() <= 1e-4 * np.abs(ref).max()
    with open(os.path.join(data, "segments.3"), "w") as f:            # utt1 has no segment any more
        f.write("utt0-a utt0 0.10 0.60\n")
    with pytest.raises(KeyError):
        extract_feats.main([data, "test", feat])
