"""Shared test helpers: regenerate the golden cases' inputs from seeds."""
import ast
import os

import numpy as np
import torch

from oracle import stft_oracle
from ssb200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
H = 600


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, "ref_%s.npz" % name))
    case = ast.literal_eval(str(z["case"]))
    return z, case


def case_features(case):
    """Per-utterance magnitude features [(T_b,F)] for mix and each source (oracle STFT)."""
    mixes, srcs = [], [[] for _ in range(case["S"])]
    waves = []
    for b, n in enumerate(case["lens"]):
        mix, ss = synth.mixture(1000 * case["seed"] + b, n, case["S"], case.get("sr", 8000))
        waves.append((mix, ss))
        mixes.append(stft_oracle.stft_mag(mix, case["n_fft"], case["hop"]).T.copy())
        for i, s in enumerate(ss):
            srcs[i].append(stft_oracle.stft_mag(s, case["n_fft"], case["hop"]).T.copy())
    return mixes, srcs, waves


def pad_time_major(feats, dtype=torch.float64):
    """list of (T_b,F) (descending T_b) -> (T,B,F) zero padded, lens."""
    lens = [f.shape[0] for f in feats]
    T, B, F = max(lens), len(feats), feats[0].shape[1]
    out = torch.zeros(T, B, F, dtype=dtype)
    for b, f in enumerate(feats):
        out[:lens[b], b] = torch.from_numpy(f).to(dtype)
    return out, lens


def case_state_dict(case, dtype=torch.float64):
    """Weights exactly as the reference SepDNN draws them under manual_seed(seed)
    (module construction order: LSTM, Linear, BatchNorm1d; uPIT.py:115-119)."""
    F = case["n_fft"] // 2 + 1
    torch.manual_seed(case["seed"])
    if case["arch"] == "RSH":
        lstm = torch.nn.LSTM(2 * F, H, num_layers=2, bidirectional=True)
        lin = torch.nn.Linear(2 * H, F)
    else:
        lstm = torch.nn.LSTM(F, H, num_layers=2, bidirectional=True)
        lin = torch.nn.Linear(2 * H, F * case["S"])
    bn = torch.nn.BatchNorm1d(2 * H)
    sd = {}
    for pre, mod in (("blstm.", lstm), ("lin.", lin), ("bn.", bn)):
        for k, v in mod.state_dict().items():
            sd[pre + k] = v.detach().clone().to(dtype) if v.is_floating_point() else v.clone()
    return sd


def case_hidden(case, seed_offset=100, dtype=torch.float64):
    B = len(case["lens"])
    torch.manual_seed(seed_offset + case["seed"])
    h0 = torch.randn(4, B, H)
    c0 = torch.randn(4, B, H)
    return h0.to(dtype), c0.to(dtype)


def grad_stat(name, g):
    g = g.detach().double().cpu().numpy().ravel()
    rng = np.random.default_rng(sum(map(ord, name)))
    r = rng.standard_normal(g.size)
    idx = np.linspace(0, g.size - 1, num=min(256, g.size)).astype(np.int64)
    return np.array([np.linalg.norm(g), float(g @ r)]), g[idx]


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-300))
