"""CPU oracle for the BLSTM mask estimator + uPIT / RSH losses.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl
reference` legs may import this module; the product path never does.

A restatement, in explicit torch CPU tensor arithmetic (any dtype, float64 by
default for a tight comparison), of what the reference computes in

  archs/uPIT.py:97-147   SepDNN: LSTM(F->600, 2 layers, bidirectional) from given
                         (h0,c0) -> zero-padded (B,T,1200) -> BatchNorm1d(1200) over
                         (B,T) *including padding* -> Linear(1200->F*S) -> sigmoid
  archs/uPIT.py:157-206  compute_loss: masked = mask*tile(mix,S); per permutation
                         (itertools.permutations order) sum of squared error per
                         utterance; min/argmin over permutations;
                         loss = sum_b(min)/S / (sum_b(len_b)*F)
  archs/RSH.py:141-187   SepDNN: LSTM(2F->600,...) -> BN -> Linear(1200->F) -> sigmoid
  archs/RSH.py:197-259   compute_loss: S passes, greedy not-yet-used source
                         assignment, residual attention update, state carry-over

It does not call torch.nn.LSTM / BatchNorm1d / MSELoss: the LSTM cell, the packed
sequence semantics (reverse direction starts at each utterance's own last frame,
outputs beyond len_b are zero), the BN statistics and the loss are written out.

PINNING: tests/golden/ref_*.npz hold outputs of the UNMODIFIED reference modules
(/root/reference/archs/uPIT.py, RSH.py) run in the build container by
tests/golden/make_golden.py; tests/test_oracle_upit.py checks this restatement
against them (loss, norm, per-permutation losses, argmin indices, every
gradient tensor's norm / projection / head values).
"""
import itertools

import torch

H = 600  # archs/uPIT.py:115, archs/RSH.py:155 (hard-coded)


def _lstm_dir(x, lens, w_ih, w_hh, b_ih, b_hh, h0, c0, reverse):
    """One direction of one layer over a length-sorted padded batch.

    x: (T,B,I) zero beyond lens; lens: python list, descending.
    Returns out (T,B,H) (zero beyond lens), h_n, c_n (state at each utterance's
    last processed frame) -- torch.nn.LSTM on a PackedSequence (uPIT.py:132).
    Gate order i,f,g,o.
    """
    T, B, _ = x.shape
    gx = x @ w_ih.t() + (b_ih + b_hh)
    h, c = h0, c0
    outs = [None] * T
    steps = range(T - 1, -1, -1) if reverse else range(T)
    for t in steps:
        n = sum(1 for L in lens if L > t)
        g = gx[t, :n] + h[:n] @ w_hh.t()
        gi, gf, gg, go = g.chunk(4, dim=1)
        c_new = torch.sigmoid(gf) * c[:n] + torch.sigmoid(gi) * torch.tanh(gg)
        h_new = torch.sigmoid(go) * torch.tanh(c_new)
        if n < B:
            outs[t] = torch.cat([h_new, torch.zeros(B - n, h_new.shape[1], dtype=x.dtype)], 0)
            h = torch.cat([h_new, h[n:]], 0)
            c = torch.cat([c_new, c[n:]], 0)
        else:
            outs[t] = h_new
            h, c = h_new, c_new
    return torch.stack(outs, 0), h, c


def blstm(x, lens, params, h0, c0, num_layers=2):
    """params: dict with torch.nn.LSTM names (weight_ih_l0, weight_hh_l0_reverse, ...).
    h0,c0: (2*num_layers,B,H), index = 2*layer + direction (torch.nn.LSTM order).
    Returns out (T,B,2H), h_n, c_n."""
    hn, cn = [], []
    inp = x
    for l in range(num_layers):
        outs = []
        for d, suf in enumerate(["", "_reverse"]):
            o, h, c = _lstm_dir(inp, lens,
                                params["weight_ih_l%d%s" % (l, suf)], params["weight_hh_l%d%s" % (l, suf)],
                                params["bias_ih_l%d%s" % (l, suf)], params["bias_hh_l%d%s" % (l, suf)],
                                h0[2 * l + d], c0[2 * l + d], reverse=(d == 1))
            outs.append(o)
            hn.append(h)
            cn.append(c)
        inp = torch.cat(outs, dim=2)
    return inp, torch.stack(hn, 0), torch.stack(cn, 0)


def batchnorm_train(x, weight, bias, eps=1e-5):
    """BatchNorm1d in training mode over all (T,B) positions, padding included
    (uPIT.py:138).  x: (T,B,C).  Returns y, batch mean, biased batch var."""
    n = x.shape[0] * x.shape[1]
    mean = x.sum((0, 1)) / n
    var = ((x - mean) ** 2).sum((0, 1)) / n
    y = (x - mean) / torch.sqrt(var + eps) * weight + bias
    return y, mean, var


def batchnorm_eval(x, weight, bias, running_mean, running_var, eps=1e-5):
    return (x - running_mean) / torch.sqrt(running_var + eps) * weight + bias


def sepdnn_forward(x, lens, sd, h0, c0, training=True, num_layers=2):
    """SepDNN.forward (uPIT.py:129-147 / RSH.py:169-187).

    x: (T,B,I) time-major padded, sd: dict keyed like the reference state_dict
    ('blstm.weight_ih_l0', 'bn.weight', 'lin.weight', ...).
    Returns mask (T,B,F*S) (values at padded positions are what the reference
    computes there: sigmoid(lin(bn(0)))), (h_n,c_n), (batch mean, biased var)."""
    lp = {k[len("blstm."):]: v for k, v in sd.items() if k.startswith("blstm.")}
    out, hn, cn = blstm(x, lens, lp, h0, c0, num_layers)
    if training:
        y, mean, var = batchnorm_train(out, sd["bn.weight"], sd["bn.bias"])
    else:
        y = batchnorm_eval(out, sd["bn.weight"], sd["bn.bias"], sd["bn.running_mean"], sd["bn.running_var"])
        mean = var = None
    z = y @ sd["lin.weight"].t() + sd["lin.bias"]
    return torch.sigmoid(z), (hn, cn), (mean, var)


def bn_running_update(running_mean, running_var, mean, var_biased, n, momentum=0.1):
    """What BatchNorm1d.forward does to its buffers in training mode."""
    unbiased = var_biased * (n / (n - 1.0))
    return ((1 - momentum) * running_mean + momentum * mean,
            (1 - momentum) * running_var + momentum * unbiased)


def upit_loss(mask, mix, sources, lens):
    """uPIT.py:178-206.  mask (T,B,F*S), mix (T,B,F), sources list of S (T,B,F).
    Returns loss (0-d), norm (0-d), losses (P,B), indices (B,)."""
    S = len(sources)
    Fd = mix.shape[2]
    masked = mask * torch.cat([mix] * S, dim=2)
    perms = list(itertools.permutations(range(S)))
    losses = torch.stack([
        ((masked - torch.cat([sources[i] for i in perm], dim=2)) ** 2).sum((0, 2))
        for perm in perms])
    min_losses, indices = torch.min(losses, 0)
    loss = min_losses.sum() / S
    norm = torch.tensor(float(sum(lens)) * Fd, dtype=mask.dtype)
    return loss / norm, norm, losses, indices


def upit_compute_loss(sd, mix, sources, lens, h0, c0, training=True, num_layers=2):
    """End-to-end uPIT compute_loss on time-major padded inputs."""
    mask, _, stats = sepdnn_forward(mix, lens, sd, h0, c0, training, num_layers)
    loss, norm, losses, indices = upit_loss(mask, mix, sources, lens)
    return loss, norm, losses, indices, mask, stats


def rsh_compute_loss(sd, mix, sources, lens, h0, c0, training=True, relu=True):
    """archs/RSH.py:197-259 for ONE sub-batch of equal speaker count S=len(sources).

    mix (T,B,F); combo = [mix || ones] on valid frames, zero on padding (the
    collator packs combo, so padding is zero in both halves, RSH.py:107-109,222).
    Returns loss_sum (un-normalised: sum over passes of sum_b(min)/S), norm_sum,
    per-pass losses list [(S,B)], per-pass indices list [(B,)], per-pass masks."""
    T, B, Fd = mix.shape
    S = len(sources)
    valid = torch.zeros(T, B, 1, dtype=mix.dtype)
    for b, L in enumerate(lens):
        valid[:L, b] = 1
    combo = torch.cat([mix, valid.expand(T, B, Fd)], dim=2)
    h, c = h0, c0
    usage = [[] for _ in range(S)]
    loss = 0
    norm = 0
    all_losses, all_idx, all_masks = [], [], []
    for _ in range(S):
        mask, (h, c), _ = sepdnn_forward(combo, lens, sd, h, c, training)
        masked = mask * mix
        losses = torch.stack([((masked - s) ** 2).sum((0, 2)) for s in sources])
        sel = losses.clone()
        for si in range(S):
            for b in usage[si]:
                sel[si, b] = float("inf")
        min_losses, indices = torch.min(sel, 0)
        for b in range(B):
            usage[int(indices[b])].append(b)
        loss = loss + min_losses.sum() / S
        norm = norm + float(sum(lens)) * Fd
        all_losses.append(sel)
        all_idx.append(indices)
        all_masks.append(mask)
        resid = combo - torch.cat([torch.zeros_like(mask), mask], dim=2)
        if relu:
            resid = torch.relu(resid)          # RSH.py:256 (compute_loss); :280 has no relu (compute_masks)
        combo = resid * valid                  # pack_padded_sequence drops padded frames (:257)
    return loss, norm, all_losses, all_idx, all_masks
