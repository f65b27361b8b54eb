"""CPU baseline of the uPIT train step.  TEST / BENCH INFRASTRUCTURE ONLY (bench.py cpu_baseline leg and
`--impl reference`; tests check it against oracle/upit_oracle.py).  The product path never imports it.

The reference's own files cannot travel to the GPU box (/root/reference is absent there and its sources may not be
copied), so this module issues the SAME torch library calls the reference's step makes on a CPU --

  archs/uPIT.py:115-119,129-147   nn.LSTM(F, 600, 2 layers, bidirectional) on a PackedSequence -> pad ->
                                  BatchNorm1d(1200) over (B,1200,T) -> Linear -> sigmoid
  archs/uPIT.py:178-206           mask * tiled mixture, squared error per permutation, min over permutations
  steps/train_qsub.py:120-122     backward, clip_grad_norm_(0.25), Adam(lr=1e-3).step()

-- so that its timing is what the reference's CPU path costs ("port", not the unmodified reference).
"""
import itertools

import torch
import torch.nn as nn
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

H = 600


class CpuNet(nn.Module):
    def __init__(self, feat_dim, num_spk):
        super().__init__()
        self.blstm = nn.LSTM(feat_dim, H, num_layers=2, bidirectional=True)
        self.lin = nn.Linear(2 * H, feat_dim * num_spk)
        self.bn = nn.BatchNorm1d(2 * H)
        self.num_spk = num_spk

    def forward(self, mix, lens, h0, c0):
        """mix (T,B,F) zero padded, lens descending.  Returns masks (T,B,F*S)."""
        packed = pack_padded_sequence(mix, torch.as_tensor(lens, dtype=torch.int64))
        out, _ = self.blstm(packed, (h0, c0))
        out, _ = pad_packed_sequence(out, total_length=mix.shape[0])          # (T,B,2H), zeros beyond lens
        y = self.bn(out.permute(1, 2, 0)).permute(2, 0, 1)                     # statistics over (B,T) incl. padding
        return torch.sigmoid(self.lin(y))


def upit_loss(mask, mix, sources, lens):
    S = len(sources)
    masked = mask * torch.cat([mix] * S, dim=2)
    losses = torch.stack([((masked - torch.cat([sources[i] for i in perm], dim=2)) ** 2).sum((0, 2))
                          for perm in itertools.permutations(range(S))])
    best, idx = torch.min(losses, 0)
    norm = float(sum(lens)) * mix.shape[2]
    return best.sum() / S / norm, norm, idx


class CpuTrainer:
    def __init__(self, feat_dim, num_spk, state_dict=None, dtype=torch.float32):
        self.net = CpuNet(feat_dim, num_spk).to(dtype)
        if state_dict is not None:
            self.net.load_state_dict(state_dict)
        self.opt = torch.optim.Adam(self.net.parameters(), lr=1e-3)

    def step(self, mix, sources, lens, h0, c0):
        self.net.zero_grad()
        mask = self.net(mix, lens, h0, c0)
        loss, norm, idx = upit_loss(mask, mix, sources, lens)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(self.net.parameters(), 0.25)
        self.opt.step()
        return float(loss.detach()), norm, idx
