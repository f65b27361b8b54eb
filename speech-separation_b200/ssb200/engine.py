"""Host orchestration of the BLSTM mask estimator step: which kernel runs on which buffer, on which stream.

Mirrors what `SepDNN.forward` + `compute_loss` + `loss.backward()` do in the reference
(archs/uPIT.py:129-147,157-206; archs/RSH.py:169-187,197-259; steps/train_qsub.py:120), but every
arithmetic step is a libssb200 kernel; torch only owns the device buffers and the streams.

Layouts: activations time-major padded, row = t*B + b.  LSTM gate columns interleaved (dir*4H + 4u + gate).

Streaming.  A recurrence kernel occupies 100 (forward) / 80 (backward, K-split) of the 148 SMs for milliseconds and
reports per direction how many frames are finished (`progress`, include/ssb200.h).  The GEMM that consumes its output
does not wait for the kernel to end: it is cut into time chunks x direction halves, each launched on a side stream
behind a stream-ordered wait on the progress counter and capped to the idle SMs --

  forward   x-projection of layer l+1:  gx[l+1][rows] (+)= out[l][rows, dir half] . W_ih[l+1][:, dir half]^T
  backward  dX of layer l:              dout[l-1][rows] (+)= dG[l][rows, dir half] . W_ih[l][dir half, :]
            dW_hh[l], (l = 0: dW_ih[0]) accumulated over row blocks as the gate gradients appear;
            dW_ih[l >= 1] and the head's weight gradients ride under the NEXT recurrence

-- so that only the chunk the recurrence finishes last is left for the main stream.  Weight gradients use the "TN"
GEMM form (ss_gemm_sb16_tn): activations and gate gradients are read as they were written, no transposed copies.
"""
import ctypes
import os

import torch

from . import ops
from ._lib import LL, call, lib, stream_ptr
from .ops import SB16, up8

H = 600          # archs/uPIT.py:115, archs/RSH.py:155 (hard-coded in the reference)
UNIT4 = 1
# stream GEMMs under the recurrences (module switch for A/B tests and profiling; results agree either way)
STREAM = os.environ.get("SSB_STREAM", "1") != "0"
# Time chunks of the streamed GEMMs.  Measured at configs[2] with the 2.7 / 3.1 ms recurrences of this round (ms per train
# step): all three at 16 / 12 / 10 / 8 / 6 / 4 / 3 -> 17.15 / 16.23 / 15.54 / 15.22 / 15.00 / 14.90 / 15.43; around (4, 4, 4):
# forward 4 / 6 / 8 -> 14.85 / 14.88 / 14.81, dX 3 / 4 / 6 -> 14.88 / 14.85 / 14.83, dW 3 / 4 / 6 -> 15.06 / 14.85 / 15.01.
N_CHUNKS = 4         # backward: dX of a layer
N_CHUNKS_FWD = 8     # forward: x-projection of the next layer
N_CHUNKS_DW = 4      # backward: dW_hh (and dW_ih of layer 0), accumulated over row blocks
FWD_STREAM_FRACTION = 0.64   # forward: stream the x-projection halves that are ready within this fraction of the recurrence
# backward: a gated job whose inputs are only ready after this fraction of the recurrence runs AFTER it, on the main stream
# at full width, instead of finishing late on the 68 idle SMs (first entry: layer 0, whose side stream starts with the
# whole dW_ih of the layer above; second: the other layers)
# Measured (configs[2], 3.1 ms recurrences): layer 0 cut at 1.0 / 0.6 / 0.45 / 0.3 -> 15.58 / 15.25 / 15.28 / 15.48 ms per step;
# other layers 1.0 / 0.85 / 0.7 -> 15.28 / 15.35 / 15.61; forward 0.55 / 0.64 / 0.75 -> 15.38 / 15.28 / 15.60.
BWD_STREAM_FRACTION = (0.6, 1.0)
MIN_STREAM_T = 48
SPLIT_WIDE_BATCH = True      # backward recurrence of 64 < B <= 128 as two half-batch launches of the K-split kernel
# The same for the forward recurrence (ss_lstm_fwd takes a sub-batch view too) does NOT pay: two launches at 5.2 us per
# time step against ~9.5 for one pass of the generic 128-row kernel (batch 100: 32.2 instead of 31.3 ms per step).
SPLIT_WIDE_BATCH_FWD = False
GRAPH_REFRESH = os.environ.get("SSB_GRAPH_REFRESH", "1") != "0"


def _sb_view(sb, row0=0, rows=None, col0=0, cols=None):
    """Rows [row0, +rows) x columns [col0, +cols) of an SB16 as a new SB16 sharing storage (pitch and plane stride kept)."""
    v = SB16.__new__(SB16)
    v.rows = sb.rows - row0 if rows is None else rows
    v.cols = sb.cols - col0 if cols is None else cols
    v.ld, v.plane = sb.ld, sb.plane
    v.buf = sb.buf[row0 * sb.ld + col0:]
    return v


def _split_into(x, out, row0=0, outT=None, col0=0, rowperm=0, hidden=0, col_scale=None):
    """ss_split of fp32 (R,C) into rows [row0..) of `out` and/or columns [col0..) of `outT`."""
    R, C = x.shape
    o_ptr = out.buf[row0 * out.ld:] if out is not None else None
    t_ptr = outT.buf[col0:] if outT is not None else None
    call("split", x, LL(x.stride(0)), R, C, rowperm, hidden, col_scale,
         o_ptr, LL(out.ld if out is not None else 0), LL(out.plane if out is not None else 0),
         t_ptr, LL(outT.ld if outT is not None else 0), LL(outT.plane if outT is not None else 0), stream_ptr())


class Weights:
    """Tensor-core operand copies of the parameters (rebuilt after every optimizer step)."""

    def __init__(self, in_dim, out_dim, num_layers, device):
        self.in_dim, self.out_dim, self.L = in_dim, out_dim, num_layers
        self.wih, self.wihT, self.whh_r, self.whhT, self.bias = [], [], [], [], []
        for l in range(num_layers):
            I = in_dim if l == 0 else 2 * H
            self.wih.append(SB16(8 * H, I, device))        # rows dir*4H + 4u+g
            self.wihT.append(SB16(I, 8 * H, device))       # [I][dir*4H + 4u+g]
            self.whh_r.append(SB16(8 * H, H, device))
            self.whhT.append(SB16(2 * H, 4 * H, device))   # row dir*H + j, col 4u+g
            self.bias.append(torch.empty(8 * H, dtype=torch.float32, device=device))
        self.linT = SB16(2 * H, out_dim, device)           # natural W_lin^T for the dY GEMM
        self.lin_folded = SB16(out_dim, 2 * H, device)     # W_lin * diag(bn scale), per batch
        self.bias_folded = torch.empty(out_dim, dtype=torch.float32, device=device)
        self._graphs = {}
        self.graph_replays = 0          # replays so far; each stands for launches_per_refresh() kernel launches

    def launches_per_refresh(self, need_grad_copies=True):
        return self.L * 2 * (3 + (1 if need_grad_copies else 0)) + (1 if need_grad_copies else 0)

    def refresh(self, p, need_grad_copies=True):
        """p: dict name -> fp32 cuda tensor with the reference's state_dict names.  The ~25 tiny split / bias launches
        are launch-latency bound (1.3 ms of host-paced gaps per train step), so they are captured once per set of
        parameter addresses into a CUDA graph and replayed (one launch) afterwards."""
        key = (bool(need_grad_copies),) + tuple(t.data_ptr() for t in p.values())
        g = self._graphs.get(key)
        if g is not None:
            g.replay()
            self.graph_replays += 1
            return
        self._refresh_eager(p, need_grad_copies)
        if not GRAPH_REFRESH:
            return
        if len(self._graphs) >= 4:
            self._graphs.clear()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._refresh_eager(p, need_grad_copies)
        self._graphs[key] = g

    def _refresh_eager(self, p, need_grad_copies):
        for l in range(self.L):
            for d, suf in enumerate(("", "_reverse")):
                w_ih = p["blstm.weight_ih_l%d%s" % (l, suf)]
                w_hh = p["blstm.weight_hh_l%d%s" % (l, suf)]
                _split_into(w_ih, self.wih[l], row0=d * 4 * H, outT=self.wihT[l] if need_grad_copies else None,
                            col0=d * 4 * H, rowperm=UNIT4, hidden=H)
                _split_into(w_hh, self.whh_r[l], row0=d * 4 * H, rowperm=UNIT4, hidden=H)
                if need_grad_copies:
                    _split_into(w_hh, None, outT=_sb_view(self.whhT[l], d * H, H), rowperm=UNIT4, hidden=H)
                call("bias_prep", p["blstm.bias_ih_l%d%s" % (l, suf)], p["blstm.bias_hh_l%d%s" % (l, suf)], 4 * H, H,
                     self.bias[l][d * 4 * H:], stream_ptr())
        if need_grad_copies:
            _split_into(p["lin.weight"], None, outT=self.linT)


def _ws_bytes(B):
    fn = lib().ss_lstm_ws_bytes
    fn.restype = ctypes.c_longlong
    return int(fn(int(B), H))


class Workspace:
    """All activation / gradient buffers of one model step, sized by CAPACITY (rows = T*B, batch B): a shuffled loader
    yields a new T_max almost every batch (steps/train_qsub.py:81), so `bind(T, B)` only re-points the views; nothing is
    allocated or cleared unless a batch exceeds the capacity (SepDNN.workspace then grows it)."""

    def __init__(self, cap_rows, cap_B, in_dim, out_dim, n_src, num_layers, device, training=True):
        self.cap_rows, self.cap_B = int(cap_rows), int(cap_B)
        self.in_dim, self.out_dim, self.L, self.device = in_dim, out_dim, num_layers, device
        R, Bc = self.cap_rows, self.cap_B
        f32 = dict(dtype=torch.float32, device=device)
        self.x0 = SB16(R, in_dim, device)
        self.gx = [torch.empty(R, 8 * H, **f32) for _ in range(num_layers)]
        self.out = [SB16(R, 2 * H, device) for _ in range(num_layers)]
        self._hn = torch.empty(2 * num_layers * Bc * H, **f32)
        self._cn = torch.empty(2 * num_layers * Bc * H, **f32)
        self.lstm_ws = torch.empty(max(_ws_bytes(Bc), _ws_bytes(min(Bc, 64))), dtype=torch.uint8, device=device)
        self.coef = torch.empty(5, 2 * H, **f32)
        self.bn_partial = torch.empty(2 * max(1, lib().ss_bn_parts(R)) * 2 * H, **f32)
        self.ldz = (out_dim + 3) // 4 * 4
        self.z = torch.empty(R, self.ldz, **f32)
        # ss_pit_chunks(T) * B * 16 floats, and chunks(T) * B <= chunks(T * B) + B
        self.pit_partial = torch.empty((lib().ss_pit_chunks(R) + Bc + 1) * 16, **f32)
        self._losses = torch.empty(24 * Bc, **f32)
        self._idx = torch.empty(Bc, dtype=torch.int32, device=device)
        self._assign = torch.empty(Bc * 4, dtype=torch.int32, device=device)
        self._used = torch.zeros(Bc, dtype=torch.int32, device=device)
        self.stats = torch.zeros(4, **f32)
        # progress counters of the recurrence kernels: [fwd dir0, fwd dir1, bwd dir0, bwd dir1]; values only grow
        # data-parallel mode B: {sum x, sum x^2 | B_total, T_max} and the backward pair {sum dy, sum dy*x}, local + global
        self.bn_sums = torch.zeros(3, 2 * (2 * H) + 2, **f32)
        self.progress = torch.zeros(4, dtype=torch.int32, device=device)
        self.prog_base = 0
        self.training = training
        if training:
            self.cs = [torch.empty(R, 2 * H, **f32) for _ in range(num_layers)]
            self.dz = SB16(R, out_dim, device)
            self.dy = torch.empty(R, 2 * H, **f32)            # dY -> dout of the top layer (in place)
            self.dout = torch.empty(R, 2 * H, **f32)          # dout of lower layers
            # one gradient buffer per layer: the upper layer's weight-gradient GEMMs still read theirs while the lower
            # layer's recurrence writes its own (see backward())
            self.dG = [SB16(R, 8 * H, device) for _ in range(num_layers)]
            self.G = torch.empty(out_dim, 2 * H, **f32)
            self.bn_bwd = torch.empty(2, 2 * H, **f32)
            self.cs_partial = [torch.empty(max(1, lib().ss_colsum_parts(R)) * 8 * H, **f32) for _ in range(num_layers + 1)]
            self.db_l = [torch.empty(8 * H, **f32) for _ in range(num_layers)]
            self.bwd_sync = torch.empty(256, dtype=torch.uint8, device=device)
            self.dsum = torch.empty(out_dim, **f32)
            self.dlinW = torch.empty(out_dim, 2 * H, **f32)
            self.dgamma = torch.empty(2 * H, **f32)
            self.dbeta = torch.empty(2 * H, **f32)
        self.T = self.B = self.TB = 0

    def fits(self, T, B):
        return T * B <= self.cap_rows and B <= self.cap_B

    def bind(self, T, B):
        assert self.fits(T, B)
        self.T, self.B, self.TB = int(T), int(B), int(T) * int(B)
        L2 = 2 * self.L
        self.hn = self._hn[:L2 * B * H].view(L2, B, H)
        self.cn = self._cn[:L2 * B * H].view(L2, B, H)
        self.losses = self._losses[:24 * B].view(24, B)
        self.idx = self._idx[:B]
        self.assign = self._assign[:4 * B].view(B, 4)
        self.used = self._used[:B]
        return self

    def next_base(self):
        """A fresh base for a recurrence launch's progress values (cyclic 32-bit comparison on the waiting side)."""
        self.prog_base = (self.prog_base + self.T + 8) & 0xFFFFFFFF
        return self.prog_base


def _gemm(a, b, out, bias=None, alpha=1.0, beta=0.0, rowmap=0, hidden=0, M=None, N=None, K=None, sm_limit=0):
    return ops.gemm(a, b, out=out, bias=bias, alpha=alpha, beta=beta, rowmap=rowmap, hidden=hidden, M=M, N=N, K=K,
                    sm_limit=sm_limit)


_SIDE, _SMS = {}, {}


def _dev_index(device):
    d = torch.device(device)
    return d.index if d.index is not None else torch.cuda.current_device()


def _side_stream(device):
    key = _dev_index(device)
    if key not in _SIDE:
        # high priority: its gated GEMMs are on the step's critical path, while other streams (the second slot of
        # ssb200.pipeline) only fill what is left of the idle SMs
        _SIDE[key] = torch.cuda.Stream(device=device, priority=-1)
    return _SIDE[key]


def idle_sms(device, backward, B):
    """SMs a recurrence kernel leaves free: it runs 100 CTAs forward (2 x 50), 80 backward in the K-split form
    (B <= 64), 76 in the generic backward form, one per SM."""
    key = _dev_index(device)
    if key not in _SMS:
        _SMS[key] = torch.cuda.get_device_properties(key).multi_processor_count
    used = (80 if B <= 64 else 76) if backward else 100
    return max(8, _SMS[key] - used)


def _chunks(T, n):
    """Frame ranges [(a, b)] of n nearly equal time chunks."""
    n = max(1, min(n, T))
    cuts = [round(i * T / n) for i in range(n + 1)]
    return [(cuts[i], cuts[i + 1]) for i in range(n) if cuts[i + 1] > cuts[i]]


class _Streamer:
    """Side-stream jobs gated on the progress counters of the recurrence kernel that is running on the main stream."""

    def __init__(self, ws, device, backward, nsteps, kernel_batch=None, cut=None):
        self.ws, self.device = ws, device
        self.cut = cut
        self.side = _side_stream(device)
        self.flag0 = 2 if backward else 0
        # largest value the kernel is certain to report (include/ssb200.h: every 4th step, never the last ones)
        if backward:
            self.max_reported = 4 * ((nsteps - 1) // 4) if nsteps - 1 >= 4 else -1
        else:
            self.max_reported = 4 * ((nsteps - 1) // 4) - 1 if nsteps - 1 >= 4 else -1
        self.sm_limit = idle_sms(device, backward, ws.B if kernel_batch is None else kernel_batch)
        self.jobs, self.late = [], []
        self.base = ws.next_base()
        # inputs of the gated jobs that precede the recurrence on the main stream: recorded BEFORE the recurrence is
        # launched (an event recorded after it would only fire when the kernel has ended)
        self.before = torch.cuda.Event()
        self.before.record()

    def add(self, waits, fn, order=0):
        """fn(sm_limit) may run once, for every (direction, steps) in `waits`, that many steps of the direction are
        finished; waits = () means as soon as the kernel is up.  Jobs the counters cannot announce run after the
        kernel (run_late)."""
        waits = tuple(waits) or ((0, 1),)
        ready = max(r for _, r in waits)
        if ready <= self.max_reported and (self.cut is None or ready <= self.cut):
            self.jobs.append((ready, order, waits, fn))
        else:
            self.late.append((ready, order, fn))

    def launch(self):
        """Enqueue the gated jobs on the side stream (call right AFTER the recurrence launch has succeeded)."""
        if not self.jobs:
            return
        self.side.wait_event(self.before)
        with torch.cuda.stream(self.side):
            for ready, _, waits, fn in sorted(self.jobs, key=lambda j: (j[0], j[1])):
                for direction, steps in waits:
                    call("stream_wait_value32", stream_ptr(), self.ws.progress[self.flag0 + direction:],
                         ctypes.c_uint((self.base + steps) & 0xFFFFFFFF))
                fn(self.sm_limit)

    def run_late(self):
        """After the recurrence, on the main stream: join the side stream, then the jobs the kernel finished last."""
        if self.jobs:
            torch.cuda.current_stream().wait_stream(self.side)
        for _, _, fn in sorted(self.late, key=lambda j: (j[0], j[1])):
            fn(0)


def _lstm_fwd(ws, wt, l, h0, c0, lens_dev, training, sj, b0=0, b1=None):
    """Rows b0..b1 of the batch (a sub-batch view: every pointer advanced to row / entry b0, batch_pitch = ws.B)."""
    o, B = ws.out[l], ws.B
    b1 = B if b1 is None else b1
    sub = lambda t: t[2 * l:2 * l + 2][:, b0:]                 # [2][B][H] state tensors: entry b0
    call("lstm_fwd", ws.gx[l][b0:], wt.whh_r[l].buf, LL(wt.whh_r[l].ld), LL(wt.whh_r[l].plane),
         sub(h0), sub(c0), lens_dev[b0:], _sb_view(o, b0).buf, LL(o.plane), ws.cs[l][b0:] if training else None,
         sub(ws.hn), sub(ws.cn), ws.lstm_ws, ws.progress if sj else None,
         ctypes.c_uint(sj.base if sj else 0), ws.T, b1 - b0, B, H, stream_ptr())


def forward(ws, wt, p, x, lens_dev, h0, c0, bn_train, save=True, bn_momentum=0.1, bn_eps=1e-5, dp=None):
    """x: fp32 (T*B, in_dim) padded time-major.  Leaves the head pre-activations in ws.z[:, :out_dim].
    bn_train: BatchNorm uses batch statistics and updates the running ones (model.train()); save: keep what the
    backward pass needs.  dp (data-parallel mode B, ssb200.train.ExactDP): BatchNorm statistics over all ranks' rows."""
    training = save
    T, B, TB = ws.T, ws.B, ws.TB
    st = stream_ptr()
    _split_into(x, ws.x0)
    _gemm(ws.x0, wt.wih[0], ws.gx[0], bias=wt.bias[0], M=TB)
    stream = STREAM and T >= MIN_STREAM_T
    for l in range(ws.L):
        nxt = l + 1 < ws.L
        # (optionally two half-batch launches for 64 < B <= 128: see SPLIT_WIDE_BATCH_FWD)
        halves = [(0, B)] if (B <= 64 or not SPLIT_WIDE_BATCH_FWD) else [(0, (B + 1) // 2), ((B + 1) // 2, B)]
        sj = None
        for hi_, (b0, b1) in enumerate(halves):
            if hi_ == len(halves) - 1 and stream and nxt:
                sj = _Streamer(ws, x.device, False, T)
            _lstm_fwd(ws, wt, l, h0, c0, lens_dev, training, sj, b0, b1)
        if not nxt:
            break
        if sj is None:
            _gemm(ws.out[l], wt.wih[l + 1], ws.gx[l + 1], bias=wt.bias[l + 1], M=TB)
            continue
        # x-projection of layer l+1, streamed: direction d's half of out[l] for frames [a, b) is final after b steps of
        # direction 0 (t = 0..) / T - a steps of direction 1 (t = T-1..); whichever half of a chunk comes first writes
        # (beta = 0, + bias), the other accumulates.  The 48 idle SMs cannot absorb the whole GEMM inside the recurrence's
        # 2.9 ms (profiles/r02_summary.md), so only halves that are ready in its first FWD_STREAM_FRACTION are streamed;
        # the others follow on the main stream at full width, as ONE full-K GEMM where both halves of a chunk are late.
        o, w, gx, bias = ws.out[l], wt.wih[l + 1], ws.gx[l + 1], wt.bias[l + 1]
        cut = min(sj.max_reported, int(FWD_STREAM_FRACTION * T))
        for (a, b) in _chunks(T, N_CHUNKS_FWD):
            r0, rows = a * B, (b - a) * B
            halves = sorted(((b, 0), (T - a, 1)))              # (ready, direction), earliest first
            if halves[0][0] > cut:
                def full(sm_limit, r0=r0, rows=rows):
                    _gemm(_sb_view(o, r0, rows), w, gx[r0:], bias=bias, M=rows, N=8 * H, K=2 * H, sm_limit=sm_limit)
                sj.late.append((halves[0][0], 0, full))
                continue
            for i, (ready, d) in enumerate(halves):
                def job(sm_limit, d=d, r0=r0, rows=rows, acc=(i == 1)):
                    _gemm(_sb_view(o, r0, rows, d * H, H), _sb_view(w, 0, 8 * H, d * H, H), gx[r0:], bias=None if acc else bias,
                          beta=1.0 if acc else 0.0, M=rows, N=8 * H, K=H, sm_limit=sm_limit)
                if ready <= cut:
                    sj.add(((d, ready),), job, order=i)
                else:
                    sj.late.append((ready, i, job))
        sj.launch()
        sj.run_late()
    top = ws.out[-1]
    if dp is not None and bn_train:
        C = 2 * H
        loc = ws.bn_sums[0]
        call("bn_sums", top.buf, LL(top.ld), LL(top.plane), None, TB, C, ws.bn_partial, loc, st)
        loc[2 * C:].copy_(dp.counts(B, T, x.device), non_blocking=True)
        dp.allreduce_sum(loc[:2 * C + 1])                      # sum x, sum x^2, total batch size
        dp.allreduce_max(loc[2 * C + 1:])                      # longest T: rows of the concatenated batch = B_total * T_max
        call("bn_stats_from_sums", loc, loc[2 * C:], C, p["bn.weight"], p["bn.bias"], float(bn_eps), float(bn_momentum),
             p["bn.running_mean"], p["bn.running_var"], ws.coef, st)
    else:
        call("bn_stats", top.buf, LL(top.ld), LL(top.plane), TB, 2 * H, p["bn.weight"], p["bn.bias"], float(bn_eps),
             float(bn_momentum), int(bn_train), p["bn.running_mean"], p["bn.running_var"], ws.bn_partial, ws.coef, st)
    _split_into(p["lin.weight"], wt.lin_folded, col_scale=ws.coef[0])
    call("head_fold_bias", p["lin.weight"], p["lin.bias"], ws.coef, ws.out_dim, 2 * H, wt.bias_folded, st)
    _gemm(top, wt.lin_folded, ws.z, bias=wt.bias_folded, M=TB)
    return ws.z


def pit_loss(ws, mix, src, lens_dev, F, n_out, n_src, mode=0, accumulate=False, mask_out=None, used=None, stats=None, dp=None):
    """mix fp32 (T*B,F); src fp32 (n_src,T*B,F).  Fills ws.losses/idx/assign and stats (ws.stats unless a shared
    accumulator is given, as the RSH passes do together with the shared `used` bitmasks).  dp (mode B): the loss
    normaliser is the frame count of ALL ranks (archs/uPIT.py:204 on the concatenated batch)."""
    st = stream_ptr()
    call("pit_loss_fwd", ws.z, LL(ws.ldz), mix, src, lens_dev, ws.T, ws.B, F, n_out, n_src, mask_out, ws.pit_partial, st)
    call("pit_select", ws.pit_partial, lib().ss_pit_chunks(ws.T), lens_dev, ws.B, F, n_out, n_src, mode,
         int(accumulate), ws.losses, ws.idx, ws.assign, ws.used if used is None else used,
         ws.stats if stats is None else stats, st)
    if dp is not None:
        stt = ws.stats if stats is None else stats
        dp.allreduce_sum(stt[:2])
        call("pit_stats_finalize", stt, n_src, st)


def _gemm_tn(a, a_col0, a_col_batch, b, b_col0, b_col_batch, R, blk0, n_blk, shifts, batch, c0, c1, ldc, M, N, beta,
             rowmap, hidden, sm_limit):
    call("gemm_sb16_tn", a.buf, LL(a.ld), LL(a.plane), a.cols, a_col0, a_col_batch,
         b.buf, LL(b.ld), LL(b.plane), b.cols, b_col0, b_col_batch, R, blk0[0], blk0[1], n_blk, shifts[0], shifts[1],
         batch, c0, c1, LL(ldc), M, N, 1.0, float(beta), rowmap, hidden, sm_limit, stream_ptr())


def _block_chunks(nblk, n):
    """Sizes of n nearly equal runs of 64-row blocks."""
    n = max(1, min(n, nblk))
    cuts = [round(i * nblk / n) for i in range(n + 1)]
    return [cuts[i + 1] - cuts[i] for i in range(n) if cuts[i + 1] > cuts[i]]


def backward(ws, wt, p, grads, mix, src, lens_dev, h0, c0, F, n_out, gout, dmask_extra=None, dhn=None, dcn=None,
             want_dx=False, want_d0=False, training=True, accumulate=False, hook=None, stats=None, dp=None):
    """Backward of loss -> every parameter gradient in `grads` (dict name -> fp32 tensor, natural layouts).
    Returns (dx or None, dh0 or None, dc0 or None)."""
    T, B, TB = ws.T, ws.B, ws.TB
    dev = mix.device
    O = ws.out_dim
    beta0 = 1.0 if accumulate else 0.0
    nblk = (TB + 63) // 64
    call("pit_loss_bwd", ws.z, LL(ws.ldz), mix, src, lens_dev, ws.assign, T, B, F, n_out,
         ws.stats if stats is None else stats, gout, dmask_extra,
         ws.dz.buf, LL(ws.dz.ld), LL(ws.dz.plane), stream_ptr())
    top = ws.out[-1]

    def head_weight_grads(sm_limit):
        """dlin.bias = colsum(dz); G = dz^T top; dlin.weight = s*(G - mean*dsum) + beta*dsum (needs dz only: off the
        critical path), then the bucket hook."""
        s = stream_ptr()
        call("colsum_sb16", ws.dz.buf, LL(ws.dz.ld), LL(ws.dz.plane), TB, O, ws.cs_partial[-1], ws.dsum, 0, 0, s)
        _gemm_tn(ws.dz, 0, 0, top, 0, 0, TB, (0, 0), nblk, (0, 0), 1, ws.G, None, 2 * H, O, 2 * H, 0.0, 0, 0, sm_limit)
        call("head_dw_fix", ws.G, ws.dsum, ws.coef, p["bn.bias"], O, 2 * H, ws.dlinW, s)
        _acc(grads, "lin.weight", ws.dlinW, accumulate)
        _acc(grads, "lin.bias", ws.dsum, accumulate)
        if hook is not None:
            hook("head")

    _gemm(ws.dz, wt.linT, ws.dy, M=TB, K=O)
    if dp is not None and training:
        C = 2 * H
        loc, glob = ws.bn_sums[1], ws.bn_sums[2]
        call("bn_sums", top.buf, LL(top.ld), LL(top.plane), ws.dy, TB, C, ws.bn_partial, loc, stream_ptr())
        glob[:2 * C].copy_(loc[:2 * C], non_blocking=True)
        dp.allreduce_sum(glob[:2 * C])                         # sum dy, sum dy*x over all ranks
        call("bn_bwd_from_sums", ws.dy, top.buf, LL(top.ld), LL(top.plane), TB, C, ws.coef, loc, glob, ws.bn_sums[0][2 * C:],
             ws.dgamma, ws.dbeta, ws.bn_bwd, stream_ptr())
    else:
        call("bn_bwd", ws.dy, top.buf, LL(top.ld), LL(top.plane), TB, 2 * H, ws.coef, int(training), ws.bn_partial,
             ws.dgamma, ws.dbeta, ws.bn_bwd, stream_ptr())
    _acc(grads, "bn.weight", ws.dgamma, accumulate)
    _acc(grads, "bn.bias", ws.dbeta, accumulate)
    dout = ws.dy
    dx = dh0 = dc0 = None
    if want_d0:
        dh0 = torch.zeros(2 * ws.L, B, H, dtype=torch.float32, device=dev)
        dc0 = torch.zeros(2 * ws.L, B, H, dtype=torch.float32, device=dev)

    def g_pair(kind, l):
        g0, g1 = grads["blstm.weight_%s_l%d" % (kind, l)], grads["blstm.weight_%s_l%d_reverse" % (kind, l)]
        assert g0.stride(0) == g1.stride(0) and g0.stride(1) == 1
        return g0, g1

    def dw_chunk(l, kind, blk0, n_blk, beta, sm_limit):
        """Rows of 64-row blocks [blk0[z], +n_blk) of direction z's weight gradient: dW[z] (+)= dG[l][:, z]^T . X."""
        g0, g1 = g_pair(kind, l)
        if kind == "ih":
            xin = ws.x0 if l == 0 else ws.out[l - 1]
            _gemm_tn(ws.dG[l], 0, 4 * H, xin, 0, 0, TB, blk0, n_blk, (0, 0), 2, g0, g1, g0.stride(0), 4 * H, xin.cols,
                     beta, UNIT4, H, sm_limit)
        else:   # carried state of frame t: out[t-1] (direction 0) / out[t+1] (direction 1); the h0 term: dwhh_edge
            _gemm_tn(ws.dG[l], 0, 4 * H, ws.out[l], 0, H, TB, blk0, n_blk, (-B, B), 2, g0, g1, g0.stride(0), 4 * H, H,
                     beta, UNIT4, H, sm_limit)

    def dw_finish(l, sm_limit=0):
        """What needs every row of dG[l]: the h0 term of dW_hh, the bias sums; then the bucket hook."""
        s = stream_ptr()
        g0, g1 = g_pair("hh", l)
        call("dwhh_edge", ws.dG[l].buf, LL(ws.dG[l].ld), LL(ws.dG[l].plane), lens_dev, h0[2 * l:2 * l + 2], B, H, g0, g1,
             LL(g0.stride(0)), s)
        db = ws.db_l[l]
        call("colsum_sb16", ws.dG[l].buf, LL(ws.dG[l].ld), LL(ws.dG[l].plane), TB, 8 * H, ws.cs_partial[l], db, UNIT4, H, s)
        for d, suf in enumerate(("", "_reverse")):
            _acc(grads, "blstm.bias_ih_l%d%s" % (l, suf), db[d * 4 * H:(d + 1) * 4 * H], accumulate)
            _acc(grads, "blstm.bias_hh_l%d%s" % (l, suf), db[d * 4 * H:(d + 1) * 4 * H], accumulate)
        if hook is not None:
            hook("l%d" % l)

    stream = STREAM and T >= MIN_STREAM_T
    nsteps = T + (1 if want_d0 else 0)
    deferred = [head_weight_grads]                             # rides under the next recurrence (or runs at the end)
    for l in range(ws.L - 1, -1, -1):
        # Batch rows are independent chains: 64 < B <= 128 runs as two halves through the K-split kernel (9.2 us per time
        # step each) instead of one pass of the generic kernel (31 us); the streamed GEMMs ride under the second half,
        # when the rows of the first are final already.
        halves = [(0, B)] if (B <= 64 or not SPLIT_WIDE_BATCH) else [(0, (B + 1) // 2), ((B + 1) // 2, B)]
        sj = None
        for hi_, (b0, b1) in enumerate(halves):
            if hi_ == len(halves) - 1 and stream:
                sj = _Streamer(ws, dev, True, nsteps, b1 - b0, cut=int(BWD_STREAM_FRACTION[0 if l == 0 else 1] * nsteps))
            sub = lambda t: None if t is None else t[2 * l:2 * l + 2][:, b0:]      # [2][B][H] state tensors: entry b0
            call("lstm_bwd", ws.gx[l][b0:], ws.cs[l][b0:], sub(c0), dout[b0:], sub(dhn), sub(dcn), lens_dev[b0:],
                 wt.whhT[l].buf, LL(wt.whhT[l].ld), LL(wt.whhT[l].plane), _sb_view(ws.dG[l], b0).buf, LL(ws.dG[l].plane),
                 sub(dh0) if want_d0 else None, sub(dc0) if want_d0 else None, ws.bwd_sync,
                 ws.progress[2:] if sj else None, ctypes.c_uint(sj.base if sj else 0), T, b1 - b0, B, H, stream_ptr())
        need_dx = l > 0 or want_dx
        dxbuf = None
        if need_dx:
            if l > 0:   # ping-pong: the running recurrence still reads `dout` while the streamed dX writes the other buffer
                dxbuf = ws.dout if dout is not ws.dout else ws.dy
            else:
                dx = dxbuf = torch.empty(TB, ws.in_dim, dtype=torch.float32, device=dev)
        if sj is None:
            for fn in deferred:
                fn(0)
            deferred = []
            if need_dx:
                _gemm(ws.dG[l], wt.wihT[l], dxbuf, M=TB, K=8 * H)
            dw_chunk(l, "hh", (0, 0), nblk, beta0, 0)
            dw_chunk(l, "ih", (0, 0), nblk, beta0, 0)
            dw_finish(l)
        else:
            # (1) what the previous layer left behind: needs nothing from this recurrence
            for i, fn in enumerate(deferred):
                sj.add((), fn, order=-10 + i)
            deferred = []
            # (2) dX, on the critical path: direction 0 walks t = T-1..0, direction 1 walks t = 0..T-1
            if need_dx:
                wT, dG = wt.wihT[l], ws.dG[l]
                for (a, b) in _chunks(T, N_CHUNKS):
                    r0, rows = a * B, (b - a) * B
                    ready = (T - a, b)
                    first = 0 if ready[0] <= ready[1] else 1
                    for d in (0, 1):
                        def job(sm_limit, d=d, r0=r0, rows=rows, acc=(d != first)):
                            _gemm(_sb_view(dG, r0, rows, d * 4 * H, 4 * H), _sb_view(wT, 0, wT.rows, d * 4 * H, 4 * H),
                                  dxbuf[r0:], beta=1.0 if acc else 0.0, M=rows, N=wT.rows, K=4 * H, sm_limit=sm_limit)
                        sj.add(((d, ready[d]),), job, order=0)
            # (3) weight gradients that only need this layer's gate gradients, accumulated over row blocks in the order
            # the two directions produce them (direction 0 from the last block down, direction 1 from the first up)
            sizes = _block_chunks(nblk, N_CHUNKS_DW)
            done = 0
            for ci, n in enumerate(sizes):
                blk0 = (nblk - done - n, done)
                done += n
                # direction 0 has passed block j once T - floor(64 j / B) steps are done; direction 1 once
                # ceil(64 (j + n) / B) are (capped at T)
                waits = ((0, T - (blk0[0] * 64) // B), (1, min(T, -(-(blk0[1] + n) * 64 // B))))
                kinds = ("hh", "ih") if l == 0 else ("hh",)
                for kind in kinds:
                    def job(sm_limit, kind=kind, blk0=blk0, n=n, beta=(beta0 if ci == 0 else 1.0)):
                        dw_chunk(l, kind, blk0, n, beta, sm_limit)
                    sj.add(waits, job, order=1)
            sj.launch()
            sj.run_late()
            if l > 0:
                def rest(sm_limit, l=l):
                    dw_chunk(l, "ih", (0, 0), nblk, beta0, sm_limit)
                    dw_finish(l, sm_limit)
                deferred = [rest]
            else:
                dw_finish(l)
        dout = dxbuf
    for fn in deferred:
        fn(0)
    return dx, dh0, dc0


def _acc(grads, name, val, accumulate):
    g = grads[name]
    call("axpby", val, g, LL(g.numel()), 1.0, 1.0 if accumulate else 0.0, stream_ptr())
