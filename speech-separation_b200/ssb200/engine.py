"""Host orchestration of the BLSTM mask estimator step: which kernel runs on which buffer.

Mirrors what `SepDNN.forward` + `compute_loss` + `loss.backward()` do in the reference
(archs/uPIT.py:129-147,157-206; archs/RSH.py:169-187,197-259; steps/train_qsub.py:120), but every
arithmetic step is a libssb200 kernel; torch only owns the device buffers.

Layouts: activations time-major padded, row = t*B + b.  LSTM gate columns interleaved (dir*4H + 4u + gate).
"""
import torch

from . import ops
from ._lib import LL, call, lib, stream_ptr
from .ops import SB16, up8

H = 600          # archs/uPIT.py:115, archs/RSH.py:155 (hard-coded in the reference)
UNIT4 = 1


def _sb_view(sb, row0, rows):
    """Rows [row0,row0+rows) of an SB16 as a new SB16 sharing storage (plane stride unchanged)."""
    v = SB16.__new__(SB16)
    v.rows, v.cols, v.ld, v.plane = rows, sb.cols, sb.ld, sb.plane
    v.buf = sb.buf[row0 * sb.ld:]
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

    def refresh(self, p, need_grad_copies=True):
        """p: dict name -> fp32 cuda tensor with the reference's state_dict names."""
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


class Workspace:
    """All activation / gradient buffers for one (T, B) shape."""

    def __init__(self, T, B, in_dim, out_dim, n_src, num_layers, device, training=True):
        self.T, self.B, self.in_dim, self.out_dim, self.L = T, B, in_dim, out_dim, num_layers
        TB = T * B
        self.TB = TB
        f32 = dict(dtype=torch.float32, device=device)
        self.x0 = SB16(TB, in_dim, device)
        self.gx = [torch.empty(TB, 8 * H, **f32) for _ in range(num_layers)]
        self.out = [SB16(TB, 2 * H, device) for _ in range(num_layers)]
        self.hn = torch.empty(2 * num_layers, B, H, **f32)
        self.cn = torch.empty(2 * num_layers, B, H, **f32)
        lib().ss_lstm_ws_bytes.restype = __import__("ctypes").c_longlong
        self.lstm_ws = torch.empty(int(lib().ss_lstm_ws_bytes(B, H)), dtype=torch.uint8, device=device)
        self.coef = torch.empty(5, 2 * H, **f32)
        self.bn_partial = torch.empty(2 * max(1, lib().ss_bn_parts(TB)) * 2 * H, **f32)
        self.ldz = (out_dim + 3) // 4 * 4
        self.z = torch.empty(TB, self.ldz, **f32)
        self.pit_partial = torch.empty(max(1, lib().ss_pit_chunks(T)) * B * 16, **f32)
        self.losses = torch.empty(24, B, **f32)
        self.idx = torch.empty(B, dtype=torch.int32, device=device)
        self.assign = torch.empty(B, 4, dtype=torch.int32, device=device)
        self.used = torch.zeros(B, dtype=torch.int32, device=device)
        self.stats = torch.zeros(4, **f32)
        self.training = training
        if training:
            self.x0T = SB16(in_dim, TB, device)
            self.outT = [SB16(2 * H, TB, device) for _ in range(num_layers)]
            self.hprevT = [SB16(2 * H, TB, device) for _ in range(num_layers)]
            self.cs = [torch.empty(TB, 2 * H, **f32) for _ in range(num_layers)]
            self.dz = SB16(TB, out_dim, device)
            self.dzT = SB16(out_dim, TB, device)
            self.dy = torch.empty(TB, 2 * H, **f32)           # dY -> dout of the top layer (in place)
            self.dout = torch.empty(TB, 2 * H, **f32)         # dout of lower layers
            # one gradient buffer per layer: the upper layer's weight-gradient GEMMs still read theirs while the lower
            # layer's recurrence writes its own (see backward())
            self.dG = [SB16(TB, 8 * H, device) for _ in range(num_layers)]
            self.dGT = [SB16(8 * H, TB, device) for _ in range(num_layers)]
            self.G = torch.empty(out_dim, 2 * H, **f32)
            self.bn_bwd = torch.empty(2, 2 * H, **f32)
            self.cs_partial = [torch.empty(max(1, lib().ss_colsum_parts(TB)) * 8 * H, **f32) for _ in range(num_layers + 1)]
            self.db_l = [torch.empty(8 * H, **f32) for _ in range(num_layers)]
            self.start_flag = torch.zeros(1, dtype=torch.int32, device=device)
            self.start_seq = 0
            self.bwd_sync = torch.empty(256, dtype=torch.uint8, device=device)


def _gemm(a, b, out, bias=None, alpha=1.0, beta=0.0, rowmap=0, hidden=0, M=None, N=None, K=None):
    return ops.gemm(a, b, out=out, bias=bias, alpha=alpha, beta=beta, rowmap=rowmap, hidden=hidden, M=M, N=N, K=K)


def forward(ws, wt, p, x, lens_dev, h0, c0, bn_train, save=True, bn_momentum=0.1, bn_eps=1e-5):
    """x: fp32 (T*B, in_dim) padded time-major.  Leaves the head pre-activations in ws.z[:, :out_dim].
    bn_train: BatchNorm uses batch statistics and updates the running ones (model.train()); save: keep what the
    backward pass needs."""
    training = save
    T, B, TB = ws.T, ws.B, ws.TB
    st = stream_ptr()
    _split_into(x, ws.x0, outT=ws.x0T if training else None)
    inp = ws.x0
    for l in range(ws.L):
        _gemm(inp, wt.wih[l], ws.gx[l], bias=wt.bias[l])
        o = ws.out[l]
        inline_T = training and not OVERLAP                    # transposed copies inside the call, on this stream
        call("lstm_fwd", ws.gx[l], wt.whh_r[l].buf, LL(wt.whh_r[l].ld), LL(wt.whh_r[l].plane),
             h0[2 * l:2 * l + 2], c0[2 * l:2 * l + 2], lens_dev, o.buf, LL(o.plane),
             ws.outT[l].buf if inline_T else None, ws.hprevT[l].buf if inline_T else None,
             LL(ws.outT[l].ld if inline_T else 0), LL(ws.outT[l].plane if inline_T else 0),
             ws.cs[l] if training else None, ws.hn[2 * l:2 * l + 2], ws.cn[2 * l:2 * l + 2], ws.lstm_ws, T, B, H, st)
        if training and OVERLAP:
            # only the backward pass reads outT / hprevT: the side stream makes them while the next GEMM and
            # recurrence run (backward() joins the side stream before it needs them)
            side, ev = _side_stream(x.device), torch.cuda.Event()
            ev.record()
            side.wait_event(ev)
            with torch.cuda.stream(side):
                oT, hT = ws.outT[l], ws.hprevT[l]
                call("transpose_sb16", o.buf, LL(o.ld), LL(o.plane), TB, 2 * H, oT.buf, LL(oT.ld), LL(oT.plane), stream_ptr())
                call("hprev_transpose", o.buf, LL(o.ld), LL(o.plane), T, B, H, lens_dev, h0[2 * l:2 * l + 2], hT.buf,
                     LL(hT.ld), LL(hT.plane), stream_ptr())
        inp = o
    top = ws.out[-1]
    call("bn_stats", top.buf, LL(top.ld), LL(top.plane), TB, 2 * H, p["bn.weight"], p["bn.bias"], float(bn_eps),
         float(bn_momentum), int(bn_train), p["bn.running_mean"], p["bn.running_var"], ws.bn_partial, ws.coef, st)
    _split_into(p["lin.weight"], wt.lin_folded, col_scale=ws.coef[0])
    call("head_fold_bias", p["lin.weight"], p["lin.bias"], ws.coef, ws.out_dim, 2 * H, wt.bias_folded, st)
    _gemm(top, wt.lin_folded, ws.z, bias=wt.bias_folded)
    return ws.z


def pit_loss(ws, mix, src, lens_dev, F, n_out, n_src, mode=0, accumulate=False, mask_out=None, used=None, stats=None):
    """mix fp32 (T*B,F); src fp32 (n_src,T*B,F).  Fills ws.losses/idx/assign and stats (ws.stats unless a shared
    accumulator is given, as the RSH passes do together with the shared `used` bitmasks)."""
    st = stream_ptr()
    call("pit_loss_fwd", ws.z, LL(ws.ldz), mix, src, lens_dev, ws.T, ws.B, F, n_out, n_src, mask_out, ws.pit_partial, st)
    call("pit_select", ws.pit_partial, lib().ss_pit_chunks(ws.T), lens_dev, ws.B, F, n_out, n_src, mode,
         int(accumulate), ws.losses, ws.idx, ws.assign, ws.used if used is None else used,
         ws.stats if stats is None else stats, st)


def backward(ws, wt, p, grads, mix, src, lens_dev, c0, F, n_out, gout, dmask_extra=None, dhn=None, dcn=None,
             want_dx=False, want_d0=False, training=True, accumulate=False, hook=None, stats=None):
    """Backward of loss -> every parameter gradient in `grads` (dict name -> fp32 tensor, natural layouts).
    Returns (dx or None, dh0 or None, dc0 or None)."""
    T, B, TB = ws.T, ws.B, ws.TB
    st = stream_ptr()
    O = ws.out_dim
    beta = 1.0 if accumulate else 0.0
    call("pit_loss_bwd", ws.z, LL(ws.ldz), mix, src, lens_dev, ws.assign, T, B, F, n_out,
         ws.stats if stats is None else stats, gout, dmask_extra,
         ws.dz.buf, LL(ws.dz.ld), LL(ws.dz.plane), ws.dzT.buf, LL(ws.dzT.ld), LL(ws.dzT.plane), st)
    top, topT = ws.out[-1], ws.outT[-1]
    # head: dlin.bias = colsum(dz); G = dz^T x; dW = s*(G - mean*dsum) + beta*dsum; dY = dz W
    dsum = ws.dsum if hasattr(ws, "dsum") else None
    if dsum is None:
        dsum = ws.dsum = torch.empty(O, dtype=torch.float32, device=mix.device)
        ws.dlinW = torch.empty(O, 2 * H, dtype=torch.float32, device=mix.device)
        ws.dgamma = torch.empty(2 * H, dtype=torch.float32, device=mix.device)
        ws.dbeta = torch.empty(2 * H, dtype=torch.float32, device=mix.device)
        ws.db = torch.empty(8 * H, dtype=torch.float32, device=mix.device)
    def head_weight_grads():
        """dlin.weight, dlin.bias (needs dz only: off the critical path), then the bucket hook."""
        s = stream_ptr()
        call("colsum_sb16", ws.dz.buf, LL(ws.dz.ld), LL(ws.dz.plane), TB, O, ws.cs_partial[-1], dsum, 0, 0, s)
        _gemm(ws.dzT, topT, ws.G, K=TB)
        call("head_dw_fix", ws.G, dsum, ws.coef, p["bn.bias"], O, 2 * H, ws.dlinW, s)
        _acc(grads, "lin.weight", ws.dlinW, accumulate)
        _acc(grads, "lin.bias", dsum, accumulate)
        if hook is not None:
            hook("head")

    _gemm(ws.dz, wt.linT, ws.dy, K=O)
    call("bn_bwd", ws.dy, top.buf, LL(top.ld), LL(top.plane), TB, 2 * H, ws.coef, int(training), ws.bn_partial,
         ws.dgamma, ws.dbeta, ws.bn_bwd, st)
    _acc(grads, "bn.weight", ws.dgamma, accumulate)
    _acc(grads, "bn.bias", ws.dbeta, accumulate)
    dout = ws.dy
    dx = dh0 = dc0 = None
    if want_d0:
        dh0 = torch.zeros(2 * ws.L, B, H, dtype=torch.float32, device=mix.device)
        dc0 = torch.zeros(2 * ws.L, B, H, dtype=torch.float32, device=mix.device)

    def weight_grads(l):
        """dW_ih, dW_hh (both directions per launch: rows d*4H.. of dGT; d*H.. of hprevT), biases; then the bucket hook."""
        s = stream_ptr()
        dGT, xT = ws.dGT[l], (ws.x0T if l == 0 else ws.outT[l - 1])
        if overlap:
            dG = ws.dG[l]
            call("transpose_sb16", dG.buf, LL(dG.ld), LL(dG.plane), TB, 8 * H, dGT.buf, LL(dGT.ld), LL(dGT.plane), s)
        for B_op, b_rows, N, kind in ((xT, 0, xT.rows, "ih"), (ws.hprevT[l], H, H, "hh")):
            g0, g1 = grads["blstm.weight_%s_l%d" % (kind, l)], grads["blstm.weight_%s_l%d_reverse" % (kind, l)]
            assert g0.stride(0) == g1.stride(0) and g0.stride(1) == 1
            call("gemm_sb16_x2", dGT.buf, LL(dGT.ld), LL(dGT.plane), 4 * H, B_op.buf, LL(B_op.ld), LL(B_op.plane),
                 b_rows, g0, g1, LL(g0.stride(0)), 4 * H, N, TB, 1.0, float(beta), UNIT4, H, s)
        db = ws.db_l[l]
        call("colsum_sb16", ws.dG[l].buf, LL(ws.dG[l].ld), LL(ws.dG[l].plane), TB, 8 * H, ws.cs_partial[l], db, UNIT4, H, s)
        for d, suf in enumerate(("", "_reverse")):
            _acc(grads, "blstm.bias_ih_l%d%s" % (l, suf), db[d * 4 * H:(d + 1) * 4 * H], accumulate)
            _acc(grads, "blstm.bias_hh_l%d%s" % (l, suf), db[d * 4 * H:(d + 1) * 4 * H], accumulate)
        if hook is not None:
            hook("l%d" % l)

    # The K-split backward recurrence occupies 80 of the 148 SMs for ~5 ms; the weight gradients of layer l+1 are not
    # on the critical path, so they run on a side stream, capped to the idle SMs, once layer l's recurrence is
    # resident (it raises ws.start_flag at its start; the side stream waits on that value without holding an SM).
    overlap = OVERLAP and B <= 64
    side = _side_stream(mix.device) if OVERLAP else None
    main = torch.cuda.current_stream()
    if side is not None:
        main.wait_stream(side)                                 # the forward pass's transposed copies
    deferred = []                                              # work that waits for the next recurrence to be resident

    def defer(fn):
        ev = torch.cuda.Event()
        ev.record(main)                                        # its inputs are complete at this point of the main stream
        side.wait_event(ev)
        deferred.append(fn)

    if overlap:
        defer(head_weight_grads)
    else:
        head_weight_grads()
    for l in range(ws.L - 1, -1, -1):
        if deferred:
            ws.start_seq += 1
            call("lstm_bwd_start_flag", ws.start_flag, ws.start_seq)
        call("lstm_bwd", ws.gx[l], ws.cs[l], c0[2 * l:2 * l + 2], dout,
             dhn[2 * l:2 * l + 2] if dhn is not None else None, dcn[2 * l:2 * l + 2] if dcn is not None else None,
             lens_dev, wt.whhT[l].buf, LL(wt.whhT[l].ld), LL(wt.whhT[l].plane), ws.dG[l].buf, LL(ws.dG[l].plane),
             None if overlap else ws.dGT[l].buf, LL(ws.dGT[l].ld), LL(ws.dGT[l].plane),
             dh0[2 * l:2 * l + 2] if want_d0 else None, dc0[2 * l:2 * l + 2] if want_d0 else None, ws.bwd_sync, T, B, H,
             stream_ptr())
        if deferred:
            with torch.cuda.stream(side):
                call("stream_wait_value32", stream_ptr(), ws.start_flag, ws.start_seq)
                call("gemm_sm_limit", IDLE_SMS)
                for fn in deferred:
                    fn()
                call("gemm_sm_limit", 0)
            deferred = []
        if overlap and l > 0:
            defer(lambda l=l: weight_grads(l))
        else:
            weight_grads(l)
        if l > 0:
            _gemm(ws.dG[l], wt.wihT[l], ws.dout, K=8 * H)
            dout = ws.dout
        elif want_dx:
            dx = torch.empty(TB, ws.in_dim, dtype=torch.float32, device=mix.device)
            _gemm(ws.dG[0], wt.wihT[0], dx, K=8 * H)
    if side is not None:
        main.wait_stream(side)
    return dx, dh0, dc0


OVERLAP = not __import__("os").environ.get("SSB_NO_OVERLAP")
IDLE_SMS = 68                                                  # 148 SMs - the 80 CTAs of the K-split backward recurrence
_SIDE = {}


def _side_stream(device):
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    if key not in _SIDE:
        _SIDE[key] = torch.cuda.Stream(device=device)
    return _SIDE[key]


def _acc(grads, name, val, accumulate):
    g = grads[name]
    call("axpby", val, g, LL(g.numel()), 1.0, 1.0 if accumulate else 0.0, stream_ptr())
