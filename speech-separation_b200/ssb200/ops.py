"""Host-side wrappers (torch tensors in, torch tensors out) over the C ABI.  No arithmetic happens here:
torch only allocates device memory and supplies the stream."""
import os

import numpy as np
import torch

from ._lib import LL, call, lib, stream_ptr


def _small(a, dtype, device):
    """Offset tables -> device through pinned memory: a pageable host->device copy makes the driver wait for everything
    already queued on the stream, which would serialise the host side of batch i+1 behind the kernels of batch i."""
    t = torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=dtype)))
    if torch.device(device).type != "cuda":
        return t
    return t.pin_memory().to(device, non_blocking=True)


def _i64(a, device):
    return _small(a, np.int64, device)


def _i32(a, device):
    return _small(a, np.int32, device)


def up8(n):
    return (int(n) + 7) // 8 * 8


class Ragged:
    """Offsets of a ragged batch of utterances (lengths in samples)."""

    def __init__(self, n_samples, n_fft, hop, device, align=1):
        """align > 1 starts every utterance at a multiple of `align` samples of the flat buffer (gaps are never
        read); align=4 gives the tensor-core STFT its 16-byte loads.  Only ss_stft_dft256 takes the lengths
        explicitly, so an aligned layout is accepted by `stft` on that path alone."""
        n = np.asarray(n_samples, dtype=np.int64)
        if n.size and n.min() <= n_fft // 2:
            raise ValueError("every utterance must be longer than n_fft//2 samples (reflect padding)")
        self.n_fft, self.hop, self.F = n_fft, hop, n_fft // 2 + 1
        self.n_samples = n
        self.T = 1 + n // hop
        self.align = int(align)
        self.sig_off = np.concatenate([[0], np.cumsum((n + align - 1) // align * align)])
        self._dft_tiles = None
        self.frame_off = np.concatenate([[0], np.cumsum(self.T)])
        tf = lib().ss_stft_tile_frames()
        self.stft_tiles = np.concatenate([[0], np.cumsum((self.T + tf - 1) // tf)]).astype(np.int32)
        self.L = hop * (self.T - 1)
        th = lib().ss_istft_tile_hops() * hop
        self.istft_tiles = np.concatenate([[0], np.cumsum((self.L + th - 1) // th)]).astype(np.int32)
        self.device = device
        self.d_sig_off = _i64(self.sig_off, device)
        self.d_frame_off = _i64(self.frame_off, device)
        self.d_stft_tiles = _i32(self.stft_tiles, device)
        self.d_istft_tiles = _i32(self.istft_tiles, device)

    def dft_tiles(self):
        """Tables of ss_stft_dft256 (include/ssb200.h): the utterance table (four int64 per utterance: first sample,
        first output element, N | T << 32, first stream row) and, per 128-frame tile of the row stream, the utterance
        its first row belongs to.  Utterance u owns T_u + 3 stream rows of 64 samples (its reflect-padded signal)."""
        if self._dft_tiles is None:
            tf = lib().ss_stft_dft256_tile_frames()
            row_off = np.concatenate([[0], np.cumsum(self.T + 3)]).astype(np.int64)
            total = int(row_off[-1])
            n_tiles = (total + tf - 1) // tf
            u0 = (np.searchsorted(row_off, np.arange(n_tiles, dtype=np.int64) * tf, side="right") - 1).astype(np.int32)
            tab = np.stack([self.sig_off[:-1], self.frame_off[:-1] * self.F, self.n_samples | (self.T << 32), row_off[:-1]],
                           axis=1).astype(np.int64)
            self._dft_tiles = (torch.as_tensor(np.ascontiguousarray(tab), device=self.device),
                               torch.as_tensor(u0, device=self.device), n_tiles, total)
        return self._dft_tiles

    @property
    def n_utts(self):
        return len(self.n_samples)

    @property
    def total_frames(self):
        return int(self.frame_off[-1])


STFT_COMPLEX, STFT_TF, STFT_LOG = 1, 2, 4
# 256/64 magnitude features on the tensor cores (csrc/stft_mma.cu) instead of the FFT kernel.  Parity-green (2e-6 against
# the oracle) but slower on configs[1]: 17.1 vs 13.3 ms (its MMAs alone take 11.1 ms, profiles/r02_summary.md section E),
# so it is opt-in.
STFT_DFT256 = os.environ.get("SSB_STFT_DFT256", "0") == "1"
ISTFT_TF, ISTFT_F32 = 1, 2


def stft(sig, rag, complex_out=False, layout_tf=False, log=False, out=None):
    """sig: flat float32 cuda tensor holding all utterances back to back.  Returns a flat tensor whose block
    for utterance u is [frame_off[u]*F, frame_off[u+1]*F) shaped (F,T_u) (or (T_u,F) with layout_tf)."""
    assert sig.is_cuda and sig.dtype == torch.float32 and sig.is_contiguous()
    assert sig.numel() == int(rag.sig_off[-1])
    n = rag.total_frames * rag.F
    if out is None:
        out = torch.empty(n, dtype=torch.complex64 if complex_out else torch.float32, device=sig.device)
    flags = (STFT_COMPLEX if complex_out else 0) | (STFT_TF if layout_tf else 0) | (STFT_LOG if log else 0)
    if (flags & ~STFT_LOG) == 0 and rag.n_fft == 256 and rag.hop == 64 and STFT_DFT256:
        tab, u0, n_tiles, total_rows = rag.dft_tiles()
        call("stft_dft256", sig, tab, rag.n_utts, u0, n_tiles, LL(total_rows), out, flags, stream_ptr())
        return out
    if rag.align != 1:
        raise ValueError("an aligned Ragged layout is only read by the tensor-core 256/64 magnitude path (STFT_DFT256)")
    call("stft", sig, rag.d_sig_off, rag.d_frame_off, rag.d_stft_tiles, rag.n_utts, int(rag.stft_tiles[-1]),
         rag.n_fft, rag.hop, flags, out, stream_ptr())
    return out


def cmvn(feat, rag, layout_tf=False, eps=1e-8):
    """In-place per-utterance mean-variance normalisation of (log-)magnitude features (an extension, see ss_cmvn)."""
    assert feat.dtype == torch.float32 and feat.numel() == rag.total_frames * rag.F
    call("cmvn", feat, rag.d_frame_off, rag.n_utts, rag.F, int(bool(layout_tf)), float(eps), stream_ptr())
    return feat


class _HostStream:
    """Cached plan of stft_host: utterance chunks, their sub-corpora and two device buffers of each kind."""

    def __init__(self, rag, chunks, complex_out):
        tot = rag.total_frames
        cuts = np.searchsorted(rag.frame_off, np.arange(1, chunks) * (tot / chunks))
        self.bounds = np.unique(np.concatenate([[0], cuts, [rag.n_utts]])).astype(np.int64)
        self.subs = [Ragged(rag.n_samples[a:b], rag.n_fft, rag.hop, rag.device, align=rag.align)
                     for a, b in zip(self.bounds[:-1], self.bounds[1:])]
        n_in = max(int(r.sig_off[-1]) for r in self.subs)
        n_out = max(r.total_frames for r in self.subs) * rag.F
        odt = torch.complex64 if complex_out else torch.float32
        self.d_in = [torch.empty(n_in, dtype=torch.float32, device=rag.device) for _ in range(2)]
        self.d_out = [torch.empty(n_out, dtype=odt, device=rag.device) for _ in range(2)]
        self.streams = [torch.cuda.Stream(device=rag.device) for _ in range(3)]


def stft_host(h_sig, rag, h_out, chunks=8, complex_out=False, layout_tf=False):
    """Pinned host samples -> pinned host features (the whole job of steps/extract_feats.py:85-90 minus the file
    IO): the corpus is cut into `chunks` runs of utterances and host->device copy, kernel and device->host copy of
    consecutive runs overlap on three streams (PCIe is full duplex; the copies dominate end to end).  Returns after
    the last copy has landed."""
    assert h_sig.is_pinned() and h_out.is_pinned() and h_sig.numel() == int(rag.sig_off[-1])
    key = (int(chunks), bool(complex_out))
    plans = rag.__dict__.setdefault("_host_plans", {})
    if key not in plans:
        plans[key] = _HostStream(rag, chunks, complex_out)
    pl = plans[key]
    s_in, s_cmp, s_out = pl.streams
    cur = torch.cuda.current_stream()
    for st in pl.streams:
        st.wait_stream(cur)
    ev_cmp, ev_out = {}, {}
    for c, sub in enumerate(pl.subs):
        b = c & 1
        a0, a1 = int(rag.sig_off[pl.bounds[c]]), int(rag.sig_off[pl.bounds[c + 1]])
        f0, f1 = int(rag.frame_off[pl.bounds[c]]) * rag.F, int(rag.frame_off[pl.bounds[c + 1]]) * rag.F
        d_in, d_out = pl.d_in[b][:a1 - a0], pl.d_out[b][:f1 - f0]
        with torch.cuda.stream(s_in):
            if c >= 2:
                s_in.wait_event(ev_cmp[c - 2])                 # the kernel that read this buffer is done
            d_in.copy_(h_sig[a0:a1], non_blocking=True)
            ev_in = torch.cuda.Event()
            ev_in.record(s_in)
        with torch.cuda.stream(s_cmp):
            s_cmp.wait_event(ev_in)
            if c >= 2:
                s_cmp.wait_event(ev_out[c - 2])                # the copy that read this output buffer is done
            stft(d_in, sub, complex_out=complex_out, layout_tf=layout_tf, out=d_out)
            ev_cmp[c] = torch.cuda.Event()
            ev_cmp[c].record(s_cmp)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_cmp[c])
            h_out[f0:f1].copy_(d_out, non_blocking=True)
            ev_out[c] = torch.cuda.Event()
            ev_out[c].record(s_out)
    s_out.synchronize()
    return h_out


def stft_views(flat, rag, layout_tf=False):
    """Per-utterance views of a flat STFT buffer."""
    F = rag.F
    outs = []
    for u in range(rag.n_utts):
        blk = flat[int(rag.frame_off[u]) * F:int(rag.frame_off[u + 1]) * F]
        outs.append(blk.view(int(rag.T[u]), F) if layout_tf else blk.view(F, int(rag.T[u])))
    return outs


def istft_masked(spec, mask, rag, n_src, layout_tf=False, out_f32=False, out=None):
    """spec: flat complex64 (ss_stft COMPLEX output); mask: flat float32, n_src blocks per utterance.
    Returns (flat wav tensor int16|float32, wav_off numpy[int64] of length n_utts+1)."""
    assert spec.is_cuda and spec.dtype == torch.complex64 and mask.dtype == torch.float32
    assert mask.numel() == rag.total_frames * rag.F * n_src
    wav_off = np.concatenate([[0], np.cumsum(rag.L * n_src)])
    if out is None:
        out = torch.empty(int(wav_off[-1]), dtype=torch.float32 if out_f32 else torch.int16, device=spec.device)
    flags = (ISTFT_TF if layout_tf else 0) | (ISTFT_F32 if out_f32 else 0)
    call("istft_masked", spec, mask, rag.d_frame_off, _i64(wav_off, spec.device), rag.d_istft_tiles, rag.n_utts, n_src,
         int(rag.istft_tiles[-1]), rag.n_fft, rag.hop, flags, out, stream_ptr())
    return out, wav_off


def ragged_to_padded(flat, rag, T=None, out=None):
    """(T_u, F) blocks (ss_stft LAYOUT_TF; float32 or complex64 -> |X|) -> time-major padded [T*B][F] model input."""
    T = int(rag.T.max()) if T is None else int(T)
    if out is None:
        out = torch.empty(T * rag.n_utts, rag.F, dtype=torch.float32, device=flat.device)
    call("ragged_to_padded", flat, int(flat.dtype == torch.complex64), rag.d_frame_off, T, rag.n_utts, rag.F, out,
         stream_ptr())
    return out


def masks_to_ragged(z, ldz, rag, T, n_src, out=None):
    """sigmoid of the head pre-activations [T*B][ldz] -> the (T_u, F) mask blocks istft_masked(layout_tf=True) reads."""
    if out is None:
        out = torch.empty(rag.total_frames * n_src * rag.F, dtype=torch.float32, device=z.device)
    call("masks_to_ragged", z, LL(ldz), rag.d_frame_off, int(T), rag.n_utts, rag.F, int(n_src), out, stream_ptr())
    return out


class SB16:
    """Split-bf16 matrix [rows][cols] with pitch `ld` (elements): planes hi, lo in one bf16 buffer."""

    def __init__(self, rows, cols, device, ld=None, buf=None):
        self.rows, self.cols = int(rows), int(cols)
        self.ld = up8(cols) if ld is None else int(ld)
        self.plane = self.rows * self.ld
        self.buf = torch.zeros(2 * self.plane, dtype=torch.bfloat16, device=device) if buf is None else buf
        assert self.buf.numel() >= 2 * self.plane

    def float(self):
        v = self.buf[:2 * self.plane].view(2, self.rows, self.ld)[:, :, :self.cols].float()
        return v[0] + v[1]


def split(x, rowperm=0, hidden=0, col_scale=None, want=True, want_t=False, out=None, out_t=None):
    """fp32 matrix (R,C) -> SB16 (R,C) and/or transposed SB16 (C,R)."""
    assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    R, C = x.shape
    if want and out is None:
        out = SB16(R, C, x.device)
    if want_t and out_t is None:
        out_t = SB16(C, R, x.device)
    call("split", x, LL(x.stride(0)), R, C, rowperm, hidden, col_scale,
         out.buf if out is not None else None, LL(out.ld if out is not None else 0), LL(out.plane if out is not None else 0),
         out_t.buf if out_t is not None else None, LL(out_t.ld if out_t is not None else 0),
         LL(out_t.plane if out_t is not None else 0), stream_ptr())
    return out, out_t


def gemm(a, b, out=None, bias=None, alpha=1.0, beta=0.0, rowmap=0, hidden=0, M=None, N=None, K=None, sm_limit=0):
    """out[M,N] = alpha * a @ b^T + bias + beta*out, a: SB16 (M,K), b: SB16 (N,K); sm_limit > 0 caps the grid."""
    M = a.rows if M is None else M
    N = b.rows if N is None else N
    K = a.cols if K is None else K
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=a.buf.device)
    assert out.dtype == torch.float32 and out.stride(1) == 1
    call("gemm_sb16", a.buf, LL(a.ld), LL(a.plane), b.buf, LL(b.ld), LL(b.plane), out, LL(out.stride(0)), M, N, K, bias,
         float(alpha), float(beta), rowmap, hidden, int(sm_limit), stream_ptr())
    return out


def gemm_tn(a, b, M, N, outs=None, a_col0=0, b_col0=0, a_col_batch=0, b_col_batch=0, R=None, blk0=(0, 0), n_blk=None,
            shifts=(0, 0), batch=1, beta=0.0, rowmap=0, hidden=0, sm_limit=0):
    """Weight-gradient form: outs[z][m][n] = sum_r a[r][a_col0 + z*a_col_batch + m] * b[r + shifts[z]][b_col0 + z*b_col_batch
    + n] (+ beta*outs[z]) over the 64-row blocks [blk0[z], +n_blk) of the R reduction rows; a, b: SB16 with rows = r."""
    R = a.rows if R is None else R
    n_blk = (R + 63) // 64 if n_blk is None else n_blk
    if outs is None:
        outs = [torch.empty(M, N, dtype=torch.float32, device=a.buf.device) for _ in range(batch)]
    o0, o1 = outs[0], outs[1] if batch == 2 else None
    call("gemm_sb16_tn", a.buf, LL(a.ld), LL(a.plane), a.cols, a_col0, a_col_batch, b.buf, LL(b.ld), LL(b.plane), b.cols,
         b_col0, b_col_batch, R, blk0[0], blk0[1], n_blk, shifts[0], shifts[1], batch, o0, o1, LL(o0.stride(0)), M, N, 1.0,
         float(beta), rowmap, hidden, int(sm_limit), stream_ptr())
    return outs


def colsum(a, rowmap=0, hidden=0, R=None):
    R = a.rows if R is None else R
    parts = lib().ss_colsum_parts(R)
    partial = torch.empty(max(parts, 1) * a.cols, dtype=torch.float32, device=a.buf.device)
    out = torch.empty(a.cols, dtype=torch.float32, device=a.buf.device)
    call("colsum_sb16", a.buf, LL(a.ld), LL(a.plane), R, a.cols, partial, out, rowmap, hidden, stream_ptr())
    return out
