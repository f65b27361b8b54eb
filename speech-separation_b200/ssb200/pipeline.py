"""On-device pipelines around the model: waveforms in, loss or separated
waveforms out, with no .npz round trip between the stages.

The reference hands data from stage to stage through files: steps/extract_feats.py:86-90 writes (F, T) arrays,
archs/uPIT.py:66-79 re-reads and transposes them, compute_masks (archs/uPIT.py:209-225) copies every utterance's
mask to the host and compresses it, steps/reconstruct_sources.py:32-42 re-reads masks and complex features.  Here the
same arithmetic runs back to back on the device: ss_stft(LAYOUT_TF) -> ss_ragged_to_padded -> BLSTM engine ->
ss_masks_to_ragged -> ss_istft_masked, one host->device copy of the samples and one device->host copy of the int16
audio.  The file-based CLIs (ssb200.extract_feats / reconstruct_sources) remain the drop-in for run_train.sh /
run_eval.sh; tests check that both routes give the same numbers.
"""
import numpy as np
import torch

from . import ops
from .uPIT import _to_device_async, run_forward


def _order(wavs, hop):
    """Collator order (archs/uPIT.py:40, ssb200/uPIT.py:Collator): descending number of frames."""
    return [int(i) for i in np.argsort(np.array([1 + len(w) // hop for w in wavs]))[::-1]]


_PINNED, _IN_FLIGHT = {}, {}


def _pinned(tag, n, dtype):
    """Grow-only pinned staging buffers: cudaHostAlloc per call costs more than the copies it feeds."""
    buf = _PINNED.get(tag)
    if buf is None or buf.numel() < n or buf.dtype != dtype:
        buf = torch.empty(max(n, 1 << 16), dtype=dtype).pin_memory()
        _PINNED[tag] = buf
    return buf[:n]


def _upload(wavs, order, device, tag="in"):
    n = sum(len(wavs[i]) for i in order)
    if tag in _IN_FLIGHT:
        _IN_FLIGHT[tag].synchronize()                          # the last copy out of this staging buffer has finished
    stage = _pinned(tag, n, torch.float32)
    np.concatenate([np.asarray(wavs[i], dtype=np.float32) for i in order], out=stage.numpy())
    dev = stage.to(device, non_blocking=True)
    _IN_FLIGHT[tag] = torch.cuda.Event()
    _IN_FLIGHT[tag].record()
    return dev


def _batch(rag, mix_padded, device):
    lens = rag.T.astype(np.int32)
    return {"device_batch": True, "T": int(rag.T.max()), "B": rag.n_utts, "lens": lens,
            "lens_dev": _to_device_async(lens, np.int32, device), "mix": mix_padded}


def train_batch(mix_wavs, src_wavs, n_fft=512, hop=128, device="cuda"):
    """f2: mixture and source waveforms -> the padded device batch compute_loss / loss_and_grads / Trainer.step take
    in place of the collator's PackedSequences.  src_wavs[i][s] is source s of mixture i.
    Returns (batch, order): batch column b holds mixture order[b]."""
    order = _order(mix_wavs, hop)
    rag = ops.Ragged([len(mix_wavs[i]) for i in order], n_fft, hop, device)
    T, S = int(rag.T.max()), len(src_wavs[0])
    mix = ops.stft(_upload(mix_wavs, order, device), rag, layout_tf=True)
    batch = _batch(rag, ops.ragged_to_padded(mix, rag, T), device)
    src = torch.empty(S, T * rag.n_utts, rag.F, dtype=torch.float32, device=device)
    for s in range(S):
        for i in order:
            assert len(src_wavs[i][s]) == len(mix_wavs[i])
        # (a second staging buffer: the previous upload may still be in flight)
        spec = ops.stft(_upload([w[s] for w in src_wavs], order, device, tag="in%d" % (s & 1)), rag, layout_tf=True, out=mix)
        ops.ragged_to_padded(spec, rag, T, out=src[s])
    batch["src"] = src
    return batch, order


_PENDING = {}


_SLOT_STREAMS = {}


def _slot_stream(device, slot):
    """Every slot has a stream (and, in the model, a workspace) of its own: batch i+1's STFT / first GEMM and batch i-1's
    masks / iSTFT / copy-out then run on the SMs batch i's recurrence leaves idle instead of queueing behind it."""
    dev = torch.device(device)
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), slot)
    if key not in _SLOT_STREAMS:
        _SLOT_STREAMS[key] = torch.cuda.Stream(device=dev)
    return _SLOT_STREAMS[key]


def submit(model, wavs, n_fft=512, hop=128, device=None, slot=0):
    """First half of `separate`: uploads the waveforms and enqueues STFT, model, masks, iSTFT and the copy of the int16
    audio into pinned memory on the slot's stream, without waiting.  Returns a handle for `collect`.  Two slots used
    alternately let the host side of batch i+1 (ordering, concatenation, h0/c0 draw, launches) AND the parts of its
    device work that do not need 100 SMs run while the device works on batch i -- the evaluation counterpart of
    Trainer.prefetch."""
    if slot in _PENDING:
        raise RuntimeError("pipeline.submit: slot %d still holds a batch that was not collected" % slot)
    device = model.lin.weight.device if device is None else device
    S, F = model.num_spk, n_fft // 2 + 1
    assert F == model.feat_dim and model.out_dim == S * F
    order = _order(wavs, hop)
    # operand copies of the parameters: rebuilt (if they changed) on the caller's stream, which every slot stream follows
    model.weights(model.param_dict(), need_grad_copies=False)
    stream = _slot_stream(device, slot)
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        rag = ops.Ragged([len(wavs[i]) for i in order], n_fft, hop, device)
        spec = ops.stft(_upload(wavs, order, device, tag="in%d" % slot), rag, complex_out=True, layout_tf=True)
        batch = _batch(rag, ops.ragged_to_padded(spec, rag), device)
        model.hidden = model.init_hidden(batch["B"])
        with torch.no_grad():
            run_forward(model, batch, model.hidden, save=False, slot=slot)
            ws = batch["ws"]
            mask = ops.masks_to_ragged(ws.z, ws.ldz, rag, batch["T"], S)
            wav, off = ops.istft_masked(spec, mask, rag, S, layout_tf=True)
        stage = _pinned("out%d" % slot, wav.numel(), wav.dtype)
        stage.copy_(wav, non_blocking=True)                    # the only device->host copy
        done = torch.cuda.Event()
        done.record()
    h = {"slot": slot, "done": done, "stage": stage, "off": off, "order": order, "L": rag.L, "S": S, "n": len(wavs),
         "keep": (spec, mask, wav, batch)}                     # device buffers stay referenced until the copy is done
    _PENDING[slot] = h
    return h


def collect(h):
    """Second half of `separate`: waits for the batch of `h` and returns its waveforms in the caller's order."""
    h["done"].synchronize()
    host = h["stage"].numpy().copy()
    _PENDING.pop(h["slot"], None)
    h["keep"] = None
    out, off, S = [None] * h["n"], h["off"], h["S"]
    for u, i in enumerate(h["order"]):
        L = int(h["L"][u])
        out[i] = [host[off[u] + s * L: off[u] + (s + 1) * L] for s in range(S)]
    return out


def separate(model, wavs, n_fft=512, hop=128, device=None):
    """f3: list of mixture waveforms -> list of [n_spk] int16 waveforms (what run_eval.sh's feature extraction,
    compute_masks and reconstruct_sources.py produce through files), uPIT-style models (mask columns [s*F, (s+1)*F))."""
    return collect(submit(model, wavs, n_fft, hop, device))
