"""Train step around the hot path: what steps/train_qsub.py:116-122 does per batch --

    loss, norm = m.compute_loss(model, epoch, batch); loss.backward()
    torch.nn.utils.clip_grad_norm_(model.parameters(), 0.25); optimizer.step()        (Adam, lr 1e-3)

-- as one call, with (new) the data-parallel gradient exchange between the backward pass and the clip.
Parameters, gradients and Adam moments live in flat fp32 buckets ordered by when the backward pass finishes them
(layer 0 | layer 1 | head), so each bucket is allreduced over NCCL on a side stream while the next layer's backward
recurrence still runs, and clip + Adam are two kernels over the whole flat buffer (libssb200 ss_grad_norm /
ss_clip_adam) instead of ~120 per-tensor launches.

DP semantics (mode A of SURVEY.md 8e): every rank computes the reference loss on its own shard of utterances with
its own BatchNorm batch statistics; gradients are averaged over ranks before the clip; parameters stay identical on
all ranks because they start identical (broadcast from rank 0) and see identical averaged gradients.
"""
import torch

from ._lib import LL, call, lib, stream_ptr


def bucket_of(name):
    """Gradient bucket a parameter belongs to, by the reference's state_dict names."""
    if name.startswith("blstm."):
        import re
        k = int(re.search(r"_l(\d+)", name).group(1))
        return "l0" if k == 0 else "l1"        # layers >= 1 share a bucket: it is complete when layer 1's backward is
    return "head"


BUCKET_ORDER = ("l0", "l1", "head")


def flat_layout(named_shapes, align=64):
    """named_shapes: [(name, numel)] -> ({name: (offset, numel)}, {bucket: (offset, numel)}, total).
    Buckets are contiguous; every tensor starts on an `align`-float boundary."""
    offs, buckets, pos = {}, {}, 0
    for bk in BUCKET_ORDER:
        start = pos
        for name, n in named_shapes:
            if bucket_of(name) != bk:
                continue
            offs[name] = (pos, n)
            pos += (n + align - 1) // align * align
        buckets[bk] = (start, pos - start)
    return offs, buckets, pos


def shard_utterances(lengths, world, rank):
    """Length-bucketed sharding of a global batch: utterances sorted by descending length are dealt
    in snake order (0..w-1, w-1..0, ...), so every rank sees a similar T_max and total frame count.  Returns the
    indices for `rank`, longest first."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    out = []
    for g in range(0, len(order), world):
        grp = order[g:g + world]
        k = rank if (g // world) % 2 == 0 else world - 1 - rank
        if k < len(grp):
            out.append(grp[k])
    return out


def allreduce_buckets(flat, buckets, group=None, tags=BUCKET_ORDER):
    """Sum-allreduce the listed buckets of a flat tensor (host-side logic shared by the NCCL path and the CPU/gloo
    tests).  Returns the async work handles."""
    import torch.distributed as dist
    works = []
    for tag in tags:
        o, n = buckets[tag]
        if n:
            works.append(dist.all_reduce(flat[o:o + n], op=dist.ReduceOp.SUM, group=group, async_op=True))
    return works


EAGER_REFRESH = True

class ExactDP:
    """Data-parallel mode B: the small collectives that make N ranks reproduce the reference's arithmetic
    on the concatenated batch -- sum of (BatchNorm sums, batch size), max of T, sum of (loss numerator, normaliser), sum
    of the BatchNorm backward sums.  Stream-ordered NCCL calls on the current stream; nothing is read back."""

    def __init__(self, group=None):
        self.group = group

    def counts(self, B, T, device):
        return torch.tensor([float(B), float(T)], dtype=torch.float32).pin_memory().to(device, non_blocking=True)

    def allreduce_sum(self, t):
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)

    def allreduce_max(self, t):
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)


class _Staged(dict):
    """A collated batch whose PackedSequence data already travels to (or sits on) the device; .event marks the copy."""
    event = None


class Trainer:
    """Owns the flat parameter / gradient / Adam buckets of one SepDNN and runs whole train steps."""

    def __init__(self, model, arch, lr=1e-3, clip=0.25, betas=(0.9, 0.999), eps=1e-8, group=None, host_rng=False,
                 exact=False):
        """host_rng: draw the random initial state (h0, c0) of every batch on the CPU generator and copy it over,
        exactly as the reference's init_hidden does (archs/uPIT.py:121-127); default draws the same N(0,1) state on
        the device generator, which keeps ~2 ms of host work and a blocking copy out of every step.
        exact: data-parallel mode B -- loss normaliser and BatchNorm statistics over all ranks' utterances, gradients
        summed (not averaged): the N-rank step IS the reference step on the concatenated batch (uPIT arch only)."""
        import torch.distributed as dist
        self.model, self.arch = model, arch
        self.host_rng = host_rng
        self.lr, self.clip, self.betas, self.eps = float(lr), float(clip), betas, float(eps)
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.group = group
        self.exact = bool(exact) and self.world > 1
        if self.exact and not hasattr(arch, "loss_and_grads"):
            raise ValueError("exact data-parallel mode needs the arch's loss_and_grads")
        self.dp = ExactDP(group) if self.exact else None
        named = [(k, p) for k, p in model.named_parameters()]
        dev = named[0][1].device
        if dev.type != "cuda":
            raise RuntimeError("ssb200.train.Trainer needs the model on a CUDA device (no CPU fallback)")
        self.offs, self.buckets, total = flat_layout([(k, p.numel()) for k, p in named])
        self.n = total
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(total, dtype=torch.float32, device=dev)
        self.grads = {}
        for k, p in named:
            o, n = self.offs[k]
            self.flat_p[o:o + n].copy_(p.detach().reshape(-1))
            p.data = self.flat_p[o:o + n].view_as(p)          # the module's parameters now alias the flat bucket
            self.grads[k] = self.flat_g[o:o + n].view_as(p)
        if self.world > 1:
            dist.broadcast(self.flat_p, src=0, group=group)
            for buf in (model.bn.running_mean, model.bn.running_var):
                dist.broadcast(buf, src=0, group=group)
        import ctypes
        self.partial = torch.empty(max(1, lib().ss_optim_parts(ctypes.c_longlong(total))), dtype=torch.float32,
                                   device=dev)
        self.norm = torch.zeros(1, dtype=torch.float32, device=dev)
        self.steps = 0
        self.comm = torch.cuda.Stream(device=dev) if self.world > 1 else None
        self.copy_stream = None
        self._works = []

    # -- gradient exchange: called by the engine as each bucket's last kernel has been enqueued --
    def _bucket_ready(self, tag):
        if self.world == 1 or tag not in self.buckets:         # (a 3-layer net reports 'l2' before 'l1': same bucket)
            return
        import torch.distributed as dist
        ev = torch.cuda.Event()
        ev.record()
        self.comm.wait_event(ev)
        o, n = self.buckets[tag]
        with torch.cuda.stream(self.comm):
            dist.all_reduce(self.flat_g[o:o + n], op=dist.ReduceOp.SUM, group=self.group)

    def prefetch(self, batch_sample):
        """Start the host->device copy of a collated batch (pinned PackedSequences) on a copy stream and return the
        staged batch; pass it to step() later.  Lets the copy of batch i+1 overlap the kernels of batch i, the usual
        DataLoader(pin_memory=True) + non_blocking pattern.  Batches that are not plain dicts are returned as is."""
        if not isinstance(batch_sample, dict):
            return batch_sample
        from torch.nn.utils.rnn import PackedSequence
        if self.copy_stream is None:
            self.copy_stream = torch.cuda.Stream()
        main = torch.cuda.current_stream()
        staged = _Staged()
        with torch.cuda.stream(self.copy_stream):
            for k, v in batch_sample.items():
                if isinstance(v, PackedSequence):
                    data = v.data.to(device=self.flat_p.device, non_blocking=True)
                    data.record_stream(main)
                    staged[k] = PackedSequence(data, v.batch_sizes)
                else:
                    staged[k] = v
            staged.event = torch.cuda.Event()
            staged.event.record(self.copy_stream)
        return staged

    def step(self, batch_sample):
        """One optimisation step on a collated batch.  Returns device scalars (loss, norm) of this rank's shard."""
        if getattr(batch_sample, "event", None) is not None:
            torch.cuda.current_stream().wait_event(batch_sample.event)
        kw = {"dp": self.dp} if self.exact else {}
        loss, norm = self.arch.loss_and_grads(self.model, batch_sample, self.grads, hook=self._bucket_ready,
                                              hidden=None if self.host_rng else self._device_hidden, **kw)
        loss, norm = loss.clone(), norm.clone()
        if self.world > 1:
            torch.cuda.current_stream().wait_stream(self.comm)
        self.apply_update()
        return loss, norm

    def grad_exchange_check(self, batch_sample):
        """Diagnostic for N > 1 (bench.py `dp_grad_check`, outside any timed region): one backward pass WITHOUT the
        exchange gives this rank's gradient; all ranks' gradients are gathered, and the bucketed allreduce the real step
        uses must deliver exactly their sum (the clip kernel then scales by 1/world in mode A).  Parameters and optimizer
        state are not touched.  Returns {max_abs_err, ref_max_abs, rank_spread}."""
        import torch.distributed as dist
        if self.world == 1:
            return None
        fixed = {}                                               # the same (h0, c0) for the two passes below

        def hidden(B):
            if B not in fixed:
                fixed[B] = self._device_hidden(B)
            return fixed[B]
        kw = {"dp": self.dp} if self.exact else {}
        bn = [b.clone() for b in (self.model.bn.running_mean, self.model.bn.running_var)]
        self.arch.loss_and_grads(self.model, batch_sample, self.grads, hook=None, hidden=hidden, **kw)
        local = self.flat_g.clone()
        gathered = [torch.empty_like(local) for _ in range(self.world)]
        dist.all_gather(gathered, local, group=self.group)
        want = torch.stack(gathered).sum(0)
        self.arch.loss_and_grads(self.model, batch_sample, self.grads, hook=self._bucket_ready, hidden=hidden, **kw)
        torch.cuda.current_stream().wait_stream(self.comm)
        err = float((self.flat_g - want).abs().max())
        spread = float((gathered[0] - gathered[-1]).abs().max())
        for dst, src in zip((self.model.bn.running_mean, self.model.bn.running_var), bn):
            dst.copy_(src)
        return {"max_abs_err": err, "ref_max_abs": float(want.abs().max()), "rank_spread_max_abs": spread,
                "exchange": "sum over ranks of flat buckets (l0 | l1 | head); scaled by %s in the clip kernel"
                            % ("1" if self.exact else "1/world")}

    def step_accumulate(self, batches):
        """One optimisation step over several micro-batches (gradient accumulation): every micro-batch runs forward and
        backward with its own loss normaliser and BatchNorm statistics, the gradients add up in the flat bucket, the
        exchange starts with the LAST micro-batch's buckets, and the mean over micro-batches (and ranks) is clipped and
        applied once.  This is how a fixed global batch larger than 64 x world is run (strong scaling, bench.py
        --global-batch).  uPIT arch.  Returns the last micro-batch's (loss, norm)."""
        n = len(batches)
        loss = norm = None
        for i, b in enumerate(batches):
            if getattr(b, "event", None) is not None:
                torch.cuda.current_stream().wait_event(b.event)
            last = i == n - 1
            loss, norm = self.arch.loss_and_grads(self.model, b, self.grads, hook=self._bucket_ready if last else None,
                                                  hidden=None if self.host_rng else self._device_hidden, accumulate=i > 0)
        loss, norm = loss.clone(), norm.clone()
        if self.world > 1:
            torch.cuda.current_stream().wait_stream(self.comm)
        self.apply_update(extra_scale=1.0 / n)
        return loss, norm

    def _device_hidden(self, batch_size):
        dev = self.flat_p.device
        n = 2 * self.model.num_layers
        return (torch.randn(n, batch_size, 600, device=dev), torch.randn(n, batch_size, 600, device=dev))

    def apply_update(self, extra_scale=1.0):
        """clip_grad_norm_(clip) + Adam.step() on the flat buckets (steps/train_qsub.py:121-122)."""
        st = stream_ptr()
        gscale = (1.0 if self.exact else 1.0 / self.world) * extra_scale   # mode B sums the ranks' gradients, mode A averages
        self.steps += 1
        call("grad_norm", self.flat_g, LL(self.n), gscale, self.partial, self.norm, st)
        call("clip_adam", self.flat_p, self.flat_g, self.flat_m, self.flat_v, LL(self.n), self.norm, gscale,
             self.clip, self.lr, float(self.betas[0]), float(self.betas[1]), self.eps, self.steps, st)
        self.model._wt_versions = None     # parameters changed behind autograd's back: refresh the operand copies ...
        # ... right away (one CUDA-graph replay behind the update): it then runs while the host prepares the next batch
        # instead of in front of the next step's first GEMM
        if EAGER_REFRESH and hasattr(self.model, "weights") and hasattr(self.model, "param_dict"):
            self.model.weights(self.model.param_dict(), need_grad_copies=True)
