#!/usr/bin/env python3
"""Drop-in for the reference's trainer `steps/train_qsub.py`: same command line, same files in `dirout`.

    python -m ssb200.train_qsub <arch-file> <gpu-id> <data-dir> <dirout> [--cv-data-dir D] [--train-copy-location L]
        [--model-config F] [--batch-size 100] [--start-epoch 0] [--num-epochs 200] [--learning-rate 0.001]

What it keeps from the reference (file:line in /root/reference/steps/train_qsub.py):
  * the arch-module protocol (:66, :80-84, :91-95): TrainSet/.collator, SepDNN(gpu_id, **model_config), compute_cv_loss;
  * DataLoader(batch_size, shuffle=True, collate_fn, num_workers=1) for training, unshuffled for CV (:81-84);
  * init.mdl at epoch 0, NNN.mdl every 5th epoch, final.mdl, `--start-epoch N` resumes from NNN.mdl (:104-110, :148-155) --
    `.mdl` files are plain `state_dict`s with the reference's key names, so they interchange with the reference;
  * train_stats/train_loss.txt and cv_loss.txt, one "NNN <loss>" line per epoch, appended (:98-102, :139-146);
  * the CV pass every 5th epoch under eval() / no_grad (:124-141);
  * Adam(lr) state is NOT checkpointed (the reference re-creates the optimizer on resume, :95).
What changes:
  * the step is ssb200.train.Trainer.step -- tape-free forward/backward, bucketed gradient exchange, fused global-norm
    clip (0.25) + Adam on flat buffers -- instead of compute_loss / backward / clip_grad_norm_ / optimizer.step (:117-122);
  * the epoch loss is accumulated ON THE DEVICE (sum of loss*norm and of norm, one 2-float kernel per step) and read back
    once per epoch; the reference reads two scalars back per step (:118-119), which drains the launch queue every step;
  * the next batch's host->device copy is started before the current step is enqueued (Trainer.prefetch);
  * under torchrun (WORLD_SIZE > 1) every rank trains on its own shard of each epoch (DistributedSampler), gradients are
    averaged (DP mode A) or the run reproduces the large-batch arithmetic (`--dp-exact`, mode B); rank 0 writes the files;
  * plots are skipped when `tools/plot.py` / matplotlib is not importable.
"""
import argparse
import importlib
import os
import sys

import numpy as np
import torch
from torch.utils.data import DataLoader

from ._lib import LL, call, stream_ptr


def get_args(argv=None):
    parser = argparse.ArgumentParser(description="""This script trains a separation neural network""")
    parser.add_argument("arch_file", metavar="arch-file", type=str, help="DNN architecture file")
    parser.add_argument("gpu_id", metavar="gpu-id", type=int, help="GPU ID")
    parser.add_argument("data_dir", metavar="data-dir", type=str, help="Training data directory")
    parser.add_argument("dirout", type=str, help="Output directory")
    parser.add_argument("--cv-data-dir", type=str, help="Cross validation data directory", default="")
    parser.add_argument("--train-copy-location", type=str, help="Copy training data here for I/O purposes", default="")
    parser.add_argument("--model-config", type=str, help="Config file for DNN", default="")
    parser.add_argument("--batch-size", type=int, help="Batch size", default=100)
    parser.add_argument("--start-epoch", type=int, help="Epoch to start from", default=0)
    parser.add_argument("--num-epochs", type=int, help="Total number of training epochs", default=200)
    parser.add_argument("--learning-rate", type=float, help="Learning rate", default=0.001)
    # extensions (not in the reference)
    parser.add_argument("--host-rng", action="store_true",
                        help="draw (h0, c0) on the CPU generator like the reference's init_hidden (default: device generator)")
    parser.add_argument("--dp-exact", action="store_true",
                        help="data-parallel mode B: loss normaliser and BatchNorm statistics over ALL ranks' utterances")
    parser.add_argument("--num-workers", type=int, default=1, help="DataLoader workers (reference: 1)")
    return parser.parse_args(argv)


def load_arch(name):
    """`uPIT` / `RSH` (optionally with a `_b200` suffix, the shim files of speech-separation_b200/archs/) -> the ssb200
    module; anything else is imported by bare name like the reference does (archs/ on sys.path)."""
    base = name[:-5] if name.endswith("_b200") else name
    if base in ("uPIT", "RSH"):
        return importlib.import_module("ssb200." + base)
    sys.path.append("archs")
    return __import__(name)


def load_losses(filename, loss_array):
    with open(filename, "r") as f:
        for line in f:
            split = line.rstrip().split()
            loss_array[0].append(int(split[0]))
            loss_array[1].append(float(split[1]))


def read_model_config(path):
    kwargs = {}
    for line in open(path):
        if "=" in line:
            kwargs[line.split("=")[0]] = line.rstrip().split("=")[1]
    return kwargs


class EpochLoss:
    """Device-side accumulator of (sum loss*norm, sum norm): stats[0:2] of every step are added by one tiny kernel."""

    def __init__(self, device):
        self.acc = torch.zeros(2, dtype=torch.float32, device=device)

    def add(self, stats):
        call("axpby", stats, self.acc, LL(2), 1.0, 1.0, stream_ptr())

    def result(self, group=None, world=1):
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(self.acc, group=group)
        s, n = (float(v) for v in self.acc.cpu())          # the ONE device->host read of the epoch
        self.acc.zero_()
        return s / n if n > 0 else float("nan")


def _pin(batch):
    """Collated batch (dict of PackedSequence / lists, or RSH's MultiSpkBatch of them) -> pinned host memory."""
    from torch.nn.utils.rnn import PackedSequence
    if isinstance(batch, dict):
        return {k: (PackedSequence(v.data.pin_memory(), v.batch_sizes) if isinstance(v, PackedSequence) else v)
                for k, v in batch.items()}
    if hasattr(batch, "sub_batches"):
        batch.sub_batches = [_pin(b) for b in batch.sub_batches]
    return batch


def main(argv=None):
    args = get_args(argv)
    print("Using " + args.arch_file + " DNN architecture")
    m = load_arch(args.arch_file)
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    gpu = int(os.environ.get("LOCAL_RANK", args.gpu_id)) if world > 1 else args.gpu_id
    print("Using GPU", gpu)
    torch.cuda.set_device(gpu)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", gpu))
    lead = rank == 0

    int_model_dir = args.dirout + "/intermediate_models/"
    plot_dir = args.dirout + "/train_stats/plots/"
    loss_file = args.dirout + "/train_stats/train_loss.txt"
    cv_loss_file = args.dirout + "/train_stats/cv_loss.txt"
    if lead:   # run_train.sh:50-52 creates these; harmless when they exist
        os.makedirs(int_model_dir, exist_ok=True)
        os.makedirs(plot_dir, exist_ok=True)

    print("loading datset")
    dataset = m.TrainSet(args.data_dir, args.train_copy_location)
    sampler = None
    if world > 1:
        from torch.utils.data.distributed import DistributedSampler
        sampler = DistributedSampler(dataset, num_replicas=world, rank=rank, shuffle=True, drop_last=True)
    dataloader = DataLoader(dataset, batch_size=args.batch_size, shuffle=sampler is None, sampler=sampler,
                            collate_fn=dataset.collator, num_workers=args.num_workers,
                            drop_last=world > 1)      # every rank must take the same number of steps
    cv_dataloader = None
    if args.cv_data_dir:
        cv_dataset = m.TrainSet(args.cv_data_dir)
        cv_sampler = None
        if world > 1:
            from torch.utils.data.distributed import DistributedSampler
            cv_sampler = DistributedSampler(cv_dataset, num_replicas=world, rank=rank, shuffle=False)
        cv_dataloader = DataLoader(cv_dataset, batch_size=args.batch_size, collate_fn=cv_dataset.collator, sampler=cv_sampler)

    print("initializing model")
    kwargs = read_model_config(args.model_config) if args.model_config else {}
    model = m.SepDNN(gpu, **kwargs)
    model.cuda()
    print("using lr=" + str(args.learning_rate))

    epoch_losses, epoch_cv_losses = [[], []], [[], []]
    lossF = open(loss_file, "a") if lead else None
    cv_lossF = open(cv_loss_file, "a") if (lead and args.cv_data_dir) else None
    if args.start_epoch == 0:
        if lead:
            save_model(model, int_model_dir + "init.mdl")
    else:
        sd = torch.load(int_model_dir + str(args.start_epoch).zfill(3) + ".mdl", map_location="cpu")
        model.load_state_dict(sd)
        load_losses(loss_file, epoch_losses)
        if args.cv_data_dir:
            load_losses(cv_loss_file, epoch_cv_losses)

    from . import train
    trainer = train.Trainer(model, m, lr=args.learning_rate, clip=0.25, host_rng=args.host_rng, exact=args.dp_exact)
    dev = trainer.flat_p.device
    acc = EpochLoss(dev)
    try:
        import plot
    except Exception:
        plot = None

    print("training")
    for epoch in range(args.start_epoch, args.num_epochs):
        if sampler is not None:
            sampler.set_epoch(epoch)
        it = iter(dataloader)
        nxt = next(it, None)
        staged = trainer.prefetch(_pin(nxt)) if nxt is not None else None
        while staged is not None:
            cur = staged
            nxt = next(it, None)                               # the worker collates batch i+1 while step i runs
            staged = trainer.prefetch(_pin(nxt)) if nxt is not None else None
            trainer.step(cur)
            acc.add(model.last["stats"])
        epoch_loss = acc.result(trainer.group, world)

        if cv_dataloader is not None and epoch % 5 == 4:
            model.eval()
            with torch.no_grad():
                for i_batch_cv, sample_batch_cv in enumerate(cv_dataloader):
                    m.compute_cv_loss(model, epoch, sample_batch_cv)
                    acc.add(model.last["stats"])
            model.train()
            cv_loss = acc.result(trainer.group, world)
            print("For epoch: " + str(epoch + 1).zfill(3) + " cv set loss is: " + str(cv_loss))
            if lead:
                cv_lossF.write(str(epoch + 1).zfill(3) + " " + str(cv_loss) + "\n")
                cv_lossF.flush()
            epoch_cv_losses[0].append(epoch + 1)
            epoch_cv_losses[1].append(cv_loss)

        print("For epoch: " + str(epoch + 1).zfill(3) + " loss is: " + str(epoch_loss))
        if lead:
            lossF.write(str(epoch + 1).zfill(3) + " " + str(epoch_loss) + "\n")
            lossF.flush()
        epoch_losses[0].append(epoch + 1)
        epoch_losses[1].append(epoch_loss)
        if epoch % 5 == 4 and lead:
            print("Saving model for epoch " + str(epoch + 1).zfill(3))
            save_model(model, int_model_dir + str(epoch + 1).zfill(3) + ".mdl")
            if plot is not None:
                os.makedirs(plot_dir + "epoch" + str(epoch + 1).zfill(3), exist_ok=True)
                plot.plot_loss(epoch_losses, epoch_cv_losses, plot_dir + "epoch" + str(epoch + 1).zfill(3) + "/Loss_"
                               + str(epoch_losses[0][0]).zfill(3) + "-" + str(epoch + 1).zfill(3) + ".png")
        sys.stdout.flush()

    if lead:
        save_model(model, args.dirout + "/final.mdl")
        if plot is not None and epoch_losses[0]:
            plot.plot_loss(epoch_losses, epoch_cv_losses, plot_dir + "Loss_" + str(epoch_losses[0][0]).zfill(3) + "-"
                           + str(args.num_epochs).zfill(3) + ".png")
        for f in (lossF, cv_lossF):
            if f is not None:
                f.close()
    if world > 1:
        dist.barrier()
    return model


def save_model(model, path):
    """torch.save(model.state_dict()) like the reference (:105, :150, :155), as independent CPU tensors: the module's
    parameters alias ONE flat device buffer, which torch.save would otherwise write out once per view."""
    torch.save({k: v.detach().cpu().clone() for k, v in model.state_dict().items()}, path)


if __name__ == "__main__":
    main()
