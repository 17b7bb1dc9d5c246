"""Training / evaluation drivers and the shared ``main`` of the four entry points.

Reference: ``train_model`` (/root/reference/part1/main.py:52-93 and its three variants),
``test_model`` (part1/main.py:96-111) and ``main`` (part1/main.py:114-130,
part2/part2b/main.py:169-195, part3/main.py:159-186).  Printed lines are byte-compatible:

    [{epoch+1}, {batch+1:5d}] loss: {mean of last 20:.3f}
    Total time for 1-39 iteration in ns: {int}
    Average time for 1-39 iteration in ns: {float}
    Test set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)

Quirks kept on purpose: the timer starts after the batch is fetched (Q6), iteration 0 is
excluded from the 1-39 window (Q7), the loss is read back every iteration (Q8), the test
loss is averaged over batches not samples (Q9), every rank evaluates the whole test set
(Q10), and ``sampler.set_epoch`` is still called (Q13).
"""
from __future__ import annotations

import os
import random
import time

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from . import cli
from .data import create_data_loaders
from .models.vgg import VGG11
from .optim import SGD, CrossEntropyLoss
from .parallel.comm import init_distributed_setup, test_distributed_setup
from .trainer import Trainer
from .utils.checkpoint import load_checkpoint, save_checkpoint

seed = 89395          # part1/main.py:14
GLOBAL_BATCH = 256    # part1/main.py:17, part2b/main.py:177


class DistributedDataParallel(nn.Module):
    """Same role as ``torch.nn.parallel.DistributedDataParallel(model)`` in part3/main.py:174:
    wraps the model (state_dict keys gain the ``module.`` prefix) and marks it for bucketed,
    overlapped gradient averaging.  The reducer itself lives in Trainer (mode 'ddp'): buckets
    are slices of the flat gradient arena, reduced by a peer-memory kernel on a side stream."""

    def __init__(self, module: nn.Module, bucket_cap_mb: float | None = None):
        """bucket_cap_mb=None: the Trainer's default bucket layout ("balanced": measured faster on NVLink than
        torch DDP's 1 MiB / 25 MiB default, which the reference ran with and which ``bucket_cap_mb=25`` or
        ``Trainer(bucket_layout="reference")`` selects); a number selects that cap in torch's scheme."""
        super().__init__()
        self.module = module
        self.bucket_cap_mb = bucket_cap_mb

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def conv_blocks(self):
        return self.module.conv_blocks()

    @property
    def fc1(self):
        return self.module.fc1


def _trainer_for(model, optimizer, mode, rank, world_size, batch_size, **kw) -> Trainer:
    tr = getattr(model, "_ddpb_trainer", None)
    if tr is not None and tr.mode == mode and tr.batch_size >= batch_size:
        return tr
    g = optimizer.param_groups[0] if optimizer is not None else dict(lr=0.1, momentum=0.9, weight_decay=1e-4)
    inner = model.module if isinstance(model, DistributedDataParallel) else model
    if isinstance(model, DistributedDataParallel) and model.bucket_cap_mb is not None:
        kw.setdefault("bucket_layout", "reference")
        kw.setdefault("bucket_bytes", int(model.bucket_cap_mb * 1024 * 1024))
    tr = Trainer(inner, mode=mode, batch_size=batch_size, lr=g["lr"], momentum=g["momentum"],
                 weight_decay=g["weight_decay"], rank=rank, world_size=world_size, optimizer=optimizer, **kw)
    if isinstance(optimizer, SGD) and getattr(tr, "engine", None) is not None:
        optimizer.bind(tr.engine)
    object.__setattr__(model, "_ddpb_trainer", tr)
    return tr


def _batches(loader):
    """(data, target, aug) triples: our loaders' raw uint8 view when they have one, else whatever
    (data, target) pairs the loader yields (e.g. a torch DataLoader)."""
    if hasattr(loader, "raw"):
        yield from loader.raw()
    else:
        for data, target in loader:
            yield data, target, None


def train_model(model, *args, rank: int | None = None, world_size: int | None = None, mode: str | None = None,
                max_iters: int | None = None, **trainer_kw):
    """One epoch.  Accepts both positional orders of the reference:

        train_model(model, train_loader, optimizer, criterion, epoch)                      part1, part3
        train_model(model, rank, world_size, train_loader, optimizer, criterion, epoch)    part2a, part2b

    (/root/reference/part1/main.py:52, part3/main.py:98, part2/part2a/main.py:118-119,
    part2/part2b/main.py:106-107).  Without explicit rank / world_size the initialised process
    group is asked, as part3's DDP does.  ``mode`` defaults to 'ddp' for a
    DistributedDataParallel-wrapped model, 'single' for one rank, and 'allreduce' otherwise."""
    if len(args) == 6 and isinstance(args[0], int):
        rank, world_size, train_loader, optimizer, criterion, epoch = args
    elif len(args) == 4:
        train_loader, optimizer, criterion, epoch = args
    elif len(args) == 6:       # legacy ddpb200 order: (..., epoch, rank, world_size)
        train_loader, optimizer, criterion, epoch, rank, world_size = args
    else:
        raise TypeError("train_model(model, [rank, world_size,] train_loader, optimizer, criterion, epoch)")
    if world_size is None:
        world_size = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    if mode is None:
        mode = "ddp" if isinstance(model, DistributedDataParallel) else ("single" if world_size == 1 else "allreduce")
    trainer = _trainer_for(model, optimizer, mode, rank, world_size, train_loader.batch_size, **trainer_kw)
    device = trainer.device
    running_loss = 0.0
    total_time = 0
    # The loop keeps the reference's shape (timer after the batch is fetched, loss read every iteration,
    # Q6/Q8) but is software pipelined on CUDA: while step i computes, the host gathers batch i+1 and
    # its host->device copy runs on the copy stream (Trainer.prefetch); the loss comes back through
    # pinned memory (Trainer.last_loss) without draining the stream.
    batches = _batches(train_loader)
    nxt = next(batches, None)
    if nxt is not None and device.type == "cuda":
        trainer.prefetch(*nxt)
    batch_idx = -1
    while nxt is not None:
        batch_idx += 1
        if max_iters is not None and batch_idx >= max_iters:
            break
        data, target, aug = nxt
        start_time = time.perf_counter_ns()
        if device.type != "cuda":
            data, target = data.to(device), target.to(device)
        # zero_grad / forward / loss / backward / [sync] / optimizer.step, fused
        trainer.step(data, target, aug)
        nxt = next(batches, None)
        if nxt is not None and device.type == "cuda" and (max_iters is None or batch_idx + 1 < max_iters):
            trainer.prefetch(*nxt)
        running_loss += trainer.last_loss()
        if batch_idx % 20 == 19:
            print(f"[{epoch + 1}, {batch_idx + 1:5d}] loss: {running_loss / 20:.3f}")
            running_loss = 0.0
        if 0 < batch_idx < 40:
            total_time += time.perf_counter_ns() - start_time
        if batch_idx == 39:
            print(f"Total time for 1-39 iteration in ns: {total_time}")
            print(f"Average time for 1-39 iteration in ns: {total_time / 39.0}")
    return None


def test_model(model, test_loader, criterion, max_iters: int | None = None):
    trainer = getattr(model, "_ddpb_trainer", None)
    if trainer is None:
        trainer = _trainer_for(model, None, "single", 0, 1, test_loader.batch_size)
    model.eval()
    test_loss = 0.0
    correct = 0
    batches = 0
    for batch_idx, (data, target, _aug) in enumerate(_batches(test_loader)):
        if max_iters is not None and batch_idx >= max_iters:
            break
        loss, ok = trainer.evaluate(data, target)
        test_loss += float(loss.item())
        correct += int(ok.item())
        batches += 1
    n_batches = len(test_loader) if max_iters is None else max(batches, 1)
    test_loss /= n_batches
    total = len(test_loader.dataset)
    print("Test set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n".format(
        test_loss, correct, total, 100.0 * correct / total))
    return test_loss, correct


def seed_everything(s: int = seed) -> None:
    random.seed(s)
    torch.manual_seed(s)
    np.random.seed(s)


def pick_device(arg: str | None, local_rank: int = 0) -> torch.device:
    if arg in (None, "cuda") and torch.cuda.is_available():
        return torch.device("cuda", local_rank % torch.cuda.device_count())
    if arg == "cuda":
        raise SystemExit("--device cuda requested but no CUDA device is visible")
    if arg not in (None, "cpu"):
        return torch.device(arg)
    torch.set_num_threads(4)  # part1/main.py:16
    return torch.device("cpu")


def run(part: str, argv=None) -> None:
    """Shared main of part1 / part2a / part2b / part3."""
    mode = {"part1": "single", "part2a": "gather_scatter", "part2b": "allreduce", "part3": "ddp"}[part]
    distributed = part != "part1"
    args = cli.parse_all(distributed, argv)
    if distributed and args.spawn:
        return _spawn(part, args, argv)
    rank, world_size = (args.rank, args.size) if distributed else (0, 1)
    if distributed and world_size is None:
        raise SystemExit("--num-nodes is required (README.md:10)")
    device = pick_device(args.device, rank)
    if distributed:
        init_distributed_setup(args.master_ip, args.master_port, rank, world_size, backend=args.backend,
                               device=device if device.type == "cuda" else None)
        test_distributed_setup()
    seed_everything()
    batch_size = int(args.batch_size / world_size)   # part2b/main.py:177
    criterion = CrossEntropyLoss().to(device)
    root = args.data_root if args.data_root is not None else ("./../data" if part in ("part1", "part3") else "./../../data")
    train_loader, test_loader = create_data_loaders(rank, world_size, batch_size, root=root,
                                                    synthetic=True if args.synthetic else None,
                                                    train_size=args.train_size, test_size=args.test_size,
                                                    augment=not args.device_augment)
    model = VGG11()
    model.to(device)
    blob = load_checkpoint(args.resume, model) if args.resume else None
    net = DistributedDataParallel(model) if part == "part3" else model
    optimizer = SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)
    kw = dict(use_graph=not args.no_graph, device_augment=args.device_augment)
    if distributed:
        kw.update(granularity=args.sync_granularity, use_multicast=False if args.no_multicast else "auto")
    if blob is not None and "optimizer" in blob:
        # the momentum buffers live in the engine's arena: bind first, then restore them, so a
        # resumed run continues exactly where the saved one stopped
        _trainer_for(net, optimizer, mode, rank, world_size, train_loader.batch_size, **kw)
        optimizer.load_state_dict(blob["optimizer"])
    for epoch in range(args.epochs):
        train_loader.sampler.set_epoch(epoch)
        train_model(net, train_loader, optimizer, criterion, epoch, rank=rank, world_size=world_size, mode=mode,
                    max_iters=args.max_iters, **kw)
        test_model(net, test_loader, criterion, max_iters=args.max_iters)
    if args.save and rank == 0:
        save_checkpoint(args.save, net, optimizer)
    tr = getattr(net, "_ddpb_trainer", None)
    if tr is not None:
        tr.close()
    if distributed:
        dist.destroy_process_group()


def _spawn(part, args, argv):
    """Single-host convenience launcher: one process per rank (the reference has none and is
    started by hand on every node, README.md:8-10)."""
    import subprocess
    import sys

    base = [a for a in (argv if argv is not None else sys.argv[1:]) if a != "--spawn"]
    script = os.path.abspath(sys.argv[0])
    procs = [subprocess.Popen([sys.executable, script, *base, "--rank", str(r)]) for r in range(args.size)]
    rc = 0
    for p in procs:
        rc = rc or p.wait()
    if rc:
        raise SystemExit(rc)
