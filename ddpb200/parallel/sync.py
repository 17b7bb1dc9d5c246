"""Gradient-synchronisation strategies (the reference's "L3" layer).

Three algorithms, same semantics as the reference:

  gather_scatter  part2a: every gradient is gathered on rank 0, averaged there and sent
                  back (/root/reference/part2/part2a/main.py:97-115)
  allreduce       part2b: per-parameter all_reduce(SUM) followed by a divide
                  (/root/reference/part2/part2b/main.py:97-103)
  ddp             part3: DistributedDataParallel -- gradients are grouped into buckets in
                  the order backward produces them and each bucket is all-reduced
                  asynchronously while backward continues (/root/reference/part3/main.py:174)

On B200 each strategy is a peer-memory kernel launched by :class:`ddpb200.trainer.Trainer`
(csrc/comm.cu).  The functions in this file are the host-side restatements over
``torch.distributed`` used when the model lives on the CPU (Gloo), which keeps every
layer above the kernels testable on a machine without GPUs; they are also the oracle
for the multi-GPU tests.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist

MIB = 1024 * 1024


# ---------------------------------------------------------------------------------------
# part2a / part2b restatements (CPU / any torch.distributed backend)
# ---------------------------------------------------------------------------------------
def sync_gradients_gather_scatter(model: torch.nn.Module, rank: int, world_size: int, root: int = 0) -> None:
    """Mean of every gradient computed on ``root`` and scattered back (part2a)."""
    for param in model.parameters():
        g = param.grad.data
        buf = g.contiguous()
        if rank == root:
            gathered = [torch.zeros_like(buf) for _ in range(world_size)]
            dist.gather(buf, gathered, dst=root)
            mean = torch.mean(torch.stack(gathered), dim=0)
            dist.scatter(buf, [mean for _ in range(world_size)], src=root)
        else:
            dist.gather(buf, dst=root)
            dist.scatter(buf, src=root)
        if buf.data_ptr() != g.data_ptr():
            g.copy_(buf)


def sync_gradients_allreduce(model: torch.nn.Module, world_size: int) -> None:
    """all_reduce(SUM) then divide, one collective per parameter (part2b)."""
    for param in model.parameters():
        g = param.grad.data
        buf = g.contiguous()
        dist.all_reduce(buf, op=dist.ReduceOp.SUM)
        buf /= world_size
        if buf.data_ptr() != g.data_ptr():
            g.copy_(buf)


def verify_param_shapes_across_ranks(model: torch.nn.Module) -> None:
    """What DistributedDataParallel's constructor checks before it broadcasts the parameters
    (torch `_verify_param_shape_across_processes`): every rank must hold the same
    list of parameter shapes.  Raises RuntimeError on every rank if one differs."""
    import hashlib

    desc = ";".join(f"{n}:{tuple(p.shape)}:{p.dtype}" for n, p in model.named_parameters())
    digest = hashlib.sha1(desc.encode()).hexdigest()
    seen = [None] * dist.get_world_size()
    dist.all_gather_object(seen, digest)
    if any(d != seen[0] for d in seen):
        bad = [r for r, d in enumerate(seen) if d != seen[0]]
        raise RuntimeError(f"parameter shapes differ across ranks (ranks {bad} disagree with rank 0); "
                           "every rank must build the same model")


# ---------------------------------------------------------------------------------------
# bucket layout shared by the GPU reducer and the CPU restatement
# ---------------------------------------------------------------------------------------
@dataclass
class Bucket:
    first_param: int   # parameter indices (model.parameters() order), inclusive
    last_param: int
    numel: int
    ready_after_group: int  # the bucket is complete once this parameter group's backward is done


def plan_buckets(group_sizes: list[tuple[int, int, int]], first_cap_bytes: int = MIB,
                 cap_bytes: int = 25 * MIB, tail_bytes: int = 0) -> list[Bucket]:
    """Assigns parameter groups to buckets the way torch's Reducer does after its first
    iteration: walk the groups in gradient-ready order (last layer first) and close a
    bucket when it EXCEEDS its cap -- 1 MiB for the first bucket, 25 MiB afterwards
    (SURVEY.md §2.4, C8).  ``group_sizes`` lists (first_param, last_param, numel) per
    layer-sized group in model order.  For VGG-11 this yields the reference's three
    buckets of 9.46 / 27.16 / 0.31 MB.
    """
    buckets: list[Bucket] = []
    cur: list[int] = []
    cur_bytes = 0
    cap = first_cap_bytes
    # tail_bytes > 0: the first (input-side) groups, up to that many bytes in total, are kept
    # out of the walk and form the LAST bucket -- the one that is reduced at full width after
    # backward -- so that it stays small whatever the cap of the other buckets is.
    tail_end = 0
    acc = 0
    while tail_bytes > 0 and tail_end < len(group_sizes) - 1 and acc + group_sizes[tail_end][2] * 4 <= tail_bytes:
        acc += group_sizes[tail_end][2] * 4
        tail_end += 1
    for gi in range(len(group_sizes) - 1, tail_end - 1, -1):
        cur.append(gi)
        cur_bytes += group_sizes[gi][2] * 4
        if cur_bytes > cap:
            buckets.append(_close(group_sizes, cur))
            cur, cur_bytes, cap = [], 0, cap_bytes
    if cur:
        buckets.append(_close(group_sizes, cur))
    if tail_end > 0:
        buckets.append(_close(group_sizes, list(range(tail_end))))
    return buckets


def _close(group_sizes, members: list[int]) -> Bucket:
    lo, hi = min(members), max(members)
    return Bucket(group_sizes[lo][0], group_sizes[hi][1], sum(group_sizes[g][2] for g in members), lo)


def param_groups_of(model) -> list[tuple[int, int, int]]:
    """One group per conv block (conv w, conv b, bn w, bn b) plus one for the classifier."""
    params = list(model.parameters())
    index = {id(p): i for i, p in enumerate(params)}
    groups = []
    for blk in model.conv_blocks():
        idx = [index[id(t)] for t in (blk.conv.weight, blk.conv.bias, blk.bn.weight, blk.bn.bias)]
        groups.append((min(idx), max(idx), sum(params[i].numel() for i in idx)))
    fc = [index[id(model.fc1.weight)], index[id(model.fc1.bias)]]
    groups.append((min(fc), max(fc), sum(params[i].numel() for i in fc)))
    return groups


# ---------------------------------------------------------------------------------------
# part3 restatement for CPU: bucketed, asynchronous all-reduce driven by autograd hooks
# ---------------------------------------------------------------------------------------
class HostBucketReducer:
    """DDP-style reducer over torch.distributed for CPU tensors (oracle + CPU path)."""

    def __init__(self, model, world_size: int, first_cap_bytes: int = MIB, cap_bytes: int = 25 * MIB):
        self.params = list(model.parameters())
        self.world = world_size
        self.buckets = plan_buckets(param_groups_of(model), first_cap_bytes, cap_bytes)
        self._owner = {}
        for bi, b in enumerate(self.buckets):
            for pi in range(b.first_param, b.last_param + 1):
                self._owner[pi] = bi
        self._pending = [0] * len(self.buckets)
        self._work = []
        self.enabled = True
        for pi, p in enumerate(self.params):
            p.register_post_accumulate_grad_hook(self._make_hook(pi))
        self.reset()

    def reset(self) -> None:
        self._pending = [b.last_param - b.first_param + 1 for b in self.buckets]
        self._work = []

    def _make_hook(self, pi: int):
        def hook(_p):
            if not self.enabled:
                return
            bi = self._owner[pi]
            self._pending[bi] -= 1
            if self._pending[bi] == 0:
                b = self.buckets[bi]
                members = self.params[b.first_param:b.last_param + 1]
                flat = torch.cat([q.grad.reshape(-1) for q in members]) / self.world  # DDP pre-divides
                self._work.append((dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=True), flat, members))
        return hook

    def finalize(self) -> None:
        """Waits for all buckets and copies the averaged slices back into .grad."""
        for work, flat, members in self._work:
            work.wait()
            o = 0
            for q in members:
                q.grad.copy_(flat[o:o + q.numel()].view_as(q.grad))
                o += q.numel()
        self.reset()
