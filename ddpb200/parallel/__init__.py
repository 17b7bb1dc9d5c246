"""Process-group bootstrap, peer-memory buffers and the gradient-sync strategies.

``sync_gradients_2a`` / ``sync_gradients_2b`` keep the reference's function-level API
(``sync_gradients(model, rank, world_size)`` at part2/part2a/main.py:97-115 and
``sync_gradients(model, world_size)`` at part2/part2b/main.py:97-103): call them after
``loss.backward()`` and before ``optimizer.step()``.  On CUDA they launch the peer-memory
kernel over the model's gradient arena; on CPU they run the torch.distributed restatement.
"""
from __future__ import annotations

import torch.distributed as dist

from . import sync as _sync
from .comm import init_distributed_setup, test_distributed_setup  # noqa: F401


def _cuda_sync_context(model):
    """(trainer, engine) owning the model's gradient arena, creating them on first use."""
    inner = getattr(model, "module", model)
    tr = getattr(model, "_ddpb_trainer", None) or getattr(inner, "_ddpb_trainer", None)
    if tr is None or tr.symm is None:
        from ..trainer import Trainer

        eng = getattr(inner, "_engine", None)
        batch = eng.max_batch if eng is not None else 256
        tr = Trainer(inner, mode="allreduce", batch_size=batch, rank=dist.get_rank(), world_size=dist.get_world_size(),
                     use_graph=False, fuse_sgd=False)
        object.__setattr__(model, "_ddpb_trainer", tr)
    return tr


def sync_gradients_2a(model, rank: int, world_size: int) -> None:
    """Mean of every gradient computed on rank 0 and sent back (part2a)."""
    p0 = next(model.parameters())
    if not p0.is_cuda:
        return _sync.sync_gradients_gather_scatter(model, rank, world_size)
    tr = _cuda_sync_context(model)
    a = tr.engine.arena
    tr.symm.reduce_root_bcast(0, a.numel, root=0, blocks=tr.comm_blocks, use_mc=tr.use_mc)


def sync_gradients_2b(model, world_size: int) -> None:
    """all_reduce(SUM) then divide by world_size for every parameter (part2b)."""
    p0 = next(model.parameters())
    if not p0.is_cuda:
        return _sync.sync_gradients_allreduce(model, world_size)
    tr = _cuda_sync_context(model)
    a = tr.engine.arena
    tr.symm.allreduce_avg(0, a.numel, blocks=tr.comm_blocks, use_mc=tr.use_mc)
