"""Trainer: one object that performs the reference's training iteration in any of its four modes.

The reference spells the iteration out in ``train_model``: ``zero_grad`` -> forward ->
loss -> backward -> [sync] -> ``optimizer.step()`` (/root/reference/part1/main.py:71-77,
part2/part2a/main.py:137-144, part2/part2b/main.py:125-132, part3/main.py:116-122).
``Trainer.step(data, target)`` is that whole sequence:

  mode            reference      gradient sync here
  'single'        part1          none
  'gather_scatter' part2a        reduce-to-root + broadcast kernel
  'allreduce'     part2b         in-place two-shot all-reduce kernel (per parameter or fused)
  'ddp'           part3          same kernel per bucket, on a side stream, started as soon as
                                 the bucket's last gradient is enqueued (overlaps backward)

On CUDA the iteration is a CUDA graph over the hand-written kernels (engine.py,
csrc/*.cu) and the sync kernels also apply the SGD update to their range; on CPU it
runs the stock torch modules with the restatements in parallel/sync.py over Gloo.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn.functional as F

from .parallel import sync as S

MODES = ("single", "gather_scatter", "allreduce", "ddp")


class Trainer:
    def __init__(self, model, mode: str = "single", batch_size: int = 256, lr: float = 0.1, momentum: float = 0.9,
                 weight_decay: float = 1e-4, rank: int = 0, world_size: int = 1, granularity: str = "fused",
                 use_multicast: bool | str = "auto", use_graph: bool = True, comm_blocks: int | None = None,
                 bucket_layout: str = "balanced", first_bucket_bytes: int | None = None,
                 bucket_bytes: int | None = None, symm_backend: str | None = None,
                 fuse_sgd: bool = True, overlap_wgrad: bool = True, comm_threads: int | None = None, optimizer=None,
                 overlap_sgd: bool = True, debug_poison_grads: bool = False, device_augment: bool = False,
                 augment_seed: int = 0, augment_pad: int = 4):
        if mode not in MODES:
            raise ValueError(f"mode must be one of {MODES}")
        if granularity not in ("fused", "param", "param_launches"):
            raise ValueError("granularity must be 'fused', 'param' or 'param_launches'")
        self.model, self.mode = model, mode
        self.batch_size = batch_size
        self.lr, self.momentum, self.weight_decay = lr, momentum, weight_decay
        self.rank, self.world = rank, world_size
        self.granularity = granularity
        self.use_graph = use_graph
        self.fuse_sgd = fuse_sgd
        self.overlap_wgrad = overlap_wgrad
        # The SGD update of a layer only needs that layer's gradients and that nobody reads its
        # weights any more in this step (its dgrad is enqueued).  With overlap_sgd the update of each
        # parameter group ('single') / each reduced bucket ('ddp') runs on a side stream while the
        # rest of backward continues, instead of as one 30 us pass over all 9.2 M parameters at the end.
        self.overlap_sgd = overlap_sgd
        # Race detector for the overlapped modes (tests): every iteration starts by filling the whole
        # gradient arena with NaN.  The backward kernels overwrite all of it, so a reduction or an SGD
        # update that ran BEFORE its gradients were final would pick up NaNs and poison the parameters.
        self.debug_poison_grads = debug_poison_grads
        # device_augment: uint8 batches given WITHOUT crop/flip draws get them drawn on the GPU inside the
        # captured step (Philox keyed by augment_seed and the rank, counter = step number kept on the
        # device), so the host sends 3 KiB of pixels per image and nothing else.
        self.device_augment = device_augment
        self.augment_seed = (int(augment_seed) * 0x9E3779B97F4A7C15 + rank) & 0xFFFFFFFFFFFFFFFF
        self.augment_pad = augment_pad
        self._aug_step_cpu = 0      # CPU path: the step counter the device keeps in _aug_step
        # CTA size of the reductions that overlap backward (mode 'ddp'): 32 CTAs of 256 threads -- the same
        # number of threads as round 1's 16 x 512, spread over twice as many SMs with half the footprint
        # each (2 GPUs, balanced buckets: 0.754 ms per step against 0.772 with 16 x 512, profiles/r2_call23_*)
        self.comm_threads = comm_threads if comm_threads is not None else (256 if mode == "ddp" else 512)
        self.device = next(model.parameters()).device
        self.distributed = world_size > 1 and mode != "single"
        self.iteration = 0
        # ``optimizer``: the user's optimizer object.  Its param_groups[0] stays the source of truth
        # for lr / momentum / weight_decay (checked before every step, so an LR scheduler works);
        # on CPU it is also the object that is stepped, so its state_dict carries the momentum.
        self._hyper_source = optimizer
        self._graphs: dict[str, torch.cuda.CUDAGraph] = {}
        # Bucket layout of mode 'ddp'.  "balanced" (default): {fc, L7} {L6..L4} {L3, L2} {L1, L0} -- ~20 MiB
        # middle buckets and a small last one.  "reference" = torch DDP's defaults, which the reference ran
        # with (1 MiB first bucket, 25 MiB cap: 9.46 / 27.16 / 0.31 MB for VGG-11): its 27 MB bucket is only
        # complete after layer 2's dgrad, so its reduction and SGD pass pile up on the memory-bound end of
        # backward (profiles/r2_call09_2gpu_ddp_timeline_reference.txt).  "per_layer" closes a bucket after
        # every parameter group.  2 GPUs, ms per step: balanced 0.761-0.772, reference 0.800-0.807,
        # per_layer 0.830 (profiles/r2_call23_*; in round 1, with CAS barriers and a cooperative BatchNorm
        # backward, reference was the best of the three).  Explicit byte caps override the layout.
        if bucket_layout not in ("per_layer", "reference", "balanced"):
            raise ValueError("bucket_layout must be 'per_layer', 'reference' or 'balanced'")
        tail_bytes = 0
        if bucket_layout == "balanced":
            # small last bucket (first layers), ~20 MiB middle buckets: {fc,L7} {L6..L4} {L3,L2} {L1,L0}
            first_bucket_bytes = S.MIB if first_bucket_bytes is None else first_bucket_bytes
            bucket_bytes = 20 * S.MIB if bucket_bytes is None else bucket_bytes
            tail_bytes = S.MIB
        if first_bucket_bytes is None:
            first_bucket_bytes = 0 if bucket_layout == "per_layer" else S.MIB
        if bucket_bytes is None:
            bucket_bytes = 0 if bucket_layout == "per_layer" else 25 * S.MIB
        self.bucket_layout = bucket_layout
        if self.distributed and self.mode == "ddp":
            if self.device.type == "cuda":
                torch.cuda.set_device(self.device)      # object collectives of NCCL use the current device
            S.verify_param_shapes_across_ranks(model)   # DDP constructor, SURVEY.md C7
        self.groups = S.param_groups_of(model)
        self.buckets = S.plan_buckets(self.groups, first_bucket_bytes, bucket_bytes, tail_bytes)
        if self.device.type == "cuda":
            self._init_cuda(use_multicast, comm_blocks, symm_backend)
        else:
            self._init_cpu()

    # =====================================================================================
    # CUDA path
    # =====================================================================================
    def _init_cuda(self, use_multicast, comm_blocks, symm_backend):
        from .engine import VGGEngine
        from .parallel.comm import SymmetricBuffer

        torch.cuda.set_device(self.device)
        n_params = sum((p.numel() + 31) // 32 * 32 for p in self.model.parameters())
        self.symm = None
        grad_storage = None
        if self.distributed:
            self.symm = SymmetricBuffer(n_params, self.device, prefer=symm_backend)
            grad_storage = self.symm.payload
        self.engine = VGGEngine(self.model, self.batch_size, grad_storage)
        self.model._engine = self.engine
        if use_multicast == "auto":
            use_multicast = bool(self.symm is not None and self.symm.has_multicast)
        self.use_mc = bool(use_multicast and self.symm is not None and self.symm.has_multicast)
        sms = torch.cuda.get_device_properties(self.device).multi_processor_count
        # 2a / 2b reductions own the GPU (full width).  'ddp': CTAs of the reductions that overlap backward
        # (2 GPUs, reference buckets, 512 threads: 8 / 16 / 32 CTAs -> 0.951 / 0.807 / 0.773 ms, profiles/r2_call23_*)
        self.comm_blocks = comm_blocks if comm_blocks is not None else (32 if self.mode == "ddp" else min(sms, 128))
        self.comm_stream = torch.cuda.Stream(device=self.device, priority=-1) if self.mode == "ddp" else None
        # wgrad runs on its own branch, concurrently with dgrad / the next BatchNorm backward
        self.wgrad_stream = torch.cuda.Stream(device=self.device) if self.overlap_wgrad else None
        self.update_stream = torch.cuda.Stream(device=self.device) if self.overlap_sgd else None
        B = self.batch_size
        hw = self.engine.image_hw
        # Two staging sets alternate between steps, so the host->device copy of batch i+1 (issued by
        # prefetch() on the copy stream) runs while step i computes.  The fp32 staging buffer is only
        # allocated when a float batch shows up.  Set 2 belongs to evaluate().
        self._stage = [dict(u8=torch.zeros(B, hw, hw, 3, dtype=torch.uint8, device=self.device),
                            aug=torch.zeros(B, 4, dtype=torch.int8, device=self.device), f32=None,
                            labels=torch.zeros(B, dtype=torch.int64, device=self.device),
                            ready=None, free=None) for _ in range(3)]
        self._slot = 0              # staging set the NEXT step reads
        self._pending = None        # (slot, kind, with_aug, N, data object) filled by prefetch()
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._capture_stream = torch.cuda.Stream(device=self.device)
        # every step's loss is published to pinned host memory by a one-thread kernel right after the
        # classifier (sequence number + value in one 8-byte word): last_loss() polls it instead of
        # issuing a D2H copy and draining the stream
        self._loss_host = torch.zeros(1, dtype=torch.int64).pin_memory()
        self._loss_seq = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._steps_published = 0
        self._aug_step = torch.zeros(1, dtype=torch.int32, device=self.device)   # Philox counter of the draws
        from .ops import kernels as K

        if os.environ.get("DDPB_PDL") is None:
            # programmatic dependent launch: +0.6% on one GPU.  Round 1 had to switch it off in 'ddp' mode
            # (early-resident CTAs kept the overlapped reductions off the SMs while a cooperative
            # BatchNorm backward waited for all of them); with the two-kernel BatchNorm backward it helps
            # there too (2 GPUs: 0.795 ms per step with it, 0.815 without; profiles/r2_call08_*)
            K.set_pdl(True)
        if self.distributed and self.mode == "ddp":
            self._broadcast_parameters()

    @property
    def in_u8(self):
        return self._stage[self._slot]["u8"]

    @property
    def in_aug(self):
        return self._stage[self._slot]["aug"]

    def _broadcast_parameters(self):
        """DDP's constructor-time parameter broadcast from rank 0, via the arena."""
        a = self.engine.arena
        a.grad.copy_(a.master)
        torch.cuda.synchronize(self.device)
        self.symm.broadcast(0, a.numel, root=0, blocks=self.comm_blocks, use_mc=self.use_mc)
        torch.cuda.synchronize(self.device)
        a.master.copy_(a.grad)
        a.grad.zero_()
        self.engine.repack(full=True)
        torch.cuda.synchronize(self.device)

    def _sgd_args(self):
        a = self.engine.arena
        return (a.master, a.momentum, self.engine.shadow, self.lr, self.momentum, self.weight_decay)

    def _publish_loss(self):
        """Enqueued right after the classifier's backward kernel (which also reduces the loss): a
        one-thread kernel on the weight-gradient branch stores the loss into pinned host memory."""
        from .ops import kernels as K

        main = torch.cuda.current_stream()
        side = self.wgrad_stream
        if side is None:
            K.publish_scalar(self.engine.loss, self._loss_seq, self._loss_host)
            return
        ev = torch.cuda.Event()
        ev.record(main)
        side.wait_event(ev)
        with torch.cuda.stream(side):
            K.publish_scalar(self.engine.loss, self._loss_seq, self._loss_host)

    def _sgd_slice(self, off: int, n: int):
        """Plain (local) SGD update of arena range [off, off + n) on the current stream."""
        from .ops import kernels as K

        a, eng = self.engine.arena, self.engine
        K.sgd_step(a.master[off:off + n], a.grad[off:off + n], a.momentum[off:off + n], self.lr, self.momentum,
                   self.weight_decay, 1.0, shadow=eng.shadow[off:off + n])

    def _sync_range(self, off, n, slot_base, stream=None):
        sgd = self._sgd_args() if self.fuse_sgd else None
        if self.mode == "gather_scatter":
            self.symm.reduce_root_bcast(off, n, root=0, blocks=self.comm_blocks, slot_base=slot_base,
                                        use_mc=self.use_mc, sgd=sgd, stream=stream)
        else:
            self.symm.allreduce_avg(off, n, blocks=self.comm_blocks, slot_base=slot_base, use_mc=self.use_mc, sgd=sgd,
                                    stream=stream)

    def _enqueue_iteration(self, kind: str, with_aug: int, B: int | None = None, slot: int = 0,
                           one_bucket: bool = False):
        """Everything between 'inputs are on the device' and 'parameters are updated'.
        ``one_bucket`` (mode 'ddp'): reduce all gradients as ONE bucket after backward, which is what
        torch's Reducer does in iteration 0 before it has seen the gradient-ready order."""
        eng, a = self.engine, self.engine.arena
        B = self.batch_size if B is None else B
        st = self._stage[slot]
        eng.labels = st["labels"]
        # the input kernel also refreshes the first layer's padded filter pack; every other conv
        # operand is the SGD-maintained shadow
        eng.mark("step begin", reset=True)
        if self.debug_poison_grads:
            for i in range(len(a.params)):      # (not the alignment padding, which no kernel writes)
                a.flat_grad(i).fill_(float("nan"))
        if kind == "u8":
            if with_aug == 2:       # draws made on the device (device_augment)
                from .ops import kernels as K

                K.draw_augmentation(st["aug"], self._aug_step, seed=self.augment_seed, pad=self.augment_pad, n=B)
            eng.load_input_u8(st["u8"][:B], st["aug"][:B] if with_aug else None, pack=True)
        else:
            eng.load_input_f32(st["f32"][:B], pack=True)
        eng.forward(B, with_loss=True, defer_loss=True)
        if not self.distributed:
            if self.overlap_sgd and self.update_stream is not None:
                # per parameter group: update as soon as the group's gradients are final and its dgrad
                # (the last reader of its weights in this step) is enqueued
                main = torch.cuda.current_stream()
                upd = self.update_stream

                def after_layer(gi, side_events):
                    first, last, _ = self.groups[gi]
                    off, n = a.range_of(first, last)
                    ev = torch.cuda.Event()
                    ev.record(main)
                    upd.wait_event(ev)
                    for sev in side_events:           # the group's weight gradient (side branch)
                        upd.wait_event(sev)
                    with torch.cuda.stream(upd):
                        self._sgd_slice(off, n)
                        eng.mark(f"sgd group {gi} (side)")

                eng.backward(B, on_loss_ready=self._publish_loss, after_layer=after_layer, side_stream=self.wgrad_stream, reduce_loss=True)
                done = torch.cuda.Event()
                done.record(upd)
                main.wait_event(done)
                eng.mark("sgd joined")
            else:
                eng.backward(B, on_loss_ready=self._publish_loss, side_stream=self.wgrad_stream, reduce_loss=True)
                eng.sgd(self.lr, self.momentum, self.weight_decay)
        elif self.mode == "ddp" and one_bucket:
            eng.backward(B, on_loss_ready=self._publish_loss, side_stream=self.wgrad_stream, reduce_loss=True)
            wide = min(torch.cuda.get_device_properties(self.device).multi_processor_count, 128)
            self.symm.allreduce_avg(0, a.numel, blocks=wide, slot_base=1024, use_mc=self.use_mc,
                                    sgd=self._sgd_args() if self.fuse_sgd else None)
            if not self.fuse_sgd:
                eng.sgd(self.lr, self.momentum, self.weight_decay)
        elif self.mode == "ddp":
            # Early buckets: reduction only, few CTAs, on the side stream while backward continues.
            # Last bucket (the first layers, ready when backward ends): one full-width kernel that
            # also applies the SGD update to the WHOLE arena, so no separate optimizer launch.
            main = torch.cuda.current_stream()
            ready = {b.ready_after_group: bi for bi, b in enumerate(self.buckets)}
            last_bi = len(self.buckets) - 1
            early_sgd = self.overlap_sgd and self.fuse_sgd

            def after_layer(li, side_events):
                bi = ready.get(li)
                if bi is None or bi == last_bi:
                    return
                b = self.buckets[bi]
                off, n = a.range_of(b.first_param, b.last_param)
                ev = torch.cuda.Event()
                ev.record(main)
                self.comm_stream.wait_event(ev)
                for sev in side_events:           # the bucket's last weight gradient (side branch)
                    self.comm_stream.wait_event(sev)
                with torch.cuda.stream(self.comm_stream):
                    self.symm.allreduce_avg(off, n, blocks=self.comm_blocks, slot_base=bi * self.comm_blocks,
                                            use_mc=self.use_mc, sgd=None, threads=self.comm_threads)
                    eng.mark(f"bucket {bi} reduced (comm)")
                    if early_sgd:
                        # the bucket's mean is final on this rank once its reduction kernel has ended:
                        # update it now (a local kernel, no peer waits), not in the tail after backward
                        self._sgd_slice(off, n)
                        eng.mark(f"bucket {bi} sgd (comm)")

            eng.backward(B, on_loss_ready=self._publish_loss, after_layer=after_layer, side_stream=self.wgrad_stream, reduce_loss=True)
            done = torch.cuda.Event()
            done.record(self.comm_stream)
            main.wait_event(done)
            eng.mark("comm joined")
            b = self.buckets[last_bi]
            off, n = a.range_of(b.first_param, b.last_param)
            wide = min(torch.cuda.get_device_properties(self.device).multi_processor_count, 128)
            if early_sgd and n <= self.symm.SMALL_MAX_FLOATS:
                # small tail bucket (the first layers): one-shot reduction, mean and SGD update from
                # registers -- no remote stores, no local grid barrier
                self.symm.allreduce_multi([(off, n)], blocks=wide, slot_base=1024, use_mc=self.use_mc,
                                          sgd=self._sgd_args())
            else:
                self.symm.allreduce_avg(off, n, blocks=wide, slot_base=1024, use_mc=self.use_mc,
                                        sgd=self._sgd_args() if self.fuse_sgd else None,
                                        sgd_range=(off, n) if early_sgd else (0, a.numel))
            eng.mark("last bucket reduced + sgd")
            if not self.fuse_sgd:
                eng.sgd(self.lr, self.momentum, self.weight_decay)
        else:
            eng.backward(B, on_loss_ready=self._publish_loss, side_stream=self.wgrad_stream, reduce_loss=True)
            if self.granularity == "param" and self.mode == "allreduce":
                # the reference's 34 per-parameter all-reduces (part2b/main.py:101-103) as ONE launch over a
                # table of 34 ranges: the 32 small tensors one-shot, the big filters two-shot
                self.symm.allreduce_multi([a.range_of(i, i) for i in range(len(a.params))], blocks=self.comm_blocks,
                                          use_mc=self.use_mc, sgd=self._sgd_args() if self.fuse_sgd else None)
            elif self.granularity in ("param", "param_launches"):
                # one launch per parameter tensor, like the reference's python loop
                for i in range(len(a.params)):
                    off, n = a.range_of(i, i)
                    self._sync_range(off, n, slot_base=0)
            else:
                self._sync_range(0, a.numel, slot_base=0)
            if not self.fuse_sgd:
                eng.sgd(self.lr, self.momentum, self.weight_decay)

    def _copy_into(self, slot: int, data: torch.Tensor, target: torch.Tensor, aug):
        """Enqueues the copies of one batch into staging set ``slot`` on the CURRENT stream."""
        st = self._stage[slot]
        N = data.shape[0]
        if N > self.batch_size or N < 1:
            raise ValueError(f"Trainer was built for batches of up to {self.batch_size}, got {N}")
        if data.dtype == torch.uint8:
            st["u8"][:N].copy_(data, non_blocking=True)
            if aug is not None:
                st["aug"][:N].copy_(aug, non_blocking=True)
            kind = "u8"
        else:
            if st["f32"] is None:
                hw = self.engine.image_hw
                for t in self._stage:
                    t["f32"] = torch.zeros(self.batch_size, 3, hw, hw, dtype=torch.float32, device=self.device)
            st["f32"][:N].copy_(data, non_blocking=True)
            kind = "f32"
        st["labels"][:N].copy_(target, non_blocking=True)
        return kind

    def _aug_kind(self, kind: str, aug) -> int:
        """0: no augmentation, 1: the caller's draws, 2: draws made on the device inside the step."""
        if aug is not None:
            return 1
        return 2 if (self.device_augment and kind == "u8") else 0

    def prefetch(self, data: torch.Tensor, target: torch.Tensor, aug: torch.Tensor | None = None) -> None:
        """Starts the host->device copy of the NEXT step's batch on the copy stream, into the staging
        set the running step does not use; the next ``step()`` (called with the same tensors or with
        no arguments) only waits for that copy.  Call it right after ``step()``: the copy then runs
        while the GPU computes.  Pinned host tensors make the copy asynchronous."""
        if self.device.type != "cuda":
            self._pending = (None, data, target, aug)
            return
        slot = self._slot
        st = self._stage[slot]
        cs = self.copy_stream
        if st["free"] is not None:
            cs.wait_event(st["free"])      # the step that last read this set has finished
        with torch.cuda.stream(cs):
            kind = self._copy_into(slot, data, target, aug)
            ready = torch.cuda.Event()
            ready.record(cs)
        st["ready"] = ready
        self._pending = (slot, kind, self._aug_kind(kind, aug), data.shape[0], data)

    def set_hyperparameters(self, lr: float | None = None, momentum: float | None = None,
                            weight_decay: float | None = None) -> None:
        """Changes the SGD hyper-parameters of the following steps.  On CUDA they are kernel
        arguments baked into the captured step, so a change drops the captured graphs (the next
        step runs eagerly and is captured again: about a millisecond, once per change -- fine for
        an epoch-wise schedule; for a per-step schedule build the Trainer with use_graph=False)."""
        new = (self.lr if lr is None else float(lr), self.momentum if momentum is None else float(momentum),
               self.weight_decay if weight_decay is None else float(weight_decay))
        if new != (self.lr, self.momentum, self.weight_decay):
            self.lr, self.momentum, self.weight_decay = new
            self._graphs.clear()
            opt = getattr(self, "optimizer", None)
            if self.device.type != "cuda" and opt is not None and opt is not self._hyper_source:
                for g in opt.param_groups:
                    g["lr"], g["momentum"], g["weight_decay"] = new

    def _refresh_hyper(self) -> None:
        src = self._hyper_source
        if src is not None:
            g = src.param_groups[0]
            self.set_hyperparameters(g["lr"], g.get("momentum", 0.0), g.get("weight_decay", 0.0))

    def step(self, data: torch.Tensor | None = None, target: torch.Tensor | None = None,
             aug: torch.Tensor | None = None) -> torch.Tensor:
        """One training iteration.  ``data`` is either raw uint8 NHWC pixels (plus optional
        per-image crop/flip parameters ``aug``) or a normalised fp32 NCHW batch, on the host
        (ideally pinned) or on the device; without arguments (or with the very tensors given to
        ``prefetch``) the prefetched batch is used.  Returns the mean loss as a 0-d device tensor
        that is overwritten by the next call; ``last_loss()`` gives the same number on the host
        without a copy or a stream synchronisation."""
        self._refresh_hyper()
        if self.device.type != "cuda":
            if data is None:
                _, data, target, aug = self._pending
            self._pending = None
            return self._step_cpu(data, target, aug)
        main = torch.cuda.current_stream()
        pend, self._pending = self._pending, None
        if pend is not None and (data is None or data is pend[4]):
            slot, kind, with_aug, N, _ = pend
            main.wait_event(self._stage[slot]["ready"])
        else:
            if data is None:
                raise ValueError("step() without arguments needs a preceding prefetch()")
            slot = self._slot
            kind = self._copy_into(slot, data, target, aug)
            with_aug, N = self._aug_kind(kind, aug), data.shape[0]
        key = f"{kind}:{int(with_aug)}:{slot}"
        first = self.iteration == 0 and self.mode == "ddp"
        if not self.use_graph or N != self.batch_size:
            # ragged last batch of an epoch (or graphs disabled): plain launches
            self._enqueue_iteration(kind, with_aug, N, slot, one_bucket=first)
        else:
            g = self._graphs.get(key)
            if g is None:
                g = self._capture(key, kind, with_aug, slot)
            else:
                g.replay()
        free = torch.cuda.Event()
        free.record(main)
        self._stage[slot]["free"] = free
        self._slot = slot ^ 1
        self._steps_published += 1
        self.iteration += 1
        return self.engine.loss

    def last_loss(self) -> float:
        """Mean loss of the most recent ``step()`` on the host.  The step itself stores it into pinned
        host memory right after the classifier, so this returns as soon as that store has landed --
        typically while the step's backward pass is still running -- instead of draining the
        stream like ``loss.item()`` (/root/reference/part1/main.py:80 reads the loss every iteration)."""
        if self.device.type != "cuda":
            return float(self._cpu_last_loss)
        import struct
        import time

        want = self._steps_published & 0xFFFFFFFF
        view = self._loss_host.numpy()
        t0 = None
        while True:
            word = int(view[0]) & 0xFFFFFFFFFFFFFFFF
            if (word >> 32) == want:
                return struct.unpack("<f", struct.pack("<I", word & 0xFFFFFFFF))[0]
            if t0 is None:
                t0 = time.perf_counter()
            elif time.perf_counter() - t0 > 120.0:
                raise RuntimeError("last_loss(): the step did not publish its loss within 120 s")

    def _capture(self, key, kind, with_aug, slot):
        # The first call runs eagerly (it also performs one-time kernel attribute set-up) and
        # IS the iteration; the graph is captured afterwards from a snapshot-free sequence: the
        # capture itself executes nothing.
        self._enqueue_iteration(kind, with_aug, None, slot, one_bucket=self.iteration == 0 and self.mode == "ddp")
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self._capture_stream):
            self._enqueue_iteration(kind, with_aug, None, slot)
        self._graphs[key] = g
        return g

    def evaluate(self, data: torch.Tensor, target: torch.Tensor):
        """Forward + loss + top-1 count for one batch (any size up to batch_size); BatchNorm uses
        batch statistics here too, as the reference's model does in eval mode
        (track_running_stats=False, model.py:24).  Returns (mean loss, number correct) tensors."""
        if self.device.type != "cuda":
            return self._evaluate_cpu(data, target)
        eng = self.engine
        N = data.shape[0]
        if N > self.batch_size:
            raise ValueError("evaluation batch larger than the engine was sized for")
        # pack=True: the first layer's padded filter pack is refreshed by the input kernel; after a
        # training step it still holds the filters from BEFORE that step's update
        kind = self._copy_into(2, data, target, None)      # evaluation has its own staging set
        st = self._stage[2]
        eng.labels = st["labels"]
        if kind == "u8":
            eng.load_input_u8(st["u8"][:N], None, pack=True)
        else:
            eng.load_input_f32(st["f32"][:N], pack=True)
        eng.forward(N, with_loss=True, want_dlogits=False)
        return eng.loss.clone(), eng.correct.clone()

    def kernel_launches_per_step(self) -> int:
        """Number of ddpb200 kernels one ``step`` enqueues (for bench.py's gpu_launches)."""
        if self.device.type != "cuda":
            return 0
        n = 1 + self.engine.launches_per_train_step()  # input prep (+ first-layer filter pack)
        n += 1                                          # the one-thread kernel that publishes the loss to the host
        if not self.distributed:
            n += len(self.groups) if (self.overlap_sgd and self.update_stream is not None) else 1
        elif self.mode == "ddp":
            early = self.overlap_sgd and self.fuse_sgd
            n += len(self.buckets) + (len(self.buckets) - 1 if early else 0) + (0 if self.fuse_sgd else 1)
        else:
            per_param = self.granularity == "param_launches" or (self.granularity == "param" and self.mode != "allreduce")
            n += (len(self.engine.arena.params) if per_param else 1) + (0 if self.fuse_sgd else 1)
        return n

    def close(self):
        if getattr(self, "symm", None) is not None:
            self._graphs.clear()
            self.symm.close()
            self.symm = None

    # =====================================================================================
    # CPU path (stock modules + torch.distributed): plumbing tests and the oracle
    # =====================================================================================
    def _init_cpu(self):
        # step the caller's optimizer when there is one (its state_dict then holds the momentum)
        self.optimizer = self._hyper_source if self._hyper_source is not None else torch.optim.SGD(
            self.model.parameters(), lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay)
        self.reducer = None
        if self.distributed and self.mode == "ddp":
            for p in self.model.parameters():  # constructor-time broadcast, as DDP does
                dist.broadcast(p.data, src=0)
            self.reducer = S.HostBucketReducer(self.model, self.world)
            self.buckets = self.reducer.buckets

    @staticmethod
    def _to_float_batch(data: torch.Tensor, aug: torch.Tensor | None = None) -> torch.Tensor:
        if data.dtype == torch.uint8:
            from .ops.reference import augment_normalize_u8

            return augment_normalize_u8(data, aug)   # crop / flip / normalise, as the device kernel does
        return data

    def _step_cpu(self, data, target, aug=None):
        if aug is None and self.device_augment and data.dtype == torch.uint8:
            # the draws the device kernel makes for this seed and step counter (Philox restated in ops/reference.py)
            from .ops.reference import draw_augmentation

            aug = draw_augmentation(data.shape[0], self.augment_pad, self.augment_seed, self._aug_step_cpu)
            self._aug_step_cpu += 1
        data = self._to_float_batch(data, aug)
        self.optimizer.zero_grad()
        loss = F.cross_entropy(self.model(data), target)
        loss.backward()
        if self.distributed:
            if self.mode == "gather_scatter":
                S.sync_gradients_gather_scatter(self.model, self.rank, self.world)
            elif self.mode == "allreduce":
                S.sync_gradients_allreduce(self.model, self.world)
            else:
                self.reducer.finalize()
        self.optimizer.step()
        self.iteration += 1
        self._cpu_last_loss = float(loss.detach())
        return loss.detach()

    def _evaluate_cpu(self, data, target):
        data = self._to_float_batch(data)
        with torch.no_grad():
            if self.reducer is not None:
                self.reducer.enabled = False
            out = self.model(data)
            if self.reducer is not None:
                self.reducer.enabled = True
            return F.cross_entropy(out, target), out.argmax(1).eq(target).sum()
