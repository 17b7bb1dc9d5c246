"""Execution engine: runs a ddpb200 VGG on one B200 with the hand-written kernels.

The reference executes ``model(data)``, ``criterion``, ``loss.backward()`` and
``optimizer.step()`` as ~100 separate ATen calls per iteration
(/root/reference/part1/main.py:71-77).  Here the whole iteration is a fixed launch
sequence over preallocated NHWC bf16 buffers:

    input prep -> [conv fprop(+stats) -> BN/ReLU/pool] x L -> FC+CE
               -> FC bwd -> [BN bwd -> wgrad -> dgrad] x L -> gradient sync + SGD -> repack

Design points:
  * parameters, gradients and momentum each live in ONE flat fp32 arena; ``param.data`` and
    ``param.grad`` are views into them, so the torch-facing API and the 34-key state_dict
    are unchanged while sync/SGD kernels see contiguous ranges ("bucket" = arena slice);
  * conv filters are stored channels_last, i.e. physically [C_out][3][3][C_in], which is
    the GEMM operand order -- the repack to bf16 is an elementwise cast;
  * gradients are written by the kernels (no zero_grad pass needed);
  * the sequence is captured once into a CUDA graph and replayed.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn

from .ops import kernels as K

ALIGN = 32  # arena alignment of every parameter, in floats (128 bytes)


# ---------------------------------------------------------------------------------------
# flat arenas
# ---------------------------------------------------------------------------------------
class ParamArena:
    """Re-homes a module's parameters (and their grads) into flat fp32 arenas."""

    def __init__(self, params: list[nn.Parameter], device: torch.device, grad_storage: torch.Tensor | None = None):
        self.params = params
        self.offsets: list[int] = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
        self.numel = off
        self.master = torch.zeros(off, dtype=torch.float32, device=device)
        self.momentum = torch.zeros(off, dtype=torch.float32, device=device)
        if grad_storage is None:
            grad_storage = torch.zeros(off, dtype=torch.float32, device=device)
        assert grad_storage.numel() >= off and grad_storage.dtype == torch.float32
        self.grad = grad_storage[:off]
        for p, o in zip(params, self.offsets):
            view = self._view(self.master, p, o)
            view.copy_(p.data)
            p.data = view
            p.grad = self._view(self.grad, p, o)

    @staticmethod
    def _view(flat: torch.Tensor, p: torch.Tensor, off: int) -> torch.Tensor:
        n = p.numel()
        if p.dim() == 4:  # conv filter (K,C,R,S): keep it physically [K][R][S][C]
            k, c, r, s = p.shape
            return flat[off:off + n].view(k, r, s, c).permute(0, 3, 1, 2)
        return flat[off:off + n].view(p.shape)

    def flat_master(self, idx: int) -> torch.Tensor:
        o = self.offsets[idx]
        return self.master[o:o + self.params[idx].numel()]

    def flat_grad(self, idx: int) -> torch.Tensor:
        o = self.offsets[idx]
        return self.grad[o:o + self.params[idx].numel()]

    def range_of(self, first_idx: int, last_idx: int) -> tuple[int, int]:
        """(offset, count) in floats covering parameters first_idx..last_idx inclusive, padding included."""
        lo = self.offsets[first_idx]
        hi = self.offsets[last_idx + 1] if last_idx + 1 < len(self.offsets) else self.numel
        return lo, hi - lo


@dataclass
class _Layer:
    H: int
    W: int
    c_in: int       # real input channels
    c_in_eff: int   # GEMM K per tap (64 for the first layer's gathered operand)
    c_out: int
    taps: int
    pool: bool
    p_w: int        # indices into the parameter list
    p_b: int
    p_g: int
    p_beta: int


class VGGEngine:
    """Owns buffers + launch sequences for one model instance on one device."""

    def __init__(self, model, max_batch: int, grad_storage: torch.Tensor | None = None, image_hw: int = 32,
                 share_from: "VGGEngine | None" = None):
        """``share_from``: an existing engine of the SAME model whose parameter / gradient / momentum
        arenas, bf16 shadow and first-layer filter pack this engine adopts instead of re-homing the
        parameters again.  Used when a larger batch than the first engine was sized for shows up
        (get_engine): only activation and scratch buffers are new, so a Trainer, a bound optimizer
        or captured CUDA graphs that still use the old engine keep seeing the same parameters and
        momentum."""
        self.model = model
        params = list(model.parameters())
        self.device = params[0].device
        assert self.device.type == "cuda", "VGGEngine needs the model on a CUDA device"
        self.max_batch = max_batch
        if share_from is not None:
            assert share_from.model is model and grad_storage is None
            self.arena = share_from.arena
        else:
            self.arena = ParamArena(params, self.device, grad_storage)
        index = {id(p): i for i, p in enumerate(params)}
        dev = self.device

        self.layers: list[_Layer] = []
        hw = image_hw
        for bi, blk in enumerate(model.conv_blocks()):
            first = bi == 0
            self.layers.append(_Layer(hw, hw, blk.c_in, 64 if first else blk.c_in, blk.c_out, 1 if first else 9,
                                      blk.pool, index[id(blk.conv.weight)], index[id(blk.conv.bias)],
                                      index[id(blk.bn.weight)], index[id(blk.bn.bias)]))
            if first:
                assert blk.c_in == 3 and blk.c_out % 64 == 0
            if blk.pool:
                hw //= 2
        assert hw == 1, "feature map must be 1x1 before the classifier"
        self.p_fcw, self.p_fcb = index[id(model.fc1.weight)], index[id(model.fc1.bias)]
        self.feat = model.fc1.in_features
        self.classes = model.fc1.out_features

        B = max_batch
        bf = dict(dtype=torch.bfloat16, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        # --- activations and their gradients
        self.operand0 = torch.empty(B * image_hw * image_hw, 64, **bf)
        self.conv_out, self.act, self.dconv, self.dact = [], [], [], []
        for L in self.layers:
            P = B * L.H * L.W
            Po = P // 4 if L.pool else P
            self.conv_out.append(torch.empty(P, L.c_out, **bf))
            self.act.append(torch.empty(Po, L.c_out, **bf))
            self.dconv.append(torch.empty(P, L.c_out, **bf))
            self.dact.append(torch.empty(Po, L.c_out, **bf))
        # --- per-channel statistics: one arena so a single memset clears all of it
        csum = sum(2 * L.c_out + 32 for L in self.layers)      # every slot carries 32 extra words
        self.stat_arena = torch.zeros(2 * csum, **f32)
        self.stats, self.bsums, self.saved = [], [], []
        o = 0
        for L in self.layers:
            self.stats.append(self.stat_arena[o:o + 2 * L.c_out + 32])   # sums + grid-barrier word
            o += 2 * L.c_out + 32
        for L in self.layers:
            self.bsums.append(self.stat_arena[o:o + 2 * L.c_out + 32])   # sums + grid-barrier word
            o += 2 * L.c_out + 32
            self.saved.append(torch.empty(2 * L.c_out, **f32))
        # --- bf16 filter operands.  The arena keeps filters as [C_out][3][3][C_in], which is the
        # K-major GEMM operand of fprop AND (read MN-major, taps reversed) of dgrad, so one bf16
        # "shadow" of the whole arena -- refreshed by the SGD kernels -- serves both.  Only the
        # first layer needs a real repack (27 values per filter, zero padded to a 64-wide row).
        if share_from is not None:
            self.shadow, self.w_f, self.pack_plan = share_from.shadow, share_from.w_f, share_from.pack_plan
        else:
            self.shadow = torch.zeros(self.arena.numel, **bf)
            self.w_f = []
            for li, L in enumerate(self.layers):
                if li == 0:
                    self.w_f.append(torch.zeros(L.c_out, 64, **bf))
                else:
                    o = self.arena.offsets[L.p_w]
                    self.w_f.append(self.shadow[o:o + L.c_out * 9 * L.c_in].view(L.c_out, 9 * L.c_in))
            L0 = self.layers[0]
            self.pack_plan = K.PackPlan([dict(master=self.arena.flat_master(L0.p_w), fprop=self.w_f[0], dgrad=None,
                                              K=L0.c_out, C=L0.c_in, taps=9, fprop_ld=64, flat=True)])
        # --- classifier / loss
        self.labels = torch.zeros(B, dtype=torch.int64, device=dev)
        self.logits = torch.empty(B, self.classes, **f32)
        self.dlogits = torch.empty(B, self.classes, **f32)
        self.row_loss = torch.empty(B, **f32)
        self.row_correct = torch.empty(B, dtype=torch.int32, device=dev)
        self.loss = torch.zeros((), **f32)
        self.correct = torch.zeros((), dtype=torch.int32, device=dev)
        # --- split reduction scratch, shared by all layers (launches are stream ordered).  The
        # planners are cost based, so a ragged last batch (any N <= max_batch) can want MORE
        # scratch than the full batch: size for the maximum over every N.  (The native side also
        # checks the capacity it is given and never writes past it.)
        need_w = need_f = 1
        for li, L in enumerate(self.layers):
            cols = 27 if li == 0 else 9 * L.c_in
            for n in range(1, B + 1):
                need_w = max(need_w, K.conv_wgrad_scratch_elems(n, L.H, L.W, L.c_in_eff, L.c_out, L.taps, cols, cols))
                need_f = max(need_f, K.conv_fprop_scratch_elems(n, L.H, L.W, L.c_in_eff, L.c_out, L.taps))
                if li > 0:
                    need_f = max(need_f, K.conv_fprop_scratch_elems(n, L.H, L.W, L.c_out, L.c_in, 9))
        self.wgrad_scratch = torch.empty(need_w, **f32)
        self.fprop_scratch = torch.empty(need_f, **f32)
        # sync words of the in-kernel split-K sums (zeroed once, left zero by every launch; fprop and dgrad
        # run one after the other on the main stream, so one buffer serves both)
        self.split_sync = torch.zeros(K.conv_split_sync_words(), dtype=torch.int32, device=dev)
        # --- partial sums of the BatchNorm backward (one buffer, the layers run one after the other)
        need_p = 1
        for L in self.layers:
            for n in {B, max(1, B // 2), max(1, B // 4), 1} | set(range(1, B + 1, max(1, B // 16))):
                need_p = max(need_p, K.bn_bwd_partials_elems(n, L.H, L.W, L.c_out, L.pool))
        self.bn_partials = torch.empty(need_p + 8192, **f32)
        self.image_hw = image_hw
        self._tl = None
        self._tl_labels = []
        if share_from is None:
            self.repack(full=True)
            _register_engine(self)

    # -- diagnostics ----------------------------------------------------------------------------
    def enable_timeline(self, slots: int = 256):
        """Every mark() from now on stamps %globaltimer into a device buffer (bench/step_timeline.py).
        Must be called before the step is captured in a graph."""
        self._tl = torch.zeros(slots, dtype=torch.int64, device=self.device)
        self._tl_labels = []

    def mark(self, label: str, reset: bool = False):
        if self._tl is None:
            return
        if reset:
            self._tl_labels = []
        i = len(self._tl_labels)
        if i < self._tl.numel():
            K.stamp(self._tl[i:i + 1])
            self._tl_labels.append(label)

    def timeline(self):
        """[(label, microseconds since the first mark)] of the last executed step."""
        t = self._tl[:len(self._tl_labels)].cpu().tolist()
        return [(lab, (v - t[0]) / 1e3) for lab, v in zip(self._tl_labels, t)]

    # -- pieces ---------------------------------------------------------------------------------
    def repack(self, stream=None, full: bool = False):
        """Refreshes the bf16 filter operands from the fp32 master arena.  The training step only
        needs the first layer's padded pack (the SGD kernels keep the shadow current); ``full``
        also recasts the whole shadow -- used after parameters were changed from outside
        (initialisation, load_state_dict, a foreign optimizer)."""
        if full:
            self.shadow.copy_(self.arena.master)
        self.pack_plan.run(stream)

    def _pack_kw(self, pack: bool) -> dict:
        if not pack:
            return {}
        return dict(w0_master=self.arena.flat_master(self.layers[0].p_w), w0_packed=self.w_f[0])

    def load_input_f32(self, x_nchw: torch.Tensor, pack: bool = False):
        """``pack``: the same launch refreshes the first layer's padded filter pack (= repack())."""
        K.first_layer_operand_f32(x_nchw.contiguous(), out=self.operand0, **self._pack_kw(pack))

    def load_input_u8(self, x_nhwc: torch.Tensor, aug: torch.Tensor | None, pack: bool = False):
        K.first_layer_operand_u8(x_nhwc, aug, out=self.operand0, **self._pack_kw(pack))

    def forward(self, N: int, with_loss: bool, want_dlogits: bool = True, defer_loss: bool = False):
        """operand0[:N*H*W] -> logits[:N] (and loss / dlogits / correct when with_loss).
        ``defer_loss``: leave the reduction of the per-row losses to backward(reduce_loss=True)."""
        a = self.arena
        self.stat_arena.zero_()
        self.mark("fwd begin")
        x = self.operand0
        for li, L in enumerate(self.layers):
            # conv (+bias, +statistics) and, inside the same persistent kernel after a grid-wide
            # barrier, BatchNorm + ReLU (+pool); split-K layers fall back to the separate kernel
            K.conv_bn_relu_pool_fwd(x, self.w_f[li], a.flat_master(L.p_b), a.flat_master(L.p_g),
                                    a.flat_master(L.p_beta), N, L.H, L.W, L.c_in_eff, L.c_out, L.taps, L.pool,
                                    conv_out=self.conv_out[li], act=self.act[li], stats=self.stats[li],
                                    saved=self.saved[li], scratch=self.fprop_scratch, split_sync=self.split_sync)
            self.mark(f"L{li} conv+bn fwd")
            x = self.act[li]
        feat = x[:N]
        if with_loss:
            K.fc_ce_fwd(feat, a.flat_master(self.p_fcw).view(self.classes, self.feat), a.flat_master(self.p_fcb),
                        self.labels[:N], logits=self.logits, row_loss=self.row_loss,
                        dlogits=self.dlogits if want_dlogits else None, row_correct=self.row_correct,
                        loss=None if defer_loss else self.loss, correct=None if defer_loss else self.correct)
        else:
            # logits only: labels are not read meaningfully, loss outputs are discarded
            K.fc_ce_fwd(feat, a.flat_master(self.p_fcw).view(self.classes, self.feat), a.flat_master(self.p_fcb),
                        self.labels[:N], logits=self.logits, row_loss=self.row_loss)
        self.mark("fc fwd")

    def backward(self, N: int, after_layer=None, side_stream=None, reduce_loss: bool = False, on_loss_ready=None):
        """dlogits[:N] -> every gradient in the arena.

        ``reduce_loss``: the classifier's backward kernel also writes ``loss`` / ``correct`` from
        the per-row values forward(defer_loss=True) left behind.

        ``after_layer(li, events)`` is called once layer li's gradients (and everything after it,
        classifier included) are enqueued -- the hook the bucketed reducer uses to start a bucket
        while the rest of backward still runs; ``events`` must have completed before the bucket's
        gradients are final.

        ``side_stream``: wgrad(l) and dgrad(l) only share their input, so with a side stream the
        weight gradient is a second branch of the captured graph next to the data gradient and the
        next BatchNorm backward.  How much the branches really overlap depends on the kernels leaving
        each other room on the SMs (measured in DESIGN.md par. 7)."""
        a = self.arena
        main = torch.cuda.current_stream()
        side_events = []
        last = len(self.layers) - 1
        feat = self.act[last][:N]
        K.fc_bwd(feat, a.flat_master(self.p_fcw).view(self.classes, self.feat), self.dlogits[:N],
                 dw=a.flat_grad(self.p_fcw), db=a.flat_grad(self.p_fcb), dx=self.dact[last],
                 **(dict(row_loss=self.row_loss[:N], row_correct=self.row_correct[:N], loss=self.loss,
                         correct=self.correct) if reduce_loss else {}))
        self.mark("fc bwd")
        if on_loss_ready is not None:
            on_loss_ready()             # ``loss`` / ``correct`` are final from here on (reduce_loss)
        if after_layer is not None:
            after_layer(last + 1, [])   # the classifier is parameter group len(layers)
        for li in range(last, -1, -1):
            L = self.layers[li]
            x_in = self.operand0 if li == 0 else self.act[li - 1]
            K.bn_relu_pool_bwd(self.conv_out[li], self.dact[li], self.saved[li], a.flat_master(L.p_g),
                               a.flat_master(L.p_beta), N, L.H, L.W, L.c_out, L.pool, sums=self.bsums[li],
                               dx=self.dconv[li], dgamma=a.flat_grad(L.p_g), dbeta=a.flat_grad(L.p_beta),
                               dbias=a.flat_grad(L.p_b), partials=self.bn_partials)
            self.mark(f"L{li} bn bwd")
            cols = 9 * L.c_in
            if side_stream is not None:
                fork = torch.cuda.Event()
                fork.record(main)
                side_stream.wait_event(fork)
                with torch.cuda.stream(side_stream):
                    K.conv_wgrad(self.dconv[li], x_in, N, L.H, L.W, L.c_in_eff, L.c_out, L.taps,
                                 out=a.flat_grad(L.p_w), cols_valid=cols, out_ld=cols, scratch=self.wgrad_scratch)
                    self.mark(f"L{li} wgrad (side)")
                    done = torch.cuda.Event()
                    done.record(side_stream)
                side_events.append(done)
            else:
                K.conv_wgrad(self.dconv[li], x_in, N, L.H, L.W, L.c_in_eff, L.c_out, L.taps, out=a.flat_grad(L.p_w),
                             cols_valid=cols, out_ld=cols, scratch=self.wgrad_scratch)
                self.mark(f"L{li} wgrad")
            if li > 0:
                K.conv_dgrad(self.dconv[li], self.w_f[li], N, L.H, L.W, L.c_in, L.c_out, out=self.dact[li - 1],
                             scratch=self.fprop_scratch, split_sync=None)   # two-kernel split-K is faster here (no tail to fuse)
                self.mark(f"L{li} dgrad")
            if after_layer is not None:
                after_layer(li, side_events[-1:] if side_stream is not None else [])
        if side_stream is not None and side_events:
            main.wait_event(side_events[-1])   # the side stream is in order: the last event covers all
            self.mark("bwd joined")

    def sgd(self, lr: float, momentum: float, weight_decay: float, grad_scale: float = 1.0):
        a = self.arena
        K.sgd_step(a.master, a.grad, a.momentum, lr, momentum, weight_decay, grad_scale, shadow=self.shadow)
        self.mark("sgd")

    def launches_per_train_step(self, splits_reduce: bool = True) -> int:
        """How many of OUR kernels one fwd+bwd enqueues (excluding sync/SGD/repack/input prep)."""
        n = 0
        B = self.max_batch
        for li, L in enumerate(self.layers):
            n += 1                  # fprop (BatchNorm forward fused in, also when split-K is summed in the kernel)
            _, _, splits, fused = K.conv_plan_ex(B, L.H, L.W, L.c_in_eff, L.c_out, L.taps, with_tail=True, with_sync=True)
            if splits > 1 and not fused:
                n += 2              # split-K finalize + separate BatchNorm forward
            if li > 0 and K.conv_plan_ex(B, L.H, L.W, L.c_out, L.c_in, 9)[2] > 1:
                n += 1              # dgrad is called without sync words: its split-K is summed by the finalize kernel
            out_pix = B * L.H * L.W // (4 if L.pool else 1)
            owner = L.c_out // 8 >= 32 and out_pix * (5 if L.pool else 2) * 16 <= 48 * 1024
            n += 1 if owner else 2  # BatchNorm backward: channel-owner kernel, or reduce + apply
            n += 1                  # wgrad
            if K.conv_wgrad_scratch_elems(B, L.H, L.W, L.c_in_eff, L.c_out, L.taps):
                n += 1              # split-K reduce
            if li > 0:
                n += 1              # dgrad
        return n + 1 + 1            # fc fwd, fc bwd (+ loss reduction)


# ---------------------------------------------------------------------------------------
# nn.Module integration: model(x) on CUDA, differentiable through one autograd node
# ---------------------------------------------------------------------------------------
class _EngineFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, model, *params):
        eng: VGGEngine = model._engine
        N = x.shape[0]
        eng.load_input_f32(x.float())
        eng.forward(N, with_loss=False)
        ctx.eng, ctx.N = eng, N
        return eng.logits[:N].clone()

    @staticmethod
    def backward(ctx, dlogits):
        eng, N = ctx.eng, ctx.N
        eng.dlogits[:N].copy_(dlogits)
        eng.backward(N)   # (single stream on the eager autograd path)
        # gradients were written straight into param.grad (arena views); nothing to hand to autograd
        return (None, None) + (None,) * len(eng.arena.params)


_ENGINES: "weakref.WeakValueDictionary[int, VGGEngine]" = None  # storage data_ptr -> engine


def _register_engine(eng: "VGGEngine") -> None:
    global _ENGINES
    import weakref

    if _ENGINES is None:
        _ENGINES = weakref.WeakValueDictionary()
    _ENGINES[eng.arena.master.data_ptr()] = eng


def engine_of_parameter(p: torch.Tensor):
    """The engine whose master arena holds ``p`` (None if p is an ordinary tensor)."""
    if _ENGINES is None:
        return None
    for base, eng in list(_ENGINES.items()):
        n = eng.arena.master.numel() * 4
        if base <= p.data_ptr() < base + n:
            return eng
    return None


def get_engine(model, batch: int, grad_storage: torch.Tensor | None = None) -> VGGEngine:
    """The model's engine, sized for at least ``batch`` images.  A larger batch than the current
    engine holds gets a new engine that SHARES the parameter / gradient / momentum arenas of the
    old one (only activations are reallocated): a Trainer or optimizer bound to the old engine
    keeps training the very tensors ``model.parameters()`` and ``state_dict()`` show."""
    eng = model._engine
    if eng is None:
        eng = VGGEngine(model, batch, grad_storage)
        model._engine = eng
    elif eng.max_batch < batch:
        eng = VGGEngine(model, batch, image_hw=eng.image_hw, share_from=eng)
        model._engine = eng
    return eng


def engine_forward(model, x: torch.Tensor) -> torch.Tensor:
    if x.dim() != 4 or x.shape[1] != 3:
        raise ValueError(f"expected a [N,3,H,W] batch, got {tuple(x.shape)}")
    eng = get_engine(model, x.shape[0])
    for p, o in zip(eng.arena.params, eng.arena.offsets):
        if p.grad is None:  # a stock optimizer's zero_grad(set_to_none=True) detached it
            p.grad = ParamArena._view(eng.arena.grad, p, o)
    if torch.is_grad_enabled() and any(p.requires_grad for p in eng.arena.params):
        eng.repack(full=True)  # parameters may have been updated by any optimizer since the last call
        return _EngineFn.apply(x, model, *eng.arena.params)
    eng.repack(full=True)
    eng.load_input_f32(x.float())
    eng.forward(x.shape[0], with_loss=False)
    return eng.logits[:x.shape[0]].clone()
