"""Tensor-level wrappers over the sm_100a kernels (one function per extern "C" entry point).

Shapes follow the kernels' native layout: activations are NHWC bf16 stored as
``[N*H*W, C]`` matrices; conv filters are ``[C_out, taps*C_in]`` bf16.  Each wrapper names
the reference op it replaces.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _native as nat

BN_EPS = 1e-5  # nn.BatchNorm2d default, /root/reference/part1/model.py:24


def _s(stream):
    return nat.stream_ptr(stream)


def _numel(t) -> int:
    """Capacity (in elements) the native side may use of an optional fp32 scratch tensor."""
    if t is None:
        return 0
    assert t.dtype == torch.float32 and t.is_contiguous()
    return t.numel()


def conv_force(pair: int = -1, block_n: int = -1, splits: int = -1) -> None:
    """Tuning/testing hook: pin the conv launch planner's choice (-1 = automatic)."""
    nat.load().ddpb_conv_force(pair, block_n, splits)


def conv_tail_from_tmem(on: bool | int) -> None:
    """Tuning hook: fused BatchNorm tail of the persistent conv kernel from TMEM (1, default: with one tile per
    CTA; 2: wherever the shape allows) or from the stored tiles in L2 (0)."""
    nat.load().ddpb_conv_tail_from_tmem(int(on))


def conv_plan(N, H, W, C_in, C_out, taps=9, block_n=0) -> tuple[int, int, int]:
    """(cta_pair, block_n, splits) the planner picks for this fprop-shaped problem."""
    out = (C.c_int * 3)()
    nat.load().ddpb_conv_plan(N, H, W, C_in, C_out, taps, block_n, out)
    return tuple(out)


_SPLIT_SYNC: dict = {}


def conv_split_sync(device) -> torch.Tensor:
    """The zeroed sync words the conv entry points use to add up split-K partial sums inside the kernel: one
    cached buffer per device (the kernels leave it zero).  Launches that may run concurrently on several
    streams need their own (``torch.zeros(conv_split_sync_words(), dtype=torch.int32)``)."""
    key = torch.device(device)
    if key.index is None:
        key = torch.device("cuda", torch.cuda.current_device())
    buf = _SPLIT_SYNC.get(key)
    if buf is None:
        buf = _SPLIT_SYNC[key] = torch.zeros(conv_split_sync_words(), dtype=torch.int32, device=key)
    return buf


def conv_split_sync_words() -> int:
    return int(nat.load().ddpb_conv_split_sync_words())


def conv_fuse_split(on: bool) -> None:
    """Tuning hook: False = split-K partial sums always go through the second (finalize) kernel."""
    nat.load().ddpb_conv_fuse_split(int(on))


def _sync_arg(split_sync, device):
    if split_sync is True:
        return conv_split_sync(device)
    if split_sync is False or split_sync is None:
        return None
    assert split_sync.dtype == torch.int32 and split_sync.numel() >= conv_split_sync_words() and split_sync.is_cuda
    return split_sync


def conv_plan_ex(N, H, W, C_in, C_out, taps=9, block_n=0, *, with_tail=False, with_sync=False) -> tuple[int, int, int, int]:
    """(kernel, block_n, splits, fused split-K) for a given call situation: ``with_tail`` = the fused
    conv + BatchNorm entry point, ``with_sync`` = sync words passed.  Host only (works without a GPU)."""
    out = (C.c_int * 4)()
    nat.load().ddpb_conv_plan_ex(N, H, W, C_in, C_out, taps, block_n, int(with_tail), int(with_sync), out)
    return tuple(out)


def conv_fprop_scratch_elems(N, H, W, C_in, C_out, taps=9, block_n=0) -> int:
    """fp32 elements of split-K scratch conv_fprop wants for this shape (0 = no split)."""
    return int(nat.load().ddpb_conv_fprop_scratch_elems(N, H, W, C_in, C_out, taps, block_n))


def conv_fprop(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None, N: int, H: int, W: int, C_in: int,
               C_out: int, taps: int = 9, *, out: torch.Tensor | None = None, stats: torch.Tensor | None = None,
               block_n: int = 0, scratch: torch.Tensor | None | bool = True, split_sync=True,
               stream=None) -> torch.Tensor:
    """O1 (and O2 when fed dY and the rotated filters): y[P,C_out] = conv3x3(x) (+bias).

    ``stats`` (fp32 [2, C_out]) is accumulated with per-channel sum / sum of squares of y.
    ``scratch``: fp32 buffer for split-K on small problems; True = allocate if the launcher
    wants one, False/None = never split.
    """
    nat.require_cuda(x, w)
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    assert x.numel() >= N * H * W * C_in and w.numel() == C_out * taps * C_in
    if out is None:
        out = torch.empty(N * H * W, C_out, dtype=torch.bfloat16, device=x.device)
    if scratch is True:
        need = conv_fprop_scratch_elems(N, H, W, C_in, C_out, taps, block_n)
        scratch = torch.empty(need, dtype=torch.float32, device=x.device) if need else None
    elif scratch is False:
        scratch = None
    rc = nat.load().ddpb_conv_fprop(nat.ptr(x), nat.ptr(w), nat.ptr(bias), nat.ptr(out), nat.ptr(stats), N, H, W,
                                    C_in, C_out, taps, block_n, nat.ptr(scratch), _numel(scratch),
                                    nat.ptr(_sync_arg(split_sync, x.device)), _s(stream))
    nat.check(rc, "ddpb_conv_fprop")
    return out


def conv_bn_relu_pool_fwd(x, w, bias, gamma, beta, N, H, W, C_in, C_out, taps, pool: bool, *, conv_out=None, act=None,
                          stats=None, saved=None, scratch: torch.Tensor | None | bool = True, eps=BN_EPS,
                          split_sync=True, stream=None):
    """O1+O4+O6+O7: conv -> BatchNorm(batch stats) -> ReLU (-> 2x2 max-pool) in one call.  Returns
    (conv_out, act, saved).  ``stats``: zeroed fp32 scratch of 2*C_out + 32 elements."""
    nat.require_cuda(x, w, gamma, beta)
    dev = x.device
    P = N * H * W
    Po = P // 4 if pool else P
    if conv_out is None:
        conv_out = torch.empty(P, C_out, dtype=torch.bfloat16, device=dev)
    if act is None:
        act = torch.empty(Po, C_out, dtype=torch.bfloat16, device=dev)
    if stats is None:
        stats = torch.zeros(2 * C_out + 32, dtype=torch.float32, device=dev)
    assert stats.numel() >= 2 * C_out + 32
    if saved is None:
        saved = torch.empty(2, C_out, dtype=torch.float32, device=dev)
    if scratch is True:
        need = conv_fprop_scratch_elems(N, H, W, C_in, C_out, taps, 0)
        scratch = torch.empty(need, dtype=torch.float32, device=dev) if need else None
    elif scratch is False:
        scratch = None
    rc = nat.load().ddpb_conv_bn_relu_pool_fwd(nat.ptr(x), nat.ptr(w), nat.ptr(bias), nat.ptr(conv_out), nat.ptr(stats),
                                               nat.ptr(gamma), nat.ptr(beta), nat.ptr(act), nat.ptr(saved), N, H, W,
                                               C_in, C_out, taps, int(pool), eps, nat.ptr(scratch), _numel(scratch),
                                               nat.ptr(_sync_arg(split_sync, dev)), _s(stream))
    nat.check(rc, "ddpb_conv_bn_relu_pool_fwd")
    return conv_out, act, saved


def conv_dgrad(dy: torch.Tensor, w_fprop: torch.Tensor, N: int, H: int, W: int, C_in: int, C_out: int, *,
               out: torch.Tensor | None = None, block_n: int = 0, scratch: torch.Tensor | None | bool = True,
               split_sync=True, stream=None) -> torch.Tensor:
    """O2: dx[P, C_in] from dy[P, C_out] and the SAME bf16 filter pack fprop uses ([C_out, 9*C_in])."""
    nat.require_cuda(dy, w_fprop)
    assert dy.dtype == torch.bfloat16 and w_fprop.dtype == torch.bfloat16
    assert dy.numel() >= N * H * W * C_out and w_fprop.numel() == C_out * 9 * C_in
    if out is None:
        out = torch.empty(N * H * W, C_in, dtype=torch.bfloat16, device=dy.device)
    if scratch is True:
        need = conv_fprop_scratch_elems(N, H, W, C_out, C_in, 9, block_n)
        scratch = torch.empty(need, dtype=torch.float32, device=dy.device) if need else None
    elif scratch is False:
        scratch = None
    rc = nat.load().ddpb_conv_dgrad(nat.ptr(dy), nat.ptr(w_fprop), nat.ptr(out), N, H, W, C_in, C_out, block_n,
                                    nat.ptr(scratch), _numel(scratch), nat.ptr(_sync_arg(split_sync, dy.device)),
                                    _s(stream))
    nat.check(rc, "ddpb_conv_dgrad")
    return out


def conv_wgrad_scratch_elems(N, H, W, C_in, C_out, taps=9, cols_valid: int | None = None,
                             out_ld: int | None = None) -> int:
    """fp32 elements of scratch conv_wgrad needs; a dw with padded rows / fewer valid columns
    (``cols_valid``, ``out_ld``) is assembled from scratch even when the reduction is not split."""
    if cols_valid is None and out_ld is None:
        return int(nat.load().ddpb_conv_wgrad_scratch_elems(N, H, W, C_in, C_out, taps))
    cols_valid = taps * C_in if cols_valid is None else cols_valid
    out_ld = cols_valid if out_ld is None else out_ld
    return int(nat.load().ddpb_conv_wgrad_scratch_elems_ld(N, H, W, C_in, C_out, taps, cols_valid, out_ld))


def conv_wgrad_splits(N, H, W, C_in, C_out, taps=9) -> int:
    return int(nat.load().ddpb_conv_wgrad_splits(N, H, W, C_in, C_out, taps))


def conv_wgrad(dy: torch.Tensor, x: torch.Tensor, N: int, H: int, W: int, C_in: int, C_out: int, taps: int = 9, *,
               out: torch.Tensor | None = None, cols_valid: int | None = None, out_ld: int | None = None,
               scratch: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    """O3: dW[C_out, taps*C_in] (fp32) = sum over pixels of dy^T * shifted x."""
    nat.require_cuda(dy, x)
    assert dy.dtype == torch.bfloat16 and x.dtype == torch.bfloat16
    cols_valid = taps * C_in if cols_valid is None else cols_valid
    out_ld = cols_valid if out_ld is None else out_ld
    if out is None:
        out = torch.empty(C_out, out_ld, dtype=torch.float32, device=x.device)
    need = conv_wgrad_scratch_elems(N, H, W, C_in, C_out, taps, cols_valid, out_ld)
    if need and (scratch is None or scratch.numel() < need):
        scratch = torch.empty(need, dtype=torch.float32, device=x.device)
    rc = nat.load().ddpb_conv_wgrad(nat.ptr(dy), nat.ptr(x), nat.ptr(out), nat.ptr(scratch), _numel(scratch), N, H, W,
                                    C_in, C_out, taps, cols_valid, out_ld, _s(stream))
    nat.check(rc, "ddpb_conv_wgrad")
    return out


def channel_stats(x: torch.Tensor, C_: int, *, stats: torch.Tensor | None = None, stream=None) -> torch.Tensor:
    nat.require_cuda(x)
    if stats is None:
        stats = torch.zeros(2, C_, dtype=torch.float32, device=x.device)
    rc = nat.load().ddpb_channel_stats(nat.ptr(x), nat.ptr(stats), x.numel() // C_, C_, _s(stream))
    nat.check(rc, "ddpb_channel_stats")
    return stats


def bn_relu_pool_fwd(x, stats, gamma, beta, N, H, W, C_, pool: bool, *, out=None, saved=None, eps=BN_EPS,
                     stream=None):
    """O4+O6+O7 forward.  Returns (y, saved[2,C] = mean, rstd)."""
    nat.require_cuda(x, stats, gamma, beta)
    Po = N * (H // 2 if pool else H) * (W // 2 if pool else W)
    if out is None:
        out = torch.empty(Po, C_, dtype=torch.bfloat16, device=x.device)
    if saved is None:
        saved = torch.empty(2, C_, dtype=torch.float32, device=x.device)
    rc = nat.load().ddpb_bn_relu_pool_fwd(nat.ptr(x), nat.ptr(stats), nat.ptr(gamma), nat.ptr(beta), nat.ptr(out),
                                          nat.ptr(saved), N, H, W, C_, int(pool), eps, _s(stream))
    nat.check(rc, "ddpb_bn_relu_pool_fwd")
    return out, saved


def stamp(slot: torch.Tensor, stream=None) -> None:
    """Diagnostics: a one-thread kernel writes %globaltimer (ns) into ``slot`` (int64[1])."""
    assert slot.is_cuda and slot.dtype == torch.int64 and slot.numel() >= 1
    nat.check(nat.load().ddpb_stamp(nat.ptr(slot), _s(stream)), "ddpb_stamp")


def publish_scalar(value: torch.Tensor, dev_seq: torch.Tensor, host_slot: torch.Tensor, stream=None) -> None:
    """Stores (++dev_seq << 32 | bits of value) into ``host_slot`` (pinned int64[1]) from the device."""
    assert value.is_cuda and value.dtype == torch.float32 and dev_seq.is_cuda and dev_seq.dtype == torch.int32
    assert host_slot.is_pinned() and host_slot.dtype == torch.int64
    nat.check(nat.load().ddpb_publish_scalar(nat.ptr(value), nat.ptr(dev_seq), nat.ptr(host_slot), _s(stream)),
              "ddpb_publish_scalar")


def sm_count() -> int:
    return int(nat.load().ddpb_sm_count())


def set_pdl(on: bool) -> None:
    """Programmatic dependent launch between consecutive kernels (default on; env DDPB_PDL=0 turns it
    off): a kernel's prologue overlaps the previous kernel's tail.  Takes effect for launches (and
    graph captures) made afterwards."""
    nat.load().ddpb_set_pdl(1 if on else 0)


def get_pdl() -> bool:
    return bool(nat.load().ddpb_get_pdl())


def bn_bwd_set_fused(mode: bool | int) -> None:
    """3 (default): two ordinary kernels, fixed-order sums (deterministic, no float atomics);
    2/True: fused cooperative kernel, up to two CTAs per SM; 1: same with one CTA per SM;
    0/False: the two-pass kernels with float atomics of round 1."""
    nat.load().ddpb_bn_bwd_set_fused(2 if mode is True else int(mode))


def bn_bwd_partials_elems(N, H, W, C_, pool: bool) -> int:
    return int(nat.load().ddpb_bn_bwd_partials_elems(N, H, W, C_, int(pool)))


def bn_relu_pool_bwd(x, dy, saved, gamma, beta, N, H, W, C_, pool: bool, *, sums=None, dx=None, dgamma=None,
                     dbeta=None, dbias=None, partials: torch.Tensor | None | bool = True, eps=BN_EPS, stream=None):
    """O5+O6+O7 backward.  ``sums`` (fp32, 2*C + 32 elements) must be zero on entry.  ``partials``:
    fp32 scratch of bn_bwd_partials_elems() for the deterministic default path (True = allocate,
    None/False = use the cooperative one-kernel variant)."""
    nat.require_cuda(x, dy, saved, gamma, beta)
    dev = x.device
    if sums is None:
        sums = torch.zeros(2 * C_ + 32, dtype=torch.float32, device=dev)
    assert sums.numel() >= 2 * C_ + 32, "bn backward scratch needs 2*C + 32 zeroed floats"
    if partials is True:
        need = bn_bwd_partials_elems(N, H, W, C_, pool)
        partials = torch.empty(need, dtype=torch.float32, device=dev) if need else None
    elif partials is False:
        partials = None
    if dx is None:
        dx = torch.empty(N * H * W, C_, dtype=torch.bfloat16, device=dev)
    if dgamma is None:
        dgamma = torch.empty(C_, dtype=torch.float32, device=dev)
    if dbeta is None:
        dbeta = torch.empty(C_, dtype=torch.float32, device=dev)
    rc = nat.load().ddpb_bn_relu_pool_bwd(nat.ptr(x), nat.ptr(dy), nat.ptr(saved), nat.ptr(gamma), nat.ptr(beta),
                                          nat.ptr(sums), nat.ptr(dx), nat.ptr(dgamma), nat.ptr(dbeta),
                                          nat.ptr(dbias), nat.ptr(partials), _numel(partials), N, H, W, C_, int(pool),
                                          eps, _s(stream))
    nat.check(rc, "ddpb_bn_relu_pool_bwd")
    return dx, dgamma, dbeta


def fc_ce_fwd(x, w, b, labels, *, logits=None, row_loss=None, dlogits=None, row_correct=None, loss=None,
              correct=None, stream=None):
    """O9+O10(+O15) forward: logits, mean CE loss, dlogits, correct count."""
    nat.require_cuda(x, w, b, labels)
    assert labels.dtype == torch.int64 and x.dtype == torch.bfloat16
    B, F = x.shape
    K = w.shape[0]
    dev = x.device
    if row_loss is None:
        row_loss = torch.empty(B, dtype=torch.float32, device=dev)
    rc = nat.load().ddpb_fc_ce_fwd(nat.ptr(x), nat.ptr(w), nat.ptr(b), nat.ptr(labels), nat.ptr(logits),
                                   nat.ptr(row_loss), nat.ptr(dlogits), nat.ptr(row_correct), nat.ptr(loss),
                                   nat.ptr(correct), B, F, K, _s(stream))
    nat.check(rc, "ddpb_fc_ce_fwd")
    return row_loss


def fc_bwd(x, w, dlogits, *, dw=None, db=None, dx=None, want_dx=True, row_loss=None, row_correct=None, loss=None,
           correct=None, stream=None):
    """O9 backward.  With ``row_loss`` (+``row_correct``) and ``loss``/``correct`` the same launch also
    reduces the forward's per-row losses to the mean loss / correct count."""
    nat.require_cuda(x, w, dlogits)
    B, F = x.shape
    K = w.shape[0]
    dev = x.device
    if dw is None:
        dw = torch.empty(K, F, dtype=torch.float32, device=dev)
    if db is None:
        db = torch.empty(K, dtype=torch.float32, device=dev)
    if dx is None and want_dx:
        dx = torch.empty(B, F, dtype=torch.bfloat16, device=dev)
    rc = nat.load().ddpb_fc_bwd(nat.ptr(x), nat.ptr(w), nat.ptr(dlogits), nat.ptr(dw), nat.ptr(db), nat.ptr(dx), B, F,
                                K, nat.ptr(row_loss), nat.ptr(row_correct), nat.ptr(loss), nat.ptr(correct),
                                _s(stream))
    nat.check(rc, "ddpb_fc_bwd")
    return dw, db, dx


def sgd_step(p, g, m, lr, momentum, weight_decay, grad_scale=1.0, *, shadow=None, stream=None):
    """O11+O12 over flat fp32 arenas, in place; ``shadow`` (bf16) receives the rounded parameters."""
    nat.require_cuda(p, g, m)
    assert p.dtype == g.dtype == m.dtype == torch.float32 and p.numel() == g.numel() == m.numel()
    assert shadow is None or (shadow.dtype == torch.bfloat16 and shadow.numel() == p.numel())
    rc = nat.load().ddpb_sgd_step(nat.ptr(p), nat.ptr(g), nat.ptr(m), nat.ptr(shadow), p.numel(), lr, momentum,
                                  weight_decay, grad_scale, _s(stream))
    nat.check(rc, "ddpb_sgd_step")


class PackPlan:
    """Pre-marshalled argument arrays for ddpb_pack_weights (built once per model)."""

    def __init__(self, layers):
        """layers: list of dict(master, fprop, dgrad|None, K, C, taps, fprop_ld, flat)."""
        n = len(layers)
        self.n = n
        self._keep = layers
        VP = C.c_void_p * n
        IA = C.c_int * n
        self.master = VP(*[l["master"].data_ptr() for l in layers])
        self.fprop = VP(*[l["fprop"].data_ptr() for l in layers])
        self.dgrad = VP(*[(l["dgrad"].data_ptr() if l["dgrad"] is not None else None) for l in layers])
        self.K = IA(*[l["K"] for l in layers])
        self.C = IA(*[l["C"] for l in layers])
        self.taps = IA(*[l["taps"] for l in layers])
        self.ld = IA(*[l["fprop_ld"] for l in layers])
        self.flat = IA(*[int(l["flat"]) for l in layers])

    def run(self, stream=None):
        rc = nat.load().ddpb_pack_weights(self.n, self.master, self.fprop, self.dgrad, self.K, self.C, self.taps,
                                          self.ld, self.flat, _s(stream))
        nat.check(rc, "ddpb_pack_weights")


def _pack_args(w0_master, w0_packed):
    if w0_master is None:
        return None, None, 0
    assert w0_master.dtype == torch.float32 and w0_packed.dtype == torch.bfloat16 and w0_packed.shape[1] == 64
    assert w0_master.numel() == w0_packed.shape[0] * 27
    return nat.ptr(w0_master), nat.ptr(w0_packed), w0_packed.shape[0]


def first_layer_operand_f32(x_nchw: torch.Tensor, *, out=None, w0_master=None, w0_packed=None,
                            stream=None) -> torch.Tensor:
    """O16 tail + im2col for layer 0 from a normalised fp32 NCHW batch.  ``w0_master`` (fp32
    [C_out*27]) / ``w0_packed`` (bf16 [C_out][64]): also refresh the first layer's filter pack."""
    nat.require_cuda(x_nchw)
    assert x_nchw.dtype == torch.float32 and x_nchw.is_contiguous() and x_nchw.shape[1] == 3
    N, _, H, W = x_nchw.shape
    if out is None:
        out = torch.empty(N * H * W, 64, dtype=torch.bfloat16, device=x_nchw.device)
    rc = nat.load().ddpb_first_layer_operand_f32(nat.ptr(x_nchw), nat.ptr(out), N, H, W,
                                                 *_pack_args(w0_master, w0_packed), _s(stream))
    nat.check(rc, "ddpb_first_layer_operand_f32")
    return out


_MEAN3 = (C.c_float * 3)(125.3 / 255.0, 123.0 / 255.0, 113.9 / 255.0)  # part1/main.py:20
_STD3 = (C.c_float * 3)(63.0 / 255.0, 62.1 / 255.0, 66.7 / 255.0)      # part1/main.py:21


def draw_augmentation(aug: torch.Tensor, step: torch.Tensor, *, seed: int, pad: int = 4, n: int | None = None,
                      stream=None) -> torch.Tensor:
    """Draws RandomCrop(32, pad) offsets and RandomHorizontalFlip bits for ``n`` images on the device
    (Philox4x32-10, counter = (step, image)) into ``aug`` (int8 [N][4]); ``step`` (int32 [1], device) is read
    and incremented, so a captured launch draws anew on every replay.  The reference draws them on CPU
    DataLoader workers (/root/reference/part1/main.py:23-26)."""
    nat.require_cuda(aug)
    assert aug.dtype == torch.int8 and aug.dim() == 2 and aug.shape[1] == 4 and aug.is_contiguous()
    assert step.dtype == torch.int32 and step.numel() == 1 and step.is_cuda
    n = aug.shape[0] if n is None else n
    assert 1 <= n <= aug.shape[0]
    nat.check(nat.load().ddpb_draw_augmentation(nat.ptr(aug), n, pad, seed & 0xFFFFFFFFFFFFFFFF, nat.ptr(step),
                                                _s(stream)), "ddpb_draw_augmentation")
    return aug


def synthetic_batch(n: int, step: torch.Tensor, *, seed: int, hw: int = 32, classes: int = 10, out=None,
                    stream=None):
    """A synthetic CIFAR-shaped batch generated on the device: (uint8 [n][hw][hw][3], int64 [n]).  Class
    templates depend on ``seed`` only; labels and noise on (seed, step).  Increments ``step``."""
    nat.require_cuda(step)
    assert step.dtype == torch.int32 and step.numel() == 1
    if out is None:
        out = (torch.empty(n, hw, hw, 3, dtype=torch.uint8, device=step.device),
               torch.empty(n, dtype=torch.int64, device=step.device))
    img, lab = out
    assert img.dtype == torch.uint8 and img.is_contiguous() and img.shape == (n, hw, hw, 3)
    assert lab.dtype == torch.int64 and lab.is_contiguous() and lab.shape == (n,)
    nat.check(nat.load().ddpb_synthetic_batch(nat.ptr(img), nat.ptr(lab), n, hw, hw, classes,
                                              seed & 0xFFFFFFFFFFFFFFFF, nat.ptr(step), _s(stream)),
              "ddpb_synthetic_batch")
    return img, lab


def first_layer_operand_u8(x_nhwc_u8: torch.Tensor, aug: torch.Tensor | None = None, *, out=None, pad: int = 4,
                           w0_master=None, w0_packed=None, stream=None) -> torch.Tensor:
    """O16: crop/flip/normalise raw uint8 NHWC pixels and emit the layer-0 GEMM operand."""
    nat.require_cuda(x_nhwc_u8)
    assert x_nhwc_u8.dtype == torch.uint8 and x_nhwc_u8.is_contiguous() and x_nhwc_u8.shape[-1] == 3
    N, H, W, _ = x_nhwc_u8.shape
    if aug is not None:
        assert aug.dtype == torch.int8 and aug.shape == (N, 4) and aug.is_cuda
    if out is None:
        out = torch.empty(N * H * W, 64, dtype=torch.bfloat16, device=x_nhwc_u8.device)
    rc = nat.load().ddpb_first_layer_operand_u8(nat.ptr(x_nhwc_u8), nat.ptr(aug), nat.ptr(out), N, H, W, _MEAN3,
                                                _STD3, pad, *_pack_args(w0_master, w0_packed), _s(stream))
    nat.check(rc, "ddpb_first_layer_operand_u8")
    return out
