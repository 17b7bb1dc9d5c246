"""Plain PyTorch fp32 restatements of every kernel in ddpb200.ops.kernels.

They serve two purposes: (1) the oracle in the numerics tests ("each sm_100a
kernel ... against the same op on CPU in fp32"), built from the very modules the reference
uses (nn.Conv2d / BatchNorm2d(track_running_stats=False) / ReLU / MaxPool2d / Linear /
CrossEntropyLoss, /root/reference/part1/model.py:11-46); (2) the CPU execution path of
the framework, so the plumbing is testable on a host without a GPU.

All functions take and return the kernels' layouts (NHWC matrices, [C_out, taps, C_in]
filters) but compute in fp32 with ATen.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5


def nhwc_to_nchw(x2d: torch.Tensor, N: int, H: int, W: int) -> torch.Tensor:
    return x2d.reshape(N, H, W, -1).permute(0, 3, 1, 2).contiguous()


def nchw_to_nhwc2d(x: torch.Tensor) -> torch.Tensor:
    N, C, H, W = x.shape
    return x.permute(0, 2, 3, 1).reshape(N * H * W, C).contiguous()


def filters_to_oihw(w: torch.Tensor, C_out: int, taps: int, C_in: int) -> torch.Tensor:
    """[C_out, taps*C_in] (tap-major, channel fastest) -> (C_out, C_in, 3, 3)."""
    assert taps == 9
    return w.reshape(C_out, 3, 3, C_in).permute(0, 3, 1, 2).contiguous()


def conv_fprop(x2d, w, bias, N, H, W, C_in, C_out, taps=9):
    x2d, w = x2d.float(), w.float()
    if taps == 1:
        y = x2d @ w.reshape(C_out, C_in).t()
        return y + bias.float() if bias is not None else y
    y = F.conv2d(nhwc_to_nchw(x2d, N, H, W), filters_to_oihw(w, C_out, taps, C_in),
                 None if bias is None else bias.float(), stride=1, padding=1)
    return nchw_to_nhwc2d(y)


def conv_dgrad(dy2d, w, N, H, W, C_in, C_out):
    """dX for y = conv3x3(x, w): conv_transpose of dy with the same filters."""
    dx = F.conv_transpose2d(nhwc_to_nchw(dy2d.float(), N, H, W), filters_to_oihw(w.float(), C_out, 9, C_in),
                            stride=1, padding=1)
    return nchw_to_nhwc2d(dx)


def rotate_filters_for_dgrad(w: torch.Tensor, C_out: int, C_in: int) -> torch.Tensor:
    """Wd[c][t][k] = W[k][8-t][c]: dgrad becomes a plain fprop with these filters."""
    return w.reshape(C_out, 9, C_in).flip(1).permute(2, 1, 0).contiguous().reshape(C_in, 9 * C_out)


def conv_wgrad(dy2d, x2d, N, H, W, C_in, C_out, taps=9):
    dy2d, x2d = dy2d.float(), x2d.float()
    if taps == 1:
        return dy2d.t() @ x2d
    x = nhwc_to_nchw(x2d, N, H, W)
    dy = nhwc_to_nchw(dy2d, N, H, W)
    dw = torch.nn.grad.conv2d_weight(x, (C_out, C_in, 3, 3), dy, stride=1, padding=1)  # OIHW
    return dw.permute(0, 2, 3, 1).reshape(C_out, 9 * C_in).contiguous()


def channel_stats(x2d):
    x = x2d.float()
    return torch.stack([x.sum(0), (x * x).sum(0)])


def bn_relu_pool_fwd(x2d, gamma, beta, N, H, W, pool: bool, eps=BN_EPS):
    x = nhwc_to_nchw(x2d.float(), N, H, W)
    y = F.batch_norm(x, None, None, gamma.float(), beta.float(), training=True, momentum=0.0, eps=eps)
    y = F.relu(y)
    if pool:
        y = F.max_pool2d(y, 2, 2)
    return nchw_to_nhwc2d(y)


def bn_relu_pool_bwd(x2d, dy2d, gamma, beta, N, H, W, pool: bool, eps=BN_EPS):
    """Returns dx2d, dgamma, dbeta via autograd over the reference modules."""
    x = nhwc_to_nchw(x2d.float(), N, H, W).requires_grad_(True)
    g = gamma.float().clone().requires_grad_(True)
    b = beta.float().clone().requires_grad_(True)
    y = F.batch_norm(x, None, None, g, b, training=True, momentum=0.0, eps=eps)
    y = F.relu(y)
    if pool:
        y = F.max_pool2d(y, 2, 2)
    Ho, Wo = y.shape[2], y.shape[3]
    y.backward(nhwc_to_nchw(dy2d.float(), N, Ho, Wo))
    return nchw_to_nhwc2d(x.grad), g.grad, b.grad


def fc_ce(x, w, b, labels):
    """Returns logits, mean loss, dlogits (grad of the mean loss), number correct."""
    logits = x.float() @ w.float().t() + b.float()
    loss = F.cross_entropy(logits, labels)
    dlogits = (torch.softmax(logits, 1) - F.one_hot(labels, logits.shape[1]).float()) / x.shape[0]
    correct = (logits.argmax(1) == labels).sum()
    return logits, loss, dlogits, correct


def fc_bwd(x, w, dlogits):
    return dlogits.t() @ x.float(), dlogits.sum(0), dlogits @ w.float()


def sgd_step(p, g, m, lr, momentum, weight_decay, grad_scale=1.0):
    """In-place torch.optim.SGD rule (dampening 0, no Nesterov); m starts at zero."""
    gg = g * grad_scale + weight_decay * p
    m.mul_(momentum).add_(gg)
    p.add_(m, alpha=-lr)


def first_layer_operand(x_nchw: torch.Tensor) -> torch.Tensor:
    """[N,3,H,W] fp32 -> [N*H*W, 64] with column tap*3+c, zero padded past 27."""
    N, C, H, W = x_nchw.shape
    cols = F.unfold(x_nchw.float(), 3, padding=1)            # [N, C*9, H*W], row index c*9+tap
    cols = cols.reshape(N, C, 9, H * W).permute(0, 3, 2, 1)  # [N, HW, tap, c]
    out = torch.zeros(N * H * W, 64, dtype=torch.float32, device=x_nchw.device)
    out[:, :27] = cols.reshape(N * H * W, 27)
    return out


CIFAR_MEAN = torch.tensor([125.3, 123.0, 113.9]) / 255.0  # part1/main.py:20
CIFAR_STD = torch.tensor([63.0, 62.1, 66.7]) / 255.0      # part1/main.py:21


def augment_normalize_u8(x_nhwc_u8: torch.Tensor, aug: torch.Tensor | None, pad: int = 4) -> torch.Tensor:
    """RandomCrop(pad)+flip with GIVEN offsets, ToTensor, Normalize -> fp32 NCHW (part1/main.py:23-28)."""
    N, H, W, _ = x_nhwc_u8.shape
    x = x_nhwc_u8.permute(0, 3, 1, 2).float()
    if aug is not None:
        # one gather for the whole batch: row / column index of every output pixel in the padded image
        xp = F.pad(x, (pad, pad, pad, pad))
        a = aug.to(torch.int64)
        rows = a[:, 0:1] + torch.arange(H, device=x.device)                      # [N, H]
        cols = torch.arange(W, device=x.device).expand(N, W)
        cols = torch.where(a[:, 2:3] != 0, W - 1 - cols, cols) + a[:, 1:2]       # flip, then shift
        n_idx = torch.arange(N, device=x.device)[:, None, None]
        x = xp.permute(0, 2, 3, 1)[n_idx, rows[:, :, None], cols[:, None, :]].permute(0, 3, 1, 2)
    mean = CIFAR_MEAN.to(x.device).view(1, 3, 1, 1)
    std = CIFAR_STD.to(x.device).view(1, 3, 1, 1)
    return (x / 255.0 - mean) / std


# ---- Philox4x32-10: the counter-based generator of the device-side augmentation draws ---------------
_PHILOX_M0, _PHILOX_M1 = 0xD2511F53, 0xCD9E8D57
_PHILOX_W0, _PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(ctr, key):
    """Philox4x32 with 10 rounds (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3") on numpy
    arrays: ``ctr`` = four uint32 arrays of one shape, ``key`` = two uint32 scalars or arrays.  Restates
    ``philox4x32_10`` of csrc/data_prep.cu word for word; checked against the published known-answer vectors
    in tests/test_cpu_plumbing.py."""
    import numpy as np

    c = [np.asarray(v, dtype=np.uint64) & 0xFFFFFFFF for v in ctr]
    k0 = np.asarray(key[0], dtype=np.uint64) & 0xFFFFFFFF
    k1 = np.asarray(key[1], dtype=np.uint64) & 0xFFFFFFFF
    for _ in range(10):
        p0 = c[0] * np.uint64(_PHILOX_M0)          # 32 x 32 -> 64 bit products
        p1 = c[2] * np.uint64(_PHILOX_M1)
        c = [(p1 >> np.uint64(32)) ^ c[1] ^ k0, p1 & np.uint64(0xFFFFFFFF),
             (p0 >> np.uint64(32)) ^ c[3] ^ k1, p0 & np.uint64(0xFFFFFFFF)]
        k0 = (k0 + np.uint64(_PHILOX_W0)) & np.uint64(0xFFFFFFFF)
        k1 = (k1 + np.uint64(_PHILOX_W1)) & np.uint64(0xFFFFFFFF)
    return [v.astype(np.uint32) for v in c]


def draw_augmentation(n: int, pad: int, seed: int, step: int) -> torch.Tensor:
    """The crop / flip parameters ``ddpb_draw_augmentation`` draws for ``n`` images at step counter ``step``:
    int8 [n][4] = {crop y, crop x in [0, 2 pad], flip, 0}; counter = (image, step, "AUG!", 0), key = seed."""
    import numpy as np

    seed &= 0xFFFFFFFFFFFFFFFF
    img = np.arange(n, dtype=np.uint64)
    zeros = np.zeros(n, dtype=np.uint64)
    r = philox4x32_10([img, zeros + (step & 0xFFFFFFFF), zeros + 0x41554721, zeros], (seed & 0xFFFFFFFF, seed >> 32))
    span = np.uint64(2 * pad + 1)
    out = np.zeros((n, 4), dtype=np.int8)
    out[:, 0] = ((r[0].astype(np.uint64) * span) >> np.uint64(32)).astype(np.int8)
    out[:, 1] = ((r[1].astype(np.uint64) * span) >> np.uint64(32)).astype(np.int8)
    out[:, 2] = (r[2] >> np.uint32(31)).astype(np.int8)
    return torch.from_numpy(out)
