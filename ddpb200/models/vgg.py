"""VGG family for 3x32x32 inputs and 10 classes.

Same public surface and the same ``state_dict`` as the reference's ``model.py``
(/root/reference/part1/model.py:1-50): ``VGG11()`` returns a module whose parameters are
``layers.<i>.weight/bias`` (Conv2d and BatchNorm2d(track_running_stats=False) inside one
flat ``nn.Sequential``) and ``fc1.weight/bias`` -- 34 tensors, 9,231,114 elements, no
buffers.  Modules are created in the reference's order, so the same
``torch.manual_seed`` yields bit-identical initial weights.

What differs is how it runs.  On a CUDA device ``forward`` does not walk the Sequential:
it hands the batch to a :class:`ddpb200.engine.VGGEngine`, which executes every conv
block as hand-written sm_100a kernels (tcgen05 implicit GEMM + fused BN/ReLU/pool) and is
differentiable through one autograd node.  On CPU the stock modules run, which is the
oracle the tests compare against.

The reference only exports VGG11 but carries configs for 13/16/19 as dead data
(model.py:3-8); they are all live here.
"""
from __future__ import annotations

import torch
import torch.nn as nn

_cfg = {
    "VGG11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "VGG19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512,
              "M"],
}


def _make_layers(cfg) -> nn.Sequential:
    layers: list[nn.Module] = []
    in_channels = 3
    for item in cfg:
        if item == "M":
            layers.append(nn.MaxPool2d(kernel_size=2, stride=2))
        else:
            layers.append(nn.Conv2d(in_channels, item, kernel_size=3, stride=1, padding=1, bias=True))
            layers.append(nn.BatchNorm2d(item, track_running_stats=False))
            layers.append(nn.ReLU(inplace=True))
            in_channels = item
    return nn.Sequential(*layers)


class ConvBlock:
    """One conv -> BN -> ReLU (-> pool) group of the Sequential, as the engine sees it."""

    def __init__(self, conv: nn.Conv2d, bn: nn.BatchNorm2d, pool: bool, conv_idx: int, bn_idx: int):
        self.conv, self.bn, self.pool = conv, bn, pool
        self.conv_idx, self.bn_idx = conv_idx, bn_idx
        self.c_in, self.c_out = conv.in_channels, conv.out_channels


class _VGG(nn.Module):
    """VGG for 3x32x32 input, 10 classes (reference: model.py:30-46)."""

    def __init__(self, name: str):
        super().__init__()
        self.name = name
        self.layers = _make_layers(_cfg[name])
        self.fc1 = nn.Linear(512, 10)
        self._engine = None  # created lazily on first CUDA forward

    # -- structure helpers -----------------------------------------------------------------
    def conv_blocks(self) -> list[ConvBlock]:
        blocks, mods = [], list(self.layers)
        i = 0
        while i < len(mods):
            if isinstance(mods[i], nn.Conv2d):
                pool = i + 3 < len(mods) and isinstance(mods[i + 3], nn.MaxPool2d)
                blocks.append(ConvBlock(mods[i], mods[i + 1], pool, i, i + 1))
                i += 4 if pool else 3
            else:  # pragma: no cover - configs always start a block with a conv
                i += 1
        return blocks

    # -- execution ---------------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            from ..engine import engine_forward

            return engine_forward(self, x)
        y = self.layers(x)
        y = y.view(y.size(0), -1)
        return self.fc1(y)

    def _apply(self, fn, *args, **kwargs):
        # moving the module (e.g. .to("cuda")) invalidates any engine bound to the old storage
        self._engine = None
        return super()._apply(fn, *args, **kwargs)


def VGG11() -> _VGG:
    return _VGG("VGG11")


def VGG13() -> _VGG:
    return _VGG("VGG13")


def VGG16() -> _VGG:
    return _VGG("VGG16")


def VGG19() -> _VGG:
    return _VGG("VGG19")
