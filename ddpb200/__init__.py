"""ddpb200 -- B200-native data-parallel training (VGG-11 / CIFAR-10, four sync modes).

Capabilities of ruc98/Distributed-Data-Parallel-ML-Training, rebuilt for Blackwell:
hand-written sm_100a kernels (tcgen05/TMEM/TMA implicit-GEMM convolutions, fused
BN/ReLU/pool, classifier+loss, SGD) and peer-memory gradient synchronisation over NVLink.
See SURVEY.md for the reference's inventory and DESIGN.md for how this package maps to it.
"""
from .models.vgg import VGG11, VGG13, VGG16, VGG19, _VGG, _cfg, _make_layers  # noqa: F401
from .optim import SGD, CrossEntropyLoss  # noqa: F401
from .trainer import Trainer  # noqa: F401

__all__ = ["VGG11", "VGG13", "VGG16", "VGG19", "SGD", "CrossEntropyLoss", "Trainer"]
__version__ = "0.1.0"
