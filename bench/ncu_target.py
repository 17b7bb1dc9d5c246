"""Tiny driver for ncu: runs a few representative kernels (warm-up + a handful of launches).

    python bench/ncu_target.py [which ...]     which in {fprop3, fprop5, fprop7, wgrad3, wgrad7, bnbwd0, bnfwd0, fprop0}
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.ops import kernels as K  # noqa: E402

LAYERS = [(32, 3, 64, True), (16, 64, 128, True), (8, 128, 256, False), (8, 256, 256, True),
          (4, 256, 512, False), (4, 512, 512, True), (2, 512, 512, False), (2, 512, 512, True)]
N = 256
dev = "cuda"
which = sys.argv[1:] or ["fprop3", "fprop5", "wgrad3", "bnbwd0"]
for w in which:
    op, li = w[:-1], int(w[-1])
    S, Ci, Co, pool = LAYERS[li]
    H = W = S
    P = N * H * W
    Cin, taps = (64, 1) if li == 0 else (Ci, 9)
    x = torch.randn(P, Cin, device=dev).bfloat16()
    wt = torch.randn(Co, taps * Cin, device=dev).bfloat16()
    y = torch.empty(P, Co, dtype=torch.bfloat16, device=dev)
    dy = torch.randn(P, Co, device=dev).bfloat16()
    stats = torch.zeros(2, Co, device=dev)
    bias = torch.zeros(Co, device=dev)
    for it in range(3):
        if op == "fprop":
            fs = torch.empty(max(1, K.conv_fprop_scratch_elems(N, H, W, Cin, Co, taps)), device=dev)
            K.conv_fprop(x, wt, bias, N, H, W, Cin, Co, taps, out=y, stats=stats, scratch=fs)
        elif op == "wgrad":
            K.conv_wgrad(dy, x, N, H, W, Cin, Co, taps)
        elif op in ("bnbwd", "bnfwd"):
            gamma, beta = torch.ones(Co, device=dev), torch.zeros(Co, device=dev)
            st = K.channel_stats(dy, Co)
            act, saved = K.bn_relu_pool_fwd(dy, st, gamma, beta, N, H, W, Co, pool)
            if op == "bnbwd":
                K.bn_relu_pool_bwd(dy, torch.randn_like(act), saved, gamma, beta, N, H, W, Co, pool)
        torch.cuda.synchronize()
print("done", which)
