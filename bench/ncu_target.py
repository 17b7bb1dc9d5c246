"""Driver for ncu: a few launches of every hot kernel of the training step at its batch-256 shape.

    python bench/ncu_target.py [which ...]
        which in {fwdN, dgradN, wgradN, bnbwdN (N = layer 0..7), fc, sgd, input}; default: a representative set

Each op runs 3 times (2 warm-ups + the one to look at: `ncu -k regex:... -s/-c` picks launches).
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
DEFAULT = ["input", "fwd0", "fwd1", "fwd3", "fwd5", "fwd7", "dgrad1", "dgrad3", "dgrad5", "dgrad7", "wgrad0", "wgrad3",
           "wgrad5", "wgrad6", "bnbwd0", "bnbwd1", "bnbwd3", "bnbwd6", "fc", "sgd"]
which = sys.argv[1:] or DEFAULT
import os

REPS = int(os.environ.get("DDPB_NCU_REPS", "3"))
for w in which:
    if w == "fc":
        x = torch.randn(N, 512, device=dev).bfloat16()
        wt, b = torch.randn(10, 512, device=dev) * 0.05, torch.zeros(10, device=dev)
        lab = torch.randint(0, 10, (N,), device=dev)
        logits, dlog = torch.empty(N, 10, device=dev), torch.empty(N, 10, device=dev)
        rl, rc = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
        loss, cor = torch.zeros((), device=dev), torch.zeros((), dtype=torch.int32, device=dev)
        for _ in range(REPS):
            K.fc_ce_fwd(x, wt, b, lab, logits=logits, row_loss=rl, dlogits=dlog, row_correct=rc)
            K.fc_bwd(x, wt, dlog, row_loss=rl, row_correct=rc, loss=loss, correct=cor)
        continue
    if w == "sgd":
        n = 9_233_024
        p, g, m = (torch.randn(n, device=dev) for _ in range(3))
        sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
        for _ in range(REPS):
            K.sgd_step(p, g, m, 0.1, 0.9, 1e-4, shadow=sh)
        continue
    if w == "input":
        img = torch.randint(0, 256, (N, 32, 32, 3), dtype=torch.uint8, device=dev)
        aug = torch.zeros(N, 4, dtype=torch.int8, device=dev)
        for _ in range(REPS):
            K.first_layer_operand_u8(img, aug)
        continue
    op, li = w[:-1], int(w[-1])
    S, Ci, Co, pool = LAYERS[li]
    H = W = S
    P = N * H * W
    Cin, taps = (64, 1) if li == 0 else (Ci, 9)
    x = torch.randn(P, Cin, device=dev).bfloat16()
    wt = (torch.randn(Co, taps * Cin, device=dev) / (taps * Cin) ** 0.5).bfloat16()
    dy = torch.randn(P, Co, device=dev).bfloat16()
    bias, gamma, beta = torch.zeros(Co, device=dev), torch.ones(Co, device=dev), torch.zeros(Co, device=dev)
    for it in range(REPS):
        if op == "fwd":
            K.conv_bn_relu_pool_fwd(x, wt, bias, gamma, beta, N, H, W, Cin, Co, taps, pool)
        elif op == "dgrad":
            K.conv_dgrad(dy, wt, N, H, W, Ci, Co, split_sync=None)     # as the engine calls it: two-kernel split-K
        elif op == "wgrad":
            cols = 27 if li == 0 else 9 * Ci
            K.conv_wgrad(dy, x, N, H, W, Cin, Co, taps, cols_valid=cols, out_ld=cols)
        elif op == "bnbwd":
            st = K.channel_stats(dy, Co)
            act, saved = K.bn_relu_pool_fwd(dy, st, gamma, beta, N, H, W, Co, pool)
            K.bn_relu_pool_bwd(dy, torch.randn_like(act), saved, gamma, beta, N, H, W, Co, pool)
        torch.cuda.synchronize()
print("done", which)
