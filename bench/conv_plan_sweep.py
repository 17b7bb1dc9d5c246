"""Measured (tile width, split-K) sweep of the forward (conv + BatchNorm) and dgrad kernel groups of every
VGG-11 layer, against what the cost-based planner (`plan_conv`, csrc/conv_fprop.cu) picks.

    python bench/conv_plan_sweep.py [--batch 256] [--reps 20]

Every candidate is forced with `ddpb_conv_force`, replayed `reps` times inside a CUDA graph (as
bench/layer_graph_times.py does) and timed with CUDA events; a line per layer shows the planner's choice,
its time, the measured best and the table.
"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.ops import kernels as K  # noqa: E402
from layer_graph_times import LAYERS, timed  # noqa: E402

dev = "cuda"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--max-splits", type=int, default=8)
    ap.add_argument("--dgrad-two-kernels", action="store_true", help="dgrad without sync words (what the engine does)")
    ap.add_argument("--no-fuse-split", action="store_true", help="split-K summed by the second kernel, not inside the conv kernel")
    a = ap.parse_args()
    N, reps = a.batch, a.reps
    K.conv_fuse_split(not a.no_fuse_split)
    gain = dict(fwd=0.0, dgrad=0.0)
    for li, (S, Ci, Co, pool) in enumerate(LAYERS):
        H = W = S
        P = N * H * W
        Cin_eff, taps = (64, 1) if li == 0 else (Ci, 9)
        Po = P // 4 if pool else P
        x = torch.randn(P, Cin_eff, device=dev).bfloat16()
        w = (torch.randn(Co, taps * Cin_eff, device=dev) / (taps * Cin_eff) ** 0.5).bfloat16()
        bias, gamma, beta = torch.randn(Co, device=dev), torch.rand(Co, device=dev) + 0.5, torch.randn(Co, device=dev)
        conv_out = torch.empty(P, Co, dtype=torch.bfloat16, device=dev)
        act = torch.empty(Po, Co, dtype=torch.bfloat16, device=dev)
        saved = torch.empty(2 * Co, device=dev)
        stats = torch.zeros(reps + 2, 2 * Co + 32, device=dev)
        dconv = torch.randn(P, Co, device=dev).bfloat16()
        dx = torch.empty(P, max(Ci, 64), dtype=torch.bfloat16, device=dev)
        scratch = torch.empty(a.max_splits * ((P + 127) // 128) * 128 * max(Co, Ci, 64), device=dev)

        def fwd(i):
            K.conv_bn_relu_pool_fwd(x, w, bias, gamma, beta, N, H, W, Cin_eff, Co, taps, pool, conv_out=conv_out,
                                    act=act, stats=stats[i], saved=saved, scratch=scratch)

        def dgrad(i):
            K.conv_dgrad(dconv, w, N, H, W, Ci, Co, out=dx, scratch=scratch, split_sync=None if a.dgrad_two_kernels else True)

        for what, fn, shape in (("fwd", fwd, (Cin_eff, Co, taps)), ("dgrad", dgrad, (Co, Ci, 9))):
            if what == "dgrad" and li == 0:
                continue
            K.conv_force(-1, -1, -1)
            auto = K.conv_plan(N, H, W, *shape)
            zero = stats if what == "fwd" else None
            t_auto = timed(fn, reps, zero=zero)
            table = {}
            for bn in (64, 128, 256):
                if shape[1] % bn:
                    continue
                for s in range(1, a.max_splits + 1):
                    K.conv_force(-1, bn, s)
                    got = K.conv_plan(N, H, W, *shape)
                    if got[1] != bn or got[2] != s:
                        continue            # not a legal plan for this shape
                    try:
                        table[(bn, s)] = timed(fn, reps, zero=zero)
                    except RuntimeError as e:      # a forced plan the launcher refuses
                        table[(bn, s)] = float("nan")
                        print("   ", what, bn, s, str(e)[:100])
            K.conv_force(-1, -1, -1)
            ok = {k: v for k, v in table.items() if v == v}
            best = min(ok, key=ok.get)
            gain[what] += t_auto - min(ok[best], t_auto)
            print(f"L{li} {what:5s} {S:2d}x{S:<2d} {shape[0]:3d}->{shape[1]:3d} | planner bn {auto[1]:3d} x{auto[2]} {t_auto:6.1f} us | "
                  f"best bn {best[0]:3d} x{best[1]} {ok[best]:6.1f} us")
            for bn in (64, 128, 256):
                row = [f"x{s} {table[(bn, s)]:5.1f}" for s in range(1, a.max_splits + 1) if (bn, s) in table]
                if row:
                    print(f"      bn {bn:3d}: " + "  ".join(row))
    print("possible gain over the planner (us):", {k: round(v, 1) for k, v in gain.items()})


if __name__ == "__main__":
    main()
