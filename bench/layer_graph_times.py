"""In-graph steady-state time of every per-layer kernel group of the VGG-11 step.

Each op is captured `reps` times back to back in one CUDA graph and the graph is replayed, so the
number includes the kernel-to-kernel launch gap of a captured step but no host overhead.  Buffers
that must be zero on entry (statistics, grid-barrier words) get one slot per repetition and are
cleared before every replay, outside the timed region.

    python bench/layer_graph_times.py [--batch 256] [--bn-fused 2]
"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.ops import kernels as K  # noqa: E402

LAYERS = [(32, 3, 64, True), (16, 64, 128, True), (8, 128, 256, False), (8, 256, 256, True),
          (4, 256, 512, False), (4, 512, 512, True), (2, 512, 512, False), (2, 512, 512, True)]
dev = "cuda"


def timed(fn, reps, zero=None, replays=3):
    for i in range(2):
        if zero is not None:
            zero.zero_()
        fn(i)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    if zero is not None:
        zero.zero_()
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for i in range(reps):
                fn(i)
    best = 1e30
    for _ in range(replays):
        if zero is not None:
            zero.zero_()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e) / reps * 1e3)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--bn-fused", type=int, default=3)
    ap.add_argument("--conv-mode", type=int, default=-1,
                    help="conv kernel: -1 automatic, 0 one tile per CTA, 1 CTA pair, 2 persistent, 3 persistent CTA pair")
    ap.add_argument("--block-n", type=int, default=-1, help="force the tile width of fprop / dgrad")
    ap.add_argument("--tail-l2", action="store_true", help="fused BatchNorm tail from the stored tiles, not from TMEM")
    a = ap.parse_args()
    K.conv_force(a.conv_mode, a.block_n, -1)
    K.conv_tail_from_tmem(not a.tail_l2)
    N, reps = a.batch, a.reps
    K.bn_bwd_set_fused(a.bn_fused)
    tot = dict(fwd=0.0, bn_bwd=0.0, wgrad=0.0, dgrad=0.0)
    print(f"batch {N}, pdl {K.get_pdl()}, bn fused mode {a.bn_fused}, tail from {'L2' if a.tail_l2 else 'TMEM'}, conv mode {a.conv_mode}, block_n {a.block_n}; "
          "microseconds per call in a graph")
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
        fs = torch.empty(max(1, K.conv_fprop_scratch_elems(N, H, W, Cin_eff, Co, taps)), device=dev)
        t_f = timed(lambda i: K.conv_bn_relu_pool_fwd(x, w, bias, gamma, beta, N, H, W, Cin_eff, Co, taps, pool,
                                                      conv_out=conv_out, act=act, stats=stats[i], saved=saved,
                                                      scratch=fs), reps, zero=stats)
        dact = torch.randn(Po, Co, device=dev).bfloat16()
        dconv = torch.empty(P, Co, dtype=torch.bfloat16, device=dev)
        dg, db_, dbias = torch.empty(Co, device=dev), torch.empty(Co, device=dev), torch.empty(Co, device=dev)
        part = torch.empty(max(1, K.bn_bwd_partials_elems(N, H, W, Co, pool)), device=dev)
        t_b = timed(lambda i: K.bn_relu_pool_bwd(conv_out, dact, saved, gamma, beta, N, H, W, Co, pool, sums=stats[i],
                                                 dx=dconv, dgamma=dg, dbeta=db_, dbias=dbias, partials=part), reps,
                    zero=stats)
        cols = 27 if li == 0 else 9 * Ci
        dw = torch.empty(Co, cols, device=dev)
        ws = torch.empty(max(1, K.conv_wgrad_scratch_elems(N, H, W, Cin_eff, Co, taps)), device=dev)
        t_w = timed(lambda i: K.conv_wgrad(dconv, x, N, H, W, Cin_eff, Co, taps, out=dw.view(-1), cols_valid=cols,
                                           out_ld=cols, scratch=ws), reps)
        t_d = 0.0
        if li > 0:
            dx = torch.empty(P, Ci, dtype=torch.bfloat16, device=dev)
            ds_ = torch.empty(max(1, K.conv_fprop_scratch_elems(N, H, W, Co, Ci, 9)), device=dev)
            t_d = timed(lambda i: K.conv_dgrad(dconv, w, N, H, W, Ci, Co, out=dx, scratch=ds_, split_sync=None), reps)
        print(f"L{li} {S:2d}x{S:<2d} {Ci:3d}->{Co:3d} pool={int(pool)} | fwd(conv+bn) {t_f:6.1f} | bn_bwd {t_b:6.1f} | "
              f"wgrad {t_w:6.1f} (splits {K.conv_wgrad_splits(N, H, W, Cin_eff, Co, taps)}) | dgrad {t_d:6.1f}")
        for k, v in zip(("fwd", "bn_bwd", "wgrad", "dgrad"), (t_f, t_b, t_w, t_d)):
            tot[k] += v
    print("totals:", {k: round(v, 1) for k, v in tot.items()}, "sum", round(sum(tot.values()), 1))


if __name__ == "__main__":
    main()
