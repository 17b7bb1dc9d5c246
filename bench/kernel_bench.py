"""Per-kernel timings over the VGG-11 layer shapes (S4 in SURVEY.md §8).

Times every conv fprop / dgrad / wgrad and the fused BN blocks at the given per-GPU batch
with CUDA events (warm-up, L2 flush between iterations) and prints achieved TFLOP/s and
GB/s against MEASURED_PEAKS.json.  Writes a JSON summary to --out.

    python bench/kernel_bench.py --batch 256 --out gpurun_out/kernel_bench.json
"""
import argparse
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from ddpb200.ops import kernels as K  # noqa: E402

# (H=W, C_in, C_out, pool) per conv block of VGG-11 (/root/reference/part1/model.py:4)
LAYERS = [(32, 3, 64, True), (16, 64, 128, True), (8, 128, 256, False), (8, 256, 256, True),
          (4, 256, 512, False), (4, 512, 512, True), (2, 512, 512, False), (2, 512, 512, True)]


def timeit(fn, iters=20, warmup=5, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-flush", action="store_true")
    a = ap.parse_args()
    dev = "cuda"
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    peak_tf = peaks.get("bf16_tflops", 1590.0)
    flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    N = a.batch
    rows = []
    tot = {"fprop": 0.0, "dgrad": 0.0, "wgrad": 0.0, "bn_fwd": 0.0, "bn_bwd": 0.0}
    for li, (S, Ci, Co, pool) in enumerate(LAYERS):
        H = W = S
        P = N * H * W
        Cin_eff, taps = (64, 1) if li == 0 else (Ci, 9)
        x = torch.randn(P, Cin_eff, device=dev).bfloat16()
        w = torch.randn(Co, taps * Cin_eff, device=dev).bfloat16()
        bias = torch.randn(Co, device=dev)
        stats = torch.zeros(2, Co, device=dev)
        y = torch.empty(P, Co, dtype=torch.bfloat16, device=dev)
        flops = 2.0 * P * Co * taps * Cin_eff
        fs = torch.empty(max(1, K.conv_fprop_scratch_elems(N, H, W, Cin_eff, Co, taps)), device=dev)
        t = timeit(lambda: K.conv_fprop(x, w, bias, N, H, W, Cin_eff, Co, taps, out=y, stats=stats, scratch=fs),
                   flush=flush)
        rows.append(dict(layer=li, op="fprop", us=t, tflops=flops / t / 1e6))
        tot["fprop"] += t
        dy = torch.randn(P, Co, device=dev).bfloat16()
        if li > 0:
            dx = torch.empty(P, Ci, dtype=torch.bfloat16, device=dev)
            ds_ = torch.empty(max(1, K.conv_fprop_scratch_elems(N, H, W, Co, Ci, 9)), device=dev)
            t = timeit(lambda: K.conv_dgrad(dy, w, N, H, W, Ci, Co, out=dx, scratch=ds_), flush=flush)
            rows.append(dict(layer=li, op="dgrad", us=t, tflops=flops / t / 1e6))
            tot["dgrad"] += t
        cols = 27 if li == 0 else taps * Cin_eff
        dw = torch.empty(Co, cols, dtype=torch.float32, device=dev)
        need = K.conv_wgrad_scratch_elems(N, H, W, Cin_eff, Co, taps)
        scratch = torch.empty(max(need, 1), dtype=torch.float32, device=dev)
        t = timeit(lambda: K.conv_wgrad(dy, x, N, H, W, Cin_eff, Co, taps, out=dw, cols_valid=cols, out_ld=cols,
                                        scratch=scratch), flush=flush)
        rows.append(dict(layer=li, op="wgrad", us=t, tflops=flops / t / 1e6,
                         splits=K.conv_wgrad_splits(N, H, W, Cin_eff, Co, taps)))
        tot["wgrad"] += t
        gamma, beta = torch.ones(Co, device=dev), torch.zeros(Co, device=dev)
        st = K.channel_stats(y, Co)
        Po = P // 4 if pool else P
        act = torch.empty(Po, Co, dtype=torch.bfloat16, device=dev)
        saved = torch.empty(2, Co, device=dev)
        t = timeit(lambda: K.bn_relu_pool_fwd(y, st, gamma, beta, N, H, W, Co, pool, out=act, saved=saved),
                   flush=flush)
        gb = (P + Po) * Co * 2 / 1e9
        rows.append(dict(layer=li, op="bn_fwd", us=t, gbs=gb / (t * 1e-6)))
        tot["bn_fwd"] += t
        dact = torch.randn(Po, Co, device=dev).bfloat16()
        sums = torch.zeros(2 * Co + 32, device=dev)
        dxb = torch.empty(P, Co, dtype=torch.bfloat16, device=dev)
        dg, db_, dbias = torch.empty(Co, device=dev), torch.empty(Co, device=dev), torch.empty(Co, device=dev)

        def bwd():
            sums.zero_()
            K.bn_relu_pool_bwd(y, dact, saved, gamma, beta, N, H, W, Co, pool, sums=sums, dx=dxb, dgamma=dg,
                               dbeta=db_, dbias=dbias)

        t = timeit(bwd, flush=flush)
        gb = (2 * P + 2 * Po + P) * Co * 2 / 1e9
        rows.append(dict(layer=li, op="bn_bwd", us=t, gbs=gb / (t * 1e-6)))
        tot["bn_bwd"] += t
    for r in rows:
        extra = f"{r['tflops']:8.1f} TFLOP/s ({100 * r['tflops'] / peak_tf:5.1f}% of measured cuBLAS bf16)" \
            if "tflops" in r else f"{r['gbs']:8.1f} GB/s"
        print(f"L{r['layer']} {r['op']:7s} {r['us']:9.1f} us  {extra}" + (f"  splits={r['splits']}" if 'splits' in r else ''))
    print("totals (us):", {k: round(v, 1) for k, v in tot.items()}, "sum", round(sum(tot.values()), 1))
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps(dict(batch=N, rows=rows, totals=tot, peak_tflops=peak_tf), indent=1))


if __name__ == "__main__":
    main()
