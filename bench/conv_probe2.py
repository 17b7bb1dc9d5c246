"""Phase timeline of the persistent conv kernels (modes 2 and 3), with the fused BatchNorm tail.

Per CTA the kernels stamp %globaltimer at: 0 start, 1 set-up done, 2 last accumulator complete (MMA
thread), 4 epilogue loop done, 5 past the grid barrier, 3 exit.  Printed: median / max over CTAs,
microseconds since the earliest start.  ``--probe`` also runs the producer-skip probes
(ddpb_conv_debug_mode 1 = no B loads, 2 = no A loads, 3 = neither): numerically wrong, timing only.

    python bench/conv_probe2.py [--probe] [--dgrad] [--tail-l2] [mode,block_n ...]     e.g. 2,0 3,256 3,128
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.ops import _native as nat  # noqa: E402
from ddpb200.ops import kernels as K  # noqa: E402

LAYERS = [(32, 64, 64, 1, True), (16, 64, 128, 9, True), (8, 128, 256, 9, False), (8, 256, 256, 9, True),
          (4, 256, 512, 9, False), (4, 512, 512, 9, True), (2, 512, 512, 9, False), (2, 512, 512, 9, True)]
N = 256
dev = "cuda"
lib = nat.load()
args = [a for a in sys.argv[1:] if not a.startswith("--")]
cfgs = [tuple(int(v) for v in c.split(",")) for c in (args or ["2,0", "3,256", "3,128"])]
probes = [0, 1, 2, 3] if "--probe" in sys.argv else [0]
dgrad = "--dgrad" in sys.argv
K.conv_tail_from_tmem("--tail-l2" not in sys.argv)     # fused BatchNorm tail from TMEM (default) or from the stored tiles
sms = torch.cuda.get_device_properties(0).multi_processor_count
for li, (S, Ci, Co, taps, pool) in enumerate(LAYERS):
    P = N * S * S
    if dgrad and li == 0:
        continue
    x = torch.randn(P, Ci, device=dev).bfloat16()
    w = (torch.randn(Co, taps * Ci, device=dev) / (taps * Ci) ** 0.5).bfloat16()
    bias, gamma, beta = torch.zeros(Co, device=dev), torch.ones(Co, device=dev), torch.zeros(Co, device=dev)
    dy = torch.randn(P, Co, device=dev).bfloat16()
    for mode, bn in cfgs:
        K.conv_force(mode, bn if bn else -1, -1)
        try:
            plan = K.conv_plan(N, S, S, Co, Ci, 9) if dgrad else K.conv_plan(N, S, S, Ci, Co, taps)
            if mode == 3 and plan[0] != 3:
                continue
            for probe in probes:
                buf = torch.zeros(sms * 8, dtype=torch.int64, device=dev)
                for it in range(3):
                    buf.zero_()
                    lib.ddpb_conv_debug_mode(probe)
                    lib.ddpb_conv_debug_timing(buf.data_ptr())
                    if dgrad:
                        K.conv_dgrad(dy, w, N, S, S, Ci, Co)
                    else:
                        K.conv_bn_relu_pool_fwd(x, w, bias, gamma, beta, N, S, S, Ci, Co, taps, pool)
                    lib.ddpb_conv_debug_timing(None)
                    lib.ddpb_conv_debug_mode(0)
                    torch.cuda.synchronize()
                t = buf.view(sms, 8).double().cpu()
                t = t[t[:, 0] > 0]
                t0 = t[:, 0].min()

                def col(i):
                    v = t[:, i]
                    v = v[v > 0]
                    return (float((v - t0).median()) / 1e3, float((v - t0).max()) / 1e3) if len(v) else (0.0, 0.0)

                names = [("setup", 1), ("mma done", 2), ("epi done", 4), ("barrier", 5), ("exit", 3)]
                txt = " | ".join(f"{n} {col(i)[0]:5.1f}/{col(i)[1]:5.1f}" for n, i in names)
                print(f"L{li} {'dgrad' if dgrad else 'fwd'} plan={plan} probe={probe} ctas={len(t)}: {txt}")
        finally:
            K.conv_force(-1, -1, -1)
            lib.ddpb_conv_debug_timing(None)
            lib.ddpb_conv_debug_mode(0)
