"""Per-CTA phase timing of the conv kernels (uses the ddpb_conv_debug_timing hook)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.ops import _native as nat  # noqa: E402
from ddpb200.ops import kernels as K  # noqa: E402

LAYERS = {0: (32, 64, 64, 1), 1: (16, 64, 128, 9), 3: (8, 256, 256, 9), 5: (4, 512, 512, 9), 7: (2, 512, 512, 9)}
N = 256
dev = "cuda"
lib = nat.load()
cfgs = [tuple(int(v) for v in c.split(",")) for c in ([a for a in sys.argv[1:] if not a.startswith("--")] or ["0,128,1", "2,128,1"])]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for li, (S, Ci, Co, taps) in LAYERS.items():
    P = N * S * S
    x = torch.randn(P, Ci, device=dev).bfloat16()
    w = torch.randn(Co, taps * Ci, device=dev).bfloat16()
    y = torch.empty(P, Co, dtype=torch.bfloat16, device=dev)
    stats = torch.zeros(2, Co, device=dev)
    bias = torch.zeros(Co, device=dev)
    scratch = torch.empty(8 * P * Co, device=dev)
    for pair, bn, sp in cfgs:
        if bn == 0:
            bn = K.conv_plan(N, S, S, Ci, Co, taps)[1]
        if Co % bn:
            continue
        m_tiles = (P + 127) // 128
        if pair:
            m_tiles = (m_tiles + 1) // 2 * 2
        ctas = m_tiles * (Co // bn) * sp
        if pair == 2:
            ctas = min(ctas, torch.cuda.get_device_properties(0).multi_processor_count)
        buf = torch.zeros(ctas * 4, dtype=torch.int64, device=dev)
        K.conv_force(pair, bn, sp)
        try:
            for it in range(3):
                flush.zero_() if "--cold" in sys.argv else None
                lib.ddpb_conv_debug_timing(buf.data_ptr())
                K.conv_fprop(x, w, None if "--nobias" in sys.argv else bias, N, S, S, Ci, Co, taps, out=y,
                             stats=None if "--nostats" in sys.argv else stats, scratch=scratch)
                lib.ddpb_conv_debug_timing(None)
                torch.cuda.synchronize()
        finally:
            K.conv_force(-1, -1, -1)
            lib.ddpb_conv_debug_timing(None)
        t = buf.view(ctas, 4).double().cpu()
        t0 = t[:, 0].min()
        setup = (t[:, 1] - t[:, 0])
        main = (t[:, 2] - t[:, 1]) if pair != 2 else (t[:, 3] - t[:, 1])
        epi = (t[:, 3] - t[:, 2]) if pair != 2 else torch.zeros_like(main)
        total = t[:, 3].max() - t0
        starts = t[:, 0] - t0
        print(f"L{li} pair={pair} bn={bn} sp={sp} ctas={ctas}: kernel {total / 1e3:6.1f} us | set-up {setup.mean() / 1e3:5.2f} "
              f"main {main.mean() / 1e3:6.2f} (max {main.max() / 1e3:6.2f}) epi {epi.mean() / 1e3:5.2f} us | "
              f"start spread: median {starts.median() / 1e3:5.1f} max {starts.max() / 1e3:5.1f} us")
