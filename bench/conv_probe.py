"""Times conv fprop on a few VGG layers under forced launch configurations (tuning aid)."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.ops import kernels as K  # noqa: E402

LAYERS = {1: (16, 64, 128), 2: (8, 128, 256), 3: (8, 256, 256), 4: (4, 256, 512), 5: (4, 512, 512), 7: (2, 512, 512)}
WITH_EPI = "--epi" in sys.argv   # bias + statistics epilogue, like the training step
N = 256
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
cfgs = [tuple(int(v) for v in c.split(",")) for c in ([a for a in sys.argv[1:] if not a.startswith("--")] or ["0,128,1"])]
COLD = "--warm" not in sys.argv
print("lib", os.environ.get("DDPB_LIB", "default"))
for li, (S, Ci, Co) in LAYERS.items():
    P = N * S * S
    x = torch.randn(P, Ci, device=dev).bfloat16()
    w = torch.randn(Co, 9 * Ci, device=dev).bfloat16()
    y = torch.empty(P, Co, dtype=torch.bfloat16, device=dev)
    scratch = torch.empty(8 * P * Co, device=dev)
    bias = torch.zeros(Co, device=dev)
    stats = torch.zeros(2, Co, device=dev)
    for pair, bn, sp in cfgs:
        if Co % bn:
            continue
        K.conv_force(pair, bn, sp)
        try:
            ts = []
            for it in range(12):
                if COLD:
                    flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                K.conv_fprop(x, w, bias if WITH_EPI else None, N, S, S, Ci, Co, out=y, scratch=scratch,
                             stats=stats if WITH_EPI else None)
                e.record()
                torch.cuda.synchronize()
                ts.append(s.elapsed_time(e) * 1e3)
            t = sorted(ts[2:])[len(ts[2:]) // 2]
            print(f"L{li} pair={pair} bn={bn} splits={sp}: {t:7.1f} us  {2.0 * P * Co * 9 * Ci / t / 1e6:7.1f} TFLOP/s")
        finally:
            K.conv_force(-1, -1, -1)
