"""Back-to-back timing of the small critical-path kernels (classifier, input, SGD, finalize)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from ddpb200.ops import kernels as K  # noqa: E402

dev = "cuda"
B = 256


def timed(fn, reps=40):
    """Per-call GPU time: `reps` calls captured in one CUDA graph (no host gaps), graph replayed."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps):
                fn()
    g.replay()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        g.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / (5 * reps) * 1e3


x = torch.randn(B, 512, device=dev).bfloat16()
w = torch.randn(10, 512, device=dev)
b = torch.randn(10, device=dev)
lab = torch.randint(0, 10, (B,), device=dev)
logits = torch.empty(B, 10, device=dev)
dl = torch.empty(B, 10, device=dev)
rl = torch.empty(B, device=dev)
rc = torch.empty(B, dtype=torch.int32, device=dev)
loss = torch.zeros((), device=dev)
cor = torch.zeros((), dtype=torch.int32, device=dev)
dw, db, dx = torch.empty(10, 512, device=dev), torch.empty(10, device=dev), torch.empty(B, 512, dtype=torch.bfloat16, device=dev)
print("fc_ce_fwd (+loss reduce)", timed(lambda: K.fc_ce_fwd(x, w, b, lab, logits=logits, row_loss=rl, dlogits=dl, row_correct=rc, loss=loss, correct=cor)))
print("fc_ce_fwd only          ", timed(lambda: K.fc_ce_fwd(x, w, b, lab, logits=logits, row_loss=rl, dlogits=dl, row_correct=rc)))
print("fc_bwd                  ", timed(lambda: K.fc_bwd(x, w, dl, dw=dw, db=db, dx=dx)))
raw = torch.randint(0, 256, (B, 32, 32, 3), dtype=torch.uint8, device=dev)
aug = torch.zeros(B, 4, dtype=torch.int8, device=dev)
op = torch.empty(B * 1024, 64, dtype=torch.bfloat16, device=dev)
print("first_layer_operand_u8  ", timed(lambda: K.first_layer_operand_u8(raw, aug, out=op)))
n = 9231360
p, g, m = torch.randn(n, device=dev), torch.randn(n, device=dev), torch.zeros(n, device=dev)
sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
print("sgd_step (+shadow)      ", timed(lambda: K.sgd_step(p, g, m, 0.1, 0.9, 1e-4, shadow=sh)))
z = torch.zeros(9000, device=dev)
print("memset 36 KB            ", timed(lambda: z.zero_()))
