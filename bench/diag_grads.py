"""Per-parameter agreement of the kernel pipeline's gradients with the fp32 module oracle."""
import copy
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import ddpb200
from ddpb200.ops.reference import augment_normalize_u8

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
torch.manual_seed(89395)
ref = ddpb200.VGG11()
mine = copy.deepcopy(ref).to("cuda")
g = torch.Generator().manual_seed(0)
img = torch.randint(0, 256, (n, 32, 32, 3), dtype=torch.uint8, generator=g)
lab = torch.randint(0, 10, (n,), generator=g)
x = augment_normalize_u8(img, None)
out = ref(x)
loss = torch.nn.functional.cross_entropy(out, lab)
loss.backward()
out2 = mine(x.cuda())
loss2 = torch.nn.functional.cross_entropy(out2, lab.cuda())
loss2.backward()
print("loss", loss.item(), loss2.item(), "logits rel", ((out2.cpu() - out).norm() / out.norm()).item())
rp = dict(ref.named_parameters())
for name, p in mine.named_parameters():
    a, b = p.grad.double().cpu().flatten(), rp[name].grad.double().flatten()
    cos = float(a @ b / (a.norm() * b.norm() + 1e-30))
    print(f"{name:20s} n={a.numel():8d} cos={cos:+.5f} |g|={a.norm():.3e} |ref|={b.norm():.3e} max|g|={a.abs().max():.3e}")
