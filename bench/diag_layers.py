"""Layer-by-layer agreement of engine buffers with a precision-matched ATen emulation."""
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import ddpb200
from ddpb200.ops.reference import augment_normalize_u8, nchw_to_nhwc2d

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
n = 64
torch.manual_seed(11)
model = ddpb200.VGG11().cuda()
g = torch.Generator().manual_seed(3)
img = torch.randint(0, 256, (n, 32, 32, 3), dtype=torch.uint8, generator=g)
x = augment_normalize_u8(img, None).cuda()
with torch.no_grad():
    out = model(x)
eng = model._engine
y = x.bfloat16().float()


def rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-30)).item()


with torch.no_grad():
    for li, blk in enumerate(model.conv_blocks()):
        L = eng.layers[li]
        P = n * L.H * L.W
        w = blk.conv.weight.bfloat16().float()
        c = F.conv2d(y, w, blk.conv.bias, padding=1)
        c_r = c.bfloat16().float()
        mine_c = eng.conv_out[li][:P].float()
        ref_c = nchw_to_nhwc2d(c_r)
        mean = ref_c.mean(0)
        var = ref_c.var(0, unbiased=False)
        s = eng.stats[li].view(2, -1)
        m_mean = s[0] / P
        m_var = s[1] / P - m_mean * m_mean
        z = F.relu(F.batch_norm(c_r, None, None, blk.bn.weight, blk.bn.bias, training=True, eps=1e-5))
        if blk.pool:
            z = F.max_pool2d(z, 2, 2)
        z = z.bfloat16().float()
        Po = P // 4 if blk.pool else P
        print(f"L{li}: conv rel {rel(mine_c, ref_c):.2e}  mean abs err {(m_mean - mean).abs().max():.2e} "
              f"var rel err {((m_var - var).abs() / var).max():.2e}  act rel {rel(eng.act[li][:Po], nchw_to_nhwc2d(z)):.2e} "
              f"saved rstd rel {((eng.saved[li].view(2, -1)[1] - (var + 1e-5).rsqrt()).abs() * (var + 1e-5).sqrt()).max():.2e}")
        # continue from MY activations so errors do not accumulate across layers
        y = eng.act[li][:Po].float().view(n, L.H // (2 if blk.pool else 1), L.W // (2 if blk.pool else 1), -1).permute(0, 3, 1, 2).contiguous()
