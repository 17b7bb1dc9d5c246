"""Kernels that are in the library but not on the default path.  They only run when
DDPB_TEST_EXPERIMENTAL=1: a pipeline bug in a tcgen05 kernel ends in the 2 s watchdog trap, which
kills the CUDA context of the whole pytest process, so these stay out of the default GPU suite
until they have passed on hardware (status in DESIGN.md par. 7)."""
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("DDPB_TEST_EXPERIMENTAL") != "1",
                                 reason="set DDPB_TEST_EXPERIMENTAL=1 to run the experimental kernels")]

DEV = "cuda"


def _rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-30))


# (N, H, W, C_in, C_out), BLOCK_N -- one pair tile per cluster, an odd number of 128-pixel tiles,
# several tiles per cluster (accumulator ring wraps: 4 buffers at 128 columns, 2 at 256)
PAIR_PERSIST_CASES = [((8, 8, 8, 128, 256), 128), ((8, 8, 8, 128, 256), 256), ((3, 8, 8, 128, 128), 128),
                      ((512, 16, 16, 64, 128), 128), ((1024, 8, 8, 128, 256), 256), ((256, 4, 4, 256, 512), 256)]


@pytest.mark.parametrize("case", PAIR_PERSIST_CASES)
def test_persistent_cta_pair_conv(case):
    """ddpb_conv_force(3, ...): persistent cta_group::2 kernel, fprop (bias + statistics) and dgrad."""
    from ddpb200.ops import kernels as K
    from ddpb200.ops import reference as R

    (N, H, W, Ci, Co), bn = case
    torch.manual_seed(0)
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    b = torch.randn(Co, device=DEV)
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    try:
        K.conv_force(3, bn, 1)
        assert tuple(K.conv_plan(N, H, W, Ci, Co, 9))[:2] == (3, bn)
        stats = torch.zeros(2, Co, device=DEV)
        y = K.conv_fprop(x, w, b, N, H, W, Ci, Co, stats=stats)
        dx = K.conv_dgrad(dy, w, N, H, W, Ci, Co) if Ci % bn == 0 else None
        torch.cuda.synchronize()
    finally:
        K.conv_force(-1, -1, -1)
    assert _rel_err(y, R.conv_fprop(x, w, b, N, H, W, Ci, Co)) < 1e-2
    assert torch.allclose(stats, R.channel_stats(y), rtol=2e-3, atol=1e-2 * (N * H * W) ** 0.5)
    if dx is not None:
        assert _rel_err(dx, R.conv_dgrad(dy, w, N, H, W, Ci, Co)) < 1e-2


@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("case", [((16, 16, 16, 64, 128), 128), ((37, 8, 8, 128, 256), 256), ((256, 8, 8, 256, 256), 256),
                                  ((64, 4, 4, 256, 512), 128)])
def test_persistent_cta_pair_conv_with_batchnorm_tail(case, pool):
    """Mode 3 with the BatchNorm + ReLU (+pool) tail after the grid barrier (not yet run on hardware
    at the end of round 1)."""
    from ddpb200.ops import kernels as K
    from ddpb200.ops import reference as R

    (N, H, W, Ci, Co), bn = case
    torch.manual_seed(1)
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    b = torch.randn(Co, device=DEV) * 0.1
    gamma = torch.rand(Co, device=DEV) + 0.5
    beta = torch.randn(Co, device=DEV) * 0.2
    try:
        K.conv_force(3, bn, 1)
        conv_out, act, saved = K.conv_bn_relu_pool_fwd(x, w, b, gamma, beta, N, H, W, Ci, Co, 9, pool)
        torch.cuda.synchronize()
    finally:
        K.conv_force(-1, -1, -1)
    assert _rel_err(conv_out, R.conv_fprop(x, w, b, N, H, W, Ci, Co)) < 1e-2
    assert _rel_err(act, R.bn_relu_pool_fwd(conv_out, gamma, beta, N, H, W, pool)) < 1e-2
    cf = conv_out.float()
    assert torch.allclose(saved[0], cf.mean(0), rtol=1e-3, atol=1e-3)
    assert torch.allclose(saved[1], (cf.var(0, unbiased=False) + 1e-5).rsqrt(), rtol=2e-3)


@pytest.mark.parametrize("shape", [(256, 8, 8, 128, 256), (256, 8, 8, 256, 256), (256, 4, 4, 256, 512),
                                   (256, 4, 4, 512, 512), (256, 2, 2, 512, 512), (3, 8, 8, 128, 256)])
def test_cta_pair_wgrad(shape):
    """ddpb_conv_wgrad_set_mode(2): weight gradient with cta_group::2 (not yet run on hardware at the
    end of round 1); deterministic, same tolerance as the default kernel."""
    from ddpb200.ops import kernels as K
    from ddpb200.ops import reference as R

    N, H, W, Ci, Co = shape
    torch.manual_seed(2)
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    ref = R.conv_wgrad(dy, x, N, H, W, Ci, Co)
    try:
        K.conv_wgrad_set_mode(2)
        dw = K.conv_wgrad(dy, x, N, H, W, Ci, Co)
        again = K.conv_wgrad(dy, x, N, H, W, Ci, Co)
        torch.cuda.synchronize()
    finally:
        K.conv_wgrad_set_mode(0)
    assert _rel_err(dw, ref) < 2e-3
    assert torch.equal(dw, again)

