"""Numerics of every sm_100a kernel against a plain PyTorch fp32 restatement of the same op
(ddpb200.ops.reference, built from the modules of /root/reference/part1/model.py).

Inputs are rounded to bf16 first, so both sides see identical operands; the remaining
difference is accumulation order and the bf16 rounding of outputs.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

from ddpb200.ops import kernels as K
from ddpb200.ops import reference as R

DEV = "cuda"

# (N, H, W, C_in, C_out): every spatial size VGG-11 sees on CIFAR plus partial tiles
CONV_SHAPES = [
    (4, 32, 32, 64, 64),
    (4, 16, 16, 64, 128),
    (8, 8, 8, 128, 256),
    (3, 8, 8, 128, 128),      # N not a multiple of the box's image count
    (32, 4, 4, 256, 512),
    (64, 2, 2, 512, 512),
    (37, 2, 2, 512, 512),     # ragged last tile
]


def _rel_err(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12)).item()


@pytest.fixture(autouse=True)
def _no_tf32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    yield


@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_fprop_bias_stats(shape):
    N, H, W, Ci, Co = shape
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    b = torch.randn(Co, device=DEV)
    stats = torch.zeros(2, Co, device=DEV)
    y = K.conv_fprop(x, w, b, N, H, W, Ci, Co, stats=stats)
    torch.cuda.synchronize()
    ref = R.conv_fprop(x, w, b, N, H, W, Ci, Co)
    assert _rel_err(y, ref) < 1e-2
    ref_stats = R.channel_stats(y)   # statistics are of the stored (rounded) tensor
    assert torch.allclose(stats, ref_stats, rtol=2e-3, atol=1e-2 * (N * H * W) ** 0.5)


@pytest.mark.parametrize("block_n", [64, 128, 256])
def test_conv_fprop_block_n(block_n):
    N, H, W, Ci, Co = 8, 8, 8, 128, 256
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    y = K.conv_fprop(x, w, None, N, H, W, Ci, Co, block_n=block_n)
    assert _rel_err(y, R.conv_fprop(x, w, None, N, H, W, Ci, Co)) < 1e-2


def test_conv_fprop_single_tap_first_layer():
    N, H, W = 8, 32, 32
    img = torch.randn(N, 3, H, W, device=DEV)
    op = K.first_layer_operand_f32(img)
    ref_op = R.first_layer_operand(img).bfloat16()
    assert torch.equal(op, ref_op)
    w27 = torch.randn(64, 27, device=DEV) / 27 ** 0.5
    w = torch.zeros(64, 64, device=DEV)
    w[:, :27] = w27
    w = w.bfloat16()
    y = K.conv_fprop(op, w, None, N, H, W, 64, 64, taps=1)
    # oracle: the actual 3-channel convolution
    w_oihw = w[:, :27].float().reshape(64, 3, 3, 3).permute(0, 3, 1, 2)
    ref = torch.nn.functional.conv2d(img.bfloat16().float(), w_oihw, padding=1)
    assert _rel_err(y, R.nchw_to_nhwc2d(ref)) < 1e-2


@pytest.mark.parametrize("shape", CONV_SHAPES[1:])
def test_conv_dgrad_via_rotated_filters(shape):
    N, H, W, Ci, Co = shape
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Co) ** 0.5).bfloat16()
    wd = R.rotate_filters_for_dgrad(w, Co, Ci)
    dx = K.conv_fprop(dy, wd, None, N, H, W, Co, Ci)
    assert _rel_err(dx, R.conv_dgrad(dy, w, N, H, W, Ci, Co)) < 1e-2


# every legal tile / split combination, fprop and dgrad
BASE_CASES = [((8, 8, 8, 128, 256), 128, 1), ((8, 8, 8, 128, 256), 256, 1), ((8, 8, 8, 256, 256), 256, 2),
              ((3, 8, 8, 128, 128), 128, 1), ((4, 16, 16, 64, 128), 128, 1), ((32, 4, 4, 256, 512), 256, 3),
              ((37, 2, 2, 512, 512), 128, 4), ((2, 32, 32, 64, 128), 128, 1)]


PERSIST_CASES = BASE_CASES + [((4, 32, 32, 64, 64), 64, 1), ((64, 8, 8, 128, 128), 128, 1),
                              ((300, 4, 4, 256, 256), 128, 2), ((8, 32, 32, 64, 128), 64, 1)]

# persistent CTA-pair kernel (cta_group::2; no split-K, tiles of 128 / 256 columns): one pair tile per cluster, an
# odd number of 128-pixel tiles, several tiles per cluster (accumulator ring wraps: 4 buffers at 128 columns, 2 at 256)
PAIR_PERSIST_CASES = [((8, 8, 8, 128, 256), 128, 1), ((8, 8, 8, 128, 256), 256, 1), ((3, 8, 8, 128, 128), 128, 1),
                      ((512, 16, 16, 64, 128), 128, 1), ((1024, 8, 8, 128, 256), 256, 1), ((256, 4, 4, 256, 512), 256, 1)]


@pytest.mark.parametrize("case", [(0, c) for c in BASE_CASES] + [(2, c) for c in PERSIST_CASES] +
                         [(3, c) for c in PAIR_PERSIST_CASES])
def test_conv_kernel_variants_fprop_and_dgrad(case):
    """mode 0 = one tile per CTA, mode 2 = persistent kernel (the default), mode 3 = persistent CTA pair
    (cta_group::2, opt-in: measured no faster, DESIGN.md par. 7); all against the oracle."""
    mode, ((N, H, W, Ci, Co), bn, splits) = case
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    b = torch.randn(Co, device=DEV)
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    try:
        K.conv_force(mode, bn, splits)
        assert tuple(K.conv_plan(N, H, W, Ci, Co, 9))[:2] == (mode, bn)
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


@pytest.mark.parametrize("shape", CONV_SHAPES[1:])
def test_conv_dgrad_from_fprop_pack(shape):
    """dgrad reads the very same [C_out][9*C_in] pack as fprop (MN-major B operand, taps reversed)."""
    N, H, W, Ci, Co = shape
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Co) ** 0.5).bfloat16()
    dx = K.conv_dgrad(dy, w, N, H, W, Ci, Co)
    assert _rel_err(dx, R.conv_dgrad(dy, w, N, H, W, Ci, Co)) < 1e-2


@pytest.mark.parametrize("shape", CONV_SHAPES + [(64, 32, 32, 64, 128)])
def test_conv_wgrad(shape):
    N, H, W, Ci, Co = shape
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    dw = K.conv_wgrad(dy, x, N, H, W, Ci, Co)
    ref = R.conv_wgrad(dy, x, N, H, W, Ci, Co)
    assert _rel_err(dw, ref) < 2e-3   # fp32 output: only the summation order differs


@pytest.mark.parametrize("shape", [(256, 8, 8, 128, 256), (256, 8, 8, 256, 256), (256, 4, 4, 256, 512),
                                   (256, 4, 4, 512, 512), (256, 2, 2, 512, 512), (100, 16, 16, 64, 128)])
def test_conv_wgrad_training_shapes_deterministic(shape):
    """The training shapes at batch 256 (split reductions of 1..32 partial tiles): fixed summation
    order, so two launches agree bitwise."""
    N, H, W, Ci, Co = shape
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    ref = R.conv_wgrad(dy, x, N, H, W, Ci, Co)
    dw = K.conv_wgrad(dy, x, N, H, W, Ci, Co)
    assert _rel_err(dw, ref) < 2e-3
    assert torch.equal(dw, K.conv_wgrad(dy, x, N, H, W, Ci, Co))


def test_conv_wgrad_first_layer_compact():
    N, H, W = 16, 32, 32
    op = R.first_layer_operand(torch.randn(N, 3, H, W, device=DEV)).bfloat16()
    dy = torch.randn(N * H * W, 64, device=DEV).bfloat16()
    dw = K.conv_wgrad(dy, op, N, H, W, 64, 64, taps=1, cols_valid=27, out_ld=27)
    ref = R.conv_wgrad(dy, op, N, H, W, 64, 64, taps=1)[:, :27]
    assert dw.shape == (64, 27)
    assert _rel_err(dw, ref) < 2e-3


@pytest.mark.parametrize("fused", [3, 2, 0])
@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("geom", [(8, 32, 32, 64), (8, 8, 8, 256), (16, 2, 2, 512), (64, 16, 16, 128), (5, 4, 4, 512),
                                  (256, 2, 2, 512)])
def test_bn_relu_pool_fwd_bwd(geom, pool, fused):
    """fused=3 (default): two ordinary kernels with fixed-order sums; 2: single cooperative kernel (slice
    parked in shared memory); 0: two passes with float atomics."""
    N, H, W, C = geom
    K.bn_bwd_set_fused(fused)
    x = (torch.randn(N * H * W, C, device=DEV) * 1.5 + 0.3).bfloat16()
    gamma = torch.rand(C, device=DEV) + 0.5
    beta = torch.randn(C, device=DEV) * 0.2
    stats = K.channel_stats(x, C)
    assert torch.allclose(stats, R.channel_stats(x), rtol=1e-4, atol=1e-2)
    y, saved = K.bn_relu_pool_fwd(x, stats, gamma, beta, N, H, W, C, pool)
    ref_y = R.bn_relu_pool_fwd(x, gamma, beta, N, H, W, pool)
    assert _rel_err(y, ref_y) < 1e-2
    Po = y.shape[0]
    dy = torch.randn(Po, C, device=DEV).bfloat16()
    dbias = torch.full((C,), 7.0, device=DEV)
    sums = torch.zeros(2 * C + 32, device=DEV)
    dx, dgamma, dbeta = K.bn_relu_pool_bwd(x, dy, saved, gamma, beta, N, H, W, C, pool, dbias=dbias, sums=sums)
    rdx, rdg, rdb = R.bn_relu_pool_bwd(x, dy, gamma, beta, N, H, W, pool)
    assert _rel_err(dx, rdx) < 2e-2
    assert _rel_err(dgamma, rdg) < 5e-3
    assert _rel_err(dbeta, rdb) < 5e-3
    if fused == 3:
        # fixed summation order: bitwise reproducible, and the ticket counters are left at zero so the
        # same scratch serves the next launch (CUDA-graph replay) without a memset
        assert int(sums[2 * C:].view(torch.int32).abs().sum()) == 0
        dx2, dg2, db2 = K.bn_relu_pool_bwd(x, dy, saved, gamma, beta, N, H, W, C, pool, sums=sums)
        assert torch.equal(dx, dx2) and torch.equal(dgamma, dg2) and torch.equal(dbeta, db2)
    K.bn_bwd_set_fused(3)
    # conv-bias gradient = column sums of dx: mathematically zero (SURVEY.md §2.5 numerics note)
    assert dbias.abs().max().item() < 1e-2 * dy.float().abs().sum(0).max().item()


@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("shape", [(8, 32, 32, 64, 64), (16, 16, 16, 64, 128), (37, 8, 8, 128, 256), (64, 4, 4, 256, 512),
                                   (96, 2, 2, 512, 512), (256, 2, 2, 512, 512)])
def test_conv_bn_relu_pool_fused_forward(shape, pool):
    """conv + BatchNorm + ReLU (+pool) in one call (BatchNorm tail inside the persistent kernel
    after a grid barrier, or the split-K fallback) against the ATen composition."""
    N, H, W, Ci, Co = shape
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    b = torch.randn(Co, device=DEV) * 0.1
    gamma = torch.rand(Co, device=DEV) + 0.5
    beta = torch.randn(Co, device=DEV) * 0.2
    conv_out, act, saved = K.conv_bn_relu_pool_fwd(x, w, b, gamma, beta, N, H, W, Ci, Co, 9, pool)
    torch.cuda.synchronize()
    ref_conv = R.conv_fprop(x, w, b, N, H, W, Ci, Co)
    assert _rel_err(conv_out, ref_conv) < 1e-2
    ref_act = R.bn_relu_pool_fwd(conv_out, gamma, beta, N, H, W, pool)   # BatchNorm of the stored conv output
    assert _rel_err(act, ref_act) < 1e-2
    cf = conv_out.float()
    assert torch.allclose(saved[0], cf.mean(0), rtol=1e-3, atol=1e-3)
    assert torch.allclose(saved[1], (cf.var(0, unbiased=False) + 1e-5).rsqrt(), rtol=2e-3)


@pytest.mark.parametrize("B", [32, 256, 85])
def test_fc_cross_entropy_fwd_bwd(B):
    F_, Kc = 512, 10
    x = torch.randn(B, F_, device=DEV).bfloat16()
    w = torch.randn(Kc, F_, device=DEV) / F_ ** 0.5
    b = torch.randn(Kc, device=DEV) * 0.1
    labels = torch.randint(0, Kc, (B,), device=DEV)
    logits = torch.empty(B, Kc, device=DEV)
    dlogits = torch.empty(B, Kc, device=DEV)
    row_correct = torch.empty(B, dtype=torch.int32, device=DEV)
    loss = torch.zeros((), device=DEV)
    correct = torch.zeros((), dtype=torch.int32, device=DEV)
    K.fc_ce_fwd(x, w, b, labels, logits=logits, dlogits=dlogits, row_correct=row_correct, loss=loss,
                correct=correct)
    rl, rloss, rdl, rc = R.fc_ce(x, w, b, labels)
    assert torch.allclose(logits, rl, rtol=1e-4, atol=1e-4)
    assert abs(loss.item() - rloss.item()) < 1e-4
    assert torch.allclose(dlogits, rdl, rtol=1e-3, atol=1e-6)
    assert correct.item() == rc.item()
    dw, db, dx = K.fc_bwd(x, w, dlogits)
    rdw, rdb, rdx = R.fc_bwd(x, w, rdl)
    assert torch.allclose(dw, rdw, rtol=1e-3, atol=1e-5)
    assert torch.allclose(db, rdb, rtol=1e-3, atol=1e-6)
    assert _rel_err(dx, rdx) < 1e-2
    # the training step's arrangement: forward leaves the per-row losses, backward reduces them
    row_loss = K.fc_ce_fwd(x, w, b, labels, dlogits=dlogits, row_correct=row_correct)
    loss2 = torch.full((), -1.0, device=DEV)
    correct2 = torch.full((), -1, dtype=torch.int32, device=DEV)
    dw2, db2, _ = K.fc_bwd(x, w, dlogits, row_loss=row_loss, row_correct=row_correct, loss=loss2, correct=correct2)
    assert torch.equal(loss2, loss) and torch.equal(correct2, correct)
    assert torch.equal(dw2, dw) and torch.equal(db2, db)


def test_fc_generic_class_count():
    B, F_, Kc = 64, 256, 7
    x = torch.randn(B, F_, device=DEV).bfloat16()
    w = torch.randn(Kc, F_, device=DEV) / F_ ** 0.5
    b = torch.randn(Kc, device=DEV) * 0.1
    labels = torch.randint(0, Kc, (B,), device=DEV)
    logits = torch.empty(B, Kc, device=DEV)
    dlogits = torch.empty(B, Kc, device=DEV)
    loss = torch.zeros((), device=DEV)
    K.fc_ce_fwd(x, w, b, labels, logits=logits, dlogits=dlogits, loss=loss)
    rl, rloss, rdl, _ = R.fc_ce(x, w, b, labels)
    assert torch.allclose(logits, rl, rtol=1e-4, atol=1e-4)
    assert abs(loss.item() - rloss.item()) < 1e-4
    dw, db, dx = K.fc_bwd(x, w, dlogits)
    rdw, rdb, rdx = R.fc_bwd(x, w, rdl)
    assert torch.allclose(dw, rdw, rtol=1e-3, atol=1e-5)
    assert torch.allclose(db, rdb, rtol=1e-3, atol=1e-6)
    assert _rel_err(dx, rdx) < 1e-2


def test_sgd_matches_torch_optim():
    n = 1_000_003
    p = torch.randn(n + 1, device=DEV)[:n + 1 - (n + 1) % 4]
    n = p.numel()
    p0 = p.clone()
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SGD([ref_p], lr=0.1, momentum=0.9, weight_decay=1e-4)
    m = torch.zeros_like(p)
    shadow = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    for _ in range(3):
        g = torch.randn(n, device=DEV)
        ref_p.grad = g.clone()
        opt.step()
        K.sgd_step(p, g, m, 0.1, 0.9, 1e-4, shadow=shadow)
    assert torch.allclose(p, ref_p.data, rtol=1e-5, atol=1e-6)
    assert torch.equal(shadow, p.bfloat16())


def test_pack_weights_layouts():
    Kc, Cc = 128, 64
    master = torch.randn(Kc, 9, Cc, device=DEV)
    fprop = torch.empty(Kc, 9 * Cc, dtype=torch.bfloat16, device=DEV)
    dgrad = torch.empty(Cc, 9 * Kc, dtype=torch.bfloat16, device=DEV)
    m0 = torch.randn(64, 27, device=DEV)
    f0 = torch.full((64, 64), 5.0, dtype=torch.bfloat16, device=DEV)
    plan = K.PackPlan([
        dict(master=m0, fprop=f0, dgrad=None, K=64, C=3, taps=9, fprop_ld=64, flat=1),
        dict(master=master, fprop=fprop, dgrad=dgrad, K=Kc, C=Cc, taps=9, fprop_ld=9 * Cc, flat=0),
    ])
    plan.run()
    assert torch.equal(fprop, master.reshape(Kc, -1).bfloat16())
    assert torch.equal(dgrad, R.rotate_filters_for_dgrad(master.bfloat16(), Kc, Cc))
    assert torch.equal(f0[:, :27], m0.bfloat16()) and f0[:, 27:].abs().max().item() == 0


def test_first_layer_operand_u8_augmentation():
    N = 16
    raw = torch.randint(0, 256, (N, 32, 32, 3), dtype=torch.uint8, device=DEV)
    aug = torch.stack([torch.randint(0, 9, (N,)), torch.randint(0, 9, (N,)), torch.randint(0, 2, (N,)),
                       torch.zeros(N, dtype=torch.long)], 1).to(torch.int8).to(DEV)
    for a in (None, aug):
        op = K.first_layer_operand_u8(raw, a)
        ref = R.first_layer_operand(R.augment_normalize_u8(raw, a)).bfloat16()
        assert (op.float() - ref.float()).abs().max().item() < 2e-2


@pytest.mark.parametrize("H,W", [(32, 32), (16, 16), (8, 64), (12, 20)])
def test_first_layer_operand_shapes_and_filter_pack(H, W):
    """Tiled kernel (rows divide 256 pixels) and the general one give the same operand; the extra
    CTA writes the zero-padded first-layer filters."""
    N = 5
    img = torch.randn(N, 3, H, W, device=DEV)
    m0 = torch.randn(64 * 27, device=DEV)
    f0 = torch.full((64, 64), 5.0, dtype=torch.bfloat16, device=DEV)
    op = K.first_layer_operand_f32(img, w0_master=m0, w0_packed=f0)
    assert torch.equal(op, R.first_layer_operand(img).bfloat16())
    assert torch.equal(f0[:, :27], m0.view(64, 27).bfloat16()) and f0[:, 27:].abs().max().item() == 0
    raw = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device=DEV)
    aug = torch.stack([torch.randint(0, 9, (N,)), torch.randint(0, 9, (N,)), torch.randint(0, 2, (N,)),
                       torch.zeros(N, dtype=torch.long)], 1).to(torch.int8).to(DEV)
    f0.fill_(5.0)
    op = K.first_layer_operand_u8(raw, aug, w0_master=m0, w0_packed=f0)
    ref = R.first_layer_operand(R.augment_normalize_u8(raw, aug)).bfloat16()
    assert (op.float() - ref.float()).abs().max().item() < 2e-2
    assert torch.equal(f0[:, :27], m0.view(64, 27).bfloat16())


def test_conv_split_k_never_overruns_a_small_scratch():
    """the split-K plan is cost based, so a caller's scratch can be smaller than what
    the planner wants for another batch size.  The launcher must then fall back to a plan without
    split-K (same result), and a FORCED split that does not fit must be an error, never a write past
    the buffer (checked with a canary behind the capacity the call was given)."""
    N, H, W, Ci, Co = 64, 2, 2, 512, 512
    want = K.conv_fprop_scratch_elems(N, H, W, Ci, Co, 9)
    assert want > 0, "shape is meant to split K"
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    ref = R.conv_fprop(x, w, None, N, H, W, Ci, Co)
    backing = torch.full((want,), 12345.0, device=DEV)
    small = backing[:want // 4]                       # capacity the call is told about
    y = K.conv_fprop(x, w, None, N, H, W, Ci, Co, scratch=small)
    dx = K.conv_dgrad(dy, w, N, H, W, Ci, Co, scratch=small)
    gamma, beta = torch.ones(Co, device=DEV), torch.zeros(Co, device=DEV)
    conv_out, act, _ = K.conv_bn_relu_pool_fwd(x, w, None, gamma, beta, N, H, W, Ci, Co, 9, True, scratch=small)
    torch.cuda.synchronize()
    assert _rel_err(y, ref) < 1e-2 and _rel_err(conv_out, ref) < 1e-2
    assert _rel_err(dx, R.conv_dgrad(dy, w, N, H, W, Ci, Co)) < 1e-2
    assert bool((backing[want // 4:] == 12345.0).all()), "wrote past the scratch capacity"
    try:
        K.conv_force(-1, -1, 4)
        with pytest.raises(RuntimeError, match="scratch"):
            K.conv_fprop(x, w, None, N, H, W, Ci, Co, scratch=small)
    finally:
        K.conv_force(-1, -1, -1)


def test_conv_wgrad_padded_dw_goes_through_scratch_even_without_a_split():
    """with one split and C_out >= 128 a DENSE dw is written directly and needs no
    scratch, but a dw with padded rows is assembled from scratch: the size query must say so and a
    too-small buffer must be refused."""
    from ddpb200.ops import _native as nat

    N, H, W, Ci, Co = 64, 2, 2, 512, 512
    assert K.conv_wgrad_splits(N, H, W, Ci, Co) == 1 and K.conv_wgrad_scratch_elems(N, H, W, Ci, Co) == 0
    ld = 9 * Ci + 64
    need = K.conv_wgrad_scratch_elems(N, H, W, Ci, Co, 9, 9 * Ci, ld)
    assert need == Co * 9 * Ci
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    out = torch.full((Co, ld), 7.0, device=DEV)
    K.conv_wgrad(dy, x, N, H, W, Ci, Co, out=out, cols_valid=9 * Ci, out_ld=ld)
    ref = R.conv_wgrad(dy, x, N, H, W, Ci, Co)
    assert _rel_err(out[:, :9 * Ci], ref) < 2e-3 and bool((out[:, 9 * Ci:] == 7.0).all())
    small = torch.empty(need // 2, device=DEV)
    rc = nat.load().ddpb_conv_wgrad(nat.ptr(dy), nat.ptr(x), nat.ptr(out), nat.ptr(small), small.numel(), N, H, W, Ci,
                                    Co, 9, 9 * Ci, ld, nat.stream_ptr(None))
    assert rc != 0 and b"scratch" in nat.load().ddpb_last_error()


@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("case", [((16, 16, 16, 64, 128), 128), ((37, 8, 8, 128, 256), 256), ((256, 8, 8, 256, 256), 256),
                                  ((64, 4, 4, 256, 512), 128)])
def test_persistent_cta_pair_conv_with_batchnorm_tail(case, pool):
    """Mode 3 (persistent cta_group::2 kernel, opt-in) with the BatchNorm + ReLU (+pool) tail after the grid barrier."""
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


@pytest.mark.parametrize("pool", [False, True])
@pytest.mark.parametrize("shape", [(256, 16, 16, 64, 128), (256, 8, 8, 128, 256), (256, 8, 8, 256, 256), (256, 4, 4, 256, 512),
                                   (256, 4, 4, 512, 512), (64, 16, 16, 64, 128), (37, 8, 8, 128, 256), (5, 4, 4, 256, 512),
                                   (3, 2, 2, 512, 512), (8, 32, 32, 64, 64)])
def test_batchnorm_tail_from_tmem_equals_the_tail_from_l2(shape, pool):
    """The fused BatchNorm tail reads the accumulators back from TMEM when every tile of a CTA still has its
    buffer (several tiles per CTA at 16x16, ragged last tiles, pooling across lanes for W = 2..16; 32x32 and
    recycled buffers fall back to the stored tiles).  Same rounding points as the pass over the stored tiles:
    the two agree to the run-to-run noise of the statistics, and both match the ATen composition."""
    N, H, W, Ci, Co = shape
    torch.manual_seed(2)
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    b = torch.randn(Co, device=DEV) * 0.1
    gamma = (torch.rand(Co, device=DEV) + 0.5) * torch.where(torch.rand(Co, device=DEV) < 0.3, -1.0, 1.0)
    beta = torch.randn(Co, device=DEV) * 0.2
    out = {}
    try:
        for mode in (2, 0):           # 2: from TMEM wherever the shape allows (also several tiles per CTA)
            K.conv_tail_from_tmem(mode)
            K.conv_force(-1, -1, 1)       # no split-K: the tail runs inside the kernel
            conv_out, act, saved = K.conv_bn_relu_pool_fwd(x, w, b, gamma, beta, N, H, W, Ci, Co, 9, pool)
            torch.cuda.synchronize()
            out[mode] = (conv_out.clone(), act.clone(), saved.clone())
    finally:
        K.conv_tail_from_tmem(1)
        K.conv_force(-1, -1, -1)
    assert torch.equal(out[0][0], out[2][0])                       # the conv output itself is deterministic
    assert torch.allclose(out[0][2], out[2][2], rtol=1e-5, atol=1e-6)
    diff = (out[0][1].float() - out[2][1].float()).abs()
    # (a last-bit difference of the statistics moves a few outputs by one bf16 ulp; a wrong tail would move most of them)
    assert float(diff.max()) <= 2 ** -6 * float(out[0][1].float().abs().max()) and float((diff > 0).float().mean()) < 1e-2
    for mode in (0, 2):
        assert _rel_err(out[mode][1], R.bn_relu_pool_fwd(out[mode][0], gamma, beta, N, H, W, pool)) < 1e-2


@pytest.mark.parametrize("case", [((256, 2, 2, 512, 512), 128, 4), ((256, 2, 2, 512, 512), 256, 8), ((256, 2, 2, 512, 512), 64, 2),
                                  ((96, 2, 2, 512, 512), 128, 8), ((64, 4, 4, 256, 512), 128, 4), ((37, 4, 4, 512, 256), 256, 7),
                                  ((8, 8, 8, 128, 256), 128, 3), ((3, 2, 2, 512, 512), 128, 5)])
def test_split_k_summed_inside_the_persistent_kernel(case):
    """Split-K with sync words: the warp that draws the last ticket of a tile sums the slices in split order, so
    the output does not depend on which CTA finished last (bitwise repeatable), the BatchNorm tail stays fused,
    and the sync words are zero again afterwards.  Oracle: ATen; cross-check: the two-kernel path."""
    (N, H, W, Ci, Co), bn, splits = case
    torch.manual_seed(4)
    x = torch.randn(N * H * W, Ci, device=DEV).bfloat16()
    w = (torch.randn(Co, 9 * Ci, device=DEV) / (9 * Ci) ** 0.5).bfloat16()
    dy = torch.randn(N * H * W, Co, device=DEV).bfloat16()
    b = torch.randn(Co, device=DEV) * 0.1
    gamma, beta = torch.rand(Co, device=DEV) + 0.5, torch.randn(Co, device=DEV) * 0.2
    sync = torch.zeros(K.conv_split_sync_words(), dtype=torch.int32, device=DEV)
    scratch = torch.full((8 * ((N * H * W + 127) // 128) * 128 * max(Ci, Co),), float("nan"), device=DEV)
    ref = R.conv_fprop(x, w, b, N, H, W, Ci, Co)
    ref_dx = R.conv_dgrad(dy, w, N, H, W, Ci, Co)
    try:
        K.conv_force(-1, bn, splits)
        assert K.conv_plan(N, H, W, Ci, Co, 9)[1:] == (bn, splits)
        runs = []
        for _ in range(3):
            y = K.conv_fprop(x, w, b, N, H, W, Ci, Co, scratch=scratch, split_sync=sync)
            conv_out, act, saved = K.conv_bn_relu_pool_fwd(x, w, b, gamma, beta, N, H, W, Ci, Co, 9, True, scratch=scratch,
                                                           split_sync=sync)
            dx = K.conv_dgrad(dy, w, N, H, W, Ci, Co, scratch=scratch, split_sync=sync) if Ci % bn == 0 else None
            torch.cuda.synchronize()
            assert int(sync.abs().sum()) == 0
            runs.append((y.clone(), conv_out.clone(), None if dx is None else dx.clone(), act.clone()))
        for r in runs[1:]:
            assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][0])      # same bits every time, both entry points
            assert r[2] is None or torch.equal(r[2], runs[0][2])
        y, conv_out, dx, act = runs[0]
        assert _rel_err(y, ref) < 1e-2
        assert dx is None or _rel_err(dx, ref_dx) < 1e-2
        assert _rel_err(act, R.bn_relu_pool_fwd(conv_out, gamma, beta, N, H, W, True)) < 1e-2
        cf = conv_out.float()
        assert torch.allclose(saved[0], cf.mean(0), rtol=1e-3, atol=1e-3)
        # the two-kernel path (no sync words) sums in another order: equal up to fp32 rounding before the bf16 store
        y2 = K.conv_fprop(x, w, b, N, H, W, Ci, Co, scratch=scratch, split_sync=None)
        diff = (y2.float() - y.float()).abs()
        assert float((diff > 0).float().mean()) < 2e-2 and float(diff.max()) <= 2 ** -7 * float(y.float().abs().max())
    finally:
        K.conv_force(-1, -1, -1)
