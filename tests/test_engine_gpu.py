"""Step-level integration on one B200: the kernel pipeline against the reference modules.

Oracle (SURVEY.md §4 "Step integration"): the same VGG-11 built from stock torch modules
(what /root/reference/part1/model.py builds), fp32 on the CPU, torch.optim.SGD.
The kernel path computes in bf16 with fp32 accumulation, so comparisons use cosine
similarity / relative norms rather than elementwise equality.
"""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu

import ddpb200
from ddpb200.ops.reference import augment_normalize_u8

DEV = "cuda"


def _cos(a, b):
    a, b = a.double().flatten().cpu(), b.double().flatten().cpu()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def _relnorm(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _make_pair(seed=89395):
    torch.manual_seed(seed)
    ref = ddpb200.VGG11()            # on CPU this IS the stock module stack
    mine = copy.deepcopy(ref).to(DEV)
    return ref, mine


def _batch(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    img = torch.randint(0, 256, (n, 32, 32, 3), dtype=torch.uint8, generator=g)
    lab = torch.randint(0, 10, (n,), generator=g)
    return img, lab


@pytest.mark.parametrize("n", [64, 24])
def test_forward_backward_matches_fp32_modules(n):
    ref, mine = _make_pair()
    img, lab = _batch(n)
    x = augment_normalize_u8(img, None)
    # oracle
    ref.zero_grad()
    out = ref(x)
    loss = torch.nn.functional.cross_entropy(out, lab)
    loss.backward()
    # kernels, through the nn.Module + autograd surface
    out2 = mine(x.to(DEV))
    loss2 = torch.nn.functional.cross_entropy(out2, lab.to(DEV))
    loss2.backward()
    assert _relnorm(out2, out) < 0.05
    assert abs(loss2.item() - loss.item()) < 0.03 * abs(loss.item())
    sd_ref = dict(ref.named_parameters())
    for name, p in mine.named_parameters():
        g, gr = p.grad, sd_ref[name].grad
        assert g.shape == gr.shape
        if name.endswith("bias") and name.startswith("layers") and gr.dim() == 1 and "layers." in name and \
                isinstance(dict(ref.named_modules())[name.rsplit(".", 1)[0]], torch.nn.Conv2d):
            # conv bias feeding BatchNorm: the true gradient is rounding noise
            assert g.abs().max().item() < 1e-2
            continue
        # bf16 activations and gradients through 8 blocks: measured 0.947 (first conv) .. 1.000 (classifier)
        assert _cos(g, gr) > 0.92, name
        assert 0.9 < float(g.norm() / gr.norm()) < 1.1, name


def test_trainer_step_matches_sgd_oracle_and_graph_replay():
    ref, mine = _make_pair()
    opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    tr = ddpb200.Trainer(mine, mode="single", batch_size=64, lr=0.01)
    p0 = {k: v.detach().clone() for k, v in ref.named_parameters()}
    losses_ref, losses = [], []
    for it in range(3):                       # it=0 eager, it=1 captures, it=2 replays the graph
        img, lab = _batch(64, seed=it)
        x = augment_normalize_u8(img, None)
        opt.zero_grad()
        l = torch.nn.functional.cross_entropy(ref(x), lab)
        l.backward()
        opt.step()
        losses_ref.append(l.item())
        losses.append(tr.step(img.pin_memory(), lab.pin_memory()).item())
    assert abs(losses[0] - losses_ref[0]) < 0.03 * losses_ref[0]
    sd = dict(mine.named_parameters())
    for name, p in ref.named_parameters():
        if p.dim() == 4 or name == "fc1.weight":
            # direction of the total update after three steps
            assert _cos(sd[name].detach().cpu() - p0[name], p.detach() - p0[name]) > 0.85, name
    assert all(l == l for l in losses)


def test_loss_decreases_on_learnable_data():
    from ddpb200.data import synthetic_cifar10

    ds = synthetic_cifar10(True, size=2048)
    torch.manual_seed(1)
    model = ddpb200.VGG11().to(DEV)
    tr = ddpb200.Trainer(model, mode="single", batch_size=128, lr=0.05)
    losses = []
    for ep in range(3):
        for s in range(0, 2048, 128):
            losses.append(tr.step(ds.images[s:s + 128], ds.labels[s:s + 128]).item())
    assert sum(losses[-4:]) / 4 < 0.6 * sum(losses[:4]) / 4, (losses[:4], losses[-4:])
    loss, correct = tr.evaluate(ds.images[:128], ds.labels[:128])
    assert int(correct) > 40


def test_state_dict_roundtrip_and_layout(tmp_path):
    from ddpb200.utils.checkpoint import load_checkpoint, save_checkpoint

    torch.manual_seed(3)
    model = ddpb200.VGG11().to(DEV)
    tr = ddpb200.Trainer(model, mode="single", batch_size=32)
    img, lab = _batch(32)
    tr.step(img, lab)
    keys = list(model.state_dict().keys())
    assert len(keys) == 34 and keys[0] == "layers.0.weight" and keys[-1] == "fc1.bias"
    assert model.state_dict()["layers.0.weight"].shape == (64, 3, 3, 3)
    save_checkpoint(tmp_path / "ck.pt", model)
    stock = ddpb200.VGG11()                       # CPU, stock modules
    load_checkpoint(tmp_path / "ck.pt", stock)
    x = augment_normalize_u8(img, None)
    assert _relnorm(model(x.to(DEV)), stock(x)) < 0.05
    # and back: loading must keep the arena views alive
    before = tr.engine.arena.master.data_ptr()
    load_checkpoint(tmp_path / "ck.pt", model)
    assert model.layers[0].weight.data_ptr() == before
    tr.step(img, lab)


def test_our_sgd_and_eager_api_like_the_reference_loop():
    """zero_grad / forward / criterion / backward / optimizer.step, spelled as in part1/main.py:71-77."""
    ref, mine = _make_pair(seed=5)
    optimizer = ddpb200.SGD(mine.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    criterion = ddpb200.CrossEntropyLoss().to(DEV)
    img, lab = _batch(32)
    x = augment_normalize_u8(img, None).to(DEV)
    w0 = mine.fc1.weight.detach().clone()
    for _ in range(2):
        optimizer.zero_grad()
        loss = criterion(mine(x), lab.to(DEV))
        loss.backward()
        optimizer.step()
    assert not torch.equal(w0, mine.fc1.weight.detach())
    assert loss.item() == loss.item()


# ---------------------------------------------------------------------------------------
# precision-matched oracle: fp32 ATen ops with bf16 rounding inserted exactly where the
# kernels round (conv output, block output, and the gradients flowing through both).  What
# is left is accumulation order, so agreement must be tight -- this is the test that would
# catch an indexing / layout / scaling bug that the loose fp32 comparison could hide.
# ---------------------------------------------------------------------------------------
class _RoundBoth(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.bfloat16().float()

    @staticmethod
    def backward(ctx, g):
        return g.bfloat16().float()


def _emulated_forward(model, x):
    import torch.nn.functional as F

    y = x.bfloat16().float()
    for blk in model.conv_blocks():
        w = blk.conv.weight.detach().bfloat16().float() + (blk.conv.weight - blk.conv.weight.detach())  # value rounded, grad to master
        y = F.conv2d(y, w, blk.conv.bias, padding=1)
        y = _RoundBoth.apply(y)
        y = F.batch_norm(y, None, None, blk.bn.weight, blk.bn.bias, training=True, eps=blk.bn.eps)
        y = F.relu(y)
        if blk.pool:
            y = F.max_pool2d(y, 2, 2)
        y = _RoundBoth.apply(y)
    return model.fc1(y.flatten(1))


@pytest.mark.parametrize("n", [64, 40])
def test_matches_precision_matched_oracle(n):
    """Layer by layer, feeding the oracle the ENGINE's previous activation (teacher forcing), the
    kernels agree with the precision-matched ATen emulation to ~1e-4 (measured 1e-5..1.2e-4,
    profiles/r1_call5_layerwise_agreement.txt).  End to end the two diverge by ~1%: each layer's
    rounding flips are amplified by the next BatchNorm -- a property of bf16 activations, also
    present between two ATen runs with different summation orders -- so the end-to-end check is
    loose and the per-layer one is tight."""
    import torch.nn.functional as F

    from ddpb200.ops.reference import nchw_to_nhwc2d

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(11)
    oracle = ddpb200.VGG11().to(DEV)
    mine = copy.deepcopy(oracle)
    img, lab = _batch(n, seed=3)
    x = augment_normalize_u8(img, None).to(DEV)
    lab = lab.to(DEV)
    # ---- end to end (loose)
    out = _emulated_forward(oracle, x)
    loss = torch.nn.functional.cross_entropy(out, lab)
    loss.backward()
    ref_grads = {k: v.grad.clone() for k, v in oracle.named_parameters()}
    out2 = mine(x)
    loss2 = torch.nn.functional.cross_entropy(out2, lab)
    loss2.backward()
    assert _relnorm(out2, out) < 3e-2
    assert abs(loss2.item() - loss.item()) < 1e-2
    conv_bias = {f"layers.{b.conv_idx}.bias" for b in mine.conv_blocks()}
    for name, p in mine.named_parameters():
        if name not in conv_bias:
            assert _cos(p.grad, ref_grads[name]) > 0.95, name
            assert 0.9 < float(p.grad.norm() / ref_grads[name].norm()) < 1.1, name
    # ---- per layer, teacher forced (tight)
    eng = mine._engine
    y = x.bfloat16().float()
    with torch.no_grad():
        for li, blk in enumerate(mine.conv_blocks()):
            L = eng.layers[li]
            P = n * L.H * L.W
            c = F.conv2d(y, blk.conv.weight.bfloat16().float(), blk.conv.bias, padding=1).bfloat16().float()
            assert _relnorm(eng.conv_out[li][:P].float(), nchw_to_nhwc2d(c)) < 1e-3, f"conv {li}"
            z = F.relu(F.batch_norm(c, None, None, blk.bn.weight, blk.bn.bias, training=True, eps=blk.bn.eps))
            if blk.pool:
                z = F.max_pool2d(z, 2, 2)
            Po = P // 4 if blk.pool else P
            assert _relnorm(eng.act[li][:Po].float(), nchw_to_nhwc2d(z.bfloat16().float())) < 1e-3, f"block {li}"
            Ho = L.H // 2 if blk.pool else L.H
            y = eng.act[li][:Po].float().view(n, Ho, Ho, -1).permute(0, 3, 1, 2).contiguous()
    # ---- per layer BACKWARD, teacher forced (tight): every kernel of the backward chain against the fp32
    # restatement of its op, fed with the ENGINE's own upstream gradient and stored activations, so a wrong dX /
    # dW / BatchNorm gradient cannot hide behind the loose end-to-end cosine above
    from ddpb200.ops import reference as R

    a = eng.arena
    for li in range(len(eng.layers) - 1, -1, -1):
        L = eng.layers[li]
        P = n * L.H * L.W
        Po = P // 4 if L.pool else P
        gamma, beta = a.flat_master(L.p_g), a.flat_master(L.p_beta)
        dconv_ref, dg_ref, db_ref = R.bn_relu_pool_bwd(eng.conv_out[li][:P], eng.dact[li][:Po], gamma, beta, n, L.H, L.W,
                                                       L.pool)
        assert _relnorm(eng.dconv[li][:P].float(), dconv_ref) < 6e-3, f"BatchNorm backward dx, layer {li}"
        assert _relnorm(a.flat_grad(L.p_g), dg_ref) < 2e-3 and _relnorm(a.flat_grad(L.p_beta), db_ref) < 2e-3, li
        x_in = eng.operand0 if li == 0 else eng.act[li - 1]
        dw_ref = R.conv_wgrad(eng.dconv[li][:P], x_in[:P], n, L.H, L.W, L.c_in_eff, L.c_out, L.taps)
        cols = 27 if li == 0 else 9 * L.c_in
        assert _relnorm(a.flat_grad(L.p_w).view(L.c_out, cols), dw_ref[:, :cols]) < 2e-3, f"wgrad, layer {li}"
        if li > 0:
            dx_ref = R.conv_dgrad(eng.dconv[li][:P], eng.w_f[li], n, L.H, L.W, L.c_in, L.c_out)
            assert _relnorm(eng.dact[li - 1][:P].float(), dx_ref) < 6e-3, f"dgrad, layer {li}"


@pytest.mark.parametrize("factory,batch", [(ddpb200.VGG16, 24), (ddpb200.VGG13, 40)])
def test_deeper_vgg_variants_train(factory, batch):
    """The reference carries configs for VGG13/16/19 as dead data (model.py:3-8); here they run."""
    torch.manual_seed(2)
    model = factory().to(DEV)
    tr = ddpb200.Trainer(model, mode="single", batch_size=batch, lr=0.01)
    img, lab = _batch(batch)
    l0 = tr.step(img, lab).item()
    for _ in range(5):
        l = tr.step(img, lab).item()
    assert l == l and l < l0       # memorising one batch: the loss must fall


def test_ragged_batches_and_evaluation():
    """Last batch of an epoch is smaller than batch_size (50000 = 195*256 + 80): runs without the graph."""
    torch.manual_seed(4)
    model = ddpb200.VGG11().to(DEV)
    tr = ddpb200.Trainer(model, mode="single", batch_size=64, lr=0.01)
    for n in (64, 64, 17, 64, 5):
        img, lab = _batch(n, seed=n)
        l = tr.step(img, lab).item()
        assert l == l, f"loss is NaN for a batch of {n}"
    img, lab = _batch(16, seed=1)          # the reference's last test batch has 16 images
    loss, correct = tr.evaluate(img, lab)
    # evaluation sees the parameters AFTER the last update, first layer's padded filter pack included
    eng = tr.engine
    w0 = eng.arena.flat_master(eng.layers[0].p_w).view(64, 27)
    assert torch.equal(eng.w_f[0][:, :27], w0.bfloat16()) and eng.w_f[0][:, 27:].abs().max().item() == 0
    x = augment_normalize_u8(img, None).to(DEV)
    with torch.no_grad():
        logits = model(x)
    ref_loss = torch.nn.functional.cross_entropy(logits, lab.to(DEV))
    assert abs(loss.item() - ref_loss.item()) < 2e-2 * abs(ref_loss.item())   # u8 vs f32 input path: bf16 rounding
    assert abs(int(correct) - int((logits.argmax(1).cpu() == lab).sum())) <= 1


def test_larger_eager_batch_keeps_the_trainers_parameters_and_momentum():
    """model(x) with a batch larger than the Trainer's engine used to re-home the
    parameters into a NEW arena, so the trainer kept training tensors nobody could see any more.  The
    regrown engine now shares the arenas: only activations are reallocated."""
    torch.manual_seed(6)
    model = ddpb200.VGG11().to(DEV)
    tr = ddpb200.Trainer(model, mode="single", batch_size=32, lr=0.05)
    img, lab = _batch(32)
    tr.step(img, lab)
    master_ptr = tr.engine.arena.master.data_ptr()
    mom = tr.engine.arena.momentum.clone()
    assert float(mom.abs().sum()) > 0
    with torch.no_grad():
        big = model(torch.randn(64, 3, 32, 32, device=DEV))     # twice the trainer's batch
    assert big.shape == (64, 10) and model._engine.max_batch >= 64
    assert model._engine.arena is tr.engine.arena and model._engine.arena.master.data_ptr() == master_ptr
    assert torch.equal(tr.engine.arena.momentum, mom)
    before = model.fc1.weight.detach().clone()
    tr.step(img, lab)                                           # still trains what model.parameters() shows
    torch.cuda.synchronize()
    assert not torch.equal(before, model.fc1.weight.detach())
    assert model.fc1.weight.data_ptr() >= master_ptr


def test_scratch_is_sized_for_every_batch_up_to_the_maximum():
    """a 256-batch engine must hold the split-K scratch of every N <= 256."""
    from ddpb200.ops import kernels as K

    torch.manual_seed(7)
    model = ddpb200.VGG11().to(DEV)
    tr = ddpb200.Trainer(model, mode="single", batch_size=256, lr=0.01, use_graph=False)
    eng = tr.engine
    for n in (103, 136, 192, 232):
        for li, L in enumerate(eng.layers):
            assert K.conv_fprop_scratch_elems(n, L.H, L.W, L.c_in_eff, L.c_out, L.taps) <= eng.fprop_scratch.numel()
            if li:
                assert K.conv_fprop_scratch_elems(n, L.H, L.W, L.c_out, L.c_in, 9) <= eng.fprop_scratch.numel()
    img, lab = _batch(232, seed=9)
    l = tr.step(img, lab).item()        # 1000 % 256 = 232: one of the sizes that used to overrun
    assert l == l


def test_prefetch_and_published_loss_match_the_plain_call():
    """Trainer.prefetch / step() / last_loss(): the software-pipelined loop of train_model gives the same
    losses and parameters as step(data, target, aug) + loss.item(), batch by batch."""
    batches = [tuple(t.pin_memory() for t in _batch(32, seed=s)) for s in range(5)]

    def run(pipelined):
        torch.manual_seed(11)
        model = ddpb200.VGG11().to(DEV)
        tr = ddpb200.Trainer(model, mode="single", batch_size=32, lr=0.02)
        out = []
        if pipelined:
            tr.prefetch(*batches[0])
        for i, (img, lab) in enumerate(batches):
            if pipelined:
                dev_loss = tr.step()
                if i + 1 < len(batches):
                    tr.prefetch(*batches[i + 1])
                host = tr.last_loss()
                assert host == dev_loss.item()          # same fp32 bits, read two ways
                out.append(host)
            else:
                out.append(tr.step(img, lab).item())
        torch.cuda.synchronize()
        return out, [p.detach().clone() for p in model.parameters()]

    la, pa = run(False)
    lb, pb = run(True)
    # two runs are not bitwise equal (the forward BatchNorm statistics are accumulated with float
    # atomics, DESIGN.md par. 7a), so compare within the run-to-run noise
    # (the first step starts from identical parameters; later ones amplify the last-bit differences)
    assert abs(la[0] - lb[0]) < 2e-3 * abs(la[0]), (la, lb)
    assert all(abs(a - b) < 0.15 * abs(a) for a, b in zip(la, lb)), (la, lb)
    assert _cos(torch.cat([p.flatten() for p in pa]), torch.cat([p.flatten() for p in pb])) > 0.99


def test_device_side_augmentation_draws():
    """K16 random numbers on the device: ranges, uniformity, determinism in (seed, step), a fresh set on every
    launch, and the drawn parameters fed to the input kernel give what the CPU restatement gives."""
    from ddpb200.ops import kernels as K

    n, pad = 4096, 4
    aug = torch.zeros(n, 4, dtype=torch.int8, device=DEV)
    step = torch.zeros(1, dtype=torch.int32, device=DEV)
    a0 = K.draw_augmentation(aug, step, seed=1234, pad=pad).clone()
    a1 = K.draw_augmentation(aug, step, seed=1234, pad=pad).clone()
    assert int(step) == 2
    assert int(a0[:, :2].min()) == 0 and int(a0[:, :2].max()) == 2 * pad
    assert set(a0[:, 2].unique().tolist()) == {0, 1} and int(a0[:, 3].abs().max()) == 0
    assert not torch.equal(a0, a1)                                  # the counter moved
    # every one of the 9 offsets ~ n/9 times, flips ~ n/2 (6 sigma of a binomial)
    for col, k in ((0, 9), (1, 9), (2, 2)):
        cnt = torch.bincount(a0[:, col].long(), minlength=k).float()
        sigma = (n * (1 / k) * (1 - 1 / k)) ** 0.5
        assert float((cnt - n / k).abs().max()) < 6 * sigma, cnt
    assert abs(float(torch.corrcoef(torch.stack([a0[:, 0].float(), a0[:, 1].float()]))[0, 1])) < 0.08
    step.zero_()
    assert torch.equal(K.draw_augmentation(aug, step, seed=1234, pad=pad), a0)      # same (seed, step)
    step.zero_()
    assert not torch.equal(K.draw_augmentation(aug, step, seed=1235, pad=pad), a0)  # another seed
    # a partial batch leaves the rest alone
    aug.fill_(7)
    step.zero_()
    K.draw_augmentation(aug, step, seed=1234, pad=pad, n=100)
    assert torch.equal(aug[:100], a0[:100]) and int((aug[100:] != 7).sum()) == 0
    # the draws drive the input kernel exactly as host-made ones do
    img, _ = _batch(64, seed=3)
    from ddpb200.ops import reference as R

    out = K.first_layer_operand_u8(img.to(DEV), a0[:64].contiguous())
    ref = R.first_layer_operand(augment_normalize_u8(img, a0[:64].cpu(), pad)).bfloat16()
    assert (out.float().cpu() - ref.float()).abs().max().item() < 2e-2


def test_trainer_draws_its_augmentation_inside_the_captured_step():
    """Trainer(device_augment=True): uint8 batches without draws get them from the device generator, a new
    set on every CUDA-graph replay, different per rank seed, and training still learns."""
    img, lab = _batch(64, seed=5)
    img, lab = img.pin_memory(), lab.pin_memory()
    torch.manual_seed(3)
    model = ddpb200.VGG11().to(DEV)
    tr = ddpb200.Trainer(model, mode="single", batch_size=64, lr=0.01, device_augment=True, augment_seed=9)
    seen, losses = [], []
    for _ in range(6):
        slot = tr._slot
        losses.append(tr.step(img, lab).item())
        seen.append(tr._stage[slot]["aug"].clone())
    assert int(tr._aug_step) == 6
    assert len({tuple(a.flatten().tolist()) for a in seen}) == 6     # eager, capture and replays all differ
    assert all(int(a[:, :2].max()) <= 8 and int(a.min()) >= 0 for a in seen)
    assert losses[-1] < losses[0]
    # host-made draws still win when they are given
    aug = torch.zeros(64, 4, dtype=torch.int8)
    aug[:, 0] = 4
    aug[:, 1] = 4
    slot = tr._slot
    tr.step(img, lab, aug.pin_memory())
    assert torch.equal(tr._stage[slot]["aug"].cpu(), aug) and int(tr._aug_step) == 6


def test_device_synthetic_batch():
    from ddpb200.ops import kernels as K

    step = torch.zeros(1, dtype=torch.int32, device=DEV)
    img0, lab0 = K.synthetic_batch(512, step, seed=7)
    img1, lab1 = K.synthetic_batch(512, step, seed=7)
    assert int(step) == 2 and img0.shape == (512, 32, 32, 3) and img0.dtype == torch.uint8
    assert int(lab0.min()) >= 0 and int(lab0.max()) <= 9 and len(lab0.unique()) == 10
    assert not torch.equal(lab0, lab1) and not torch.equal(img0, img1)
    step.zero_()
    img2, lab2 = K.synthetic_batch(512, step, seed=7)
    assert torch.equal(img0, img2) and torch.equal(lab0, lab2)
    # images of one class share a template: their means are closer to the class mean than to other classes'
    x = img0.float().view(512, -1) / 255
    means = torch.stack([x[lab0 == c].mean(0) for c in range(10)])
    d = torch.cdist(x, means)
    assert float((d.argmin(1) == lab0).float().mean()) > 0.95
    # noise of about sigma 0.25 (clamping trims it a little)
    resid = (x - means[lab0]).std()
    assert 0.12 < float(resid) < 0.3
