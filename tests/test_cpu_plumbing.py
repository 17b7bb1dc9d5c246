"""CPU-only tests of everything above the kernels: model / state_dict parity, CLI, data
sharding, bucket planning, printed line formats (SURVEY.md §4 "Invariants" and "CLI")."""
import io
import math
import re
from contextlib import redirect_stdout

import pytest
import torch

import ddpb200
from ddpb200 import cli
from ddpb200.data import BatchLoader, ShardSampler, create_data_loaders, synthetic_cifar10
from ddpb200.parallel import sync as S

# state_dict keys in model.parameters() order
EXPECTED_KEYS = [f"layers.{i}.{n}" for i in (0, 1, 4, 5, 8, 9, 11, 12, 15, 16, 18, 19, 22, 23, 25, 26)
                 for n in ("weight", "bias")] + ["fc1.weight", "fc1.bias"]


def test_state_dict_layout_matches_reference():
    m = ddpb200.VGG11()
    sd = m.state_dict()
    assert list(sd.keys()) == EXPECTED_KEYS
    assert sum(v.numel() for v in sd.values()) == 9_231_114
    assert sd["layers.0.weight"].shape == (64, 3, 3, 3) and sd["fc1.weight"].shape == (10, 512)
    assert len(list(m.buffers())) == 0          # track_running_stats=False: no buffers
    assert [b.pool for b in m.conv_blocks()] == [True, True, False, True, False, True, False, True]


def test_same_seed_gives_reference_initialisation():
    """Same construction order as model.py => same RNG consumption => same weights as a stock build."""
    import torch.nn as nn

    torch.manual_seed(89395)
    mine = ddpb200.VGG11()
    torch.manual_seed(89395)
    layers, cin = [], 3
    for c in [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"]:
        if c == "M":
            layers.append(nn.MaxPool2d(2, 2))
        else:
            layers += [nn.Conv2d(cin, c, 3, 1, 1, bias=True), nn.BatchNorm2d(c, track_running_stats=False),
                       nn.ReLU(inplace=True)]
            cin = c
    stock_layers = nn.Sequential(*layers)
    stock_fc = nn.Linear(512, 10)
    for (k, v), w in zip(mine.state_dict().items(), list(stock_layers.state_dict().values()) + list(stock_fc.state_dict().values())):
        assert torch.equal(v, w), k


def test_other_vgg_depths_build():
    for f, n in ((ddpb200.VGG13, 10), (ddpb200.VGG16, 13), (ddpb200.VGG19, 16)):
        m = f()
        assert len(m.conv_blocks()) == n
        assert m(torch.randn(2, 3, 32, 32)).shape == (2, 10)


def test_cli_flags_and_lazy_rank(monkeypatch):
    assert cli.parse_arguments(["--num-nodes", "4", "--rank", "2"]) == ("10.10.1.1", "4000", 2, 4)
    ip, port, rank, size = cli.parse_arguments(["--master-ip", "127.0.0.1", "--master-port", "29500", "--num-nodes",
                                                "2", "--rank", "0"])
    assert (ip, port, rank, size) == ("127.0.0.1", "29500", 0, 2) and isinstance(port, str)

    class U:
        nodename = "node3"

    monkeypatch.setattr(cli.os, "uname", lambda: U)
    assert cli.get_rank() == 3
    assert cli.parse_arguments(["--num-nodes", "4"])[2] == 3
    U.nodename = "vm"                                  # the reference dies here even with --rank (Q2)
    assert cli.parse_arguments(["--num-nodes", "4", "--rank", "1"])[2] == 1
    with pytest.raises(SystemExit):
        cli.parse_arguments(["--num-nodes", "4"])


def test_shard_sampler_matches_distributed_sampler():
    from torch.utils.data.distributed import DistributedSampler

    for n, w in ((50_000, 4), (50_000, 3), (1000, 8), (10, 4)):
        for r in range(w):
            ref = list(DistributedSampler(range(n), num_replicas=w, rank=r, shuffle=False, drop_last=False))
            assert ShardSampler(n, w, r).indices().tolist() == ref


def test_union_of_rank_batches_is_the_single_process_batch():
    """with shuffle=False the ranks' i-th batches partition part1's i-th batch."""
    w, n = 4, 2048
    per = 256 // w
    single = ShardSampler(n).indices()
    shards = [ShardSampler(n, w, r).indices() for r in range(w)]
    for i in range(n // 256):
        union = torch.cat([s[i * per:(i + 1) * per] for s in shards]).sort().values
        assert torch.equal(union, single[i * 256:(i + 1) * 256])


def test_loaders_shapes_and_determinism():
    tr, te = create_data_loaders(1, 4, 64, root=None, synthetic=True, train_size=1000, test_size=300)
    assert len(tr) == math.ceil(250 / 64) and len(te) == math.ceil(300 / 64)
    assert len(te.dataset) == 300                       # the test split is NOT sharded (Q10)
    img, lab, aug = next(iter(tr.raw()))
    assert img.dtype == torch.uint8 and img.shape == (64, 32, 32, 3) and lab.dtype == torch.int64
    assert aug.shape == (64, 4) and int(aug[:, :2].max()) <= 8 and int(aug[:, 2].max()) <= 1
    # plain iteration = what the reference's DataLoader yields (part1/main.py:65): normalised fp32 NCHW pairs,
    # same crop / flip draws as the raw view
    data, target = next(iter(tr))
    assert data.dtype == torch.float32 and data.shape == (64, 3, 32, 32) and torch.equal(target, lab)
    from ddpb200.ops.reference import augment_normalize_u8
    assert torch.equal(data, augment_normalize_u8(img, aug))
    n0 = 5
    oy, ox, flip = (int(v) for v in aug[n0, :3])
    padded = torch.nn.functional.pad(img[n0].permute(2, 0, 1).float(), (4, 4, 4, 4))[:, oy:oy + 32, ox:ox + 32]
    want = (padded.flip(-1) if flip else padded) / 255.0
    mean = torch.tensor([125.3, 123.0, 113.9]).view(3, 1, 1) / 255.0
    std = torch.tensor([63.0, 62.1, 66.7]).view(3, 1, 1) / 255.0
    assert torch.allclose(data[n0], (want - mean) / std, atol=1e-6)
    a = synthetic_cifar10(True, 128).images
    b = synthetic_cifar10(True, 128).images
    assert torch.equal(a, b)
    tr.sampler.set_epoch(3)                             # API parity (Q13)
    last = list(BatchLoader(synthetic_cifar10(False, 100), 64).raw())[-1]
    assert last[0].shape[0] == 36 and last[2] is None


def test_bucket_plan_reproduces_ddp_layout():
    """buckets of ~9.46 / 27.16 / 0.31 MB in gradient-ready order."""
    m = ddpb200.VGG11()
    groups = S.param_groups_of(m)
    assert len(groups) == 9 and sum(g[2] for g in groups) == 9_231_114
    b = S.plan_buckets(groups)
    assert [x.numel for x in b] == [2_365_962, 6_789_120, 76_032]
    assert [(x.first_param, x.last_param) for x in b] == [(28, 33), (8, 27), (0, 7)]
    assert [x.ready_after_group for x in b] == [7, 2, 0]
    fine = S.plan_buckets(groups, first_cap_bytes=0, cap_bytes=0)      # one bucket per layer group
    assert len(fine) == 9


def test_training_and_eval_print_formats():
    from ddpb200.train import test_model, train_model

    torch.manual_seed(0)
    m = ddpb200.VGG11()
    tr, te = create_data_loaders(0, 1, 8, root=None, synthetic=True, train_size=8 * 41, test_size=16)
    opt = ddpb200.SGD(m.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    out = io.StringIO()
    with redirect_stdout(out):
        train_model(m, tr, opt, ddpb200.CrossEntropyLoss(), 0)
        test_model(m, te, ddpb200.CrossEntropyLoss())
    lines = out.getvalue().splitlines()
    assert re.fullmatch(r"\[1,    20\] loss: \d+\.\d{3}", lines[0])
    assert re.fullmatch(r"\[1,    40\] loss: \d+\.\d{3}", lines[1])
    assert re.fullmatch(r"Total time for 1-39 iteration in ns: \d+", lines[2])
    assert re.fullmatch(r"Average time for 1-39 iteration in ns: \d+\.\d+", lines[3])
    assert re.fullmatch(r"Test set: Average loss: \d+\.\d{4}, Accuracy: \d+/16 \(\d+%\)", lines[-2])


def test_checkpoint_roundtrip_cpu(tmp_path):
    from ddpb200.train import DistributedDataParallel
    from ddpb200.utils.checkpoint import load_checkpoint, save_checkpoint

    torch.manual_seed(1)
    m = ddpb200.VGG11()
    save_checkpoint(tmp_path / "a.pt", m)
    wrapped = DistributedDataParallel(ddpb200.VGG11())
    assert list(wrapped.state_dict().keys()) == ["module." + k for k in EXPECTED_KEYS]
    load_checkpoint(tmp_path / "a.pt", wrapped)          # prefix is added on load
    for k, v in m.state_dict().items():
        assert torch.equal(v, wrapped.state_dict()["module." + k])


def test_entry_points_export_the_reference_names():
    """Every module-level name of the reference's scripts exists under the same
    path, so code written against them keeps importing."""
    import importlib.util
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    common = ["seed", "device", "create_data_loaders", "train_model", "test_model", "main"]
    dist_names = ["parse_arguments", "get_rank", "test_distributed_setup", "init_distributed_setup"]
    want = {
        "part1/main.py": common + ["batch_size"],
        "part2/part2a/main.py": common + dist_names + ["sync_gradients"],
        "part2/part2b/main.py": common + dist_names + ["sync_gradients"],
        "part3/main.py": common + dist_names + ["DDP"],
    }
    for rel in ["part1/model.py", "part2/part2a/model.py", "part2/part2b/model.py", "part3/model.py"]:
        want[rel] = ["_cfg", "_make_layers", "_VGG", "VGG11"]
    for rel, names in want.items():
        path = root / rel
        sys.path.insert(0, str(path.parent))
        try:
            spec = importlib.util.spec_from_file_location("entry_" + rel.replace("/", "_")[:-3], path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
        finally:
            sys.path.remove(str(path.parent))
            sys.modules.pop("model", None)   # each part has its own model.py
        missing = [n for n in names if not hasattr(mod, n)]
        assert not missing, f"{rel} lacks {missing}"
    assert mod.VGG11().fc1.out_features == 10



def test_clock_sampler_nvml_path_and_summary(monkeypatch):
    """The bench's clock sampler: NVML in-process when present (fake module here), throttle bits
    decoded into the reason names bench.py reports."""
    import sys
    import time
    import types

    from ddpb200.utils.timing import ClockSampler

    fake = types.ModuleType("pynvml")
    fake.NVML_CLOCK_SM = 1
    fake.nvmlInit = lambda: None
    fake.nvmlDeviceGetHandleByIndex = lambda i: ("h", i)
    fake.nvmlDeviceGetHandleByUUID = lambda u: ("h", u)
    fake.nvmlDeviceGetClockInfo = lambda h, k: 1965
    fake.nvmlDeviceGetMaxClockInfo = lambda h, k: 1965
    fake.nvmlDeviceGetPowerUsage = lambda h: 512000
    fake.nvmlDeviceGetCurrentClocksEventReasons = lambda h: 0x4   # software power cap
    monkeypatch.setitem(sys.modules, "pynvml", fake)
    with ClockSampler(0, period_s=0.002, uuid="abc") as c:
        time.sleep(0.1)
    out = c.summary()
    assert out["source"] == "nvml" and out["samples"] >= 5
    assert out["sm_mhz"] == 1965 and out["sm_max_mhz"] == 1965 and out["power_w_max"] == 512.0
    assert out["reasons"] == ["sw_power_cap"]


def test_reference_style_loop_and_both_train_model_orders():
    """A loop written against the reference's scripts (part1/main.py:65-80) runs unchanged on our loader,
    model, criterion and optimizer; train_model takes part1/part3's AND part2a/part2b's argument order
    (part2/part2b/main.py:106-107)."""
    from ddpb200.train import train_model

    torch.manual_seed(0)
    device = "cpu"
    model = ddpb200.VGG11()
    train_loader, _ = create_data_loaders(0, 1, 8, root=None, synthetic=True, train_size=24, test_size=8)
    optimizer = ddpb200.SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    criterion = ddpb200.CrossEntropyLoss().to(device)
    losses = []
    for batch_idx, (data, target) in enumerate(train_loader):
        data, target = data.to(device), target.to(device)
        optimizer.zero_grad()
        output = model(data)
        loss = criterion(output, target)
        loss.backward()
        optimizer.step()
        losses.append(loss.item())
    assert len(losses) == 3 and all(l == l for l in losses)
    out = io.StringIO()
    with redirect_stdout(out):
        train_model(model, train_loader, optimizer, criterion, 0)              # part1 / part3 order
        train_model(model, 0, 1, train_loader, optimizer, criterion, 0)        # part2a / part2b order
    with pytest.raises(TypeError):
        train_model(model, train_loader, optimizer)


def test_resume_restores_momentum_and_lr_changes_are_followed(tmp_path):
    """--save / --resume keep the SGD momentum: two steps straight == one step, save,
    load into a fresh model + optimizer, one more step.  And a changed param_groups[0]['lr'] is used
    by the next Trainer.step."""
    from ddpb200.utils.checkpoint import load_checkpoint, save_checkpoint

    x = torch.randint(0, 256, (8, 32, 32, 3), dtype=torch.uint8)
    y = torch.randint(0, 10, (8,))

    def make():
        torch.manual_seed(3)
        m = ddpb200.VGG11()
        o = ddpb200.SGD(m.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
        return m, o, ddpb200.Trainer(m, mode="single", batch_size=8, lr=0.05, optimizer=o)

    m1, o1, t1 = make()
    t1.step(x, y), t1.step(x, y)
    m2, o2, t2 = make()
    t2.step(x, y)
    save_checkpoint(tmp_path / "c.pt", m2, o2)
    m3, o3, t3 = make()
    load_checkpoint(tmp_path / "c.pt", m3, o3)
    t3.step(x, y)
    for a, b in zip(m1.parameters(), m3.parameters()):
        assert torch.allclose(a, b, atol=1e-6), "momentum was lost across save/resume"
    # LR schedule: the trainer follows the optimizer's param group
    before = [p.detach().clone() for p in m3.parameters()]
    o3.param_groups[0]["lr"] = 0.0
    o3.param_groups[0]["momentum"] = 0.0
    o3.param_groups[0]["weight_decay"] = 0.0
    t3.step(x, y)
    assert t3.lr == 0.0
    # lr = 0: parameters must not move (momentum buffers do not matter with lr 0)
    for a, b in zip(before, m3.parameters()):
        assert torch.equal(a, b)


def test_device_augment_option_on_the_cpu_path():
    """--device-augment / Trainer(device_augment=True): loaders stop drawing, the trainer draws (on CPU with a
    torch generator keyed by seed, rank and iteration): deterministic, and different between ranks."""
    import ddpb200
    from ddpb200.data import create_data_loaders

    assert cli.parse_all(False, ["--device-augment"]).device_augment is True
    assert cli.parse_all(False, []).device_augment is False
    train_loader, _ = create_data_loaders(0, 1, 8, root=None, synthetic=True, train_size=32, test_size=8, augment=False)
    img, lab, aug = next(iter(train_loader.raw()))
    assert aug is None and img.dtype == torch.uint8

    def run(rank, device_augment=True):
        torch.manual_seed(1)
        tr = ddpb200.Trainer(ddpb200.VGG11(), mode="single", batch_size=8, lr=0.01, device_augment=device_augment,
                             augment_seed=5, rank=rank)
        return [float(tr.step(img, lab)) for _ in range(2)]

    assert run(0) == run(0)
    assert run(0) != run(1)
    assert run(0) != run(0, device_augment=False)


def test_conv_planner_picks_the_measured_best_plans_at_batch_256():
    """The (tile width, split-K) planner is a cost model; profiles/r2_call37_conv_plan_sweep_in_kernel_split.txt is the
    measured sweep it was calibrated against (batch 256, 148 SMs).  Host-only: pins its choices for the VGG-11 layers
    to the measured best (or within 1 us of it), so a change of the cost terms that moves a layer shows up here."""
    from ddpb200.ops import kernels as K

    layers = [(32, 64, 64, 1), (16, 64, 128, 9), (8, 128, 256, 9), (8, 256, 256, 9), (4, 256, 512, 9), (4, 512, 512, 9),
              (2, 512, 512, 9), (2, 512, 512, 9)]
    fwd = [K.conv_plan_ex(256, s, s, ci, co, t, with_tail=True, with_sync=True)[1:] for s, ci, co, t in layers]
    assert fwd == [(64, 1, 0), (128, 1, 0), (256, 1, 0), (256, 1, 0), (128, 1, 0), (128, 1, 0), (128, 4, 1), (128, 4, 1)]
    dgrad = [K.conv_plan_ex(256, s, s, co, ci, 9)[1:] for s, ci, co, _ in layers[1:]]
    assert dgrad == [(64, 1, 0), (128, 1, 0), (256, 1, 0), (256, 4, 0), (128, 1, 0), (128, 4, 0), (128, 4, 0)]
    # the in-kernel sum is only planned with sync words, for the persistent kernel, with one item per CTA
    assert K.conv_plan_ex(256, 2, 2, 512, 512, 9, with_tail=True, with_sync=False)[3] == 0
    for n in (1, 24, 32, 85, 256, 1000):
        for s, ci, co, t in layers:
            kern, bn, splits, fused = K.conv_plan_ex(n, s, s, ci, co, t, with_tail=True, with_sync=True)
            tiles = (n * s * s + 127) // 128 * (co // bn)
            assert co % bn == 0 and 1 <= splits <= 8
            assert not fused or (kern == 2 and splits > 1 and tiles * splits <= K.sm_count() and tiles * 16 <= K.conv_split_sync_words())
            # the scratch query covers the plan with and without sync words / tail
            need = K.conv_fprop_scratch_elems(n, s, s, ci, co, t)
            assert need >= (splits * ((n * s * s + 127) // 128) * 128 * co if splits > 1 else 0)


def test_philox_reference_matches_the_published_vectors_and_draws_in_range():
    """ops/reference.py restates the device generator (Philox4x32-10, csrc/data_prep.cu) in numpy: the Random123
    known-answer vectors pin the algorithm; the crop / flip draws derived from it are in range, balanced,
    deterministic in (seed, step) and different across steps, seeds and images."""
    import numpy as np
    from ddpb200.ops.reference import draw_augmentation, philox4x32_10

    def kat(ctr, key):
        return [int(v) for v in philox4x32_10([np.uint32(c) for c in ctr], [np.uint32(k) for k in key])]

    assert kat([0, 0, 0, 0], [0, 0]) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    assert kat([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert kat([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0]) == \
        [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]
    a = draw_augmentation(4096, 4, 1234, 0)
    assert a.dtype == torch.int8 and a.shape == (4096, 4)
    assert int(a[:, :2].min()) == 0 and int(a[:, :2].max()) == 8 and set(a[:, 2].unique().tolist()) == {0, 1}
    assert int(a[:, 3].abs().max()) == 0
    for col, k in ((0, 9), (1, 9), (2, 2)):
        cnt = torch.bincount(a[:, col].long(), minlength=k).float()
        assert float((cnt - 4096 / k).abs().max()) < 6 * (4096 * (1 / k) * (1 - 1 / k)) ** 0.5
    assert torch.equal(a, draw_augmentation(4096, 4, 1234, 0))
    assert not torch.equal(a, draw_augmentation(4096, 4, 1234, 1))
    assert not torch.equal(a, draw_augmentation(4096, 4, 1235, 0))
    assert torch.equal(a[:100], draw_augmentation(100, 4, 1234, 0))          # an image's draw does not depend on the batch size


def test_c_header_library_and_ctypes_signatures_agree():
    """include/ddpb200.h is the C ABI: it must compile as C99 and as C++, declare exactly the `ddpb_*` symbols the
    library exports, and the ctypes signatures of ops/_native.py must have the declared number of parameters."""
    import re
    import shutil
    import subprocess
    from pathlib import Path

    from ddpb200.ops import _native as nat
    import ddpb200.parallel.comm  # noqa: F401  (registers its optional entry points)

    header = Path(__file__).resolve().parent.parent / "include" / "ddpb200.h"
    for cc, lang in (("gcc", ["-x", "c", "-std=c99"]), ("g++", ["-x", "c++"])):
        if shutil.which(cc):
            subprocess.run([cc, "-fsyntax-only", "-Wall", "-Werror", *lang, str(header)], check=True)
    text = re.sub(r"/\*.*?\*/", " ", header.read_text(), flags=re.S)
    declared = {}
    for m in re.finditer(r"\b(ddpb_\w+)\s*\(([^;{}]*?)\)\s*;", text, flags=re.S):
        params = m.group(2).strip()
        declared[m.group(1)] = 0 if params in ("", "void") else params.count(",") + 1
    nat.load()
    if shutil.which("nm"):
        out = subprocess.run(["nm", "-D", "--defined-only", str(nat.lib_path())], capture_output=True, text=True, check=True).stdout
        exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln and ln.split()[-1].startswith("ddpb_")}
        assert exported == set(declared), (sorted(exported - set(declared)), sorted(set(declared) - exported))
    signatures = {**nat._SIGNATURES, **nat._OPTIONAL}
    assert set(signatures) <= set(declared), sorted(set(signatures) - set(declared))
    wrong = {k: (len(v), declared[k]) for k, v in signatures.items() if len(v) != declared[k]}
    assert not wrong, wrong
    # and the same kinds (32-bit int, 64-bit int, float, pointer) position by position
    import ctypes as C

    def header_kind(param):
        if "*" in param:
            return "ptr"
        words = param.replace("const ", "").split()
        return {"int": "i32", "unsigned int": "i32", "long long": "i64", "unsigned long long": "i64",
                "float": "f32"}[" ".join(words[:-1])]

    def ctypes_kind(t):
        if t in (C.c_void_p, C.c_char_p) or issubclass(t, C._Pointer):
            return "ptr"
        return {C.c_int: "i32", C.c_uint: "i32", C.c_longlong: "i64", C.c_ulonglong: "i64", C.c_float: "f32"}[t]

    for m in re.finditer(r"\b(ddpb_\w+)\s*\(([^;{}]*?)\)\s*;", text, flags=re.S):
        name, params = m.group(1), m.group(2).strip()
        if name in signatures and params not in ("", "void"):
            assert [header_kind(q) for q in params.split(",")] == [ctypes_kind(t) for t in signatures[name]], name


def test_launch_count_of_a_step_follows_the_planner():
    """bench.py's gpu_launches is computed, not measured: VGGEngine.launches_per_train_step asks the planner which layers
    need a finalize / separate BatchNorm kernel.  Host only (stub engine): 8 conv+BN, 14 BatchNorm backward, 8 + 5
    wgrad (+ split reductions), 7 + 3 dgrad (+ split-K finalize), 2 classifier = 47 at batch 256."""
    from types import SimpleNamespace

    from ddpb200.engine import VGGEngine, _Layer

    layers, hw = [], 32
    for i, (ci, co, pool) in enumerate([(3, 64, True), (64, 128, True), (128, 256, False), (256, 256, True), (256, 512, False),
                                        (512, 512, True), (512, 512, False), (512, 512, True)]):
        layers.append(_Layer(hw, hw, ci, 64 if i == 0 else ci, co, 1 if i == 0 else 9, pool, 0, 0, 0, 0))
        hw = hw // 2 if pool else hw
    assert VGGEngine.launches_per_train_step(SimpleNamespace(max_batch=256, layers=layers)) == 47


def test_c_example_builds_against_the_header_and_runs(tmp_path):
    """examples/c_api_plan.c: plain C99 against include/ddpb200.h, linked with the library; the host-side planner
    entry points need no GPU."""
    import shutil
    import subprocess
    from pathlib import Path

    from ddpb200.ops import _native as nat

    if not shutil.which("gcc"):
        pytest.skip("no C compiler")
    nat.load()
    root = Path(__file__).resolve().parent.parent
    exe = tmp_path / "c_api_plan"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", f"-I{root / 'include'}", str(root / "examples" / "c_api_plan.c"), "-o",
                    str(exe), str(nat.lib_path()), f"-Wl,-rpath,{nat.lib_path().parent}"], check=True)
    out = subprocess.run([str(exe), "256"], capture_output=True, text=True, check=True).stdout
    assert "ABI 5" in out and "L7  2x2  512->512 | fwd: 128-wide tile x4 (summed in the kernel)" in out
    assert out.count("\n") == 9
