"""Headline benchmark: VGG-11 on CIFAR-10-shaped data, images/s and ms/iteration (max over ranks).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--mode ddp|allreduce|gather_scatter|single] [--scaling weak|strong]

Metric and configs are BASELINE.json's.  The default is the flagship config: part3-style
(bucketed, overlapped) data parallel with per-GPU batch 256 -- weak scaling, i.e. at N=1
this is the reference's part1 step (batch 256), at N=8 the global batch is 2048.
``--scaling strong`` keeps the global batch at 256 (per-GPU 256/N), the setting of the
reference's report.  For N>1 launch under torchrun (the driver does); a bare
``python bench.py --gpus N`` re-executes itself under torchrun.

Two numbers per run:
  value  device time of K steps (CUDA events, barrier + synchronize on both sides, max over
         ranks) with the batch already on the device;
  e2e    the same K steps through the public API from pinned host memory: every step copies
         its uint8 images + labels host->device and reads the loss back device->host.

``--impl reference`` runs the UNMODIFIED reference scripts copied to baseline/_ref
(baseline/install_reference.py) through their own ``train_model`` / ``DDP`` code path on
cuda + NCCL, on synthetic data of the same shape, timed the same way.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

PUBLISHED_IMAGES_PER_S = {1: 138.0, 2: 256 / 0.86, 3: 256 / 0.76, 4: 385.5}  # BASELINE.md §1 (part1 / DDP rows)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--mode", choices=["ddp", "allreduce", "gather_scatter", "single"], default="ddp")
    ap.add_argument("--scaling", choices=["weak", "strong"], default="weak")
    ap.add_argument("--per-gpu-batch", type=int, default=256)
    ap.add_argument("--granularity", choices=["fused", "param", "param_launches"], default="fused")
    ap.add_argument("--no-multicast", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-pdl", action="store_true", help="disable programmatic dependent launch")
    ap.add_argument("--no-overlap-wgrad", action="store_true", help="weight gradients on the main stream")
    ap.add_argument("--bn-fused", type=int, default=None, help="BatchNorm backward: 3 deterministic two-kernel (default), 1/2 cooperative fused, 0 two-pass atomics")
    ap.add_argument("--conv-mode", type=int, default=-1,
                    help="conv kernel of fprop/dgrad: -1 automatic (persistent), 1 CTA pair, 3 persistent CTA pair")
    ap.add_argument("--comm-blocks", type=int, default=None)
    ap.add_argument("--bucket-layout", choices=["per_layer", "reference", "balanced"], default="balanced")
    ap.add_argument("--comm-threads", type=int, default=None)
    ap.add_argument("--bucket-mb", type=float, default=None, help="explicit bucket cap in MiB (first bucket: min(1, cap))")
    ap.add_argument("--no-fuse-sgd", action="store_true", help="run SGD as its own full-width kernel after the sync")
    ap.add_argument("--no-overlap-sgd", action="store_true",
                    help="one SGD pass after backward instead of per-group / per-bucket updates on a side stream")
    ap.add_argument("--symm-backend", default=None)
    ap.add_argument("--no-extra", action="store_true",
                    help="skip the extra_configs block (strong-scaling runs of BASELINE.json configs 2-4)")
    return ap.parse_args()


def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def maybe_reexec_under_torchrun(a):
    if a.gpus > 1 and "WORLD_SIZE" not in os.environ:
        port = 29500 + (os.getpid() % 2000)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), str(Path(__file__).resolve()), *sys.argv[1:]]
        raise SystemExit(subprocess.call(cmd))


_JSON_FD = None


def isolate_stdout():
    """Only the result line may appear on stdout: libraries (NCCL prints its version there)
    are redirected to stderr, the JSON goes to the saved descriptor."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = json.dumps(obj) + "\n"
    if _JSON_FD is None:
        sys.stdout.write(line)
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, line.encode())


def synthetic_host_batches(n_batches, batch, seed, pinned=True):
    import torch

    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n_batches):
        img = torch.randint(0, 256, (batch, 32, 32, 3), dtype=torch.uint8, generator=g)
        lab = torch.randint(0, 10, (batch,), dtype=torch.int64, generator=g)
        aug = torch.stack([torch.randint(0, 9, (batch,), generator=g), torch.randint(0, 9, (batch,), generator=g),
                           torch.randint(0, 2, (batch,), generator=g), torch.zeros(batch, dtype=torch.int64)],
                          1).to(torch.int8)
        if pinned:
            img, lab, aug = img.pin_memory(), lab.pin_memory(), aug.pin_memory()
        out.append((img, lab, aug))
    return out


# =========================================================================================
# our arm
# =========================================================================================
def _measure_ours(a, mode, scaling, steps, warmup, granularity, world, rank, local_rank, device):
    """Builds a Trainer for one (mode, scaling) config, times ``steps`` steps twice (batch resident on the
    device / end to end from pinned host memory) and returns the numbers (max over ranks)."""
    import torch
    import torch.distributed as dist

    import ddpb200
    from ddpb200.utils.timing import ClockSampler, max_over_ranks

    per_gpu = a.per_gpu_batch if scaling == "weak" else max(1, 256 // world)
    mode = mode if world > 1 else "single"
    torch.manual_seed(89395)
    model = ddpb200.VGG11().to(device)
    trainer = ddpb200.Trainer(model, mode=mode, batch_size=per_gpu, lr=0.1, momentum=0.9, weight_decay=1e-4,
                              rank=rank, world_size=world, granularity=granularity,
                              use_multicast=False if a.no_multicast else "auto", use_graph=not a.no_graph,
                              comm_blocks=a.comm_blocks, symm_backend=a.symm_backend, bucket_layout=a.bucket_layout,
                              fuse_sgd=not a.no_fuse_sgd, comm_threads=a.comm_threads,
                              overlap_wgrad=not a.no_overlap_wgrad, overlap_sgd=not a.no_overlap_sgd,
                              first_bucket_bytes=None if a.bucket_mb is None else int(min(1.0, a.bucket_mb) * 2**20),
                              bucket_bytes=None if a.bucket_mb is None else int(a.bucket_mb * 2**20))
    if a.bn_fused is not None:
        from ddpb200.ops import kernels as _K2
        _K2.bn_bwd_set_fused(a.bn_fused)
    n_rot = 8
    host = synthetic_host_batches(n_rot, per_gpu, seed=1234 + rank)
    dev_batches = [(i.to(device), l.to(device), g.to(device)) for i, l, g in host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    def timed(batches, k, readback):
        barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        s.record()
        last = None
        if readback:
            # the public training loop (ddpb200.train.train_model): prefetch the next pinned batch on the
            # copy stream while the current step computes; every step's loss is read on the host
            trainer.prefetch(*batches[0])
        for it in range(k):
            img, lab, aug = batches[it % n_rot]
            trainer.step(img, lab, aug)
            if readback:
                if it + 1 < k:
                    trainer.prefetch(*batches[(it + 1) % n_rot])
                last = trainer.last_loss()   # device -> host: 8 bytes stored by the step into pinned memory
        e.record()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        return s.elapsed_time(e), wall, last

    for it in range(max(warmup, 3)):
        img, lab, aug = dev_batches[it % n_rot]
        trainer.step(img, lab, aug)
    first_loss = float(trainer.engine.loss.item())
    with ClockSampler(local_rank, uuid=_gpu_uuid(local_rank)) as clocks:
        dev_ms, _, _ = timed(dev_batches, steps, readback=False)
    # end to end from pinned host memory (a few extra warm-up steps for the copy path)
    for it in range(3):
        img, lab, aug = host[it % n_rot]
        trainer.step(img, lab, aug)
        trainer.last_loss()
    e2e_dev_ms, e2e_wall_ms, last_loss = timed(host, steps, readback=True)
    dev_ms = max_over_ranks(dev_ms, device)
    e2e_ms = max_over_ranks(max(e2e_dev_ms, e2e_wall_ms), device)
    eng = trainer.engine
    act_mb = sum(t.numel() * 2 for ts in (eng.conv_out, eng.act, eng.dconv, eng.dact) for t in ts) / 1e6
    out = dict(mode=mode, scaling=scaling, per_gpu=per_gpu, global_batch=per_gpu * world, steps=steps,
               warmup=max(warmup, 3), ms_step=dev_ms / steps, e2e_ms_step=e2e_ms / steps,
               launches=trainer.kernel_launches_per_step(), clocks=clocks.summary(), first_loss=first_loss,
               last_loss=last_loss, h2d=world * sum(t.numel() * t.element_size() for t in host[0]),
               ws_mb=act_mb + eng.operand0.numel() * 2 / 1e6 + 3 * eng.arena.numel * 4 / 1e6,
               sync=None if world == 1 else {"backend": trainer.symm.backend, "multicast": trainer.use_mc,
                                             "granularity": granularity,
                                             "bucket_layout": a.bucket_layout if mode == "ddp" else None,
                                             "comm_blocks": trainer.comm_blocks,
                                             "buckets_floats": [b.numel for b in trainer.buckets] if mode == "ddp" else None})
    trainer.close()
    del trainer, model
    torch.cuda.empty_cache()
    return out


def _extra_list(world):
    """BASELINE.json configs 2-4: the reference report's strong-scaling experiment (global batch 256 =
    int(256 / world) per rank, /root/reference/part2/part2b/main.py:177) in each sync mode."""
    if world == 1:
        return []
    return [("strong", "ddp", "fused"), ("allreduce_strong", "allreduce", "fused"),
            ("allreduce_param_strong", "allreduce", "param"), ("gather_scatter_strong", "gather_scatter", "fused")]


def run_ours(a):
    import torch
    import torch.distributed as dist

    rank, local_rank, world = env_rank()
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world}"
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    if a.no_pdl:
        os.environ["DDPB_PDL"] = "0"     # (the Trainer leaves an explicit setting alone)
        from ddpb200.ops import kernels as _K
        _K.set_pdl(False)
    if a.conv_mode >= 0:
        from ddpb200.ops import kernels as _K3
        _K3.conv_force(a.conv_mode, -1, -1)
    m = _measure_ours(a, a.mode, a.scaling, a.steps, a.warmup, a.granularity, world, rank, local_rank, device)
    extras = {}
    if not a.no_extra:
        for name, mode, gran in _extra_list(world):
            x = _measure_ours(a, mode, "strong", min(a.steps, 30), min(a.warmup, 5), gran, world, rank, local_rank, device)
            extras[name] = {"mode": x["mode"], "granularity": gran, "global_batch": x["global_batch"],
                            "per_gpu_batch": x["per_gpu"], "ms_per_step": round(x["ms_step"], 4),
                            "images_per_s": round(x["global_batch"] / (x["ms_step"] * 1e-3), 1),
                            "e2e_ms_per_step": round(x["e2e_ms_step"], 4),
                            "e2e_images_per_s": round(x["global_batch"] / (x["e2e_ms_step"] * 1e-3), 1),
                            "steps": x["steps"], "warmup": x["warmup"]}
    if rank == 0:
        mode = m["mode"]
        value = m["global_batch"] / (m["ms_step"] * 1e-3)
        e2e_value = m["global_batch"] / (m["e2e_ms_step"] * 1e-3)
        pub = PUBLISHED_IMAGES_PER_S.get(world)
        par = "single" if world == 1 else {"ddp": "dp", "allreduce": "dp-allreduce", "gather_scatter": "dp-gather-scatter"}[mode]
        emit({
            "metric": "VGG-11 CIFAR-10 images/sec (ms/iter max over ranks)", "value": round(value, 1),
            "unit": "images/s", "n_gpus": world, "steps": m["steps"], "warmup": m["warmup"],
            "ms_per_step": round(m["ms_step"], 4), "higher_is_better": True, "scaling": a.scaling,
            "vs_baseline": round(value / pub, 2) if pub else None, "dtype": "bf16", "data": "synthetic", "impl": "ours",
            "config": {"model": "VGG-11 (3x32x32, 10 classes, BatchNorm batch stats)", "global_batch": m["global_batch"],
                       "per_gpu_batch": m["per_gpu"], "seq_len": None, "parallelism": f"{par}{world}", "mode": mode,
                       "optimizer": "SGD lr=0.1 momentum=0.9 wd=1e-4 (inside the timed step)",
                       "sync": m["sync"],
                       "cuda_graph": not a.no_graph, "programmatic_dependent_launch": bool(_pdl_state()),
                       "conv_mode": "automatic" if a.conv_mode < 0 else a.conv_mode,
                       "l2": f"no explicit flush: working set {m['ws_mb']:.0f} MB/step > 126 MB L2; inputs rotate over 8 device batches",
                       "baseline_ref": "BASELINE.md §1: published CPU/Gloo images/s for this worker count"},
            "e2e": {"value": round(e2e_value, 1), "unit": "images/s", "ms_per_step": round(m["e2e_ms_step"], 4),
                    "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": world * 8,
                    "api": "ddpb200.Trainer.prefetch(next pinned uint8 NHWC, labels, aug) / step() / last_loss() -- "
                           "the loop of ddpb200.train.train_model; every step copies its batch host->device (copy "
                           "stream, overlapping the previous step) and stores its loss word (sequence + fp32) into "
                           "pinned host memory, which the host reads every step"},
            "gpu_launches": m["launches"] * m["steps"], "gpu_launches_per_step": m["launches"],
            "clocks": m["clocks"], "loss_first": m["first_loss"], "loss_last": m["last_loss"],
            "extra_configs": extras,
        })
    if world > 1:
        dist.destroy_process_group()


# =========================================================================================
# reference arm: the unmodified scripts from baseline/_ref, their own train_model
# =========================================================================================
def find_reference():
    for cand in (ROOT / "baseline" / "_ref", Path("/root/reference")):
        if (cand / "part1" / "main.py").exists() and (cand / "part3" / "main.py").exists():
            return cand
    return None


def import_reference_main(ref_root: Path, part: str):
    """Imports baseline/_ref/<part>/main.py as a module, unmodified -- except part2/part2a, which does not
    parse as committed (a stray '/' after the closing bracket at line 70): for that file the
    one character is dropped IN MEMORY (the copy on disk stays byte-identical; baseline/PATCHES.md)."""
    d = ref_root / part
    sys.path.insert(0, str(d))          # the script does `import model as mdl`
    sys.modules.pop("model", None)
    name = f"ref_{part.replace('/', '_')}_main"
    if part == "part2/part2a":
        import types

        src = (d / "main.py").read_text()
        lines = src.splitlines(keepends=True)
        assert lines[69].rstrip("\r\n") == "            ])/", f"unexpected line 70: {lines[69]!r}"
        lines[69] = lines[69].replace("])/", "])", 1)
        mod = types.ModuleType(name)
        mod.__file__ = str(d / "main.py")
        exec(compile("".join(lines), str(d / "main.py"), "exec"), mod.__dict__)
        return mod
    spec = importlib.util.spec_from_file_location(name, d / "main.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class TimedLoader:
    """Feeds the reference's train_model (which iterates ``enumerate(train_loader)``) W+K
    normalised fp32 CPU batches from pinned memory -- what its DataLoader(pin_memory=True)
    yields -- and brackets batches W..W+K-1 with barrier + synchronize + CUDA events."""

    def __init__(self, batches, warmup, steps, barrier):
        self.batches, self.warmup, self.steps, self.barrier = batches, warmup, steps, barrier
        self.start = self.end = None
        self.wall0 = self.wall_ms = None

    def __len__(self):
        return self.warmup + self.steps

    def __iter__(self):
        import torch

        for i in range(self.warmup + self.steps):
            if i == self.warmup:
                self.barrier()
                self.start = torch.cuda.Event(enable_timing=True)
                self.wall0 = time.perf_counter()
                self.start.record()
            yield self.batches[i % len(self.batches)]
        self.end = torch.cuda.Event(enable_timing=True)
        self.end.record()
        self.barrier()
        self.wall_ms = (time.perf_counter() - self.wall0) * 1e3


_REF_MODS = {}


def _measure_reference(a, ref_root, part, scaling, steps, warmup, world, rank, local_rank, device):
    """Runs the reference's own train_model (its DDP / sync_gradients, NCCL, cuDNN) for one config."""
    import contextlib
    import io
    import random

    import numpy as np
    import torch
    import torch.distributed as dist

    from ddpb200.ops.reference import CIFAR_MEAN, CIFAR_STD
    from ddpb200.utils.timing import ClockSampler, max_over_ranks

    per_gpu = a.per_gpu_batch if scaling == "weak" else max(1, 256 // world)
    # get_rank() is evaluated when the module's parse_arguments runs, not at import: importing is safe
    ref = _REF_MODS.get(part)
    if ref is None:
        ref = _REF_MODS[part] = import_reference_main(ref_root, part)
    ref.device = str(device)                      # the scripts read this module global (main.py:15)
    random.seed(ref.seed)
    torch.manual_seed(ref.seed)
    np.random.seed(ref.seed)
    criterion = torch.nn.CrossEntropyLoss().to(device)
    model = ref.mdl.VGG11()
    model.to(device)
    net = ref.DDP(model) if part == "part3" else model
    optimizer = ref.optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)
    mean, std = CIFAR_MEAN.view(1, 3, 1, 1), CIFAR_STD.view(1, 3, 1, 1)
    host = []
    for img, lab, _ in synthetic_host_batches(8, per_gpu, seed=1234 + rank, pinned=False):
        x = ((img.permute(0, 3, 1, 2).float() / 255.0 - mean) / std).contiguous().pin_memory()
        host.append((x, lab.pin_memory()))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    warm = max(warmup, 3)
    loader = TimedLoader(host, warm, steps, barrier)
    sink = io.StringIO()
    with ClockSampler(local_rank, uuid=_gpu_uuid(local_rank)) as clocks, contextlib.redirect_stdout(sink):
        if part in ("part1", "part3"):
            ref.train_model(net, loader, optimizer, criterion, 0)
        else:
            ref.train_model(net, rank, world, loader, optimizer, criterion, 0)
    dev_ms = loader.start.elapsed_time(loader.end)
    ms = max_over_ranks(max(dev_ms, loader.wall_ms), device)
    out = dict(part=part, per_gpu=per_gpu, global_batch=per_gpu * world, steps=steps, warmup=warm, ms_step=ms / steps,
               clocks=clocks.summary(), h2d=world * sum(t.numel() * t.element_size() for t in host[0]))
    del net, model, optimizer
    torch.cuda.empty_cache()
    return out


def run_reference(a):
    ref_root = find_reference()
    if ref_root is None:
        emit({"impl": "reference", "unavailable": "reference scripts not found under baseline/_ref or /root/reference "
                                                  "(pip cannot install them: no setup.py/pyproject.toml)"})
        return
    try:
        import torch
        import torch.distributed as dist
        import torchvision  # noqa: F401  (the reference imports it at module level)
    except Exception as e:  # noqa: BLE001
        emit({"impl": "reference", "unavailable": f"import failed: {e}"})
        return
    rank, local_rank, world = env_rank()
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    parts = {"ddp": "part3", "allreduce": "part2/part2b", "single": "part3", "gather_scatter": "part2/part2a"}
    part = "part1" if world == 1 else parts[a.mode]
    if world > 1:
        # the reference's own bootstrap (identical in 2a / 2b / 3: part2/part2b/main.py:52-58), with its
        # `backend` parameter set to nccl
        boot = import_reference_main(ref_root, "part3")
        _REF_MODS["part3"] = boot
        boot.init_distributed_setup(os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"], rank, world, "nccl")
    m = _measure_reference(a, ref_root, part, a.scaling, a.steps, a.warmup, world, rank, local_rank, device)
    extras = {}
    if not a.no_extra:
        for name, mode, gran in _extra_list(world):
            if gran == "param" and mode == "allreduce":
                continue      # the reference's part2b IS per parameter: one row ("allreduce_strong") covers it
            x = _measure_reference(a, ref_root, parts[mode], "strong", min(a.steps, 30), min(a.warmup, 5), world, rank,
                                   local_rank, device)
            extras[name] = {"script": f"baseline/_ref/{x['part']}/main.py train_model()", "global_batch": x["global_batch"],
                            "per_gpu_batch": x["per_gpu"], "ms_per_step": round(x["ms_step"], 4),
                            "images_per_s": round(x["global_batch"] / (x["ms_step"] * 1e-3), 1),
                            "e2e_ms_per_step": round(x["ms_step"], 4),
                            "e2e_images_per_s": round(x["global_batch"] / (x["ms_step"] * 1e-3), 1),
                            "steps": x["steps"], "warmup": x["warmup"],
                            "patch": "line 70: stray '/' dropped in memory (baseline/PATCHES.md)" if "part2a" in x["part"] else None}
        if "allreduce_strong" in extras:
            extras["allreduce_param_strong"] = dict(extras["allreduce_strong"])
    if rank == 0:
        ms_step = m["ms_step"]
        value = m["global_batch"] / (ms_step * 1e-3)
        pub = PUBLISHED_IMAGES_PER_S.get(world)
        emit({
            "metric": "VGG-11 CIFAR-10 images/sec (ms/iter max over ranks)", "value": round(value, 1),
            "unit": "images/s", "n_gpus": world, "steps": m["steps"], "warmup": m["warmup"], "ms_per_step": round(ms_step, 4),
            "higher_is_better": True, "scaling": a.scaling, "vs_baseline": round(value / pub, 2) if pub else None,
            "dtype": "fp32 (cuDNN TF32 convolutions, torch defaults)", "data": "synthetic", "impl": "reference",
            "config": {"model": "VGG-11 (reference model.py, unmodified)", "global_batch": m["global_batch"],
                       "per_gpu_batch": m["per_gpu"], "seq_len": None, "parallelism": f"{'single' if world == 1 else 'dp'}{world}",
                       "script": f"baseline/_ref/{part}/main.py train_model()", "backend": None if world == 1 else "nccl",
                       "l2": "no explicit flush; inputs rotate over 8 pinned host batches"},
            "e2e": {"value": round(value, 1), "unit": "images/s", "ms_per_step": round(ms_step, 4),
                    "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": world * 4,
                    "note": "the reference's loop copies each batch host->device and calls loss.item() itself, "
                            "so its only number is end to end"},
            "gpu_launches": 0, "clocks": m["clocks"], "extra_configs": extras,
        })
    if world > 1:
        dist.destroy_process_group()


def _gpu_uuid(index):
    """UUID of cuda:index (NVML indices ignore CUDA_VISIBLE_DEVICES; the UUID does not)."""
    try:
        import torch

        return str(torch.cuda.get_device_properties(index).uuid)
    except Exception:  # noqa: BLE001
        return None


def _pdl_state():
    from ddpb200.ops import kernels as _K
    return _K.get_pdl()


def main():
    a = parse()
    maybe_reexec_under_torchrun(a)
    isolate_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
