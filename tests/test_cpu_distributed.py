"""Multi-process CPU tests (gloo, 127.0.0.1): the three sync strategies against their
definitions and against each other, plus replica consistency."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent

WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.environ["DDPB_ROOT"])
import ddpb200
mode = sys.argv[1]
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
torch.set_num_threads(2)
torch.manual_seed(89395)
model = ddpb200.VGG11()
tr = ddpb200.Trainer(model, mode=mode, batch_size=4, rank=rank, world_size=world, lr=0.05)
g = torch.Generator().manual_seed(10 + rank)
img = torch.randint(0, 256, (4, 32, 32, 3), dtype=torch.uint8, generator=g)
lab = torch.randint(0, 10, (4,), generator=g)
# gradient after sync == mean over ranks of the local gradients
from ddpb200.ops.reference import augment_normalize_u8
ref = ddpb200.VGG11(); ref.load_state_dict(model.state_dict())
loss = torch.nn.functional.cross_entropy(ref(augment_normalize_u8(img, None)), lab); loss.backward()
local = torch.cat([p.grad.reshape(-1) for p in ref.parameters()])
gathered = [torch.empty_like(local) for _ in range(world)]
dist.all_gather(gathered, local)
expect = torch.stack(gathered).mean(0)
p_before = torch.cat([p.detach().reshape(-1).clone() for p in model.parameters()])
tr.step(img, lab)
got = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
assert torch.allclose(got, expect, atol=1e-6), (mode, (got - expect).abs().max())
# SGD first step: p1 = p0 - lr * (g + wd * p0)
p_after = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
assert torch.allclose(p_after, p_before - 0.05 * (expect + 1e-4 * p_before), atol=1e-6)
for _ in range(2):
    tr.step(img, lab)
flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
allp = [torch.empty_like(flat) for _ in range(world)]
dist.all_gather(allp, flat)
assert all(torch.equal(allp[0], q) for q in allp), "replicas diverged"
if rank == 0:
    torch.save(flat, os.environ["DDPB_OUT"] + f"/{mode}.pt")
    print("OK", mode)
dist.destroy_process_group()
'''


def _free_port() -> int:
    """A port nobody listens on right now (a fixed formula collided once in a while)."""
    import socket

    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as sk:
        sk.bind(("127.0.0.1", 0))
        return sk.getsockname()[1]


def _run(mode, world, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, DDPB_ROOT=str(ROOT), DDPB_OUT=str(tmp_path), OMP_NUM_THREADS="2")
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), mode]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    assert "OK " + mode in res.stdout


@pytest.mark.parametrize("mode", ["gather_scatter", "allreduce", "ddp"])
def test_sync_mode_matches_mean_of_local_gradients(mode, tmp_path):
    _run(mode, 2, tmp_path)


def test_three_modes_agree_with_each_other(tmp_path):
    import torch

    for mode in ("gather_scatter", "allreduce", "ddp"):
        _run(mode, 2, tmp_path)
    a, b, c = (torch.load(tmp_path / f"{m}.pt") for m in ("gather_scatter", "allreduce", "ddp"))
    # different summation orders: equal to rounding, not bitwise
    assert torch.allclose(a, b, atol=1e-5) and torch.allclose(b, c, atol=1e-5)


MISMATCH_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.environ["DDPB_ROOT"])
import ddpb200
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
torch.set_num_threads(2)
model = ddpb200.VGG11()
if rank == 1:   # a rank that built a different classifier
    model.fc1 = torch.nn.Linear(512, 11)
try:
    ddpb200.Trainer(model, mode="ddp", batch_size=4, rank=rank, world_size=world)
except RuntimeError as e:
    assert "parameter shapes differ" in str(e), e
    os.write(1, f"REFUSED-{rank};".encode())   # one write: the two ranks' lines must not interleave
dist.destroy_process_group()
'''


def test_ddp_refuses_ranks_with_different_models(tmp_path):
    """DistributedDataParallel's constructor verifies the parameter shapes across ranks before its
    broadcast; so does Trainer(mode='ddp')."""
    script = tmp_path / "mismatch.py"
    script.write_text(MISMATCH_WORKER)
    env = dict(os.environ, DDPB_ROOT=str(ROOT), OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    assert "REFUSED-0;" in res.stdout and "REFUSED-1;" in res.stdout

