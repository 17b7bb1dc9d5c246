"""Timeline of one captured training step: a one-thread kernel stamps %globaltimer after every
kernel group (VGGEngine.enable_timeline).  The stamps serialise the step a little (they are not
programmatic-launch capable and add ~2 us each), so read the numbers as a breakdown, not as the
step time -- bench.py measures that.

    python bench/step_timeline.py [--batch 256] [--no-overlap-wgrad]
"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import ddpb200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--no-overlap-wgrad", action="store_true")
    ap.add_argument("--mode", default="ddp", help="sync mode when launched under torchrun with several ranks")
    ap.add_argument("--bucket-layout", default="balanced")
    a = ap.parse_args()
    import os

    import torch.distributed as dist

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    model = ddpb200.VGG11().to(dev)
    tr = ddpb200.Trainer(model, mode=a.mode if world > 1 else "single", batch_size=a.batch, lr=0.1, momentum=0.9,
                         weight_decay=1e-4, overlap_wgrad=not a.no_overlap_wgrad, rank=rank, world_size=world,
                         bucket_layout=a.bucket_layout)
    tr.engine.enable_timeline()
    x = torch.randint(0, 256, (a.batch, 32, 32, 3), dtype=torch.uint8).pin_memory()
    y = torch.randint(0, 10, (a.batch,)).pin_memory()
    aug = torch.zeros(a.batch, 4, dtype=torch.int8).pin_memory()
    for _ in range(30):
        tr.step(x, y, aug)
    torch.cuda.synchronize()
    tl = tr.engine.timeline()
    if rank == 0:
        prev = {}
        print(f"{'mark':34s} {'t (us)':>9s} {'since previous mark on the same stream':>10s}")
        for lab, t in tl:
            lane = "side" if "(side)" in lab else ("comm" if "(comm)" in lab else "main")
            d = "" if lane not in prev else f"{t - prev[lane]:8.1f}"
            print(f"{lab:34s} {t:9.1f} {d}")
            prev[lane] = t
        print(f"total {max(t for _, t in tl) - tl[0][1]:.1f} us (with stamps), ranks {world}")
    tr.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
