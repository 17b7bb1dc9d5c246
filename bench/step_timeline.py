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
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = ddpb200.VGG11().to(dev)
    tr = ddpb200.Trainer(model, mode="single", batch_size=a.batch, lr=0.1, momentum=0.9, weight_decay=1e-4,
                         overlap_wgrad=not a.no_overlap_wgrad)
    tr.engine.enable_timeline()
    x = torch.randint(0, 256, (a.batch, 32, 32, 3), dtype=torch.uint8).pin_memory()
    y = torch.randint(0, 10, (a.batch,)).pin_memory()
    aug = torch.zeros(a.batch, 4, dtype=torch.int8).pin_memory()
    for _ in range(30):
        tr.step(x, y, aug)
    torch.cuda.synchronize()
    tl = tr.engine.timeline()
    main_prev = side_prev = None
    print(f"{'mark':28s} {'t (us)':>9s} {'since previous mark on the same stream':>10s}")
    for lab, t in tl:
        side = "(side)" in lab
        prev = side_prev if side else main_prev
        d = "" if prev is None else f"{t - prev:8.1f}"
        print(f"{lab:28s} {t:9.1f} {d}")
        if side:
            side_prev = t
        else:
            main_prev = t
    print(f"total {tl[-1][1] - tl[0][1]:.1f} us (with stamps)")


if __name__ == "__main__":
    main()
