"""Training-result parity over the reference's full run: ONE epoch (196 iterations of 256 images) plus
the test pass, in both arms, on the same data with the same seed and hyper-parameters.

The reference's deliverable is exactly this run (/root/reference/part1/main.py:128-130: train_model
then test_model; lr 0.1, momentum 0.9, weight decay 1e-4 at :124-125; seed 89395 at :14) and its
report compares the modes by the test loss / accuracy it prints (report p.4 Table 1).  Here:

  ours       ddpb200.train.train_model / test_model on the B200 kernels (bf16 activations, fp32 master
             weights and accumulation)
  reference  the unmodified /root/reference part1 train_model / test_model and model.py on the same GPU
             (fp32, cuDNN with torch's default TF32 convolutions)

Both read ddpb200.data.synthetic_cifar10 (learnable: class templates + noise; there is no network
for the real CIFAR-10) through the same loader, i.e. the same images, order, crop offsets and flips.

    python bench/parity_epoch.py [--impl both|ours|reference] [--train-size 50000] [--test-size 10000]
"""
import argparse
import contextlib
import io
import json
import re
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import ddpb200  # noqa: E402
from ddpb200.data import create_data_loaders  # noqa: E402

SEED = 89395
TEST_RE = re.compile(r"Test set: Average loss: ([0-9.]+), Accuracy: (\d+)/(\d+)")


class Tee(io.StringIO):
    def __init__(self, echo):
        super().__init__()
        self.echo = echo

    def write(self, s):
        if self.echo:
            sys.__stdout__.write(s)
        return super().write(s)


def seed_all():
    import random

    import numpy as np

    random.seed(SEED)
    torch.manual_seed(SEED)
    np.random.seed(SEED)


def many(a, device):
    """--repeat N: N runs per arm with seeds SEED, SEED+1, ... (model init; the data stays the same) -> the
    distribution of the end-of-epoch test loss / accuracy of each arm."""
    global SEED
    base = SEED
    out = {"ours": [], "reference": []}
    for i in range(a.repeat):
        for impl in ("ours", "reference"):
            SEED = base + i
            r = one(impl, a, device)
            out[impl].append((r["test_loss"], r["accuracy"], r["running_loss"][-1]))
    SEED = base
    summary = {}
    for impl, rows in out.items():
        tl = sorted(x[0] for x in rows)
        ac = sorted(x[1] for x in rows)
        summary[impl] = {"runs": rows, "test_loss_median": tl[len(tl) // 2], "test_loss_min": tl[0], "test_loss_max": tl[-1],
                         "accuracy_median": ac[len(ac) // 2], "accuracy_min": ac[0], "accuracy_max": ac[-1]}
    return summary


def run_ours(a, device):
    from ddpb200.train import test_model, train_model

    seed_all()
    train_loader, test_loader = create_data_loaders(0, 1, 256, root=None, synthetic=True, train_size=a.train_size,
                                                    test_size=a.test_size)
    model = ddpb200.VGG11().to(device)
    optimizer = ddpb200.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)
    criterion = ddpb200.CrossEntropyLoss().to(device)
    train_model(model, train_loader, optimizer, criterion, 0)
    test_model(model, test_loader, criterion)


def run_reference(a, device):
    import bench as headline  # repo-root bench.py: finds and imports the unmodified reference script

    ref_root = headline.find_reference()
    assert ref_root is not None, "reference scripts not found (python baseline/install_reference.py)"
    ref = headline.import_reference_main(ref_root, "part1")
    ref.device = str(device)
    seed_all()
    train_loader, test_loader = create_data_loaders(0, 1, 256, root=None, synthetic=True, train_size=a.train_size,
                                                    test_size=a.test_size)
    criterion = torch.nn.CrossEntropyLoss().to(device)
    model = ref.mdl.VGG11()
    model.to(device)
    optimizer = ref.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)
    ref.train_model(model, train_loader, optimizer, criterion, 0)
    ref.test_model(model, test_loader, criterion)


def one(impl, a, device):
    out = Tee(not a.quiet)
    if not a.quiet:
        print(f"===== {impl} =====")
    with contextlib.redirect_stdout(out):
        (run_ours if impl == "ours" else run_reference)(a, device)
    text = out.getvalue()
    m = TEST_RE.search(text)
    assert m, f"no 'Test set' line from the {impl} arm"
    losses = [float(x) for x in re.findall(r"\] loss: ([0-9.]+)", text)]
    return dict(impl=impl, test_loss=float(m.group(1)), correct=int(m.group(2)), total=int(m.group(3)),
                accuracy=100.0 * int(m.group(2)) / int(m.group(3)), running_loss=losses, log=text)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", choices=["both", "ours", "reference"], default="both")
    ap.add_argument("--train-size", type=int, default=50000)
    ap.add_argument("--test-size", type=int, default=10000)
    ap.add_argument("--quiet", action="store_true")
    ap.add_argument("--ref-fp32", action="store_true",
                    help="reference arm with cudnn.allow_tf32 = False (true fp32 convolutions instead of torch's default)")
    ap.add_argument("--spread", action="store_true",
                    help="also run the reference a second time in the other convolution precision: the difference "
                         "between its two runs is the yardstick for 'same result'")
    ap.add_argument("--json", default=None, help="write the summary here")
    ap.add_argument("--repeat", type=int, default=0, help="N seeds per arm: print the distribution instead of one pair")
    a = ap.parse_args()
    device = torch.device("cuda", 0)
    if a.repeat > 0:
        a.quiet = True
        summary = many(a, device)
        print(json.dumps(summary))
        if a.json:
            Path(a.json).write_text(json.dumps(summary, indent=1))
        return summary
    torch.backends.cudnn.allow_tf32 = not a.ref_fp32
    res = [one(i, a, device) for i in (["ours", "reference"] if a.impl == "both" else [a.impl])]
    keys = ("test_loss", "correct", "total", "accuracy", "running_loss")
    summary = {r["impl"]: {k: r[k] for k in keys} for r in res}
    if "reference" in summary:
        summary["reference"]["conv_precision"] = "fp32" if a.ref_fp32 else "tf32 (torch default)"
    if a.spread and "reference" in summary:
        torch.backends.cudnn.allow_tf32 = a.ref_fp32          # the other precision
        r2 = one("reference", a, device)
        summary["reference_other_precision"] = {k: r2[k] for k in keys}
        summary["reference_other_precision"]["conv_precision"] = "tf32 (torch default)" if a.ref_fp32 else "fp32"
        r1 = summary["reference"]
        summary["reference_spread"] = {"test_loss_rel": abs(r1["test_loss"] - r2["test_loss"]) / max(r1["test_loss"], 1e-12),
                                       "accuracy_points": abs(r1["accuracy"] - r2["accuracy"])}
    if len(res) == 2:
        o, r = res
        summary["delta"] = {"test_loss_rel": abs(o["test_loss"] - r["test_loss"]) / max(r["test_loss"], 1e-12),
                            "accuracy_points": abs(o["accuracy"] - r["accuracy"])}
    print(json.dumps(summary))
    if a.json:
        Path(a.json).write_text(json.dumps(summary, indent=1))
    return summary


if __name__ == "__main__":
    main()
