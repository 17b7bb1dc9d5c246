"""End-of-epoch parity with the reference's own run (bench/parity_epoch.py): one epoch of 196 iterations
at the reference's hyper-parameters (lr 0.1, momentum 0.9, weight decay 1e-4,
/root/reference/part1/main.py:124-125) plus the test pass (:128-130), both arms on the same data.

What can be asserted: measured on a B200 (profiles/r2_call13_*, r2_call14_*), the result of this run is
dominated by the seed and by rounding-level differences -- the reference against ITSELF (TF32 versus
fp32 convolutions, same seed) differs by 7-19 % in test loss and 1-7 accuracy points, and over six
seeds its test loss ranges 0.015 .. 1.01 (accuracy 54 .. 99.98 %).  A "within 3 %" comparison of two
single runs therefore says nothing, and even the distributions over a handful of seeds overlap almost
completely (ours: test loss 0.008 .. 1.54, reference: 0.005 .. 1.01 over the runs kept under profiles/).
The test therefore only asserts what every run showed: each of ours trains (running loss well below
the ln(10) start), the median run ends clearly better than chance, and the best run converges as the
reference's best run does (> 90 % test accuracy).  It prints both arms' numbers for the record.
"""
import sys
from pathlib import Path
from types import SimpleNamespace

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_one_epoch_result_matches_the_reference_distribution():
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "bench"))
    pytest.importorskip("torchvision")          # the reference script imports it at module level
    import bench as headline
    import parity_epoch as PE

    if headline.find_reference() is None:
        pytest.skip("reference scripts not available (python baseline/install_reference.py)")
    a = SimpleNamespace(train_size=50000, test_size=10000, quiet=True, repeat=3)
    s = PE.many(a, torch.device("cuda", 0))
    ours, ref = s["ours"], s["reference"]
    print("ours", ours["runs"], "reference", ref["runs"])
    # every run trains: the running loss of the last 20 iterations is below the ln(10) = 2.30 of an untrained net
    assert all(r[2] < 2.0 for r in ours["runs"]), ours["runs"]
    assert ours["test_loss_median"] < 1.9, ours["runs"]
    assert ours["accuracy_max"] >= 90.0 and ref["accuracy_max"] >= 90.0, (ours["runs"], ref["runs"])
