"""Entry point of part1 -- same command line and the same module-level names as
/root/reference/part1/main.py (README.md of the reference: `python main.py` for part1,
`python main.py --num-nodes N [--master-ip IP --master-port P --rank R]` on every node for the
others).  The logic lives in ddpb200; this file selects the mode and re-exports the
reference's function names so code written against the reference scripts keeps working."""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))

import model as mdl  # noqa: E402,F401  (local import, as in the reference)
from ddpb200.data import create_data_loaders  # noqa: E402,F401
from ddpb200.train import run, seed, test_model, train_model  # noqa: E402,F401

device = "cuda"  # the reference hard-codes "cpu" (main.py:15); pass --device cpu for that
batch_size = 256  # main.py:17


def main():
    run("part1")


if __name__ == "__main__":
    main()
