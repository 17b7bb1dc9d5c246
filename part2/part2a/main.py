"""Entry point of part2a -- same command line as /root/reference/part2/part2a/main.py
(README.md of the reference: `python main.py` for part1, `python main.py --num-nodes N
[--master-ip IP --master-port P --rank R]` on every node for the others).
All logic lives in ddpb200.train.run; this file only selects the mode."""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "../..")))

from ddpb200.train import run  # noqa: E402

if __name__ == "__main__":
    run("part2a")
