"""Command line of the four entry points.

The reference's distributed scripts take exactly four flags
(/root/reference/part2/part2b/main.py:20-32, README.md:12-19):

    --master-ip    str, default "10.10.1.1"
    --master-port  str, default '4000' (text: it goes into os.environ)
    --num-nodes    int, no default (required in practice)
    --rank         int, default = 5th character of the host name (``node0`` -> 0)

They are kept verbatim.  One behavioural fix: the reference evaluates the
host-name default eagerly, so it crashes on any host whose name has no digit at index 4 even
when --rank is given; here the host name is consulted only if --rank is absent.  Everything
else (--device, --data-root, --synthetic, ...) is additive and optional.
"""
from __future__ import annotations

import argparse
import os


def get_rank() -> int:
    """Rank from the computer name, e.g. ``node3`` -> 3 (part2b/main.py:35-39)."""
    name = os.uname().nodename
    try:
        return int(name[4])
    except (IndexError, ValueError) as e:
        raise SystemExit(f"cannot infer --rank from host name {name!r} (expected e.g. 'node0'); pass --rank") from e


def build_parser(distributed: bool) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser()
    if distributed:
        p.add_argument("--master-ip", dest="master_ip", default="10.10.1.1", type=str,
                       help="Speficy master ip. Default is 10.10.1.1")
        p.add_argument("--master-port", dest="master_port", default="4000", type=str,
                       help="Specify master port. Default is 4000.")
        p.add_argument("--num-nodes", dest="size", type=int,
                       help="Specify the number of nodes to distribute training over.")
        p.add_argument("--rank", dest="rank", default=None, type=int,
                       help="Specify the rank for this machine. The default takes the number from the computer name.")
        p.add_argument("--backend", default=None, help="process-group backend (default: nccl with GPUs, else gloo)")
        p.add_argument("--spawn", action="store_true",
                       help="convenience: start all --num-nodes ranks of this host from one command")
        p.add_argument("--sync-granularity", choices=["fused", "param", "param_launches"], default="fused",
                       help="'param' launches one sync kernel per parameter tensor, like the reference's loop")
        p.add_argument("--no-multicast", action="store_true", help="do not use NVLS multimem instructions")
    p.add_argument("--device", default=None, help="cuda | cpu (default: cuda when available)")
    p.add_argument("--data-root", default=None, help="directory holding cifar-10-batches-py (reference: ./../data)")
    p.add_argument("--synthetic", action="store_true", help="force the synthetic CIFAR-shaped dataset")
    p.add_argument("--epochs", type=int, default=1)
    p.add_argument("--batch-size", type=int, default=256, help="GLOBAL batch (reference: 256)")
    p.add_argument("--max-iters", type=int, default=None, help="stop each epoch after this many iterations")
    p.add_argument("--train-size", type=int, default=None, help="synthetic train-set size")
    p.add_argument("--test-size", type=int, default=None, help="synthetic test-set size")
    p.add_argument("--device-augment", action="store_true",
                   help="draw the random crop / flip parameters on the GPU inside the step (Philox) instead of on the host")
    p.add_argument("--no-graph", action="store_true", help="launch kernels one by one instead of a CUDA graph")
    p.add_argument("--save", default=None, help="write a state_dict checkpoint here after training")
    p.add_argument("--resume", default=None, help="load a checkpoint written by --save before training")
    return p


def parse_arguments(argv=None):
    """Reference-compatible: returns (master_ip, master_port, rank, size)."""
    a = build_parser(True).parse_args(argv)
    rank = a.rank if a.rank is not None else get_rank()
    return a.master_ip, a.master_port, rank, a.size


def parse_all(distributed: bool, argv=None) -> argparse.Namespace:
    a = build_parser(distributed).parse_args(argv)
    if distributed and a.rank is None and not a.spawn:
        a.rank = get_rank()
    return a
