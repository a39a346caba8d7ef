"""Command-line flags of the demo entry points.

Same seven flags, choices and defaults as the reference's argument_parser.py:6-28
(kept so that `python demo.py --backend=nccl --torchrun`, the line in
hpc_files/torchrun_configs.txt:1, works unchanged).  New knobs of this implementation
are environment variables (see README/DESIGN), never new flags.
"""
import argparse
import random

_FLAGS = (
    ("--dataloader", dict(type=str, choices=["distributed", "standard"], default="distributed")),
    ("--backend", dict(type=str, choices=["nccl", "mpi", "gloo"], default="nccl")),
    ("--torchrun", dict(action="store_true", help="Specify we are using torchrun to distribute jobs")),
    ("--use_node_rank", dict(action="store_true",
                             help="Specify whether to use the NODE_RANK environment variable to determine "
                                  "global rank of each process")),
    ("--seed", dict(type=int, default=None)),
    ("--num_workers", dict(type=int, default=0)),
    ("--dry_run", dict(action="store_true", help="Dry run (do not log to wandb)")),
)


def build_parser():
    parser = argparse.ArgumentParser()
    for flag, spec in _FLAGS:
        spec = dict(spec)
        if flag == "--seed":
            spec["default"] = random.randint(0, 2 ** 32 - 1)   # drawn per process, like the reference
        parser.add_argument(flag, **spec)
    return parser


def get_args(argv=None):
    return build_parser().parse_args(argv)
