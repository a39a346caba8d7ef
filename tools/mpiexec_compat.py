#!/usr/bin/env python3
"""`mpiexec -n N <command>` for boxes without an MPI installation.

Spawns N copies of <command> on this node with the environment an Open MPI launcher would give
them (OMPI_COMM_WORLD_RANK / OMPI_COMM_WORLD_SIZE / OMPI_COMM_WORLD_LOCAL_RANK) plus
MASTER_ADDR=127.0.0.1 and a free MASTER_PORT (what the reference's script otherwise obtains
with comm.bcast, demo_assume_started_with_mpiexec.py:43-44).  Exit code = first non-zero child.
"""
import os
import socket
import subprocess
import sys


def main():
    args = sys.argv[1:]
    n = 1
    while args and args[0].startswith("-"):
        if args[0] in ("-n", "-np"):
            n = int(args[1]); args = args[2:]
        else:
            raise SystemExit(f"mpiexec_compat: unsupported option {args[0]}")
    if not args:
        raise SystemExit("usage: mpiexec_compat.py -n N command [args...]")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(n):
        env = dict(os.environ, OMPI_COMM_WORLD_RANK=str(rank), OMPI_COMM_WORLD_SIZE=str(n),
                   OMPI_COMM_WORLD_LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        env.setdefault("TASKS_PER_NODE", str(n))
        procs.append(subprocess.Popen(args, env=env))
    codes = [p.wait() for p in procs]
    sys.exit(next((c for c in codes if c), 0))


if __name__ == "__main__":
    main()
