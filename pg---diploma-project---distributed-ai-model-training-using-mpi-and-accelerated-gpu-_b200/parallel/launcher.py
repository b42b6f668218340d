"""``mpirun -np N`` replacement.

The reference is launched with ``mpirun --mca pml ob1 --mca btl tcp,self -np 2 python
cifar10_mpi_mobilenet_224.py`` (cifar10_gpu_parallel.sh:19-20).  This image has neither
``mpirun`` nor ``srun``, so ``python -m b200dp.launch -np N -- <module or script> [args]``
forks N workers with the OpenMPI-style variables (``OMPI_COMM_WORLD_RANK`` ...) *and* the
torchrun-style ones set, which is exactly what ``bootstrap.discover_rank_world`` reads.
Failure handling the reference lacks: the first non-zero exit kills the
remaining ranks' process groups and the launcher exits with that code.
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List, Optional, Sequence


def find_free_port(addr: str = "127.0.0.1") -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind((addr, 0))
        return s.getsockname()[1]


def worker_env(rank: int, world: int, master_addr: str, master_port: int,
               base: Optional[dict] = None) -> dict:
    env = dict(os.environ if base is None else base)
    env.update({
        "OMPI_COMM_WORLD_RANK": str(rank),
        "OMPI_COMM_WORLD_SIZE": str(world),
        "OMPI_COMM_WORLD_LOCAL_RANK": str(rank),
        "RANK": str(rank),
        "WORLD_SIZE": str(world),
        "LOCAL_RANK": str(rank),
        "MASTER_ADDR": master_addr,
        "MASTER_PORT": str(master_port),
    })
    return env


def launch(nprocs: int, cmd: Sequence[str], master_addr: str = "127.0.0.1",
           master_port: int = 0, poll_s: float = 0.2, timeout_s: float = 0.0) -> int:
    """Start ``nprocs`` copies of ``cmd``; return the job's exit code (0 iff all ranks exit 0)."""
    if nprocs < 1:
        raise ValueError("nprocs must be >= 1")
    port = master_port or find_free_port(master_addr)
    procs: List[subprocess.Popen] = []
    for r in range(nprocs):
        procs.append(subprocess.Popen(list(cmd), env=worker_env(r, nprocs, master_addr, port),
                                      start_new_session=True))
    deadline = time.time() + timeout_s if timeout_s > 0 else None
    exit_code = 0
    try:
        alive = set(range(nprocs))
        while alive:
            for r in list(alive):
                rc = procs[r].poll()
                if rc is None:
                    continue
                alive.discard(r)
                if rc != 0 and exit_code == 0:
                    exit_code = rc
                    sys.stderr.write(f"[b200dp.launch] rank {r} exited with code {rc}; stopping the job\n")
                    _terminate(procs, alive)
            if deadline is not None and time.time() > deadline and alive:
                sys.stderr.write("[b200dp.launch] timeout; stopping the job\n")
                exit_code = exit_code or 124
                _terminate(procs, alive)
            if alive:
                time.sleep(poll_s)
    except KeyboardInterrupt:
        _terminate(procs, set(range(nprocs)))
        exit_code = 130
    return exit_code


def _terminate(procs: List[subprocess.Popen], alive: set) -> None:
    # Each worker is its own session leader: signal exactly the process groups we created.
    for r in list(alive):
        p = procs[r]
        if p.poll() is None:
            try:
                os.killpg(p.pid, signal.SIGTERM)
            except ProcessLookupError:
                pass
    t0 = time.time()
    while time.time() - t0 < 5.0 and any(procs[r].poll() is None for r in alive):
        time.sleep(0.05)
    for r in list(alive):
        p = procs[r]
        if p.poll() is None:
            try:
                os.killpg(p.pid, signal.SIGKILL)
            except ProcessLookupError:
                pass
        p.wait()
    alive.clear()


def main(argv: Optional[Sequence[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="b200dp.launch", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("-np", "--nprocs", type=int, default=0, help="ranks (default: number of visible GPUs, else 1)")
    ap.add_argument("--master-addr", default="127.0.0.1")
    ap.add_argument("--master-port", type=int, default=0)
    ap.add_argument("--timeout", type=float, default=0.0)
    ap.add_argument("-m", "--module", action="store_true", help="treat the first positional as a module (python -m)")
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    ns = ap.parse_args(argv)
    cmd = [c for c in ns.cmd if c != "--"] if ns.cmd and ns.cmd[0] == "--" else list(ns.cmd)
    if not cmd:
        cmd = ["-m", "b200dp.train", "--mode", "ddp"]
    elif ns.module:
        cmd = ["-m"] + cmd
    n = ns.nprocs
    if n <= 0:
        try:
            import torch
            n = max(torch.cuda.device_count(), 1)
        except Exception:
            n = 1
    full = cmd if os.path.basename(cmd[0]).startswith("python") else [sys.executable] + cmd
    return launch(n, full, ns.master_addr, ns.master_port, timeout_s=ns.timeout)


if __name__ == "__main__":
    sys.exit(main())
