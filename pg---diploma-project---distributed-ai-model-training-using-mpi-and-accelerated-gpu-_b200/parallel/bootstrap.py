"""Rank / world discovery and process-group set-up.

The reference reads ``rank``/``size`` from ``mpi4py.MPI.COMM_WORLD`` at import time
(cifar10_mpi_mobilenet_224.py:24-26) and then exports ``MASTER_ADDR``/``RANK``/... itself
(:28-35).  MPI never carries tensor data.  Here the same information is taken straight
from whichever launcher started the process -- OpenMPI, PMI/PMIx (MPICH, Slurm's pmi2),
``srun``, ``torchrun`` or our own ``b200dp.launch`` -- so the identical entry point works
under ``mpirun -np N``, ``srun -n N``, ``torchrun`` or standalone, with no mpi4py import.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Mapping, Optional, Tuple

# (rank, world, local_rank) variable families, in priority order.
_ENV_FAMILIES = (
    ("openmpi", "OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK"),
    ("pmix", "PMIX_RANK", "PMIX_SIZE", None),
    ("pmi", "PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID"),
    ("slurm", "SLURM_PROCID", "SLURM_NTASKS", "SLURM_LOCALID"),
    ("torchrun", "RANK", "WORLD_SIZE", "LOCAL_RANK"),
)


@dataclass(frozen=True)
class RankInfo:
    rank: int
    world_size: int
    local_rank: int
    source: str            # which env family supplied the numbers ("single" if none)

    @property
    def is_main(self) -> bool:
        return self.rank == 0


def discover_rank_world(env: Optional[Mapping[str, str]] = None) -> RankInfo:
    """Return this process's place in the job from the launcher's environment.

    Equivalent of ``comm.Get_rank()`` / ``comm.Get_size()`` (ref :24-26).  A family is used
    only if both its rank and size variables are set (PMIx does not always export a size;
    then the Slurm/OMPI size is used if available).
    """
    env = os.environ if env is None else env
    for name, rk, wk, lk in _ENV_FAMILIES:
        if rk in env:
            size = env.get(wk)
            if size is None:
                for _n, _r, wk2, _l in _ENV_FAMILIES:
                    if wk2 in env:
                        size = env[wk2]
                        break
            if size is None:
                continue
            rank, world = int(env[rk]), int(size)
            if not (0 <= rank < world):
                raise RuntimeError(f"{name}: rank {rank} outside world of size {world}")
            local = int(env[lk]) if lk and lk in env else None
            if local is None:
                for _n, _r, _w, lk2 in _ENV_FAMILIES:
                    if lk2 and lk2 in env:
                        local = int(env[lk2])
                        break
            return RankInfo(rank, world, rank if local is None else local, name)
    return RankInfo(0, 1, 0, "single")


@dataclass
class DistContext:
    info: RankInfo
    device: "object"      # torch.device
    backend: str
    initialized: bool

    @property
    def rank(self) -> int:
        return self.info.rank

    @property
    def world_size(self) -> int:
        return self.info.world_size

    @property
    def is_main(self) -> bool:
        return self.info.rank == 0


def setup_distributed(master_addr: str = "127.0.0.1", master_port: int = 29500,
                      backend: str = "auto", info: Optional[RankInfo] = None,
                      use_cuda: Optional[bool] = None, timeout_s: float = 600.0) -> DistContext:
    """``setup_distributed()`` of the reference (:28-45): export rendezvous variables,
    pick ``nccl`` on GPUs / ``gloo`` on CPU, ``init_process_group``, bind rank -> device
    ``rank % device_count``.  ``MASTER_ADDR``/``MASTER_PORT`` already in the environment
    (torchrun, our launcher) win over the arguments; the reference hard-codes ``localhost``,
    which limits it to one node -- here any reachable address works.
    """
    import datetime
    import torch
    import torch.distributed as dist

    info = info or discover_rank_world()
    cuda = torch.cuda.is_available() if use_cuda is None else use_cuda
    os.environ.setdefault("MASTER_ADDR", master_addr)
    os.environ.setdefault("MASTER_PORT", str(master_port))
    os.environ["RANK"] = str(info.rank)
    os.environ["WORLD_SIZE"] = str(info.world_size)
    os.environ.setdefault("LOCAL_RANK", str(info.local_rank))

    if backend == "auto":
        backend = "nccl" if cuda else "gloo"
    if cuda:
        ndev = torch.cuda.device_count()
        local = info.local_rank % max(ndev, 1)
        torch.cuda.set_device(local)
        device = torch.device(f"cuda:{local}")
    else:
        device = torch.device("cpu")

    initialized = False
    if info.world_size > 1 or backend == "gloo_force":
        kwargs = dict(backend=backend, rank=info.rank, world_size=info.world_size,
                      timeout=datetime.timedelta(seconds=timeout_s))
        if cuda and backend == "nccl":
            kwargs["device_id"] = device
        dist.init_process_group(**kwargs)
        initialized = True
    return DistContext(info=info, device=device, backend=backend, initialized=initialized)


def cleanup(ctx: Optional[DistContext] = None) -> None:
    """``cleanup()`` of the reference (:47-48)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
    if ctx is not None:
        ctx.initialized = False


def barrier(ctx: DistContext) -> None:
    """Host-visible barrier (ref :102).  The per-step device-side barrier lives in the comm kernel."""
    import torch.distributed as dist
    if ctx.initialized:
        if ctx.device.type == "cuda":
            dist.barrier(device_ids=[ctx.device.index])
        else:
            dist.barrier()


def rank_zero_first(ctx: DistContext):
    """Context manager: rank 0 runs the body first (dataset materialisation, ref :93-102), then the rest."""
    import contextlib

    @contextlib.contextmanager
    def _cm():
        if not ctx.is_main:
            barrier(ctx)
        yield
        if ctx.is_main:
            barrier(ctx)
    return _cm()
