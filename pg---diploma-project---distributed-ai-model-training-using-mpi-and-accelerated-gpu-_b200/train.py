"""``python -m b200dp.train --mode {cpu,gpu,ddp}`` -- the single entry point that replaces
cifar10_serial_mobilenet_224.py, cifar10_128batch.py and cifar10_mpi_mobilenet_224.py.

Under ``mpirun`` / ``srun`` / ``torchrun`` / ``python -m b200dp.launch`` the rank and world size come
from the launcher's environment (parallel/bootstrap.py); see scripts/*.slurm for the job files.
"""
from __future__ import annotations

import sys
from typing import Optional, Sequence

from .config import config_from_args
from .trainer import Trainer


def main(argv: Optional[Sequence[str]] = None) -> int:
    cfg = config_from_args(sys.argv[1:] if argv is None else argv)
    trainer = Trainer(cfg)
    try:
        trainer.fit()
    finally:
        trainer.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
