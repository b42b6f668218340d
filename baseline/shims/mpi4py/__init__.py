"""Stand-in for mpi4py on a box without MPI.

The reference uses MPI for exactly two calls -- `MPI.COMM_WORLD.Get_rank()` and `.Get_size()`
(cifar10_mpi_mobilenet_224.py:24-26); no tensor ever travels over MPI.  This shim answers them from the
launcher's environment (torchrun: RANK / WORLD_SIZE; OpenMPI: OMPI_COMM_WORLD_*).
"""
from . import MPI  # noqa: F401
