"""``hetu.communicator`` module paths of the reference (python/hetu/communicator/): the classes
live in :mod:`hetu_61a7_b200.parallel.comm`; these modules keep the import lines working."""
from . import mpi_comm, mpi_nccl_comm, nccl_comm          # noqa: F401
