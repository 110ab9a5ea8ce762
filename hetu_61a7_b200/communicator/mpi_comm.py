"""reference python/hetu/communicator/mpi_comm.py: host-array all-reduce over MPI
(``MPI_Communicator.allReduce`` on a CPU NDArray).  Served by the same communicator (gloo for
host tensors); the TCP ring in ``parallel.ring_comm`` is the MPI-free stand-alone variant."""
from enum import Enum

from ..parallel.comm import Communicator as MPI_Communicator
from ..parallel.comm import wrapped_mpi_nccl_init


class MPIDataType_t(Enum):
    MPI_CHAR = 0
    MPI_INT = 1
    MPI_FLOAT = 2
    MPI_DOUBLE = 3


class MPIOp_t(Enum):
    MPI_MAX = 0
    MPI_MIN = 1
    MPI_SUM = 2
    MPI_PROD = 3


def mpi_communicator():
    return wrapped_mpi_nccl_init(init_nccl=False)
