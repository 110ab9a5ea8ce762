"""reference python/hetu/communicator/mpi_nccl_comm.py: ``MPI_Communicator`` (bootstrap: rank,
size, local rank, device mapping, unique-id exchange), ``NCCL_Communicator`` (collectives, p2p,
sub-groups), ``mpi_communicator()``, ``GroupStart/GroupEnd``, the nccl enums.

One class plays both roles here (``parallel.comm.Communicator``: torch.distributed rendezvous
instead of MPI, NCCL underneath, in-kernel NVLink collectives where the symmetric heap is up)."""
from enum import Enum

from ..parallel.comm import Communicator, new_group_comm, world, wrapped_mpi_nccl_init

MPI_Communicator = Communicator
NCCL_Communicator = Communicator


class ncclDataType_t(Enum):
    ncclInt8 = 0
    ncclChar = 0
    ncclUint8 = 1
    ncclInt32 = 2
    ncclInt = 2
    ncclUint32 = 3
    ncclInt64 = 4
    ncclUint64 = 5
    ncclFloat16 = 6
    ncclHalf = 6
    ncclFloat32 = 7
    ncclFloat = 7
    ncclFloat64 = 8
    ncclDouble = 8
    ncclNumTypes = 9


class ncclRedOp_t(Enum):
    ncclSum = 0
    ncclProd = 1
    ncclMax = 2
    ncclMin = 3
    ncclNumOps = 4


def mpi_communicator(devices=None):
    """The process-wide communicator (reference: ``MPI_Communicator(devices)``)."""
    return wrapped_mpi_nccl_init(devices=devices) if devices is not None else wrapped_mpi_nccl_init()


def GroupStart():
    """NCCL group calls are batched by the executor's pipeline scheduler; kept for scripts."""
    from ..parallel import comm
    getattr(comm, 'group_start', lambda: None)()


def GroupEnd():
    from ..parallel import comm
    getattr(comm, 'group_end', lambda: None)()
