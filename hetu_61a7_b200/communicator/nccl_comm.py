"""reference python/hetu/communicator/nccl_comm.py: single-process multi-GPU communicator
(``ncclCommInitAll`` over local devices).  One process per GPU is the model here, so
``nccl_communicator(devs, n)`` returns the sub-group communicator of those devices."""
from ..parallel.comm import Communicator as NCCL_Communicator
from ..parallel.comm import new_group_comm


def nccl_communicator(devs, devs_number=None):
    from ..ndarray import gpu
    return new_group_comm([gpu(int(d)) for d in list(devs)[:devs_number]])
