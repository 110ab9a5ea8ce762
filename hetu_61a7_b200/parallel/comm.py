"""Process-group communicator: one process per B200, ``torch.distributed`` (NCCL
on GPU, gloo on CPU) for rendezvous and baseline collectives.

reference: src/communication/mpi_nccl_communication.cu:31-483 (MPI bootstrap +
NCCL wrappers) and python/hetu/communicator/mpi_nccl_comm.py:62-336.  MPI is
replaced by the launcher's env (RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT - what
``heturun`` and ``torchrun`` both export); sub-group communicators are
``dist.new_group``.  The fused in-kernel NVLink collectives live in
``parallel/symm.py`` and ``csrc/collectives.cu``; this class is the NCCL
baseline + the p2p path that BASELINE.json keeps on NCCL.

With WORLD_SIZE unset the communicator is a single-rank loopback world, so
graph passes and ops can be unit-tested without any launcher.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

_REDUCE = {'sum': dist.ReduceOp.SUM, 'max': dist.ReduceOp.MAX, 'min': dist.ReduceOp.MIN,
           'prod': dist.ReduceOp.PRODUCT, 'avg': dist.ReduceOp.AVG}


class _Val(object):
    """mimics ctypes c_int ``.value`` used by reference scripts (comm.myRank.value)."""

    def __init__(self, v):
        self.value = v


def dist_env():
    rank = int(os.environ.get('RANK', os.environ.get('OMPI_COMM_WORLD_RANK', '0')))
    world = int(os.environ.get('WORLD_SIZE', os.environ.get('OMPI_COMM_WORLD_SIZE', '1')))
    local = int(os.environ.get('LOCAL_RANK', os.environ.get('OMPI_COMM_WORLD_LOCAL_RANK', str(rank))))
    return rank, world, local


def ensure_process_group() -> bool:
    """Initialise the default group from the launcher's environment (idempotent)."""
    if dist.is_available() and dist.is_initialized():
        return True
    rank, world, local = dist_env()
    if world <= 1:
        return False
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    os.environ.setdefault('MASTER_PORT', '29511')
    use_cuda = torch.cuda.is_available() and os.environ.get('HETU_COMM_BACKEND', '') != 'gloo'
    if use_cuda:
        torch.cuda.set_device(local % torch.cuda.device_count())
        dist.init_process_group('nccl', rank=rank, world_size=world,
                                device_id=torch.device('cuda', local % torch.cuda.device_count()))
    else:
        dist.init_process_group('gloo', rank=rank, world_size=world)
    return True


class Communicator(object):
    """Collectives over a (sub)group.  Tensors are modified in place unless a
    separate output is given, matching the reference's dlarray* calls."""

    def __init__(self, group=None, ranks: Optional[Sequence[int]] = None):
        self.active = ensure_process_group()
        self.group = group
        if self.active:
            self.world_rank = dist.get_rank()
            self.world_size = dist.get_world_size()
            self.ranks = list(ranks) if ranks is not None else list(range(self.world_size))
            self.rank = self.ranks.index(self.world_rank) if self.world_rank in self.ranks else -1
            self.nrank = len(self.ranks)
        else:
            self.world_rank = self.rank = 0
            self.world_size = self.nrank = 1
            self.ranks = [0]
        _, _, local = dist_env()
        self.local_rank = local
        self.device_id = _Val(local)
        self.myRank = _Val(self.rank)
        self.nRanks = _Val(self.nrank)
        self._grouped = None

    # -- reference-compatible accessors --------------------------------------
    @property
    def dev_id(self):
        return self.local_rank

    def ncclInit(self):
        return self

    def ncclFinish(self):
        pass

    def ncclSetDevice(self, dev):
        if torch.cuda.is_available():
            torch.cuda.set_device(dev)

    def getRank(self):
        return self.rank

    def getNRanks(self):
        return self.nrank

    def world_of(self, group_rank: int) -> int:
        return self.ranks[group_rank]

    @property
    def multi(self) -> bool:
        return self.active and self.nrank > 1

    # -- collectives -----------------------------------------------------------
    def allreduce(self, t: torch.Tensor, op: str = 'sum') -> torch.Tensor:
        if self.multi:
            dist.all_reduce(t, op=_REDUCE[op], group=self.group)
        return t

    def allgather(self, t: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if not self.multi:
            if out is None:
                return t.clone()
            out.copy_(t)
            return out
        if out is None:
            out = torch.empty((t.shape[0] * self.nrank,) + tuple(t.shape[1:]), dtype=t.dtype,
                              device=t.device)
        dist.all_gather_into_tensor(out, t.contiguous(), group=self.group)
        return out

    def reducescatter(self, t: torch.Tensor, out: Optional[torch.Tensor] = None, op='sum'):
        if not self.multi:
            if out is None:
                return t.clone()
            out.copy_(t)
            return out
        if out is None:
            out = torch.empty((t.shape[0] // self.nrank,) + tuple(t.shape[1:]), dtype=t.dtype,
                              device=t.device)
        dist.reduce_scatter_tensor(out, t.contiguous(), op=_REDUCE[op], group=self.group)
        return out

    def broadcast(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        if self.multi:
            dist.broadcast(t, src=self.world_of(root), group=self.group)
        return t

    def reduce(self, t: torch.Tensor, root: int = 0, op='sum') -> torch.Tensor:
        if self.multi:
            dist.reduce(t, dst=self.world_of(root), op=_REDUCE[op], group=self.group)
        return t

    def alltoall(self, t: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Equal split of dim 0 over the group (reference _ncclAllToAll,
        mpi_nccl_communication.cu:245-277)."""
        if out is None:
            out = torch.empty_like(t)
        if not self.multi:
            out.copy_(t)
            return out
        dist.all_to_all_single(out, t.contiguous(), group=self.group)
        return out

    def send(self, t: torch.Tensor, dst: int):
        if self._grouped is not None:
            self._grouped.append(dist.P2POp(dist.isend, t.contiguous(), self.world_of(dst), self.group))
        else:
            dist.send(t.contiguous(), dst=self.world_of(dst), group=self.group)

    def recv(self, t: torch.Tensor, src: int):
        if self._grouped is not None:
            self._grouped.append(dist.P2POp(dist.irecv, t, self.world_of(src), self.group))
        else:
            dist.recv(t, src=self.world_of(src), group=self.group)
        return t

    def group_start(self):
        self._grouped = []

    def group_end(self):
        ops, self._grouped = self._grouped, None
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()

    def barrier(self):
        if self.multi:
            dist.barrier(group=self.group)

    def allgather_object(self, obj) -> list:
        if not self.multi:
            return [obj]
        out = [None] * self.nrank
        dist.all_gather_object(out, obj, group=self.group)
        return out

    # reference MPI / NCCL communicator helpers (communicator/mpi_nccl_comm.py:62-336) ------------
    # MPI is replaced by the launcher's environment + torch.distributed, so the bootstrap calls are
    # no-ops that return what the launcher already decided.
    def MPI_Init(self):
        return self

    def MPI_Finalize(self):
        pass

    def MPIGetComm(self):
        return self

    MPI_GetComm = MPIGetComm

    def allReduce(self, arr):
        """Host-side sum all-reduce of a numpy array in place (reference mpi_communication.cc:37-47)."""
        if self.multi:
            t = torch.from_numpy(arr)
            if dist.get_backend() == 'nccl':
                d = t.cuda()
                dist.all_reduce(d, group=self.group)
                t.copy_(d.cpu())
            else:
                dist.all_reduce(t, group=self.group)
        return arr

    def MPI_Comm_rank(self):
        return self.rank

    def MPI_Comm_size(self):
        return self.nrank

    def MPI_Broadcast(self, arr, root=0):
        """Host-side broadcast of a numpy array / python object from ``root`` (in place for arrays)."""
        if not self.multi:
            return arr
        box = [arr if self.rank == root else None]
        dist.broadcast_object_list(box, src=self.world_of(root), group=self.group)
        if isinstance(arr, np.ndarray) and self.rank != root:
            arr[...] = box[0]
            return arr
        return box[0]

    def getLocalRank(self):
        return self.local_rank

    def getGlobalDevice(self):
        return self.local_rank

    def getDeviceFromLocalRank(self, local_rank):
        return local_rank

    def getLocalRankFromDevice(self, device_id):
        return device_id

    def getRankFromDevice(self, hostname_or_ctx, device_id=None):
        """Rank that drives a device: on one node rank == device id (group-relative inside a sub-group)."""
        dev = hostname_or_ctx.device_id if device_id is None else device_id
        return self.ranks.index(dev) if dev in self.ranks else dev

    def ncclGroupInit(self, devices_context, stream=None):
        return new_group_comm(devices_context)

    def ncclGetUniqueId(self, senderRank=0):
        return None                 # rendezvous is torch.distributed's store

    ncclGetGroupUniqueId = ncclGetUniqueId

    def ncclCommInitRank(self):
        return self

    def ncclCommDestroy(self):
        pass

    def dlarrayNcclReduce(self, dlarray, out=None, root=0, datatype=None, reduceop=None, stream=None):
        from ..ndarray import NDArray
        t = dlarray.tensor if isinstance(dlarray, NDArray) else dlarray
        self.reduce(t, root)
        if out is not None and out is not dlarray and self.rank == root:
            (out.tensor if isinstance(out, NDArray) else out).copy_(t)

    # hierarchical all-to-all (reference mpi_nccl_communication.cu:152-244: gather to the node leader,
    # leader exchange, scatter).  Inside one NVSwitch domain the hierarchy only adds copies, so the three
    # steps collapse into one direct exchange issued by the last call.
    def dlarrayHA2AGather(self, sendarr, recvarr, datatype=None, myrank=None, num_local_gpus=None, stream=None):
        self._ha2a_src = sendarr

    def dlarrayHAllToAll(self, sendarr, recvarr, datatype=None, num_local_gpus=None, stream=None):
        pass

    def dlarrayHA2AScatter(self, sendarr, recvarr, datatype=None, myrank=None, num_local_gpus=None, stream=None):
        src = getattr(self, '_ha2a_src', None)
        self._ha2a_src = None
        self.dlarrayAllToAll(src if src is not None else sendarr, recvarr)

    # reference dlarray* spellings ------------------------------------------------
    def dlarrayNcclAllReduce(self, dlarray, out=None, datatype=None, reduceop=None, stream=None):
        from ..ndarray import NDArray
        t = dlarray.tensor if isinstance(dlarray, NDArray) else dlarray
        self.allreduce(t)
        if out is not None and out is not dlarray:
            (out.tensor if isinstance(out, NDArray) else out).copy_(t)

    def dlarrayBroadcast(self, dlarray, out=None, datatype=None, root=0, stream=None):
        from ..ndarray import NDArray
        t = dlarray.tensor if isinstance(dlarray, NDArray) else dlarray
        self.broadcast(t, root)

    def dlarrayAllGather(self, inp, out, datatype=None, stream=None):
        from ..ndarray import NDArray
        self.allgather(inp.tensor if isinstance(inp, NDArray) else inp,
                       out.tensor if isinstance(out, NDArray) else out)

    def dlarrayReduceScatter(self, inp, out, datatype=None, reduceop=None, stream=None):
        from ..ndarray import NDArray
        self.reducescatter(inp.tensor if isinstance(inp, NDArray) else inp,
                           out.tensor if isinstance(out, NDArray) else out)

    def dlarrayAllToAll(self, inp, out, datatype=None, stream=None):
        from ..ndarray import NDArray
        self.alltoall(inp.tensor if isinstance(inp, NDArray) else inp,
                      out.tensor if isinstance(out, NDArray) else out)

    def dlarraySend(self, arr, datatype, target, stream=None):
        from ..ndarray import NDArray
        self.send(arr.tensor if isinstance(arr, NDArray) else arr, target)

    def dlarrayRecv(self, arr, datatype, src, stream=None):
        from ..ndarray import NDArray
        self.recv(arr.tensor if isinstance(arr, NDArray) else arr, src)


_WORLD: Optional[Communicator] = None
_GROUPS = {}


def world() -> Communicator:
    global _WORLD
    if _WORLD is None:
        _WORLD = Communicator()
    return _WORLD


def wrapped_mpi_nccl_init(init_nccl: bool = True, devices: Optional[List[int]] = None) -> Communicator:
    """reference gpu_ops/executor.py:60-71.  Binds this process to its GPU and
    returns the world communicator."""
    comm = world()
    if torch.cuda.is_available():
        dev = comm.local_rank if devices is None else devices[comm.local_rank % len(devices)]
        torch.cuda.set_device(dev % torch.cuda.device_count())
    return comm


def get_nccl_communicate() -> Communicator:
    return world()


get_mpi_communicate = get_nccl_communicate


def new_group_comm(devices_context=None) -> Communicator:
    """Sub-group communicator over the given device group (world ranks are the
    device ids on a single node).  reference executor.py:74-83 /
    mpi_nccl_comm.py:188-250."""
    comm = world()
    if devices_context is None:
        return comm
    if isinstance(devices_context, (list, tuple)) and all(isinstance(r, int) for r in devices_context):
        ranks = sorted(devices_context)
    else:
        ranks = sorted(c.device_id for c in devices_context.flat_devices()
                       if c.device_type == 2) if hasattr(devices_context, 'flat_devices') else \
            sorted(c.device_id for c in devices_context)
    key = tuple(ranks)
    if key == tuple(range(comm.world_size)):
        return comm
    if key not in _GROUPS:
        group = dist.new_group(ranks=list(ranks)) if comm.active else None
        _GROUPS[key] = Communicator(group=group, ranks=ranks)
    return _GROUPS[key]


_P2P = {}
_PENDING = []


def make_p2p_comms():
    """Two extra world-spanning communicators, one per pipeline direction (activations forward,
    gradients backward).  Collective: every rank must call it (the placement pass does)."""
    comm = world()
    if not comm.active or _P2P:
        return
    for d in (0, 1):
        _P2P[d] = Communicator(group=dist.new_group(ranks=list(range(comm.world_size))))


def p2p_comm(direction: int = 0) -> Communicator:
    return _P2P.get(direction, world())


def isend(t: torch.Tensor, dst: int, direction: int = 0):
    comm = p2p_comm(direction)
    t = t.contiguous()
    req = dist.isend(t, dst=comm.world_of(dst), group=comm.group)
    _PENDING.append((req, t))
    if len(_PENDING) > 64:
        drain_sends(keep=32)


def recv_now(t: torch.Tensor, src: int, direction: int = 0):
    comm = p2p_comm(direction)
    dist.irecv(t, src=comm.world_of(src), group=comm.group).wait()
    return t


def drain_sends(keep: int = 0):
    while len(_PENDING) > keep:
        req, _ = _PENDING.pop(0)
        req.wait()


def device_max_ms(ms: float, device=None) -> float:
    """MAX over ranks of a device-measured duration."""
    comm = world()
    if not comm.multi:
        return ms
    t = torch.tensor([ms], dtype=torch.float64,
                     device=device if device is not None else (
                         torch.device('cuda', torch.cuda.current_device())
                         if torch.cuda.is_available() and dist.get_backend() == 'nccl' else 'cpu'))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
