"""Single-process multi-GPU communicator.

reference: python/hetu/communicator/nccl_comm.py:20-90 + src/communication/nccl_communication.cu
(``ncclCommInitAll`` over the local devices, grouped all-reduce / all-to-all where ONE process
drives every GPU: ``NCCL_Communicator(devs).all_reduce(arrays)``).

The framework's own model is one process per GPU (``parallel.comm.Communicator``); this class is
for scripts that hold one array per local device in a single process.  NCCL is reached through
``torch.cuda.nccl`` (communicators are created and cached by torch per device set); the all-to-all
is a grid of peer copies on each destination device's stream, which on an NVSwitch node is what a
grouped ncclSend/ncclRecv lowers to anyway."""
from __future__ import annotations

from typing import List, Sequence

import torch

from ..ndarray import NDArray, gpu
from ..parallel.comm import new_group_comm


def _t(a):
    return a.tensor if isinstance(a, NDArray) else a


class NCCL_Communicator(object):
    def __init__(self, devs: Sequence[int], devs_number=None):
        self.devs = [int(d) for d in list(devs)[:devs_number]]
        if len(set(self.devs)) != len(self.devs):
            raise ValueError('duplicate device ids')
        self.ndev = len(self.devs)

    def _check(self, arrays) -> List[torch.Tensor]:
        ts = [_t(a) for a in arrays]
        if len(ts) != self.ndev:
            raise ValueError('expected one array per device (%d), got %d' % (self.ndev, len(ts)))
        for t, d in zip(ts, self.devs):
            if t.device.type != 'cuda' or t.device.index != d:
                raise ValueError('array on %s, expected cuda:%d' % (t.device, d))
        return ts

    # reference spelling: dlarrayNcclAllReduce over the list of local arrays, in place
    def all_reduce(self, arrays, op='sum'):
        from torch.cuda import nccl
        ts = self._check(arrays)
        if self.ndev > 1:
            nccl.all_reduce(ts, op={'sum': 0, 'prod': 1, 'max': 2, 'min': 3}[op])
        return arrays

    ncclAllReduce = all_reduce

    def broadcast(self, arrays, root=0):
        from torch.cuda import nccl
        ts = self._check(arrays)
        if self.ndev > 1:
            nccl.broadcast(ts, root=int(root))
        return arrays

    def all_gather(self, arrays, outputs):
        """outputs[d] = concat over devices of arrays[*] along dim 0."""
        from torch.cuda import nccl
        ins, outs = self._check(arrays), self._check(outputs)
        if self.ndev > 1:
            nccl.all_gather(ins, outs)
        else:
            outs[0].copy_(ins[0].reshape(outs[0].shape))
        return outputs

    def reduce_scatter(self, arrays, outputs, op='sum'):
        from torch.cuda import nccl
        ins, outs = self._check(arrays), self._check(outputs)
        if self.ndev > 1:
            nccl.reduce_scatter(ins, outs, op={'sum': 0, 'prod': 1, 'max': 2, 'min': 3}[op])
        else:
            outs[0].copy_(ins[0].reshape(outs[0].shape))
        return outputs

    def all_to_all(self, arrays, outputs):
        """Equal split along dim 0: chunk j of arrays[i] lands in chunk i of outputs[j]."""
        ins, outs = self._check(arrays), self._check(outputs)
        n = self.ndev
        for j, dst in enumerate(outs):
            with torch.cuda.device(dst.device):
                for i, src in enumerate(ins):
                    rows = src.shape[0] // n
                    dst[i * rows:(i + 1) * rows].copy_(src[j * rows:(j + 1) * rows], non_blocking=True)
        for d in self.devs:
            torch.cuda.synchronize(d)
        return outputs

    ncclAllToAll = all_to_all


def nccl_communicator(devs, devs_number=None):
    """reference ``nccl_communicator(devs, devs_number)``: in a single process returns the
    one-process-many-GPUs communicator; under a multi-process launch the group communicator of
    those devices."""
    import os
    if int(os.environ.get('WORLD_SIZE', '1')) > 1:
        return new_group_comm([gpu(int(d)) for d in list(devs)[:devs_number]])
    return NCCL_Communicator(devs, devs_number)
