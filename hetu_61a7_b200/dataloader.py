"""Dataloaders.  reference: python/hetu/dataloader.py:11-257.

Batches are staged through a 3-deep ring of *pinned* host buffers and copied to
HBM with a non-blocking copy, so the next batch's H2D overlaps the current step.
``get_next_arr`` peeks the following batch (used to prefetch embeddings).
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np
import torch

from .graph.node import Op
from .ndarray import cpu


class Dataloader(object):
    def __init__(self, raw_data, batch_size, name='default', func=None, drop_last=True, dtype=np.float32):
        """``dtype``: the reference stores every batch as float32 (ids included, dataloader.py:16);
        pass ``np.int64`` to keep embedding ids exact beyond 2**24."""
        self.func = func if func else (lambda x: x)
        self.raw_data = np.ascontiguousarray(self.func(raw_data), dtype=dtype)
        self.dtype = torch.from_numpy(np.empty(0, dtype=dtype)).dtype
        self.batch_size = int(batch_size)
        self.drop_last = drop_last
        self.name = str(name)
        self.dp_nrank = None
        self.dp_rank = None
        self.parts = None
        self.cur_part = None
        self.slices = None
        self.queue_size = 3
        self._inited = False

    def init_states(self):
        if self._inited:
            return
        if self.dp_nrank is not None:
            n = self.raw_data.shape[0] // self.dp_nrank
            self.raw_data = self.raw_data[self.dp_rank * n:(self.dp_rank + 1) * n]
        self.samples_num = len(self.raw_data)
        if self.drop_last:
            self.batch_num = self.samples_num // self.batch_size
        else:
            self.batch_num = int(np.ceil(self.samples_num / self.batch_size))
        assert self.batch_num > 0, 'Batch size %d larger than dataset (%d samples)' % (
            self.batch_size, self.samples_num)
        self.shape = tuple([self.batch_size] + list(self.raw_data.shape[1:]))
        self.seq = np.arange(self.samples_num)
        self.index = 0
        self.batch_index = 0
        pin = torch.cuda.is_available()
        # small enough datasets are pinned once as a whole: a batch is then a VIEW of pinned memory and
        # the per-step host memcpy into the ring disappears (HETU_PIN_DATASET_MB, default 4096)
        import os
        limit = float(os.environ.get('HETU_PIN_DATASET_MB', '4096')) * (1 << 20)
        self.pinned_all = None
        if pin and not self.parts and self.raw_data.nbytes <= limit:
            self.pinned_all = torch.from_numpy(self.raw_data).pin_memory()
        self.ring = [None if self.pinned_all is not None else
                     torch.empty(self.reshape_shape(self.shape), dtype=self.dtype, pin_memory=pin)
                     for _ in range(self.queue_size)]
        self.ring_rows = [0] * self.queue_size
        self.inflight = [None] * self.queue_size
        self.rd = 0
        self.reads = 0
        # whole dataset pinned, full batches only: every batch is a fixed view - build them once, a step then costs one list
        # lookup instead of three slice / ring updates
        self._views = None
        if self.pinned_all is not None and self.drop_last:
            self._views = [self.pinned_all[b * self.batch_size:(b + 1) * self.batch_size] for b in range(self.batch_num)]
        for i in range(self.queue_size):
            self._fill(i)
        self._inited = True

    def seek(self, batch_index):
        """Continue from batch ``batch_index`` of the epoch (checkpoint resume): the pinned ring is refilled from there."""
        self.init_states()
        b = int(batch_index) % self.batch_num
        self.index = b * self.batch_size
        self.batch_index = b
        self.rd = 0
        self.reads = 0
        for i in range(self.queue_size):
            self._fill(i)

    def reshape_shape(self, shape):
        if not self.parts:
            return tuple(shape)
        shp = list(shape)
        for i, p in self.parts.items():
            shp[i] //= p
        return tuple(shp)

    def _fill(self, slot):
        res = self.index
        if res + self.batch_size > self.samples_num and self.drop_last:
            res = 0
        end = min(res + self.batch_size, self.samples_num)
        if self.inflight[slot] is not None:              # an async H2D may still be reading this slot
            self.inflight[slot].synchronize()
            self.inflight[slot] = None
        if self.pinned_all is not None:
            self.ring[slot] = self.pinned_all[res:end]
            rows = end - res
        else:
            batch = self.raw_data[res:end]
            batch = self.reshape_tensor(batch)
            rows = batch.shape[0]
            self.ring[slot][:rows].copy_(torch.from_numpy(np.ascontiguousarray(batch)))
        self.ring_rows[slot] = rows
        self.index = end if end < self.samples_num else 0
        if self.drop_last and self.index + self.batch_size > self.samples_num:
            self.index = 0

    def get_arr(self):
        """Current batch (pinned host tensor); advances the ring and refills the
        consumed slot."""
        if self._views is not None:
            out = self._views[self.batch_index]
            self.last_batch_size = self.batch_size
            self.batch_index = (self.batch_index + 1) % self.batch_num
            self.index = self.batch_index * self.batch_size
            self.reads += 1
            return out
        slot = self.rd
        out = self.ring[slot][:self.ring_rows[slot]]
        self.last_batch_size = self.ring_rows[slot]
        # copy-out protection: the consumer copies to device asynchronously, so
        # refill the slot that was consumed two reads ago (ring depth 3)
        self.rd = (self.rd + 1) % self.queue_size
        refill = (slot + self.queue_size - 1) % self.queue_size
        if self.reads >= 1:
            self._fill(refill)
        self.reads += 1
        self.batch_index = (self.batch_index + 1) % self.batch_num
        return out

    def _get_arr(self, batchind):
        """Rows of batch ``batchind`` straight from the dataset (reference dataloader.py:88-99); does not
        touch the staging ring."""
        self.init_states()
        beg = (batchind % self.batch_num) * self.batch_size
        end = min(beg + self.batch_size, self.samples_num)
        return self.reshape_tensor(self.raw_data[beg:end])

    def get_next_arr(self):
        if self._views is not None:
            return self._views[self.batch_index]
        slot = self.rd
        return self.ring[slot][:self.ring_rows[slot]]

    def mark_inflight(self, event):
        """The executor started an async H2D of the NEXT batch; its slot must not be refilled before
        ``event`` has completed."""
        if self._views is None:                          # views of the pinned dataset are never overwritten
            self.inflight[self.rd] = event

    def set_dp_rank(self, dp_rank, dp_nrank):
        if self.dp_nrank is not None:
            assert self.dp_rank == dp_rank and self.dp_nrank == dp_nrank
        self.dp_rank, self.dp_nrank = dp_rank, dp_nrank

    def set_mp_parts(self, cur_part, parts):
        if self.parts is not None:
            assert self.cur_part == cur_part and self.parts == parts
        self.cur_part, self.parts = cur_part, parts

    def set_slices(self):
        pass

    def reshape_tensor(self, tensor):
        if not self.parts:
            return tensor
        slcs = []
        for i, n in enumerate(tensor.shape):
            if i in self.parts:
                part = n // self.parts[i]
                st = part * self.cur_part.get(i, 0)
                slcs.append(slice(st, st + part))
            else:
                slcs.append(slice(0, n))
        return tensor[tuple(slcs)]

    def get_cur_shape(self):
        return tuple(self.ring[self.rd][:self.ring_rows[self.rd]].shape)


class DataloaderOp(Op):
    """Graph node fed by one Dataloader per sub-executor name."""
    graph_safe = True     # the executor stages its batch into a static device buffer

    def __init__(self, dataloaders):
        super().__init__(DataloaderOp, [], ctx=cpu(0))
        self.dataloaders: Dict[str, Dataloader] = {dl.name: dl for dl in dataloaders}
        self.name = "DataloaderOp%d(%s)" % (self.id, '_'.join(self.dataloaders.keys()))

    @property
    def desc(self):
        return self.name

    def set_dp_rank(self, dp_rank, dp_nrank):
        for dl in self.dataloaders.values():
            dl.set_dp_rank(dp_rank, dp_nrank)

    def set_mp_parts(self, cur_part, parts):
        for dl in self.dataloaders.values():
            dl.set_mp_parts(cur_part, parts)

    def get_batch_num(self, name):
        return self.dataloaders[name].batch_num

    def get_arr(self, name):
        return self.dataloaders[name].get_arr()

    def get_next_arr(self, name):
        return self.dataloaders[name].get_next_arr()

    def mark_inflight(self, name, event):
        self.dataloaders[name].mark_inflight(event)

    def get_cur_shape(self, name):
        return self.dataloaders[name].get_cur_shape()

    def compute(self, input_vals, output_val=None, stream_handle=None):
        raise AssertionError('DataloaderOp is fed by the executor')

    def gradient(self, output_grad):
        return None

    def infer_shape(self, input_shapes):
        raise NotImplementedError

    def forward_hook(self, config):
        pass

    def backward_hook(self, config):
        self.ctx = cpu(0)
        if getattr(config, 'dp_nrank', 1) and config.dp_nrank > 1 and config.shard_dataloaders:
            self.set_dp_rank(config.dp_rank, config.dp_nrank)
        for dl in self.dataloaders.values():
            dl.init_states()


class GNNDataLoaderOp(Op):
    """Feeds graph mini-batches produced by a user handler (``handler(graph)``
    returns the numpy array for the current graph).  reference dataloader.py:147-183."""
    graph = None
    nxt_graph = None
    graph_safe = False

    def __init__(self, handler, ctx=cpu(0)):
        super().__init__(GNNDataLoaderOp, [], ctx=ctx)
        self.handler = handler
        self.name = "GNNDataloaderOp%d" % self.id

    @property
    def desc(self):
        return self.name

    def get_batch_num(self, name):
        return None

    def get_arr(self, name):
        return torch.as_tensor(np.asarray(self.handler(self.graph), dtype=np.float32))

    def get_next_arr(self, name):
        return torch.as_tensor(np.asarray(self.handler(self.nxt_graph), dtype=np.float32))

    def get_cur_shape(self, name):
        return tuple(self.handler(self.graph).shape)

    def gradient(self, output_grad):
        return None

    def infer_shape(self, input_shapes):
        raise NotImplementedError

    def backward_hook(self, config):
        pass

    @classmethod
    def step(cls, graph):
        cls.graph = cls.nxt_graph
        cls.nxt_graph = graph


def dataloader_op(dataloaders):
    """``dataloaders``: list of ``Dataloader`` or of ``[raw_data, batch_size, name(, func)]``."""
    temp = []
    for dl in dataloaders:
        temp.append(dl if isinstance(dl, Dataloader) else Dataloader(*dl))
    return DataloaderOp(temp)
