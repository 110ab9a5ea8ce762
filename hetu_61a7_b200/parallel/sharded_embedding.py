"""GPU-sharded embedding table over the NVSwitch domain.

reference: embeddings live in host DRAM behind ps-lite; every step the worker pulls rows
(SparsePull / HET cache ``embedding_lookup``), copies them H2D, and pushes ``-lr * grad``
back (``SparsePush``; server does ``+=``) - python/hetu/gpu_ops/ParameterServerCommunicate.py:
38-57, ps-lite/include/ps/server/PSFHandle.h:101-164.

Here row ``r`` lives in the HBM of GPU ``r % world`` at local row ``r // world`` inside a
CUDA-IPC symmetric buffer.  The lookup kernel reads rows straight from the owner over
NVLink (``hetu_gather_rows_sharded``); the update kernel adds ``-lr * grad`` into the owner's
rows with remote atomics (``hetu_sparse_sgd_sharded``) - the PS pull, push and server-side
``+=`` are one kernel each, with ASP semantics by default and BSP when ``bsp == 0``
(a cross-GPU barrier after the update).
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _lib as L
from ..ops.embedding import _IDX_CODE


def should_shard(node, config) -> bool:
    if not (getattr(node, 'is_embed', False) and node.trainable and config.symm is not None):
        return False
    if config.ps_comm is not None:
        return False
    strategy = config.node_strategy.get(node, config.comm_mode)
    return strategy in ('Hybrid', 'PS') or config.comm_mode in ('Hybrid', 'PS')


class ShardedTable(object):
    sharded = True

    def __init__(self, engine, shape, initializer, seed, name='embedding', bsp=-1):
        self.engine = engine
        self.shape = tuple(shape)
        self.vocab, self.width = int(shape[0]), int(shape[1])
        self.world, self.rank = engine.world, engine.rank
        self.local_rows = (self.vocab + self.world - 1) // self.world
        self.bsp = bsp
        self.name = name
        nbytes = self.local_rows * self.width * 4
        self.buf = engine.buffer(('embed', name), nbytes)
        self.local = self.buf.tensor(numel=self.local_rows * self.width).view(self.local_rows, self.width)
        self.device = engine.device
        self.dtype = torch.float32
        if initializer is not None:
            a, b = initializer.args()
            L.call('init_random', L.p(self.local), L.c_sz(self.local.numel()), L.c_int(initializer.kind),
                   L.c_f32(a), L.c_f32(b), L.c_u64((seed * 1000003 + self.rank) & 0xFFFFFFFFFFFFFFFF),
                   L.stream_ptr(self.device))
        torch.cuda.synchronize(self.device)
        engine.comm.barrier()

    def numel(self):
        return self.vocab * self.width

    def lookup(self, idx: torch.Tensor) -> torch.Tensor:
        if idx.device.type == 'meta':
            return torch.empty(tuple(idx.shape) + (self.width,), device='meta')
        idx = idx.contiguous()
        if idx.dtype not in _IDX_CODE:
            idx = idx.long()
        out = torch.empty(tuple(idx.shape) + (self.width,), dtype=torch.float32, device=self.device)
        L.call('gather_rows_sharded', self.buf.ptrs, L.c_int(self.world), L.p(idx), L.c_int(_IDX_CODE[idx.dtype]),
               L.p(out), L.c_sz(idx.numel()), L.c_int(self.width), L.c_sz(self.vocab), L.stream_ptr(self.device))
        return out

    def _owner_state(self, optimizer, n_ids):
        """Accumulation shard, per-row flags and work list (symmetric: peers write them with remote atomics) and the
        optimizer's state shards (local) for the optimizers that are applied at the owner."""
        st = getattr(self, '_owner', None)
        cap = self.world * n_ids
        if st is None or st['cap'] < cap:
            eng, nrow, w = self.engine, self.local_rows, self.width
            acc = eng.buffer(('embed_acc', self.name), nrow * w * 4)
            flag = eng.buffer(('embed_flag', self.name), nrow * 4)
            lst = eng.buffer(('embed_list', self.name, cap), cap * 4)
            cnt = eng.buffer(('embed_cnt', self.name), 256)
            acc.tensor(numel=nrow * w).zero_()
            flag.tensor(dtype=torch.int32, numel=nrow).zero_()
            cnt.tensor(dtype=torch.int32, numel=64).zero_()
            old = st or {}
            st = self._owner = dict(cap=cap, acc=acc, flag=flag, list=lst, cnt=cnt,
                                    done=old.get('done', torch.zeros(1, dtype=torch.int32, device=self.device)),
                                    s1=old.get('s1', torch.zeros((nrow, w), dtype=torch.float32, device=self.device)),
                                    s2=old.get('s2'))
            torch.cuda.synchronize(self.device)
            eng.comm.barrier()
        from ..optimizer import OPT_ADAM
        if optimizer.kind >= OPT_ADAM and st['s2'] is None:
            st['s2'] = torch.zeros((self.local_rows, self.width), dtype=torch.float32, device=self.device)
        return st

    def apply_sparse_grad(self, ind: torch.Tensor, val: torch.Tensor, optimizer) -> None:
        from ..optimizer import OPT_SGD
        if optimizer.kind != OPT_SGD:
            # optimizers with state: gradients are summed at the owner (remote atomics = de-duplication across the
            # whole job), then every owner applies the step to the rows on its work list (csrc/embedding.cu)
            ind = ind.contiguous()
            val = val.contiguous()
            st = self._owner_state(optimizer, ind.numel())
            h = optimizer.hyper()
            L.call('sparse_push_sharded', st['acc'].ptrs, st['flag'].ptrs, st['list'].ptrs, st['cnt'].ptrs, L.c_int(self.world),
                   L.c_int(st['cap']), L.p(ind), L.c_int(_IDX_CODE[ind.dtype]), L.p(val), L.c_sz(ind.numel()),
                   L.c_int(self.width), L.c_sz(self.vocab), L.c_f32(1.0), L.stream_ptr(self.device))
            self.engine.barrier()                          # every rank's pushes have landed
            L.call('sparse_apply_owner', L.c_int(optimizer.kind), L.p(self.local), ctypes.c_void_p(st['acc'].local_ptr),
                   ctypes.c_void_p(st['flag'].local_ptr), ctypes.c_void_p(st['list'].local_ptr),
                   ctypes.c_void_p(st['cnt'].local_ptr), L.p(st['done']), L.p(st['s1']), L.p(st['s2']), L.c_int(self.width),
                   L.c_int(st['cap']), L.p(optimizer.lr_t), L.p(optimizer.step_t), L.c_f32(h[0]), L.c_f32(h[1]),
                   L.c_f32(h[2]), L.c_f32(h[3]), L.c_f32(float(optimizer.l2reg)), L.stream_ptr(self.device))
            self.engine.barrier()                          # rows are updated and the work lists empty before the next step
            return
        ind = ind.contiguous()
        if self.bsp == 0:
            self.engine.barrier()      # BSP: every rank has finished reading this step's rows
        L.call('sparse_sgd_sharded', self.buf.ptrs, L.c_int(self.world), L.p(ind), L.c_int(_IDX_CODE[ind.dtype]),
               L.p(val.contiguous()), L.c_sz(ind.numel()), L.c_int(self.width), L.c_sz(self.vocab),
               L.p(optimizer.lr_t), L.c_f32(1.0), L.stream_ptr(self.device))
        if self.bsp == 0:
            self.engine.barrier()

    # -- checkpointing ---------------------------------------------------------------------------
    def export_dense(self) -> torch.Tensor:
        """Full table on the host (rows fetched through the same peer-gather kernel, in chunks)."""
        out = torch.empty((self.vocab, self.width), dtype=torch.float32)
        step = 1 << 20
        for s in range(0, self.vocab, step):
            ids = torch.arange(s, min(self.vocab, s + step), device=self.device)
            out[s:s + ids.numel()] = self.lookup(ids).cpu()
        return out

    def import_dense(self, value) -> None:
        v = torch.as_tensor(np.asarray(value), dtype=torch.float32)
        mine = v[self.rank::self.world]
        self.local[:mine.shape[0]].copy_(mine.to(self.device))
        torch.cuda.synchronize(self.device)
        self.engine.comm.barrier()

    def detach(self):
        return self

    def cpu(self):
        return self.export_dense()
