"""Symmetric heap + in-kernel NVLink collectives (host side of csrc/collectives.cu).

Every rank allocates the same buffers (``cudaMalloc`` outside torch's caching
allocator), exports them with CUDA IPC and maps all peers' copies; the pointer
tables go to the kernels by value.  ``torch.distributed`` is used only to ship
the 64-byte IPC handles around.

reference counterpart: src/communication/mpi_nccl_communication.cu (NCCL calls)
and python/hetu/optimizer.py:145-163 (one all-reduce op + one update per tensor).
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import _lib as L

_PTRS16 = ctypes.c_void_p * 16


class _CudaArray(object):
    """Zero-copy torch view of raw device memory via __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {'shape': (nbytes,), 'typestr': '|u1', 'data': (ptr, False),
                                         'version': 2, 'strides': None}


class SymmBuffer(object):
    """A buffer that exists at the same size on every rank; ``local`` is this rank's
    copy as a tensor, ``ptrs`` the table of all ranks' base pointers (peer-mapped)."""

    def __init__(self, engine: 'SymmEngine', nbytes: int):
        self.engine = engine
        self.nbytes = int((nbytes + 255) // 256 * 256)
        lib = L.require_cuda_lib()
        ptr = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        L.check(lib.hetu_symm_alloc(L.c_sz(self.nbytes), ctypes.byref(ptr), handle), 'symm_alloc')
        self.local_ptr = ptr.value
        handles = engine.comm.allgather_object(handle.raw)
        addrs = []
        for r, h in enumerate(handles):
            if r == engine.rank:
                addrs.append(self.local_ptr)
            else:
                p = ctypes.c_void_p()
                L.check(lib.hetu_symm_open(ctypes.create_string_buffer(h, 64), ctypes.byref(p)), 'symm_open')
                addrs.append(p.value)
        self.addrs = addrs
        self.ptrs = _PTRS16(*(addrs + [0] * (16 - len(addrs))))
        self._holder = _CudaArray(self.local_ptr, self.nbytes)
        self.bytes_view = torch.as_tensor(self._holder, device=engine.device)
        engine.comm.barrier()

    def tensor(self, dtype=torch.float32, offset_bytes: int = 0, numel: Optional[int] = None) -> torch.Tensor:
        v = self.bytes_view[offset_bytes:]
        t = v.view(dtype)
        return t if numel is None else t[:numel]

    def peer_ptrs(self, offset_bytes: int = 0):
        return _PTRS16(*([a + offset_bytes for a in self.addrs] + [0] * (16 - len(self.addrs))))


class MulticastBuffer(object):
    """Symmetric buffer that also has an NVLS multicast mapping: a store to ``mc_ptr + off`` lands at offset ``off``
    of EVERY rank's copy (the NVSwitch replicates it), a ``multimem.ld_reduce`` there returns the sum over the ranks.
    The allocation (cuMemCreate + cuMulticastCreate / BindMem and the exchange of the handles between the
    processes) is torch's symmetric-memory allocator - plumbing only; the kernels that use the mapping are ours
    (csrc/wdl_tower.cu, csrc/collectives.cu).  ``mc_ptr`` is 0 where the fabric has no multicast support."""

    def __init__(self, engine: 'SymmEngine', nbytes: int):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as sm
        self.engine = engine
        self.nbytes = int((nbytes + 255) // 256 * 256)
        group = engine.comm.group if engine.comm.group is not None else dist.group.WORLD
        self._t = sm.empty(self.nbytes // 4, dtype=torch.float32, device=engine.device)
        self._h = sm.rendezvous(self._t, group.group_name)
        self.addrs = [int(p_) for p_ in self._h.buffer_ptrs]
        self.local_ptr = self.addrs[engine.rank]
        self.mc_ptr = int(getattr(self._h, 'multicast_ptr', 0) or 0)
        self.ptrs = _PTRS16(*(self.addrs + [0] * (16 - len(self.addrs))))
        self._t.zero_()
        torch.cuda.synchronize(engine.device)
        engine.comm.barrier()

    def tensor(self, dtype=torch.float32, offset_bytes: int = 0, numel: Optional[int] = None) -> torch.Tensor:
        v = self._t.view(torch.uint8)[offset_bytes:]
        t = v.view(dtype)
        return t if numel is None else t[:numel]

    def peer_ptrs(self, offset_bytes: int = 0):
        return _PTRS16(*([a + offset_bytes for a in self.addrs] + [0] * (16 - len(self.addrs))))


class _ArenaView(object):
    """A sub-range of the arena with the SymmBuffer interface (tensor / ptrs / peer_ptrs)."""

    def __init__(self, arena: SymmBuffer, offset: int, nbytes: int):
        self.arena, self.offset, self.nbytes = arena, offset, nbytes
        self.addrs = [a + offset for a in arena.addrs]
        self.local_ptr = arena.local_ptr + offset
        self.ptrs = _PTRS16(*(self.addrs + [0] * (16 - len(self.addrs))))

    def tensor(self, dtype=torch.float32, offset_bytes: int = 0, numel: Optional[int] = None) -> torch.Tensor:
        return self.arena.tensor(dtype, self.offset + offset_bytes, numel)

    def peer_ptrs(self, offset_bytes: int = 0):
        return _PTRS16(*([a + offset_bytes for a in self.addrs] + [0] * (16 - len(self.addrs))))


class SymmEngine(object):
    """Per-process singleton bound to the world communicator."""
    _inst: Optional['SymmEngine'] = None
    MAX_FUSED_BYTES = int(os.environ.get('HETU_SYMM_MAX_FUSED_BYTES', str(64 << 20)))

    @staticmethod
    def get(comm) -> Optional['SymmEngine']:
        if SymmEngine._inst is None:
            try:
                SymmEngine._inst = SymmEngine(comm)
            except Exception as exc:                  # no peer access etc. -> NCCL baseline
                import warnings
                warnings.warn('symmetric-heap collectives unavailable (%s); using NCCL' % exc)
                SymmEngine._inst = False
        return SymmEngine._inst or None

    def __init__(self, comm):
        assert comm.multi and comm.nrank <= 16
        self.comm = comm
        self.rank, self.world = comm.rank, comm.nrank
        self.device = torch.device('cuda', torch.cuda.current_device())
        self.flags = SymmBuffer(self, 4096)                       # uint32 flags[world] (+ padding)
        self.epoch = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.gate = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.done_ctr = torch.zeros(4, dtype=torch.int32, device=self.device)
        self._bufs: Dict[tuple, SymmBuffer] = {}
        self._parity: Dict[tuple, int] = {}
        torch.cuda.synchronize(self.device)
        comm.barrier()

    # -- buffers -----------------------------------------------------------------------------
    def buffer(self, key, nbytes: int) -> SymmBuffer:
        """Symmetric scratch keyed by (op identity, size).  Single-buffered: every collective
        kernel closes with a cross-GPU barrier, so the buffer is free again when it returns
        (needed because CUDA-graph replays bake the address in).

        With ``HETU_SYMM_ARENA_MB=<n>`` all buffers are carved out of ONE symmetric arena by the
        best-fit-with-coalescing allocator (memory_pool.BFCAllocator): a single cudaMalloc + IPC
        exchange instead of one per buffer; every rank runs the same allocation sequence, so the
        offsets agree on all GPUs."""
        k = (key, int(nbytes))
        if k not in self._bufs:
            arena_mb = int(os.environ.get('HETU_SYMM_ARENA_MB', '0'))
            if arena_mb > 0:
                if getattr(self, '_arena', None) is None:
                    from ..memory_pool import BFCAllocator
                    self._arena = SymmBuffer(self, arena_mb << 20)
                    self._arena_alloc = BFCAllocator(arena_mb << 20, 256)
                off = self._arena_alloc.alloc(int(nbytes))
                if off >= 0:
                    self._bufs[k] = _ArenaView(self._arena, off, int(nbytes))
                    return self._bufs[k]
            self._bufs[k] = SymmBuffer(self, int(nbytes))
        return self._bufs[k]

    def next_slot(self, key, nbytes: int) -> int:
        return 0

    def mc_buffer(self, key, nbytes: int) -> Optional[MulticastBuffer]:
        """Zeroed symmetric buffer with an NVLS multicast mapping, or None where that is not available
        (``HETU_NVLS=0`` switches it off)."""
        if os.environ.get('HETU_NVLS', '1') != '1' or getattr(self, '_mc_broken', False):
            return None
        k = ('mc', key, int(nbytes))
        if k not in self._bufs:
            try:
                buf = MulticastBuffer(self, int(nbytes))
            except Exception as exc:                      # old driver / no NVSwitch multicast: peer pointers remain
                import warnings
                warnings.warn('NVLS multicast buffers unavailable (%s); using peer stores' % exc)
                self._mc_broken = True
                return None
            if not buf.mc_ptr:
                self._mc_broken = True
                return None
            self._bufs[k] = buf
        return self._bufs[k]

    def _stream(self):
        return L.stream_ptr(self.device)

    # -- collectives ---------------------------------------------------------------------------
    def barrier(self):
        L.call('symm_barrier', self.flags.ptrs, L.c_int(self.rank), L.c_int(self.world), L.p(self.epoch),
               self._stream())

    # -- NVLS (in-switch) collectives ------------------------------------------------------------
    def _nvls(self, t: torch.Tensor, key, mode: int, root: int = 0, scale: float = 1.0) -> Optional[torch.Tensor]:
        """mode 0 all-reduce, 1 broadcast from root, 2 reduce to root (csrc/collectives.cu hetu_nvls_collective).
        None when the tensor does not qualify (no multicast mapping, numel % 4, not fp32)."""
        n = t.numel()
        if t.dtype != torch.float32 or n % 4 or n == 0:
            return None
        buf = self.mc_buffer((key, 'nvls'), n * 4)
        if buf is None:
            return None
        t = t.contiguous()
        local = buf.tensor(numel=n)
        if mode != 1:
            local.copy_(t.reshape(-1))                   # this rank's contribution, in its copy of the buffer
        out = torch.empty_like(t) if mode == 2 else None
        L.call('nvls_collective', ctypes.c_void_p(buf.mc_ptr), L.p(t if mode == 1 else None), L.p(out), L.c_i64(n),
               L.c_int(mode), L.c_int(root), L.c_f32(scale), self.flags.ptrs, L.c_int(self.rank), L.c_int(self.world),
               L.p(self.epoch), self._stream())
        if mode == 2:
            return out if self.rank == root else t.clone()
        return local.view(t.shape).clone()

    def broadcast(self, t: torch.Tensor, root: int = 0, key='bcast') -> Optional[torch.Tensor]:
        return self._nvls(t, (key, tuple(t.shape)), 1, root)

    def reduce(self, t: torch.Tensor, root: int = 0, key='reduce') -> Optional[torch.Tensor]:
        return self._nvls(t, (key, tuple(t.shape)), 2, root)

    def allreduce(self, t: torch.Tensor, key='allreduce') -> torch.Tensor:
        if os.environ.get('HETU_NVLS_ALLREDUCE', '1') == '1' and t.numel() * 4 >= (1 << 16):
            out = self._nvls(t, (key, tuple(t.shape)), 0)
            if out is not None:
                return out
        t = t.contiguous()
        n = t.numel()
        nbytes = n * 4
        buf = self.buffer((key, 'ar'), nbytes)
        off = self.next_slot((key, 'ar'), nbytes)
        buf.tensor(offset_bytes=off, numel=n).copy_(t.reshape(-1))
        out = torch.empty_like(t)
        L.call('symm_allreduce', buf.ptrs, L.c_sz(off), self.flags.ptrs, L.c_int(self.rank), L.c_int(self.world),
               L.p(self.epoch), L.p(self.gate), L.c_i64(n), L.p(out), self._stream())
        return out

    def reducescatter(self, t: torch.Tensor, key='rs') -> torch.Tensor:
        t = t.contiguous()
        n = t.numel()
        chunk = n // self.world
        if chunk % 4:
            return self.comm.reducescatter(t)
        nbytes = n * 4
        buf = self.buffer((key, 'rs'), nbytes)
        off = self.next_slot((key, 'rs'), nbytes)
        buf.tensor(offset_bytes=off, numel=n).copy_(t.reshape(-1))
        out = torch.empty((t.shape[0] // self.world,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        L.call('symm_reduce_scatter', buf.ptrs, L.c_sz(off), self.flags.ptrs, L.c_int(self.rank),
               L.c_int(self.world), L.p(self.epoch), L.p(self.gate), L.c_i64(chunk), L.p(out), self._stream())
        return out

    def _exchange(self, t: torch.Tensor, same_chunk: bool, key) -> torch.Tensor:
        t = t.contiguous()
        chunk = t.numel() if same_chunk else t.numel() // self.world
        nbytes = chunk * self.world * 4
        buf = self.buffer((key, 'x'), nbytes)
        off = self.next_slot((key, 'x'), nbytes)
        L.call('symm_exchange', L.p(t), buf.ptrs, L.c_sz(off), self.flags.ptrs, L.c_int(self.rank),
               L.c_int(self.world), L.p(self.epoch), L.p(self.gate), L.c_i64(chunk),
               L.c_int(int(same_chunk)), self._stream())
        recv = buf.tensor(offset_bytes=off, numel=chunk * self.world)
        if same_chunk:
            return recv.reshape((t.shape[0] * self.world,) + tuple(t.shape[1:]))
        return recv.reshape(t.shape)

    def alltoall(self, t: torch.Tensor, key='a2a') -> torch.Tensor:
        """Returns a view of the symmetric receive buffer (valid until this op's next call)."""
        if (t.numel() // self.world) % 4:
            return self.comm.alltoall(t)
        return self._exchange(t, False, (key, tuple(t.shape)))

    def allgather(self, t: torch.Tensor, key='ag') -> torch.Tensor:
        if t.numel() % 4:
            return self.comm.allgather(t)
        return self._exchange(t, True, (key, tuple(t.shape)))


class FusedDPUpdate(object):
    """Dense data-parallel step: pack local grads into the symmetric arena, then one kernel
    does barrier + cross-GPU sum + optimizer update (K1).  Falls back to symm all-reduce +
    multi-tensor step for optimizers without a fused path (Lamb) or very large models."""

    def __init__(self, engine: SymmEngine, param_ids: Sequence[int]):
        self.engine = engine
        self.ids = list(param_ids)
        self._layout = None

    def _build(self, optimizer, grads):
        offs, sizes, off = [], [], 0
        for i in self.ids:
            n = optimizer.tensors[i].numel()
            offs.append(off)
            sizes.append(n)
            off += (n + 3) // 4 * 4
        quantum = 4 * self.engine.world                 # equal, 16-byte aligned slices for the two-shot path
        self.total = (off + quantum - 1) // quantum * quantum
        self.offs, self.sizes = offs, sizes
        self.key = ('dp', id(self))
        self.buf = self.engine.buffer(self.key, self.total * 4)
        self._layout = True

    def step(self, optimizer, grads):
        from ..optimizer import OPT_LAMB
        eng = self.engine
        if self._layout is None:
            self._build(optimizer, grads)
        ids = self.ids
        dense = [grads[i] for i in ids]
        assert all(g is not None for g in dense)
        off = eng.next_slot(self.key, self.total * 4)
        arena = self.buf.tensor(offset_bytes=off, numel=self.total)
        n = len(ids)
        PT = ctypes.c_void_p * n
        SZ = ctypes.c_int64 * n
        L.call('pack_tensors', L.c_int(n), PT(*[g.contiguous().data_ptr() for g in dense]), SZ(*self.offs),
               SZ(*self.sizes), L.p(arena), L.c_f32(1.0), eng._stream())
        fused_ok = optimizer.kind != OPT_LAMB and self.total * 4 <= SymmEngine.MAX_FUSED_BYTES
        optimizer._bump_step()
        rest = [i for i in range(len(grads)) if i not in ids and grads[i] is not None]
        if fused_ok:
            h = optimizer.hyper()
            for s in range(0, n, 48):
                e = min(n, s + 48)
                sub = ids[s:e]
                m = len(sub)
                PT2 = ctypes.c_void_p * m
                SZ2 = ctypes.c_int64 * m
                st = [optimizer.states[i] for i in sub]
                if s > 0:
                    # later chunks reuse the already-synchronised arena: plain barrier-free read is
                    # safe because every rank finished packing before the first chunk's barrier
                    pass
                L.call('symm_allreduce_opt', self.buf.ptrs, L.c_sz(off), eng.flags.ptrs, L.c_int(eng.rank),
                       L.c_int(eng.world), L.p(eng.epoch), L.p(eng.gate), L.c_int(optimizer.kind), L.c_int(m),
                       PT2(*[optimizer.tensors[i].data_ptr() for i in sub]),
                       PT2(*[(x[0].data_ptr() if len(x) > 0 else 0) for x in st]),
                       PT2(*[(x[1].data_ptr() if len(x) > 1 else 0) for x in st]),
                       SZ2(*self.offs[s:e]), SZ2(*self.sizes[s:e]), L.p(optimizer.lr_t), L.p(optimizer.step_t),
                       L.c_f32(h[0]), L.c_f32(h[1]), L.c_f32(h[2]), L.c_f32(h[3]), L.c_f32(float(optimizer.l2reg)),
                       eng._stream())
        else:
            if self.total * 4 > SymmEngine.MAX_FUSED_BYTES:
                # large models: two-shot all-reduce.  The one-shot kernels make every rank read every
                # peer's whole buffer ((world-1) x N bytes inbound - fine for latency-bound small models,
                # 9.4 GB per step for BERT-large on 8 GPUs); reduce-scatter by peer loads + all-gather by
                # peer stores moves 2 (world-1)/world x N instead.
                chunk = self.total // eng.world
                red = torch.empty(chunk, dtype=torch.float32, device=eng.device)
                L.call('symm_reduce_scatter', self.buf.ptrs, L.c_sz(off), eng.flags.ptrs, L.c_int(eng.rank),
                       L.c_int(eng.world), L.p(eng.epoch), L.p(eng.gate), L.c_i64(chunk), L.p(red), eng._stream())
                out = eng.allgather(red, key=('dp_ag', id(self))).reshape(-1)
            else:
                out = torch.empty(self.total, dtype=torch.float32, device=eng.device)
                L.call('symm_allreduce', self.buf.ptrs, L.c_sz(off), eng.flags.ptrs, L.c_int(eng.rank),
                       L.c_int(eng.world), L.p(eng.epoch), L.p(eng.gate), L.c_i64(self.total), L.p(out), eng._stream())
            ps = [optimizer.tensors[i] for i in ids]
            gs = [out[o:o + sz].view(p.shape) for o, sz, p in zip(self.offs, self.sizes, ps)]
            optimizer._dense_kernel(ps, gs, [optimizer.states[i] for i in ids])
        # parameters outside the fused set (experts, sparse) take the ordinary local step
        if rest:
            saved = optimizer._bump_step
            optimizer._bump_step = lambda: None
            try:
                masked = [g if i in rest else None for i, g in enumerate(grads)]
                optimizer.update(masked)
            finally:
                optimizer._bump_step = saved
