"""Device contexts and array containers.

Storage is a ``torch.Tensor`` (HBM3e for ``gpu(i)``, pinned/pageable host memory
for ``cpu(i)``); there is no DLArray/ctypes struct layer.  The public surface
mirrors the reference (python/hetu/ndarray.py:10-618): ``DLContext``, ``cpu``,
``gpu``, ``rcpu``, ``rgpu``, ``NDArray``, ``array``, ``empty``,
``sparse_array``/``ND_Sparse_Array`` (CSR) and ``IndexedSlices``.

Indices are kept as integer tensors on device (the reference stores them as
float32, which loses precision above 2**24 rows - SURVEY.md section 0).
"""
from __future__ import annotations

import socket
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

CPU_TYPE = 1
GPU_TYPE = 2
_SELF_HOST = socket.gethostname()


class DLContext(object):
    """A (host, device type, device id) triple.  reference: ndarray.py:10-57."""

    MASK2STR = {CPU_TYPE: 'cpu', GPU_TYPE: 'gpu'}
    __slots__ = ('device_id', 'device_type', 'hostname', 'local')

    def __init__(self, device_id, device_type, hostname='localhost'):
        self.device_id = int(device_id)
        self.device_type = int(device_type)
        if hostname in ('localhost', '127.0.0.1', _SELF_HOST):
            self.hostname = _SELF_HOST
            self.local = True
        else:
            self.hostname = hostname
            self.local = False

    def __repr__(self):
        base = "%s(%d)" % (self.MASK2STR[self.device_type], self.device_id)
        return base if self.local else "%s:%s" % (self.hostname, base)

    def full_repr(self):
        return "%s:%s:%d" % (self.hostname, self.MASK2STR[self.device_type], self.device_id)

    def relocalize(self):
        self.local = self.hostname in ('localhost', _SELF_HOST)

    def _key(self):
        if self.local:
            return (self.device_type, self.device_id)
        return (self.hostname, self.device_type, self.device_id)

    def __hash__(self):
        return hash(self._key())

    def __eq__(self, other):
        return isinstance(other, DLContext) and self._key() == other._key()

    def __ne__(self, other):
        return not self.__eq__(other)

    @property
    def torch_device(self) -> torch.device:
        """Where tensors of this context live.  A gpu context on a box without a
        GPU (CI, graph-pass unit tests) degrades to host memory."""
        if self.device_type == GPU_TYPE and torch.cuda.is_available():
            return torch.device('cuda', self.device_id % max(torch.cuda.device_count(), 1))
        return torch.device('cpu')


def cpu(dev_id=0):
    return DLContext(dev_id, CPU_TYPE)


def gpu(dev_id=0):
    return DLContext(dev_id, GPU_TYPE)


def rcpu(hostname, dev_id=0):
    return DLContext(dev_id, CPU_TYPE, hostname=hostname)


def rgpu(hostname, dev_id=0):
    return DLContext(dev_id, GPU_TYPE, hostname=hostname)


def is_gpu_ctx(ctx):
    return bool(ctx) and isinstance(ctx, DLContext) and ctx.device_type == GPU_TYPE


def shape_to_stride(shape):
    ndim = len(shape)
    stride = [1] * ndim
    for i in range(ndim - 1, 0, -1):
        stride[i - 1] = stride[i] * shape[i]
    return tuple(stride)


_NP2TORCH = {
    np.dtype('float32'): torch.float32, np.dtype('float64'): torch.float32,
    np.dtype('float16'): torch.float16, np.dtype('int32'): torch.int32,
    np.dtype('int64'): torch.int64, np.dtype('bool'): torch.bool,
    np.dtype('uint8'): torch.uint8, np.dtype('int8'): torch.int8,
}


def to_tensor(value, device=None, dtype=None) -> torch.Tensor:
    """numpy / NDArray / python scalar / tensor -> torch tensor on ``device``."""
    if isinstance(value, NDArray):
        t = value.tensor
    elif isinstance(value, torch.Tensor):
        t = value
    else:
        arr = np.asarray(value)
        if arr.dtype == np.float64:
            arr = arr.astype(np.float32)
        t = torch.from_numpy(np.ascontiguousarray(arr))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if device is not None and t.device != torch.device(device):
        t = t.to(device, non_blocking=True)
    return t


_PINNED_RING = {}


def _pinned_copy(t):
    """Copies ``t`` into the next slot of a small ring of pinned host buffers (one ring per shape / dtype / device);
    eight copies may be in flight before a slot is reused."""
    import torch
    key = (tuple(t.shape), t.dtype, t.device.index)
    ring = _PINNED_RING.get(key)
    if ring is None:
        ring = _PINNED_RING[key] = {'bufs': [torch.empty(tuple(t.shape), dtype=t.dtype).pin_memory() for _ in range(8)],
                                    'events': [torch.cuda.Event() for _ in range(8)], 'turn': 0}
    i = ring['turn'] % 8
    ring['turn'] += 1
    buf, ev = ring['bufs'][i], ring['events'][i]
    buf.copy_(t, non_blocking=True)
    ev.record()                                          # on the current stream of the current device (the producer's)
    return buf, ev


class HostFuture(object):
    """Result of ``NDArray.asnumpy_async``: ``get()`` returns the numpy value once its copy has landed."""

    def __init__(self, host_tensor, event):
        self._t, self._ev = host_tensor, event

    def ready(self) -> bool:
        return self._ev is None or self._ev.query()

    def get(self) -> np.ndarray:
        if self._ev is not None:
            self._ev.synchronize()
        return self._t.numpy().copy()


class NDArray(object):
    """Thin handle over a torch tensor.  reference: ndarray.py:140-403."""

    __slots__ = ('tensor', '_ctx', '__weakref__')

    def __init__(self, tensor: torch.Tensor, ctx: Optional[DLContext] = None):
        self.tensor = tensor
        if ctx is None:
            ctx = gpu(tensor.device.index or 0) if tensor.is_cuda else cpu(0)
        self._ctx = ctx

    @property
    def shape(self) -> Tuple[int, ...]:
        return tuple(self.tensor.shape)

    @property
    def stride(self):
        return tuple(self.tensor.stride())

    @property
    def lazy(self):
        return not self.tensor.is_contiguous()

    @property
    def ctx(self) -> DLContext:
        return self._ctx

    @property
    def dtype(self):
        return self.tensor.dtype

    def __len__(self):
        return self.tensor.shape[0]

    def __setitem__(self, in_slice, value):
        if not (isinstance(in_slice, slice) and in_slice.start is None and in_slice.stop is None):
            raise ValueError('Array only support set from numpy array')
        src = to_tensor(value, self.tensor.device, self.tensor.dtype)
        if tuple(src.shape) != self.shape:
            raise ValueError('array shape do not match the shape of NDArray')
        self.tensor.copy_(src, non_blocking=True)

    def _sync_copyfrom(self, source_array, data_type=np.float32):
        self[:] = source_array
        if self.tensor.is_cuda:
            torch.cuda.current_stream(self.tensor.device).synchronize()

    def async_h2d(self, source_array, stream_handle=None, event_handle=None):
        src = source_array.tensor if isinstance(source_array, NDArray) else to_tensor(source_array)
        if stream_handle is not None and self.tensor.is_cuda:
            with torch.cuda.stream(stream_handle.handle):
                self.tensor.copy_(src, non_blocking=True)
                if event_handle is not None:
                    event_handle.record(stream_handle)
        else:
            self.tensor.copy_(src, non_blocking=True)

    async_d2h = async_h2d
    _async_copyfrom = async_h2d

    def asnumpy(self) -> np.ndarray:
        return self.tensor.detach().cpu().numpy()

    def asnumpy_async(self) -> 'HostFuture':
        """Device -> host copy that does not stall the host: the copy into a pinned staging buffer is queued on the
        current stream behind the work that produces this array, and the returned handle's ``get()`` waits for
        that copy only.  A training loop reads step i's loss while step i+1 is already queued, so the host never
        idles on the device (the reference's ``asnumpy`` is a synchronous cudaMemcpy, python/hetu/ndarray.py:265-279).
        """
        t = self.tensor.detach()
        if t.device.type != 'cuda':
            return HostFuture(t.cpu().clone(), None)
        return HostFuture(*_pinned_copy(t))

    def copyto(self, target):
        if isinstance(target, DLContext):
            return NDArray(self.tensor.to(target.torch_device, copy=True), target)
        if isinstance(target, NDArray):
            target.tensor.copy_(self.tensor, non_blocking=True)
            return target
        raise ValueError("Unsupported target type %s" % str(type(target)))

    def reshape(self, shape, target=None):
        out = NDArray(self.tensor.reshape(tuple(int(s) for s in shape)), self._ctx)
        if target is not None:
            target.tensor = out.tensor
            return target
        return out

    def broadcast_to(self, shape, target=None, add_axes=None):
        t = self.tensor
        if add_axes:
            for ax in sorted(add_axes):
                t = t.unsqueeze(ax)
        out = NDArray(t.expand(*shape), self._ctx)
        if target is not None:
            target.tensor = out.tensor
            return target
        return out

    def inplace_copy(self, target):
        target.tensor = self.tensor
        return target

    def lazy_callback(self, stream=None):
        self.tensor = self.tensor.contiguous()

    wrapped_lazy_callback = lazy_callback

    def __repr__(self):
        return 'NDArray(shape=%s, ctx=%s)' % (self.shape, self._ctx)


def array(arr, ctx, data_type=np.float32):
    """Create an NDArray from array-like on ``ctx``.  reference: ndarray.py:405-423."""
    if isinstance(arr, NDArray):
        arr = arr.tensor
    if not isinstance(arr, torch.Tensor):
        arr = np.asarray(arr, dtype=data_type)
    t = to_tensor(arr, ctx.torch_device)
    if not isinstance(arr, torch.Tensor) or t.data_ptr() == getattr(arr, 'data_ptr', lambda: 0)():
        t = t.clone() if t.device.type == 'cpu' else t
    return NDArray(t, ctx)


def empty(shape, ctx=cpu(0), dtype=torch.float32):
    """Uninitialised array.  Host arrays are pinned when a GPU is present (the
    reference's CPU allocator is cudaMallocHost: src/common/cpu_device_api.cc:22-34)."""
    dev = ctx.torch_device
    shape = tuple(int(s) for s in shape)
    if dev.type == 'cpu' and torch.cuda.is_available():
        return NDArray(torch.empty(shape, dtype=dtype, pin_memory=True), ctx)
    return NDArray(torch.empty(shape, dtype=dtype, device=dev), ctx)


def numpyasdlarrayhandle(data):
    return NDArray(to_tensor(data), cpu(0))


class ND_Sparse_Array(object):
    """CSR matrix (data, row pointer, col index).  reference: ndarray.py:460-505."""

    __slots__ = ('data', 'row', 'col', 'nrow', 'ncol', 'ctx')

    def __init__(self, data, row, col, nrow, ncol, ctx=None):
        self.data, self.row, self.col = data, row, col
        self.nrow, self.ncol = int(nrow), int(ncol)
        self.ctx = ctx or cpu(0)

    @property
    def shape(self):
        return (self.nrow, self.ncol)

    def to_torch(self) -> torch.Tensor:
        return torch.sparse_csr_tensor(self.row.tensor.long(), self.col.tensor.long(),
                                       self.data.tensor, size=self.shape)

    def to_dense(self) -> torch.Tensor:
        return self.to_torch().to_dense()


def sparse_array(values, indices, shape, ctx=cpu(0)):
    """COO triples -> CSR ND_Sparse_Array.  reference: ndarray.py:477-505."""
    import scipy.sparse
    assert len(shape) == len(indices) == 2
    mat = scipy.sparse.csr_matrix((np.asarray(values), indices), shape)
    dev = ctx.torch_device
    data = NDArray(torch.from_numpy(mat.data.astype(np.float32)).to(dev), ctx)
    row = NDArray(torch.from_numpy(mat.indptr.astype(np.int32)).to(dev), ctx)
    col = NDArray(torch.from_numpy(mat.indices.astype(np.int32)).to(dev), ctx)
    return ND_Sparse_Array(data, row, col, shape[0], shape[1], ctx)


class IndexedSlices(object):
    """Sparse row gradient: ``values[i]`` belongs to row ``indices[i]`` of a
    dense tensor of ``dense_shape``.  reference: ndarray.py:507-618.

    Deduplication runs on device (sort + segmented sum) - the reference does the
    unique on host with numpy (ndarray.py:535-552)."""

    __slots__ = ('indices', 'values', 'dense_shape', 'deduplicated', 'dedup_ind', 'dedup_val',
                 'dense_arr')

    def __init__(self, indices=None, values=None, dense_shape=None):
        self.indices = indices
        self.values = values
        self.dense_shape = tuple(dense_shape) if dense_shape is not None else None
        self.deduplicated = False
        self.dedup_ind = self.dedup_val = self.dense_arr = None

    def get_dense_shape(self):
        assert self.dense_shape is not None
        return self.dense_shape

    def get_sparse_shape(self):
        assert isinstance(self.values, (NDArray, torch.Tensor))
        return tuple(self.values.shape)

    def update(self, indices, values, dense_shape):
        self.indices, self.values = indices, values
        if self.dense_shape is not None:
            assert tuple(self.dense_shape) == tuple(dense_shape)
        else:
            self.dense_shape = tuple(dense_shape)
        self.deduplicated = False

    def _iv(self):
        ind = self.indices.tensor if isinstance(self.indices, NDArray) else self.indices
        val = self.values.tensor if isinstance(self.values, NDArray) else self.values
        width = self.dense_shape[-1]
        return ind.reshape(-1).long(), val.reshape(-1, width)

    def deduplicate(self, stream=None):
        ind, val = self._iv()
        uniq, inv = torch.unique(ind, return_inverse=True)
        acc = torch.zeros((uniq.numel(), val.shape[1]), dtype=val.dtype, device=val.device)
        acc.index_add_(0, inv, val)
        self.dedup_ind, self.dedup_val = uniq, acc
        self.deduplicated = True
        return uniq, acc

    cpu_deduplicate = deduplicate

    def free_deduplicate(self):
        self.dedup_ind = self.dedup_val = None
        self.deduplicated = False

    def to_dense(self, stream=None):
        ind, val = self._iv()
        out = torch.zeros(self.dense_shape, dtype=val.dtype, device=val.device)
        out.index_add_(0, ind, val)
        self.dense_arr = out
        return out

    def free_dense(self):
        self.dense_arr = None

    def merge(self, other: 'IndexedSlices'):
        ind, val = self._iv()
        oind, oval = other._iv()
        return IndexedSlices(torch.cat([ind, oind]), torch.cat([val, oval]), self.dense_shape)


#: reference ndarray.py exposes the ctypes ``DLArray`` struct; tensors are torch tensors here and the C ABI
#: takes raw pointers, so the name is kept as an alias of the array handle for isinstance checks
DLArray = NDArray
