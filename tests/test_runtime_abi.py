"""Native runtime C ABI (csrc/runtime.cu): DLArray alloc / copy / free, streams, events and the
workspace chunk cache, driven through ctypes without torch in the data path (pattern: reference
tests that go through ``ht.array(...).asnumpy()`` on top of DLArrayAlloc / DLArrayCopyFromTo)."""
import ctypes
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class DLContext(ctypes.Structure):
    _fields_ = [('device_id', ctypes.c_int), ('device_type', ctypes.c_int)]


class DLArray(ctypes.Structure):
    _fields_ = [('data', ctypes.c_void_p), ('ctx', DLContext), ('ndim', ctypes.c_int),
                ('shape', ctypes.POINTER(ctypes.c_longlong)), ('stride', ctypes.POINTER(ctypes.c_longlong))]


class Handle(ctypes.Structure):
    _fields_ = [('device_id', ctypes.c_int), ('handle', ctypes.c_void_p)]


@pytest.mark.gpu
def test_dlarray_stream_event_chunk_abi():
    lib = ctypes.CDLL(os.path.join(ROOT, 'hetu_61a7_b200', 'lib', 'libhetu_b200.so'))
    lib.hetu_last_error.restype = ctypes.c_char_p
    lib.hetu_find_chunk.restype = ctypes.c_void_p
    lib.hetu_find_chunk.argtypes = [ctypes.c_size_t, ctypes.c_int]
    lib.hetu_del_chunk.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.hetu_chunk_cache_bytes.restype = ctypes.c_size_t

    def ok(rc):
        assert rc == 0, lib.hetu_last_error().decode()
    shape = (ctypes.c_longlong * 2)(257, 33)
    host, dev, dev2, back = (ctypes.POINTER(DLArray)() for _ in range(4))
    ok(lib.DLArrayAlloc(shape, 2, DLContext(0, 1), ctypes.byref(host)))
    ok(lib.DLArrayAlloc(shape, 2, DLContext(0, 2), ctypes.byref(dev)))
    ok(lib.DLArrayAlloc(shape, 2, DLContext(0, 2), ctypes.byref(dev2)))
    ok(lib.DLArrayAlloc(shape, 2, DLContext(0, 1), ctypes.byref(back)))
    assert host.contents.ndim == 2 and host.contents.stride[0] == 33 and host.contents.stride[1] == 1
    n = 257 * 33
    x = np.random.RandomState(0).randn(n).astype(np.float32)
    ctypes.memmove(host.contents.data, x.ctypes.data, n * 4)
    stream, e0, e1 = ctypes.POINTER(Handle)(), ctypes.POINTER(Handle)(), ctypes.POINTER(Handle)()
    ok(lib.DLStreamCreate(0, ctypes.byref(stream)))
    ok(lib.DLEventCreate(0, ctypes.byref(e0)))
    ok(lib.DLEventCreate(0, ctypes.byref(e1)))
    ok(lib.DLEventRecord(stream, e0))
    ok(lib.DLArrayCopyFromTo(host, dev, stream))                       # H2D on the stream
    # relu through the kernel ABI on the same stream: hetu_unary(op=12 U_RELU, x, y, n, a, b, stream)
    ok(lib.hetu_unary(12, ctypes.c_void_p(dev.contents.data), ctypes.c_void_p(dev2.contents.data), ctypes.c_size_t(n),
                      ctypes.c_float(0), ctypes.c_float(0), ctypes.c_void_p(stream.contents.handle)))
    ok(lib.DLArrayCopyFromTo(dev2, back, stream))                      # D2H
    ok(lib.DLEventRecord(stream, e1))
    ok(lib.DLEventSync(e1))
    ms = ctypes.c_float(-1)
    ok(lib.DLEventElapsedTime(e0, e1, ctypes.byref(ms)))
    assert 0 <= ms.value < 1000
    y = np.empty(n, dtype=np.float32)
    ctypes.memmove(y.ctypes.data, back.contents.data, n * 4)
    np.testing.assert_array_equal(y, np.maximum(x, 0))
    ok(lib.DLArrayCopyFromTo(dev, dev2, None))                         # blocking D2D
    ok(lib.DLArrayCopyFromTo(dev2, back, None))
    ctypes.memmove(y.ctypes.data, back.contents.data, n * 4)
    np.testing.assert_array_equal(y, x)
    bad = ctypes.POINTER(DLArray)()
    ok(lib.DLArrayAlloc((ctypes.c_longlong * 1)(5), 1, DLContext(0, 2), ctypes.byref(bad)))
    assert lib.DLArrayCopyFromTo(host, bad, None) != 0 and b'elements' in lib.hetu_last_error()
    # chunk cache: a released chunk is handed out again for a request it fits
    p1 = lib.hetu_find_chunk(1 << 20, 0)
    assert p1
    ok(lib.hetu_del_chunk(p1, 0))
    assert lib.hetu_chunk_cache_bytes(0) == 1 << 20
    p2 = lib.hetu_find_chunk((1 << 20) - 4096, 0)
    assert p2 == p1 and lib.hetu_chunk_cache_bytes(0) == 0
    p3 = lib.hetu_find_chunk(1 << 10, 0)                                # far smaller: a chunk of its own
    assert p3 and p3 != p1
    ok(lib.hetu_del_chunk(p2, 0))
    ok(lib.hetu_del_chunk(p3, 0))
    assert lib.hetu_del_chunk(ctypes.c_void_p(12345), 0) != 0
    ok(lib.hetu_chunk_cache_release(0))
    for a in (host, dev, dev2, back, bad):
        ok(lib.DLArrayFree(a))
    ok(lib.DLStreamSync(stream))
    ok(lib.DLEventDestroy(e0))
    ok(lib.DLEventDestroy(e1))
    ok(lib.DLStreamDestroy(stream))
