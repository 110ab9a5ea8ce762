"""hetu_lamb_step_noalloc: the Lamb step for hosts that must not allocate GPU memory.

Who needs it: a caller that reaches the library through its C ABI and owns nothing but the
tensors it passes in - a plugin running inside another framework's callback (re-entering that
framework's allocator is off limits), a latency-bound service that bans cudaMalloc after
start-up because it synchronises the whole device and may fail under memory pressure, a thin
FFI client.  The regular entry point (hetu_multi_tensor_opt, kind = Lamb) makes such a caller
provide two floats of zeroed device scratch per tensor; this one keeps that scratch in a static
device array inside the library (768 bytes per GPU, at most 96 tensors per call).

The test drives both entry points through ctypes on identical copies of the same state, twice in
a row (the second call sees whatever the first one left in the shared array, so it also checks
that the entry point clears what it uses), and checks the over-limit call is refused.  It runs
last (file name sorts after every other test file) so a failure here cannot hide other results.
"""
import ctypes
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]

LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                        'hetu_61a7_b200', 'lib', 'libhetu_b200.so')
OPT_LAMB = 6
MAX_TENSORS = 96


def _arr(ptrs, ctype=ctypes.c_void_p):
    return (ctype * len(ptrs))(*ptrs)


def _state(shapes, device, seed):
    g = torch.Generator(device='cpu').manual_seed(seed)
    params = [torch.randn(s, generator=g).to(device) for s in shapes]
    m = [torch.zeros_like(p) for p in params]
    v = [torch.zeros_like(p) for p in params]
    return params, m, v


REGULAR_MAX = 48          # tensors per launch of the regular entry point


def _regular(lib, n, p, g, m, v, sizes, lr, step, hyper, stream, dev):
    """The regular entry point takes at most 48 tensors per call and a zeroed scratch from the caller."""
    f = ctypes.c_float
    for base in range(0, n, REGULAR_MAX):
        cnt = min(REGULAR_MAX, n - base)
        scratch = torch.zeros(2 * cnt, device=dev)    # the allocation the regular path needs
        sl = slice(base, base + cnt)
        rc = lib.hetu_multi_tensor_opt(
            ctypes.c_int(OPT_LAMB), ctypes.c_int(cnt), _arr([t.data_ptr() for t in p[sl]]),
            _arr([t.data_ptr() for t in g[sl]]), _arr([t.data_ptr() for t in m[sl]]),
            _arr([t.data_ptr() for t in v[sl]]), _arr(list(sizes[sl]), ctypes.c_longlong),
            ctypes.c_void_p(lr.data_ptr()), ctypes.c_void_p(step.data_ptr()), *[f(h) for h in hyper], f(0.0),
            ctypes.c_void_p(scratch.data_ptr()), stream)
        assert rc == 0, lib.hetu_last_error()


def test_lamb_noalloc_two_chunks_then_a_smaller_call():
    """96 tensors = both 48-tensor chunks of the static array, checked against two regular calls; then a
    17-tensor call right after it, which reuses slots the big call left non-zero."""
    assert os.path.exists(LIB_PATH), 'build the library first: python __graft_entry__.py'
    lib = ctypes.CDLL(LIB_PATH)
    lib.hetu_last_error.restype = ctypes.c_char_p
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    base_shapes = [(13, 256), (256, 256), (3584, 1), (7,), (1000, 33), (64, 64, 3), (5,), (300, 17)]
    hyper = (0.9, 0.999, 1e-7, 0.1)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    f = ctypes.c_float
    lr = torch.full((1,), 0.01, device=dev)
    for shapes in (base_shapes * 12, base_shapes * 2 + [(11, 3)]):
        n = len(shapes)
        assert n == MAX_TENSORS or n < REGULAR_MAX
        pa, ma, va = _state(shapes, dev, 3)
        pb, mb, vb = [t.clone() for t in pa], [t.clone() for t in ma], [t.clone() for t in va]
        sizes = [p.numel() for p in pa]
        step = torch.ones(1, dtype=torch.int64, device=dev)
        g = torch.Generator(device='cpu').manual_seed(7 + n)
        grads = [torch.randn(s, generator=g).to(dev) for s in shapes]
        _regular(lib, n, pa, grads, ma, va, sizes, lr, step, hyper, stream, dev)
        rc = lib.hetu_lamb_step_noalloc(
            ctypes.c_int(n), _arr([t.data_ptr() for t in pb]), _arr([t.data_ptr() for t in grads]),
            _arr([t.data_ptr() for t in mb]), _arr([t.data_ptr() for t in vb]), _arr(sizes, ctypes.c_longlong),
            ctypes.c_void_p(lr.data_ptr()), ctypes.c_void_p(step.data_ptr()), *[f(h) for h in hyper], f(0.0), stream)
        assert rc == 0, lib.hetu_last_error()
        torch.cuda.synchronize(dev)
        for a, b in zip(pa + ma + va, pb + mb + vb):
            torch.testing.assert_close(b, a, rtol=2e-4, atol=1e-5)


def test_lamb_noalloc_matches_regular_entry_point():
    assert os.path.exists(LIB_PATH), 'build the library first: python __graft_entry__.py'
    lib = ctypes.CDLL(LIB_PATH)
    lib.hetu_last_error.restype = ctypes.c_char_p
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(dev)
    shapes = [(13, 256), (256, 256), (3584, 1), (7,), (1000, 33), (64, 64, 3)] * 6 + [(5,), (2048, 17)]
    assert len(shapes) <= REGULAR_MAX
    n = len(shapes)
    pa, ma, va = _state(shapes, dev, 1)            # driven by the regular entry point
    pb, mb, vb = [t.clone() for t in pa], [t.clone() for t in ma], [t.clone() for t in va]
    lr = torch.full((1,), 0.01, device=dev)
    step = torch.zeros(1, dtype=torch.int64, device=dev)
    sizes = _arr([p.numel() for p in pa], ctypes.c_longlong)
    beta1, beta2, eps, wd = 0.9, 0.999, 1e-7, 0.1
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    f = ctypes.c_float
    for it in range(2):                             # two calls in a row
        g = torch.Generator(device='cpu').manual_seed(100 + it)
        grads = [torch.randn(s, generator=g).to(dev) for s in shapes]
        step += 1
        scratch = torch.zeros(2 * n, device=dev)    # the allocation the regular path needs
        rc = lib.hetu_multi_tensor_opt(
            ctypes.c_int(OPT_LAMB), ctypes.c_int(n), _arr([t.data_ptr() for t in pa]),
            _arr([t.data_ptr() for t in grads]), _arr([t.data_ptr() for t in ma]),
            _arr([t.data_ptr() for t in va]), sizes, ctypes.c_void_p(lr.data_ptr()),
            ctypes.c_void_p(step.data_ptr()), f(beta1), f(beta2), f(eps), f(wd), f(0.0),
            ctypes.c_void_p(scratch.data_ptr()), stream)
        assert rc == 0, lib.hetu_last_error()
        rc = lib.hetu_lamb_step_noalloc(
            ctypes.c_int(n), _arr([t.data_ptr() for t in pb]), _arr([t.data_ptr() for t in grads]),
            _arr([t.data_ptr() for t in mb]), _arr([t.data_ptr() for t in vb]), sizes,
            ctypes.c_void_p(lr.data_ptr()), ctypes.c_void_p(step.data_ptr()), f(beta1), f(beta2), f(eps),
            f(wd), f(0.0), stream)
        assert rc == 0, lib.hetu_last_error()
        torch.cuda.synchronize(dev)
        for a, b in zip(pa + ma + va, pb + mb + vb):
            torch.testing.assert_close(b, a, rtol=2e-4, atol=1e-5)
    # sanity: the step really moved the parameters
    fresh, _, _ = _state(shapes, dev, 1)
    assert any((a - b).abs().max().item() > 1e-4 for a, b in zip(fresh, pb))
    # more tensors than the static scratch holds -> refused with a message, nothing launched
    too_many = MAX_TENSORS + 1
    dummy = _arr([pb[0].data_ptr()] * too_many)
    rc = lib.hetu_lamb_step_noalloc(
        ctypes.c_int(too_many), dummy, dummy, dummy, dummy, _arr([1] * too_many, ctypes.c_longlong),
        ctypes.c_void_p(lr.data_ptr()), ctypes.c_void_p(step.data_ptr()), f(beta1), f(beta2), f(eps), f(wd),
        f(0.0), stream)
    assert rc != 0 and b'exceed' in lib.hetu_last_error()
