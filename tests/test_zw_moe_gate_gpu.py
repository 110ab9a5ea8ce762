"""Fused MoE gate kernels (csrc/moe_gate.cu) on a B200 against the composed gate.  Late-collected file."""
import pytest

import hetu as ht
from test_graph_cpu import _fused_gate_vs_composed


@pytest.mark.gpu
def test_fused_moe_gate_kernels_on_b200():
    _fused_gate_vs_composed(ht.gpu(0), 2e-5)


@pytest.mark.gpu
def test_fused_moe_gate_small_and_wide_expert_counts():
    """Thread-per-token kernel (E <= 32, three register-array widths) and warp-per-token kernel (E > 32) against the
    PyTorch statement of the same packed tensor, forward and backward."""
    import torch
    from hetu_61a7_b200.ops import moe_gate as G
    from hetu_61a7_b200 import _lib as L
    import ctypes
    dev = torch.device('cuda', 0)
    for T, E, k in ((16384, 2, 2), (1000, 7, 3), (4100, 16, 2), (777, 32, 4), (900, 48, 2)):
        g = torch.Generator().manual_seed(T + E)
        x = torch.randn(T, E, generator=g).to(dev)
        ref = G._reference_forward(x, k)
        P = torch.empty_like(ref)
        nblk = (T + 255) // 256
        blk = torch.empty(k * nblk * E, dtype=torch.int32, device=dev)
        me = torch.empty(E, device=dev)
        L.call('moe_gate_fwd', L.p(x), L.p(P), L.p(blk), L.p(me), L.c_int(T), L.c_int(E), L.c_int(k), L.stream_ptr(dev))
        torch.testing.assert_close(P[:2 * k * T], ref[:2 * k * T], rtol=0, atol=0)          # indices and slots: exact
        torch.testing.assert_close(P[2 * k * T:], ref[2 * k * T:], rtol=2e-5, atol=2e-5)
        dg = [torch.randn(T, generator=g).to(dev) if i != 1 else None for i in range(k)]     # one missing gradient
        daux = torch.randn(1, generator=g).to(dev)
        want = G._reference_backward(x, ref, dg, daux, k)
        out = torch.empty_like(x)
        ptrs = (ctypes.c_void_p * k)(*[d.data_ptr() if d is not None else None for d in dg])
        L.call('moe_gate_bwd', L.p(x), L.p(P), ptrs, L.p(daux), L.p(out), L.c_int(T), L.c_int(E), L.c_int(k), L.stream_ptr(dev))
        torch.testing.assert_close(out, want, rtol=2e-4, atol=2e-6)
