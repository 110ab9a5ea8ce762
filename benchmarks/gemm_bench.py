"""Achieved TFLOP/s of hetu_gemm at the MoE / BERT shapes (CUDA events, L2 flushed between
iterations by cycling through operands larger than L2 where the shape allows)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hetu_61a7_b200.ops.linalg import gemm

dev = torch.device('cuda')
peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(__file__), '..', 'MEASURED_PEAKS.json')))
except Exception:
    pass
shapes = [(32768, 2048, 2048, False, False), (32768, 2048, 2048, False, True), (2048, 2048, 32768, True, False),
          (16384, 2048, 2048, False, False), (4096, 4096, 1024, False, False), (4096, 1024, 4096, False, False),
          (8192, 8192, 8192, False, False), (16384, 16, 2048, False, False),
          (2048, 2048, 16384, True, False), (16384, 2, 2048, False, False), (2048, 2, 16384, True, False)]
flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
for (m, n, k, tA, tB) in shapes:
    a = torch.randn((k, m) if tA else (m, k), device=dev)
    b = torch.randn((n, k) if tB else (k, n), device=dev)
    for _ in range(3):
        c = gemm(a, b, tA, tB)
    times = []
    for _ in range(8):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        c = gemm(a, b, tA, tB)
        e.record()
        e.synchronize()
        times.append(s.elapsed_time(e))
    ms = sorted(times)[len(times) // 2]
    tf = 2.0 * m * n * k / ms / 1e9
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    torch.backends.cuda.matmul.allow_tf32 = True
    aa, bb = (a.t() if tA else a), (b.t() if tB else b)
    for _ in range(3):
        torch.matmul(aa, bb)
    flush.zero_()
    t0.record(); torch.matmul(aa, bb); t1.record(); t1.synchronize()
    ref_ms = t0.elapsed_time(t1)
    print('M%6d N%6d K%6d tA=%d tB=%d  %.3f ms  %.1f TFLOP/s (tf32)   cuBLAS tf32 %.3f ms %.1f TFLOP/s' % (
        m, n, k, tA, tB, ms, tf, ref_ms, 2.0 * m * n * k / ref_ms / 1e9))
