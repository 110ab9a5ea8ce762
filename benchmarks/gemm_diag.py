"""Diagnostic: time one shape under HETU_GEMM_DBG modes (set in the environment by the caller)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hetu_61a7_b200.ops.linalg import gemm
for (m, n, k, tA, tB) in [(32768, 2048, 2048, 0, 1), (32768, 2048, 2048, 0, 0), (32768, 2048, 2048, 1, 1), (2048, 2048, 16384, 1, 0), (2048, 2048, 16384, 0, 1)]:
    a = torch.randn((k, m) if tA else (m, k), device='cuda'); b = torch.randn((n, k) if tB else (k, n), device='cuda')
    for _ in range(3):
        gemm(a, b, bool(tA), bool(tB))
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        gemm(a, b, bool(tA), bool(tB))
    e.record(); e.synchronize()
    ms = s.elapsed_time(e) / 5
    print('dbg=%s M%d N%d K%d tA=%d tB=%d %.3f ms %.0f TFLOP/s' % (os.environ.get('HETU_GEMM_DBG', '0'), m, n, k, tA, tB, ms, 2.0 * m * n * k / ms / 1e9), flush=True)
