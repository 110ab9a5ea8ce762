"""GPU-side fixed vs per-k-block cost of hetu_gemm: 20 calls captured in a CUDA graph (no host
overhead in the timed region), one or a few output tiles, K swept."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hetu_61a7_b200.ops.linalg import gemm, set_matmul_precision
dev = torch.device('cuda')
REP = 20
for mode in ('tf32', 'fp32'):
    set_matmul_precision(mode)
    for (m, n) in ((128, 256), (128, 128), (1024, 1024)):
        for k in (32, 256, 1024, 4096):
            a = torch.randn(m, k, device=dev)
            b = torch.randn(k, n, device=dev)
            for _ in range(3):
                gemm(a, b)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(REP):
                    c = gemm(a, b)
            for _ in range(3):
                g.replay()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(10):
                g.replay()
            e.record(); e.synchronize()
            print('%s M%5d N%5d K%6d  %.2f us/gemm (graph)' % (mode, m, n, k, s.elapsed_time(e) / 10 / REP * 1000))
