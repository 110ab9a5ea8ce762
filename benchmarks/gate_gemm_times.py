import sys, torch
sys.path.insert(0, '/root/repo')
from hetu_61a7_b200.ops.linalg import gemm
dev = torch.device('cuda', 0)
for E in (2, 4, 8, 16):
    x = torch.randn(16384, 2048, device=dev); w = torch.randn(2048, E, device=dev); g = torch.randn(16384, E, device=dev)
    for name, fn in (('x@W', lambda: gemm(x, w)), ('x^T@g', lambda: gemm(x, g, True, False)), ('g@W^T', lambda: gemm(g, w, False, True))):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(20): fn()
        e.record(); torch.cuda.synchronize()
        print('E=%2d %-6s %.1f us' % (E, name, s.elapsed_time(e) / 20 * 1e3))
