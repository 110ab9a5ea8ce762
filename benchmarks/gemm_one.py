"""One GEMM shape, a few launches: the target of the ncu capture (profiles/).
usage: gemm_one.py M N K [tA tB]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hetu_61a7_b200.ops.linalg import gemm
argv = [int(v) for v in sys.argv[1:]]
m, n, k = argv[:3] if len(argv) >= 3 else (32768, 2048, 2048)
tA, tB = (bool(argv[3]), bool(argv[4])) if len(argv) >= 5 else (False, False)
a = torch.randn((k, m) if tA else (m, k), device='cuda'); b = torch.randn((n, k) if tB else (k, n), device='cuda')
for _ in range(3):
    c = gemm(a, b, tA, tB)
torch.cuda.synchronize()
ref = (a.t() if tA else a)[:256].double() @ (b.t() if tB else b).double()
print('relerr', float((c[:256].double() - ref).abs().max() / ref.abs().max()))
