"""Per-shape / per-layout error table for hetu_gemm (diagnostic, run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hetu_61a7_b200.ops.linalg import gemm

dev = torch.device('cuda')
shapes = [(128, 128, 128), (128, 128, 32), (128, 32, 32), (128, 32, 144), (256, 256, 64), (128, 256, 256),
          (8192, 32, 144), (8192, 144, 32), (512, 384, 256), (2048, 2048, 2048), (1000, 520, 264), (96, 40, 72),
          (4096, 1024, 512), (1160, 776, 200), (8192, 256, 4096), (640, 2048, 96),
          (2048, 2048, 16384), (1024, 768, 5000), (16384, 2, 2048), (2048, 2, 16384), (5000, 7, 1001), (4096, 8, 512)]
torch.manual_seed(0)
for (m, n, k) in shapes:
    for tA in (False, True):
        for tB in (False, True):
            a = torch.randn((k, m) if tA else (m, k), device=dev)
            b = torch.randn((n, k) if tB else (k, n), device=dev)
            ref = (a.t() if tA else a).double() @ (b.t() if tB else b).double()
            out = gemm(a, b, tA, tB)
            torch.cuda.synchronize()
            err = (out.double() - ref).abs().max().item() / (ref.abs().max().item() + 1e-9)
            nz = (out != 0).float().mean().item()
            # per-quadrant error map to spot layout problems
            print('M%5d N%5d K%5d tA=%d tB=%d  relerr %.3e  nonzero %.2f  %s' % (
                m, n, k, tA, tB, err, nz, 'OK' if err < 2e-3 else 'FAIL'))

# batched + epilogue on the 2-SM path
a = torch.randn(3, 1024, 256, device=dev); b = torch.randn(3, 256, 512, device=dev); bias = torch.randn(512, device=dev)
out = gemm(a, b, False, False, bias=bias, act=1)
ref = torch.relu(a.double() @ b.double() + bias.double())
err = (out.double() - ref).abs().max().item() / ref.abs().max().item()
print('batched 3x1024x512x256 bias+relu relerr %.3e %s' % (err, 'OK' if err < 2e-3 else 'FAIL'))
