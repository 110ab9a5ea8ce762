"""Phase breakdown of the fused Wide&Deep tower kernel (csrc/wdl_tower.cu) from in-kernel %globaltimer stamps.
Usage (on a GPU box): HETU_WDL_STAMPS=1 python benchmarks/wdl_tower_phases.py"""
import ctypes
import os
import sys

os.environ.setdefault('HETU_WDL_STAMPS', '1')
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from hetu_61a7_b200 import _lib as L  # noqa: E402
from hetu_61a7_b200.graph import fusion_wdl as fw  # noqa: E402

dev = torch.device('cuda', 0)
B, nd, hid, fields, width, vocab = 128, 13, 256, 26, 128, 2000000
m = dict(nd=nd, width=width, emb_cols=fields * width, fields=fields)
cls = fw._make_op_class()
op = cls.__new__(cls)
op.m, op._bufs, op.last, op.fused_launches = m, None, None, 0
g = torch.Generator().manual_seed(0)
x = torch.rand(B, nd, generator=g).to(dev)
labels = (torch.rand(B, 1, generator=g) < 0.3).float().to(dev)
E = (torch.randn(vocab, width, generator=g) * 0.05).to(dev)
Ws = [(torch.randn(nd, hid, generator=g) * 0.3).to(dev), (torch.randn(hid, hid, generator=g) * 0.08).to(dev),
      (torch.randn(hid, hid, generator=g) * 0.08).to(dev), (torch.randn(fields * width + hid, 1, generator=g) * 0.05).to(dev)]
names = ['setup+P0 (stage x, emb logits)', 'P1 h1 (+barrier)', 'P2 h2 GEMM (+barrier)', 'P3 y3 GEMM (+barrier)',
         'head (dsmem, sigmoid, bce)', 'P5 vectors + dh2 (+barrier)', 'P6a dh1 GEMM | dW4', 'P6b dW2 epilogue | dW1']
acc = [0.0] * 8
reps = 20
for it in range(reps + 3):
    ids = torch.randint(0, vocab, (B, fields), generator=g).to(dev)
    op.compute([x, ids, labels, E, *Ws])
    out = (ctypes.c_ulonglong * 9)()
    L.call('wdl_tower_stamps', out)
    if it >= 3:
        for i in range(8):
            acc[i] += (out[i + 1] - out[i]) / 1e3
print('phase breakdown of wdl_tower_kernel, mean of %d launches (us):' % reps)
for n, a in zip(names, acc):
    print('  %-34s %7.2f' % (n, a / reps))
print('  %-34s %7.2f' % ('total', sum(acc) / reps))
