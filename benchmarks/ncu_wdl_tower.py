"""Smallest program that launches the fused Wide&Deep kernels a few times (for `ncu --set full -k regex:wdl_tower`)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from hetu_61a7_b200.graph import fusion_wdl as fw  # noqa: E402

dev = torch.device('cuda', 0)
B, nd, hid, fields, width, vocab = 128, 13, 256, 26, 128, 1000000
m = dict(nd=nd, width=width, emb_cols=fields * width, fields=fields)
cls = fw._make_op_class()
op = cls.__new__(cls)
op.m, op._bufs, op._dp, op.last, op.fused_launches = m, None, None, None, 0
g = torch.Generator().manual_seed(0)
x = torch.rand(B, nd, generator=g).to(dev)
labels = (torch.rand(B, 1, generator=g) < 0.3).float().to(dev)
E = (torch.randn(vocab, width, generator=g) * 0.05).to(dev)
Ws = [(torch.randn(nd, hid, generator=g) * 0.3).to(dev), (torch.randn(hid, hid, generator=g) * 0.08).to(dev),
      (torch.randn(hid, hid, generator=g) * 0.08).to(dev), (torch.randn(fields * width + hid, 1, generator=g) * 0.05).to(dev)]
for it in range(6):
    ids = torch.randint(0, vocab, (B, fields), generator=g).to(dev)
    loss, pred = op.compute([x, ids, labels, E, *Ws])
torch.cuda.synchronize()
print('loss', float(loss))
