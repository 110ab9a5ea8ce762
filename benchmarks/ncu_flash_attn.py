"""Smallest program that launches the flash attention forward kernel (for `ncu --set full -k regex:flash_attn`)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hetu_61a7_b200 import _lib as L  # noqa: E402

dev = torch.device('cuda', 0)
B, H, S, D = 2, 16, 2048, 64
q, k, v = (torch.randn(B, H, S, D, device=dev) * 0.5 for _ in range(3))
out, lse = torch.empty_like(q), torch.empty(B, H, S, device=dev)
for _ in range(4):
    L.call('flash_attn_fwd', L.p(q), L.p(k), L.p(v), L.p(None), L.p(out), L.p(lse), L.c_int(B), L.c_int(H), L.c_int(S), L.c_int(S),
           L.c_int(D), L.c_f32(1.0 / np.sqrt(D)), L.stream_ptr(dev))
torch.cuda.synchronize()
print('out', float(out.abs().mean()))
