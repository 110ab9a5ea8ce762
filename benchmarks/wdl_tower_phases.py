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
op.m, op._bufs, op._dp, op.last, op.fused_launches = m, None, None, None, 0


class _Opt(object):          # just enough of an optimizer for the in-kernel SGD path
    lr_t = torch.full((1,), 0.01, device=dev)
    step_t = torch.zeros(1, dtype=torch.int64, device=dev)
    params = []


class _OptNode(object):
    optimizer = _Opt()
    fused_dp = None


if os.environ.get('INKERNEL_SGD', '1') == '1':
    op.opt_node = _OptNode()
g = torch.Generator().manual_seed(0)
x = torch.rand(B, nd, generator=g).to(dev)
labels = (torch.rand(B, 1, generator=g) < 0.3).float().to(dev)
E = (torch.randn(vocab, width, generator=g) * 0.05).to(dev)
Ws = [(torch.randn(nd, hid, generator=g) * 0.3).to(dev), (torch.randn(hid, hid, generator=g) * 0.08).to(dev),
      (torch.randn(hid, hid, generator=g) * 0.08).to(dev), (torch.randn(fields * width + hid, 1, generator=g) * 0.05).to(dev)]
names = ['setup+P0 (stage x, emb logits)', 'P1 h1 (+barrier)', 'P2 h2 GEMM (+barrier)', 'P3 y3 GEMM (+barrier)',
         'head (dsmem, sigmoid, bce)', 'P5 vectors + dh2 (+barrier)', 'P6a dh1 GEMM | dW4', 'P6b dW2 epilogue | dW1',
         'tail: all-reduce + SGD']
acc = [0.0] * 9
rel = [0.0] * 5     # gather start / end, sparse start / end, tower end relative to the tower's first stamp
reps = 20
for it in range(reps + 3):
    ids = torch.randint(0, vocab, (B, fields), generator=g).to(dev)
    op.compute([x, ids, labels, E, *Ws])
    if getattr(op, 'opt_node', None) is not None:
        op.apply_update(op.opt_node)
    out = (ctypes.c_ulonglong * 16)()
    L.call('wdl_tower_stamps', out)
    if it >= 3:
        for j, k in enumerate((10, 11, 12, 13)):
            rel[j] += (out[k] - out[0]) / 1e3 if out[k] not in (0, 2 ** 64 - 1) else 0.0
        rel[4] += (max(out[8], out[9]) - out[0]) / 1e3
        for i in range(9):
            acc[i] += max(0, out[i + 1] - out[i]) / 1e3 if out[i + 1] else 0.0
print('phase breakdown of wdl_tower_kernel, mean of %d launches (us):' % reps)
for n, a in zip(names, acc):
    print('  %-34s %7.2f' % (n, a / reps))
print('  %-34s %7.2f' % ('total', sum(acc) / reps))
print('relative to the tower kernel start (us): gather %.2f .. %.2f, sparse SGD %.2f .. %.2f, tower end %.2f'
      % tuple(r / reps for r in rel))


def ev_time(fn, n=200):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


ids = torch.randint(0, vocab, (B, fields), generator=g).to(dev)
inputs = [x, ids, labels, E, *Ws]
print('back-to-back launches, CUDA events (us per call):')
def full():
    op.compute(inputs)
    if getattr(op, 'opt_node', None) is not None:
        op.apply_update(op.opt_node)


print('  gather + tower + sparse (stream) %7.2f' % ev_time(full))
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    full()
print('  gather + tower + sparse (graph)  %7.2f' % ev_time(graph.replay))

# the same timeline inside graph replays (no host work between the launches)
rel = [0.0] * 6
prev_end = None
for it in range(12):
    graph.replay()
    out = (ctypes.c_ulonglong * 16)()
    L.call('wdl_tower_stamps', out)
    if it >= 2:
        rel[5] += (out[14] - out[0]) / 1e3
        if it == 11:
            print('  last replay: stamp7 %.2f, dW2 MMAs done %.2f, vector warps done %.2f, stamp8 %.2f' % tuple((out[k] - out[0]) / 1e3 for k in (7, 9, 15, 8)))
        for j, k in enumerate((10, 11, 12, 13)):
            rel[j] += (out[k] - out[0]) / 1e3
        rel[4] += (max(out[8], out[9]) - out[0]) / 1e3
print('graph replay, relative to the tower kernel start (us): gather %.2f .. %.2f, sparse SGD %.2f .. %.2f, tower end %.2f, '
      'last CTA after TMEM release %.2f' % tuple(r / 10 for r in rel[:6]))
