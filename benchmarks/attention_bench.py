"""Attention forward: the flash kernel (csrc/attention_sm100.cu) vs the three-kernel path (batched tcgen05 GEMM, row
softmax, batched GEMM) and torch's scaled_dot_product_attention (library code, fp32 inputs) - CUDA events, L2 flushed
between iterations, clocks sampled.

    python benchmarks/attention_bench.py [out.txt]
"""
import os
import subprocess
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hetu_61a7_b200 import _lib as L                                   # noqa: E402
from hetu_61a7_b200.ops.linalg import gemm                              # noqa: E402


def clocks():
    try:
        return subprocess.run(['nvidia-smi', '--query-gpu=clocks.sm,clocks.max.sm,clocks_throttle_reasons.active',
                               '--format=csv,noheader', '-i', '0'], capture_output=True, text=True, timeout=10).stdout.strip()
    except Exception as e:                                              # noqa: BLE001
        return 'n/a (%s)' % e


def timed(fn, flush, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/attention_bench.txt'
    dev = torch.device('cuda', 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    lines = ['attention forward, fp32 storage / tf32 MMA, median of 20, L2 flushed (64 keys per block, two CTAs per SM)',
             'clocks before: ' + clocks()]
    for (B, H, S, D) in [(32, 12, 128, 64), (32, 16, 128, 64), (8, 16, 512, 64), (4, 16, 1024, 64), (2, 16, 2048, 64), (1, 16, 4096, 64)]:
        q, k, v = (torch.randn(B, H, S, D, device=dev) * 0.5 for _ in range(3))
        out = torch.empty_like(q)
        lse = torch.empty(B, H, S, device=dev)
        scale = 1.0 / np.sqrt(D)

        def flash():
            L.call('flash_attn_fwd', L.p(q), L.p(k), L.p(v), L.p(None), L.p(out), L.p(lse), L.c_int(B), L.c_int(H), L.c_int(S),
                   L.c_int(S), L.c_int(D), L.c_f32(scale), L.stream_ptr(dev))

        q3, k3, v3 = (t.view(B * H, S, D) for t in (q, k, v))

        def unfused():
            p = gemm(q3, k3, False, True)
            L.call('attn_softmax', L.p(p), L.p(None), L.p(lse), L.c_i64(B * H * S), L.c_int(S), L.c_i64(H * S), L.c_f32(scale),
                   L.stream_ptr(dev))
            return gemm(p, v3)

        def sdpa():
            return torch.nn.functional.scaled_dot_product_attention(q, k, v)

        flops = 4.0 * B * H * S * S * D
        tf, tu, ts = timed(flash, flush), timed(unfused, flush), timed(sdpa, flush)
        ref = unfused().view(B, H, S, D)
        flash()
        err = (out - ref).abs().max().item() / ref.abs().max().item()
        lines.append('B %3d H %2d S %4d D %d: flash %7.3f ms (%6.1f TFLOP/s)  gemm+softmax+gemm %7.3f ms (%6.1f)  torch sdpa fp32 %7.3f ms (%6.1f)  '
                     'flash vs 3-kernel max rel diff %.1e' % (B, H, S, D, tf, flops / tf / 1e9, tu, flops / tu / 1e9, ts, flops / ts / 1e9, err))
    lines.append('clocks after: ' + clocks())
    os.makedirs(os.path.dirname(out_path) or '.', exist_ok=True)
    with open(out_path, 'w') as f:
        f.write('\n'.join(lines) + '\n')
    print('\n'.join(lines))


if __name__ == '__main__':
    main()
