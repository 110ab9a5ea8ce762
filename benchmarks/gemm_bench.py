"""Achieved TFLOP/s of the GEMM kernels at the MoE / BERT shapes: tf32 (hetu_gemm) and bf16 (hetu_gemm_f16) against
cuBLAS at the same precision.  CUDA events, L2 flushed between iterations, median of 8; every row carries the SM clock
and throttle reasons sampled by nvidia-smi while that row ran.
    python benchmarks/gemm_bench.py [out.txt]"""
import json
import os
import subprocess
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from hetu_61a7_b200.ops import linalg  # noqa: E402
from hetu_61a7_b200.ops.linalg import gemm  # noqa: E402

dev = torch.device('cuda')
Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
    'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'


class Clocks(object):
    def __init__(self):
        self.rows = []
        self.p = subprocess.Popen(['nvidia-smi', '--query-gpu=' + Q, '--format=csv,noheader,nounits', '-lms', '20', '-i', '0'],
                                  stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        threading.Thread(target=self._rd, daemon=True).start()

    def _rd(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(',')]))

    def window(self, t0, t1):
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for t, r in self.rows[-1:]]
        sm = sorted(float(r[0]) for r in rows)
        names = ['hw_slowdown', 'hw_thermal', 'sw_thermal', 'sw_power_cap']
        reasons = sorted({n for r in rows for i, n in enumerate(names) if r[2 + i].lower().startswith('active')})
        return 'sm %d/%s MHz reasons %s (%d samples)' % (sm[len(sm) // 2], rows[0][1], reasons or 'none', len(rows))


def timed(fn, flush, reps=8):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts) // 2]


def main():
    out = open(sys.argv[1], 'w') if len(sys.argv) > 1 else None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'MEASURED_PEAKS.json')))
    except Exception:
        pass

    def emit(s):
        print(s, flush=True)
        if out:
            out.write(s + '\n')
    emit('# GEMM, B200, CUDA events, L2 flushed between iterations, median of 8.  MEASURED_PEAKS: %s' % json.dumps(peaks))
    shapes = [(32768, 2048, 2048, False, False), (32768, 2048, 2048, False, True), (2048, 2048, 32768, True, False),
              (16384, 2048, 2048, False, False), (4096, 4096, 1024, False, False), (4096, 1024, 4096, False, False),
              (8192, 8192, 8192, False, False), (4096, 4096, 4096, False, True)]
    if os.environ.get('GEMM_BENCH_SHAPES') == 'bert':       # the products of one BERT-large / BERT-base layer at 4096 tokens
        shapes = []
        for h, f in ((1024, 4096), (768, 3072)):
            shapes += [(4096, h, h, False, False), (4096, f, h, False, False), (4096, h, f, False, False),
                       (4096, h, h, False, True), (4096, h, f, False, True), (4096, f, h, False, True),
                       (h, h, 4096, True, False), (h, f, 4096, True, False), (f, h, 4096, True, False)]
    flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
    clk = Clocks()
    time.sleep(0.3)
    torch.backends.cuda.matmul.allow_tf32 = True
    for (m, n, k, tA, tB) in shapes:
        a = torch.randn((k, m) if tA else (m, k), device=dev)
        b = torch.randn((n, k) if tB else (k, n), device=dev)
        aa, bb = (a.t() if tA else a), (b.t() if tB else b)
        a16, b16 = aa.bfloat16(), bb.bfloat16()
        flops = 2.0 * m * n * k
        t0 = time.time()
        linalg.set_matmul_precision('tf32')
        ms_tf32 = timed(lambda: gemm(a, b, tA, tB), flush)
        ms_cublas_tf32 = timed(lambda: torch.matmul(aa, bb), flush)
        if os.environ.get('GEMM_BENCH_SHAPES') == 'bert':
            t1 = time.time()
            emit('M%6d N%6d K%6d tA=%d tB=%d | tf32 %7.1f TFLOP/s %7.1f us (cuBLAS %7.1f, x%.2f) | %s' % (
                m, n, k, tA, tB, flops / ms_tf32 / 1e9, ms_tf32 * 1e3, flops / ms_cublas_tf32 / 1e9, ms_cublas_tf32 / ms_tf32,
                clk.window(t0, t1)))
            continue
        linalg.set_matmul_precision('bf16')
        ms_bf16 = timed(lambda: gemm(a, b, tA, tB), flush)            # includes the fp32 -> bf16 operand conversion
        ms_cublas_bf16 = timed(lambda: torch.matmul(a16, b16), flush)  # operands already bf16
        linalg.set_matmul_precision('tf32')
        t1 = time.time()
        emit('M%6d N%6d K%6d tA=%d tB=%d | tf32 %7.1f TFLOP/s (cuBLAS %7.1f, x%.2f) | bf16 incl. conversion %7.1f TFLOP/s '
             '(cuBLAS bf16 on bf16 operands %7.1f) | %s' % (
                 m, n, k, tA, tB, flops / ms_tf32 / 1e9, flops / ms_cublas_tf32 / 1e9, ms_cublas_tf32 / ms_tf32,
                 flops / ms_bf16 / 1e9, flops / ms_cublas_bf16 / 1e9, clk.window(t0, t1)))
    clk.p.terminate()


if __name__ == '__main__':
    main()
