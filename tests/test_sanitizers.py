"""Race / memory checking of the native host libraries (the reference has no sanitizer builds).
Each driver under tests/native links the library sources directly, is built with ThreadSanitizer or
Address+UndefinedBehaviour sanitizer, and must finish without a single report.  GPU kernels are checked with
compute-sanitizer through tools/sanitize.sh on a GPU box."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, 'hetu_61a7_b200')
CASES = {
    'ps_tsan': (['-fsanitize=thread', '-fopenmp'], ['tests/native/ps_stress.cc', 'csrc_ps/ps.cc', 'csrc_ps/cache.cc', 'csrc_ps/ring.cc'],
                'PS STRESS OK'),
    'ps_asan': (['-fsanitize=address,undefined', '-fno-omit-frame-pointer', '-fopenmp'],
                ['tests/native/ps_stress.cc', 'csrc_ps/ps.cc', 'csrc_ps/cache.cc', 'csrc_ps/ring.cc'], 'PS STRESS OK'),
    'graph_tsan': (['-fsanitize=thread'], ['tests/native/graph_stress.cc', 'csrc_graph/graph.cc'], 'GRAPH STRESS OK'),
    'ml_asan': (['-fsanitize=address,undefined', '-fno-omit-frame-pointer', '-fopenmp'], ['tests/native/ml_asan.cc', 'csrc_ml/ml.cc'],
                'ML ASAN OK'),
    'gemm_asan': (['-fsanitize=address,undefined', '-fno-omit-frame-pointer', '-fopenmp', '-march=x86-64-v3'],
                  ['tests/native/cpu_gemm_asan.cc', 'csrc_cpu/cpu_ops.cc'], 'CPU GEMM ASAN OK'),
    'cpu_ops_asan': (['-fsanitize=address,undefined', '-fno-omit-frame-pointer', '-fopenmp', '-march=x86-64-v3'],
                     ['tests/native/cpu_ops_asan.cc', 'csrc_cpu/cpu_ops.cc', 'csrc_cpu/bfc_allocator.cc'], 'CPU OPS ASAN OK'),
}


@pytest.mark.parametrize('name', sorted(CASES))
def test_native_library_under_sanitizer(name, tmp_path):
    if shutil.which('g++') is None:
        pytest.skip('no g++')
    flags, sources, ok = CASES[name]
    srcs = [os.path.join(ROOT, s) if s.startswith('tests/') else os.path.join(PKG, s) for s in sources]
    exe = str(tmp_path / name)
    b = subprocess.run(['g++', '-std=c++17', '-O1', '-g', '-I', os.path.join(ROOT, 'include')] + flags + srcs + ['-lpthread', '-o', exe],
                       capture_output=True, text=True, timeout=600)
    if b.returncode != 0 and ('cannot find' in b.stderr and 'san' in b.stderr):
        pytest.skip('sanitizer runtime not installed')
    assert b.returncode == 0, b.stderr[-3000:]
    env = dict(os.environ, OMP_NUM_THREADS='4', TSAN_OPTIONS='halt_on_error=0', ASAN_OPTIONS='detect_leaks=1')
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    out = r.stdout + r.stderr
    assert r.returncode == 0 and ok in out, out[-3000:]
    for bad in ('WARNING: ThreadSanitizer', 'ERROR: AddressSanitizer', 'runtime error:', 'ERROR: LeakSanitizer'):
        assert bad not in out, out[-4000:]
