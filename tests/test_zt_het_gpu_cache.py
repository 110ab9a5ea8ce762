"""GPU-resident HET cache (csrc/het_cache.cu, parallel/gpu_cache.py) in front of a real parameter server over loopback:
exactness against a numpy model under LRU / LFU / LFUOpt with and without capacity pressure, flush, staleness bound.
reference: src/hetu_cache/src/cache.cc:59-106,131-196, tests/hetu_cache/hetu_cache_test.py."""
import os
import socket
import subprocess
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]


def test_gpu_het_cache_against_the_parameter_server():
    from conftest import free_port_block
    port = free_port_block(2)
    env = dict(os.environ, DMLC_PS_ROOT_URI='127.0.0.1', DMLC_PS_ROOT_PORT=str(port), DMLC_NUM_WORKER='1',
               DMLC_NUM_SERVER='1', PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
    procs = []
    try:
        procs.append(subprocess.Popen([sys.executable, '-c', 'import hetu_61a7_b200.launcher as l; l.start_sched()'], env=env))
        time.sleep(0.5)
        procs.append(subprocess.Popen([sys.executable, '-c', 'import hetu_61a7_b200.launcher as l; l.start_server()'], env=env))
        w = subprocess.Popen([sys.executable, os.path.join(ROOT, 'tests', 'ps_gpu_cache_worker.py')], env=env,
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        out = w.communicate(timeout=500)[0]
        assert w.returncode == 0 and 'GPU HET CACHE OK' in out, out[-4000:]
    finally:
        for p in procs:
            p.terminate()
