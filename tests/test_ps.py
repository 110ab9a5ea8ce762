"""Parameter server + HET cache integration: scheduler, 2 servers and 2 workers as local
processes over loopback TCP (pattern: reference tests/pstests/test_apis.py with
pstests/local_s2_w2.yml, and tests/hetu_cache/hetu_cache_test.py)."""
import os
import socket
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def test_ps_scheduler_2servers_2workers():
    port = _free_port()
    env = dict(os.environ, DMLC_PS_ROOT_URI='127.0.0.1', DMLC_PS_ROOT_PORT=str(port), DMLC_NUM_WORKER='2',
               DMLC_NUM_SERVER='2', PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
    procs = []
    with tempfile.TemporaryDirectory() as d:
        try:
            procs.append(subprocess.Popen([sys.executable, '-c', 'import hetu_61a7_b200.launcher as l; l.start_sched()'], env=env))
            time.sleep(0.5)
            for _ in range(2):
                procs.append(subprocess.Popen([sys.executable, '-c', 'import hetu_61a7_b200.launcher as l; l.start_server()'], env=env))
            workers = [subprocess.Popen([sys.executable, os.path.join(ROOT, 'tests', 'ps_worker.py'), d], env=env,
                                        stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for _ in range(2)]
            outs = [w.communicate(timeout=300)[0] for w in workers]
            for w, o in zip(workers, outs):
                assert w.returncode == 0 and 'OK' in o, o[-3000:]
            assert any(f.endswith('.dat') for f in os.listdir(d))          # raw fp32 shards <node>_<server>.dat
        finally:
            for p in procs:
                p.terminate()
