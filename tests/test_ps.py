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
    from conftest import free_port_block
    return free_port_block(8)            # scheduler / servers / ring ranks bind base, base + 1, ...


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


def _spawn(code, env, **kw):
    return subprocess.Popen([sys.executable, '-c', code], env=env, **kw)


def test_ps_resend_survives_dropped_messages():
    """PS_DROP_MSG loses requests and replies at the server; PS_RESEND retransmits and the server's
    reply cache keeps pushes exactly-once (reference: resender.h:15-148, van.cc:334-336)."""
    port = _free_port()
    env = dict(os.environ, DMLC_PS_ROOT_URI='127.0.0.1', DMLC_PS_ROOT_PORT=str(port), DMLC_NUM_WORKER='1',
               DMLC_NUM_SERVER='1', PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
    procs = []
    try:
        procs.append(_spawn('import hetu_61a7_b200.launcher as l; l.start_sched()', env))
        time.sleep(0.5)
        procs.append(_spawn('import hetu_61a7_b200.launcher as l; l.start_server()', dict(env, PS_DROP_MSG='25')))
        w = subprocess.Popen([sys.executable, os.path.join(ROOT, 'tests', 'ps_fault_worker.py')],
                             env=dict(env, PS_RESEND='1', PS_RESEND_TIMEOUT='60'),
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        out = w.communicate(timeout=300)[0]
        assert w.returncode == 0 and 'OK' in out, out[-3000:]
    finally:
        for p in procs:
            p.terminate()


def test_ps_heartbeat_dead_node_and_recovery():
    """A server that stops beating shows up in get_dead_nodes; a replacement server that registers
    late takes over its rank and the worker finds it through the scheduler
    (reference: van.cc:132-162,567-579, postoffice.cc:201-221)."""
    port = _free_port()
    env = dict(os.environ, DMLC_PS_ROOT_URI='127.0.0.1', DMLC_PS_ROOT_PORT=str(port), DMLC_NUM_WORKER='1',
               DMLC_NUM_SERVER='1', PS_HEARTBEAT_INTERVAL='0.1', PS_HEARTBEAT_TIMEOUT='2.0',
               PS_RESEND='1', PS_RESEND_TIMEOUT='200',
               PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
    code = r'''
import os, sys, time, subprocess, numpy as np
sys.path.insert(0, %r)
from hetu_61a7_b200.parallel.ps import PSWorker, get_dead_nodes
w = PSWorker.from_env()
w.InitTensor(1, 0, 16, 1, 0, 2.0, 2.0, 1)
assert np.allclose(w.Pull(1), 2.0)
assert get_dead_nodes(2.0) == []
print('READY', flush=True)
sys.stdin.readline()                       # parent kills the server, then pokes us
deadline = time.time() + 40
dead = []
while time.time() < deadline and 8 not in dead:
    time.sleep(0.2); dead = get_dead_nodes(2.0)
assert dead == [8], dead                  # server rank 0; this worker (id 9) still beats
print('DEAD', flush=True)
sys.stdin.readline()                       # parent has started the replacement server
w.InitTensor(1, 0, 16, 1, 0, 5.0, 5.0, 1)  # state is not restored: re-initialise, as after LoadParam
assert np.allclose(w.Pull(1), 5.0)
assert get_dead_nodes(2.0) == []
print('OK', flush=True)
w.finalize()
''' % ROOT
    procs = []
    try:
        procs.append(_spawn('import hetu_61a7_b200.launcher as l; l.start_sched()', env))
        time.sleep(0.5)
        srv = _spawn('import hetu_61a7_b200.launcher as l; l.start_server()', env)
        procs.append(srv)
        w = subprocess.Popen([sys.executable, '-c', code], env=env, stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True)
        procs.append(w)

        def expect(tag):
            lines = []
            while True:
                line = w.stdout.readline()
                if not line:
                    raise AssertionError('worker ended before %s:\n%s' % (tag, ''.join(lines)[-3000:]))
                lines.append(line)
                if line.strip() == tag:
                    return
        expect('READY')
        srv.kill()
        srv.wait()
        w.stdin.write('\n'); w.stdin.flush()
        expect('DEAD')
        procs.append(_spawn('import hetu_61a7_b200.launcher as l; l.start_server()', env))
        # the replacement has taken over once the scheduler stops listing server 0 as dead (its start-up time varies with load)
        sys.path.insert(0, ROOT)
        from hetu_61a7_b200.parallel.ps import get_dead_nodes
        deadline = time.time() + 120
        while 8 in get_dead_nodes(2.0, '127.0.0.1', port):
            assert time.time() < deadline, 'replacement server never registered'
            time.sleep(0.3)
        w.stdin.write('\n'); w.stdin.flush()
        expect('OK')
        assert w.wait(timeout=60) == 0
    finally:
        for p in procs:
            if p.poll() is None:
                p.terminate()


def test_cpu_ring_allreduce_three_ranks():
    """Legacy CPU data-parallel path: TCP ring reduce-scatter + all-gather
    (reference src/communication/c_communication_nthread.cc)."""
    base = _free_port()
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''), OMP_NUM_THREADS='2')
    ws = [subprocess.Popen([sys.executable, os.path.join(ROOT, 'tests', 'ring_worker.py'), str(r), '3', str(base)], env=env,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(3)]
    outs = [w.communicate(timeout=120)[0] for w in ws]
    for w, o in zip(ws, outs):
        assert w.returncode == 0 and 'OK' in o, o[-3000:]


def test_ps_server_side_optimizers():
    """SGD / Momentum / Nesterov / AdaGrad / Adam applied by the server to pushed gradients
    (reference ps-lite include/ps/server/optimizer.h:36-357), against numpy."""
    port = _free_port()
    env = dict(os.environ, DMLC_PS_ROOT_URI='127.0.0.1', DMLC_PS_ROOT_PORT=str(port), DMLC_NUM_WORKER='1',
               DMLC_NUM_SERVER='2', PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
    procs = []
    try:
        procs.append(_spawn('import hetu_61a7_b200.launcher as l; l.start_sched()', env))
        time.sleep(0.5)
        for _ in range(2):
            procs.append(_spawn('import hetu_61a7_b200.launcher as l; l.start_server()', env))
        w = subprocess.Popen([sys.executable, os.path.join(ROOT, 'tests', 'ps_opt_worker.py')], env=env,
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        out = w.communicate(timeout=120)[0]
        assert w.returncode == 0 and 'OK' in out, out[-3000:]
    finally:
        for p in procs:
            p.terminate()
