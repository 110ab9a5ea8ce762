"""``heturun`` - cluster launcher.

reference: bin/heturun -> python/runner.py:1-262: parse ``-c cluster.yml`` (nodes: [{host,
servers, workers, chief}]) or ``-w N -s N``, write /tmp/hetu_config.yml, start the PS scheduler
and servers, then ``mpirun -np W <cmd>``; remote hosts through ssh (paramiko).

There is no MPI on the B200 boxes: ranks are spawned directly, one process per GPU, with the
rendezvous environment both ``torch.distributed`` and the symmetric-heap bootstrap read
(RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT) plus the ``DMLC_*`` variables of
the PS plane.  Remote hosts are started through ``ssh`` with the same environment.
"""
from __future__ import annotations

import argparse
import os
import shlex
import signal
import socket
import subprocess
import sys
import time

from .context import DistConfig

_children = []


def signal_handler(sig, frame):
    print("SIGINT signal caught, stop Training")
    for p in _children:
        try:
            p.terminate()
        except Exception:
            pass
    sys.exit(0)


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _spawn(cmd, env, host=None):
    if host in (None, 'localhost', '127.0.0.1', socket.gethostname()):
        p = subprocess.Popen(cmd, env=env)
    else:
        exports = ' '.join('%s=%s' % (k, shlex.quote(str(v))) for k, v in env.items()
                           if k.startswith(('DMLC_', 'RANK', 'LOCAL_RANK', 'WORLD_SIZE', 'MASTER_', 'HETU_', 'PYTHONPATH')))
        remote = 'cd %s && env %s %s' % (shlex.quote(os.getcwd()), exports, ' '.join(shlex.quote(c) for c in cmd))
        p = subprocess.Popen(['ssh', '-o', 'StrictHostKeyChecking=no', host, remote])
    _children.append(p)
    return p


def main(argv=None):
    signal.signal(signal.SIGINT, signal_handler)
    parser = argparse.ArgumentParser(prog='heturun')
    parser.add_argument('-c', '--config', default=None, help='cluster YAML (nodes: [{host, servers, workers, chief}])')
    parser.add_argument('-w', '--workers', type=int, default=0, help='shorthand: local workers')
    parser.add_argument('-s', '--servers', type=int, default=0, help='shorthand: local PS servers')
    parser.add_argument('-i', '--identify', default='', help='ssh identity file')
    parser.add_argument('--max-restarts', type=int, default=0,
                        help='relaunch ALL workers (new rendezvous, HETU_RESTART_COUNT=k) when one exits with an error; '
                             'the script resumes from its own checkpoint (Executor.save / load).  Not for PS launches.')
    parser.add_argument('command', nargs=argparse.REMAINDER)
    args = parser.parse_args(argv)
    assert args.command, 'heturun: no command given'
    if args.config is None:
        assert args.workers > 0, 'Please specify the configuration file or set the number of local workers.'
        cfg = DistConfig(num_local_servers=args.servers, num_local_workers=args.workers)
    else:
        cfg = DistConfig(args.config)
    print(cfg)
    config_path = os.environ.get('HETU_CONFIG_PATH', '/tmp/hetu_config.yml')
    cfg.save(config_path)
    base_env = dict(os.environ, HETU_LAUNCHED='1', HETU_CONFIG_PATH=config_path)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    base_env['PYTHONPATH'] = root + os.pathsep + base_env.get('PYTHONPATH', '')
    single = len(cfg.hosts) == 1
    master = '127.0.0.1' if single else cfg.chief_address
    base_env.update(MASTER_ADDR=master, MASTER_PORT=str(_free_port()), WORLD_SIZE=str(cfg.num_workers))
    if cfg.enable_PS:
        ps = cfg.make_ps_config()
        base_env.update({k: str(v) for k, v in ps.items()})
        sched_env = dict(base_env, DMLC_ROLE='scheduler')
        _spawn([sys.executable, '-c', 'import hetu_61a7_b200.launcher as l; l.start_sched()'], sched_env, cfg.chief)
        for host, n in cfg.servers.items():
            for _ in range(n):
                env = dict(base_env, DMLC_ROLE='server', DMLC_NODE_HOST='127.0.0.1' if single else host)
                _spawn([sys.executable, '-c', 'import hetu_61a7_b200.launcher as l; l.start_server()'], env, host)
        time.sleep(0.3)
    def launch_workers(restart):
        env0 = dict(base_env, HETU_RESTART_COUNT=str(restart))
        if restart:
            env0['MASTER_PORT'] = str(_free_port())               # fresh rendezvous for the new incarnation
        out, rank = [], 0
        for host, n in cfg.workers.items():
            for local in range(n):
                env = dict(env0, RANK=str(rank), LOCAL_RANK=str(local), DMLC_ROLE='worker', HETU_LOCAL_WORLD=str(n))
                out.append(_spawn(args.command, env, host))
                rank += 1
        return out

    # Failure detection on the collective plane (the reference has none: NCCLCHECK aborts one rank and the others
    # hang in their next collective).  Workers are polled; the first non-zero exit takes the rest down, and with
    # --max-restarts the whole group is started again.
    restart, code = 0, 0
    while True:
        workers = launch_workers(restart)
        code, pending = 0, list(workers)
        while pending and code == 0:
            for p in list(pending):
                rc = p.poll()
                if rc is not None:
                    pending.remove(p)
                    code = code or rc
            if pending and code == 0:
                time.sleep(0.2)
        if code != 0:
            for p in pending:
                p.terminate()
            for p in pending:
                try:
                    p.wait(timeout=10)
                except subprocess.TimeoutExpired:
                    p.kill()
            if restart < args.max_restarts and not cfg.enable_PS:
                restart += 1
                print('heturun: a worker failed with exit code %d; restarting all workers (%d of %d)' % (code, restart, args.max_restarts),
                      flush=True)
                continue
        break
    for p in _children:
        if p.poll() is None:
            p.terminate()
    return code


if __name__ == '__main__':
    sys.exit(main())
