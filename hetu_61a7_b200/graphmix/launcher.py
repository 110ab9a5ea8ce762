"""Process launcher for graph servers and workers.

reference: third_party/GraphMix/python/graphmix/launcher.py:1-135 - ``launcher(target, args,
server_init)`` reads ``args.config`` (YAML: ``env`` with GRAPHMIX_ROOT_URI / GRAPHMIX_ROOT_PORT /
GRAPHMIX_NUM_WORKER, ``launch`` with worker / server / scheduler / data), forks one process per
role, calls ``server_init(server)`` in every server and ``target(args)`` in every worker.
Graph server r listens on GRAPHMIX_ROOT_PORT + 1 + r; there is no separate scheduler process
(server 0 is the barrier point), a ``scheduler`` entry in the config is accepted and ignored."""
from __future__ import annotations

import multiprocessing as mp
import os
import signal
import sys

import yaml

from .client import Client, default_server_port
from .server import Server
from .shard import Shard


def _addresses(env, nserver):
    host = str(env.get('GRAPHMIX_ROOT_URI', '127.0.0.1'))
    port = int(env.get('GRAPHMIX_ROOT_PORT', default_server_port))
    hosts = env.get('GRAPHMIX_SERVER_HOSTS')
    hosts = [h.strip() for h in str(hosts).split(',')] if hosts else [host] * nserver
    return [(hosts[r % len(hosts)], port + 1 + r) for r in range(nserver)]


def _server_main(path, rank, addrs, nworker, server_init, env):
    os.environ.update({k: str(v) for k, v in env.items()})
    shard = Shard(path).load_graph_shard(rank)
    server = Server(shard, rank, addrs, nworker)
    if server_init is not None:
        server_init(server)
    else:
        server.is_ready()
    server.wait()


def _worker_main(target, args, rank, addrs, nworker, env):
    os.environ.update({k: str(v) for k, v in env.items()})
    os.environ['GRAPHMIX_NUM_WORKER'] = str(nworker)
    os.environ['GRAPHMIX_WORKER_RANK'] = str(rank)
    os.environ['GRAPHMIX_SERVERS'] = ','.join('%s:%d' % a for a in addrs)
    target(args)


def launcher(target, args, server_init=None, config=None):
    """Run ``target(args)`` in every worker against freshly started graph servers; returns when all
    workers are done (servers are then stopped).  ``config`` overrides ``args.config``."""
    cfg_path = config or getattr(args, 'config')
    with open(cfg_path) as f:
        cfg = yaml.safe_load(f)
    env = cfg.get('env', {}) or {}
    launch = cfg.get('launch', {}) or {}
    data = os.path.expanduser(str(launch['data']))
    meta = Shard(data).meta
    nserver_total = int(meta['num_part'])
    nworker_total = int(env.get('GRAPHMIX_NUM_WORKER', launch.get('worker', 1)))
    addrs = _addresses(env, nserver_total)
    local_servers = launch.get('server', nserver_total)
    server_ranks = list(local_servers) if isinstance(local_servers, (list, tuple)) else list(range(int(local_servers)))
    worker_base = int(launch.get('worker_rank_base', 0))
    ctx = mp.get_context('fork')
    servers = [ctx.Process(target=_server_main, args=(data, r, addrs, nworker_total, server_init, env), daemon=True)
               for r in server_ranks]
    workers = [ctx.Process(target=_worker_main, args=(target, args, worker_base + w, addrs, nworker_total, env))
               for w in range(int(launch.get('worker', 1)))]

    def _kill(*_):
        for p in servers + workers:
            if p.is_alive():
                p.terminate()
        sys.exit(1)
    old = signal.signal(signal.SIGINT, _kill)
    try:
        for p in servers + workers:
            p.start()
        for p in workers:
            p.join()
        codes = [p.exitcode for p in workers]
    finally:
        signal.signal(signal.SIGINT, old)
        for p in servers:
            if p.is_alive():
                p.terminate()
        for p in servers:
            p.join(timeout=5)
    if any(codes):
        raise RuntimeError('graphmix worker(s) failed with exit codes %s' % codes)


def start_standalone_server(data_path, port=default_server_port, server_init=None):
    """One in-process server over a single-part graph (reference tests/test_standalone.py)."""
    shard = Shard(data_path).load_graph_shard(0)
    server = Server(shard, 0, [('127.0.0.1', port)], 0)
    if server_init is not None:
        server_init(server)
    else:
        server.is_ready()
    return server
