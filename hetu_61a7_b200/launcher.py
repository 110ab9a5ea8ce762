"""Process launcher for the parameter-server plane.

reference: python/hetu/launcher.py:11-82 (``launch(target, args)`` starts scheduler, servers
and workers as local processes from a YAML with ``shared: {DMLC_*}`` and
``launch: {worker, server, scheduler}``) and ``python -m hetu.launcher cfg.yml -n N [--sched]``."""
from __future__ import annotations

import argparse
import multiprocessing
import os
import signal
import time

import yaml

_procs = []


def signal_handler(sig, frame):
    print("SIGINT signal caught, stop Training")
    for proc in _procs:
        if proc.is_alive():
            proc.terminate()
    raise SystemExit(0)


def start_sched():
    os.environ["DMLC_ROLE"] = "scheduler"
    from .parallel.ps import start_scheduler_from_env
    start_scheduler_from_env(block=True)


def start_server():
    os.environ["DMLC_ROLE"] = "server"
    from .parallel.ps import start_server_from_env
    start_server_from_env(block=True)


def start_worker(target, args):
    os.environ["DMLC_ROLE"] = "worker"
    from .graph import executor as ex
    ex.worker_init()
    try:
        target(args)
    finally:
        ex.worker_finish()


def launch(target, args):
    """``args.config`` names a YAML with ``shared`` env vars and ``launch`` process counts."""
    with open(args.config) as f:
        settings = yaml.safe_load(f.read())
    for k, v in settings.get('shared', {}).items():
        os.environ[k] = str(v)
    args.num_local_worker = int(settings["launch"]["worker"])
    ctx = multiprocessing.get_context('spawn')
    if settings["launch"].get("scheduler", 0):
        _procs.append(ctx.Process(target=start_sched))
    for _ in range(int(settings["launch"].get("server", 0))):
        _procs.append(ctx.Process(target=start_server))
    for i in range(args.num_local_worker):
        _procs.append(ctx.Process(target=start_worker, args=(target, args)))
    signal.signal(signal.SIGINT, signal_handler)
    for proc in _procs:
        proc.start()
    nworker = args.num_local_worker
    for proc in _procs[len(_procs) - nworker:]:
        proc.join()
    for proc in _procs[:len(_procs) - nworker]:
        proc.terminate()


def main():
    parser = argparse.ArgumentParser()
    parser.add_argument("config")
    parser.add_argument("-n", type=int, default=1, help='number of servers to start on this host')
    parser.add_argument("--sched", action='store_true')
    a = parser.parse_args()
    with open(a.config) as f:
        settings = yaml.safe_load(f.read())
    for k, v in (settings.get('shared', settings) or {}).items():
        if isinstance(v, (str, int, float)):
            os.environ[k] = str(v)
    ctx = multiprocessing.get_context('spawn')
    if a.sched:
        _procs.append(ctx.Process(target=start_sched))
    for _ in range(a.n):
        _procs.append(ctx.Process(target=start_server))
    signal.signal(signal.SIGINT, signal_handler)
    for p in _procs:
        p.start()
    for p in _procs:
        p.join()


if __name__ == '__main__':
    main()
