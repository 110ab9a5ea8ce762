"""Device groups, the ``ht.context`` stack, launch-config discovery and the
cluster description used by ``heturun``.

reference: python/hetu/context.py:19-319.  Differences by design:
 * one process drives exactly one B200 (rank == worker index); a group's
   ``cur_worker(rank)`` picks the unit this process executes;
 * tuple entries (one model-parallel worker) are first-class: ``is_mp`` and
   ``mp_dev_num`` are computed here (the reference references but never
   defines them - SURVEY.md section 0).
"""
from __future__ import annotations

import contextlib
import os
import re
import socket
from typing import Dict, Iterator, List, Optional, Set, Tuple, Union

from .ndarray import DLContext, cpu, gpu, is_gpu_ctx, rcpu, rgpu

ContextUnit = Union[DLContext, Tuple[DLContext, ...]]


class DeviceGroup(object):
    def __init__(self, ctxs) -> None:
        self._contexts: List[ContextUnit] = self.parse_contexts(ctxs)
        self.get_servers_n_workers()

    def get_servers_n_workers(self) -> None:
        """Split the parsed contexts into workers (GPUs, or tuples = one model-parallel worker) and servers
        (CPU contexts).  reference context.py:101-111."""
        workers, servers = [], []
        mp = 1
        for c in self._contexts:
            if isinstance(c, tuple):
                workers.append(c)
                mp = max(mp, len(c))
            elif is_gpu_ctx(c):
                workers.append(c)
            else:
                servers.append(c)
        self._workers = tuple(workers)
        self._servers = tuple(servers)
        self._mp_dev_num = mp
        self._is_mp = mp > 1
        self._hash = hash(tuple(self._contexts))

    @classmethod
    def parse_contexts(cls, ctxs) -> List[ContextUnit]:
        if isinstance(ctxs, DeviceGroup):
            return list(ctxs._contexts)
        if isinstance(ctxs, str):
            ctxs = [c for c in re.split(r';|,| +', ctxs.lower()) if c]
        if not isinstance(ctxs, list):
            ctxs = [ctxs]
        out = []
        for c in ctxs:
            if isinstance(c, tuple):
                unit = tuple(cls.str2ctx(x) for x in c)
                out.append(unit if len(unit) > 1 else unit[0])
            else:
                c = cls.str2ctx(c)
                if c is not None:
                    out.append(c)
        return out

    @classmethod
    def str2ctx(cls, c) -> DLContext:
        if isinstance(c, str):
            parts = c.lower().split(':')
            assert parts[-2] in ('cpu', 'gpu'), 'Context invalid: %s' % c
            host = 'localhost' if len(parts) == 2 else parts[0]
            idx = int(parts[-1])
            c = rcpu(host, idx) if parts[-2] == 'cpu' else rgpu(host, idx)
        assert isinstance(c, DLContext), 'Context invalid: %s' % str(c)
        return c

    def index(self, ctx) -> int:
        for i, c in enumerate(self._contexts):
            if c == ctx or (isinstance(c, tuple) and ctx in c):
                return i
        raise ValueError('%s not in %s' % (ctx, self))

    def __contains__(self, ctx) -> bool:
        try:
            self.index(ctx)
            return True
        except ValueError:
            return False

    def __getitem__(self, key: int) -> ContextUnit:
        return self._workers[key]

    def __iter__(self) -> Iterator[ContextUnit]:
        return iter(self._contexts)

    def __len__(self) -> int:
        return len(self._contexts)

    def check_mp_num(self, mp_dev_num: int) -> None:
        assert mp_dev_num == self._mp_dev_num

    is_mp = property(lambda self: self._is_mp)
    mp_dev_num = property(lambda self: self._mp_dev_num)
    worker_num = property(lambda self: len(self._workers))
    server_num = property(lambda self: len(self._servers))
    workers = property(lambda self: self._workers)
    servers = property(lambda self: self._servers)

    def flat_devices(self) -> List[DLContext]:
        out = []
        for c in self._contexts:
            out.extend(c if isinstance(c, tuple) else [c])
        return out

    def __repr__(self) -> str:
        return 'DeviceGroup(' + ''.join('%s, ' % (c,) for c in self._contexts) + ')'

    def full_repr(self) -> str:
        return str([c.full_repr() if not isinstance(c, tuple) else tuple(x.full_repr() for x in c)
                    for c in self._contexts])

    def __hash__(self) -> int:
        return self._hash

    def __eq__(self, other) -> bool:
        return isinstance(other, DeviceGroup) and self._contexts == other._contexts

    def get_sorted(self) -> 'DeviceGroup':
        return DeviceGroup(sorted(self.flat_devices(),
                                  key=lambda x: '{}:{}:{:04d}'.format(x.hostname, x.device_type, x.device_id)))

    def cur_worker(self, dp_index: int = 0) -> Optional[ContextUnit]:
        if not self._workers:
            return None
        return self._workers[dp_index % len(self._workers)]

    def get_only(self) -> DLContext:
        assert self.server_num + self.worker_num == 1
        if self.server_num == 1:
            return self._servers[0]
        res = self._workers[0]
        if isinstance(res, tuple):
            assert len(res) == 1
            res = res[0]
        return res


class ContextStack(object):
    def __init__(self) -> None:
        self._stack: List[DeviceGroup] = []

    def peek(self) -> Optional[DeviceGroup]:
        return self._stack[-1] if self._stack else None

    def push(self, ctx: DeviceGroup) -> None:
        self._stack.append(ctx)

    def pop(self) -> DeviceGroup:
        return self._stack.pop()


_default_ctx_stack = ContextStack()


def get_current_context() -> Optional[DeviceGroup]:
    return _default_ctx_stack.peek()


@contextlib.contextmanager
def context(ctx):
    """``with ht.context([...])`` - every op built inside gets this raw_ctx."""
    try:
        ctx = DeviceGroup(ctx)
        _default_ctx_stack.push(ctx)
        yield ctx
    finally:
        _default_ctx_stack.pop()


def get_launch_config_by_traverse_nodes(node_list, default_ctx: DeviceGroup):
    """Walk the graph and decide, per node, whether its parameters are
    synchronised by the parameter server ('PS': group has CPU servers and GPU
    workers) or by all-reduce ('AllReduce': >1 workers).
    reference: context.py:184-234.  Returns
    (launchMPI, launchPS, node_strategy, devices, min_worker_num)."""
    from .optimizer import OptimizerOp
    from .dataloader import DataloaderOp
    node_strategy: Dict = {}
    devices: Set[DLContext] = set(default_ctx.flat_devices())
    min_worker_num = default_ctx.worker_num
    stack = list(node_list)
    while stack:
        node = stack.pop()
        if node in node_strategy:
            continue
        rc = node.raw_ctx
        nworker = 1 if rc is None else rc.worker_num
        strategy = None
        if rc is not None and rc.server_num > 0 and nworker > 0:
            strategy = 'PS'
        elif rc is not None and nworker > 1:
            strategy = 'AllReduce'
        node_strategy[node] = strategy
        if rc is not None and not isinstance(node, (OptimizerOp, DataloaderOp)):
            devices.update(rc.flat_devices())
        stack.extend(node.inputs)
    # a group of CPU contexts only (e.g. ``with ht.context(ht.cpu(0))``) is a local CPU run, not a server without workers
    launchPS = (default_ctx.server_num > 0 and default_ctx.worker_num > 0) or any(s == 'PS' for s in node_strategy.values())
    launchMPI = ((not launchPS) and min_worker_num > 1) or any(
        s == 'AllReduce' for s in node_strategy.values())
    return launchMPI, launchPS, node_strategy, devices, min_worker_num


class DistConfig(object):
    """Cluster YAML: ``nodes: [{host, servers, workers, chief}]``.
    reference: context.py:237-319."""

    def __init__(self, file: Optional[str] = None, num_local_servers: int = 0,
                 num_local_workers: int = 1) -> None:
        import yaml
        if file is None:
            assert num_local_workers > 0, \
                'Please specify the configuration file or set the number of local workers.'
            self.settings = {'nodes': [{'host': socket.gethostname(), 'servers': num_local_servers,
                                        'workers': num_local_workers, 'chief': True}]}
        else:
            with open(file) as fr:
                self.settings = yaml.safe_load(fr.read())
        attributes = {'host', 'servers', 'workers', 'chief'}
        hosts, servers, workers, chief = [], {}, {}, None
        for node in self.settings['nodes']:
            assert set(node.keys()) <= attributes, \
                'Attributes of nodes invalid, %s / %s.' % (set(node.keys()), attributes)
            hosts.append(node['host'])
            if node.get('servers', 0):
                servers[node['host']] = node['servers']
            if node.get('workers', 0):
                workers[node['host']] = node['workers']
            if node.get('chief', False):
                assert chief is None, 'There should be only one chief.'
                chief = node['host']
        assert chief, 'There should be one chief.'
        self.num_servers = sum(servers.values())
        self.num_workers = sum(workers.values())
        self.enable_PS = self.num_servers > 0
        self.servers, self.workers = servers, workers
        self.chief, self.hosts = chief, hosts
        try:
            self.chief_address = socket.gethostbyname(socket.gethostname())
        except OSError:
            self.chief_address = '127.0.0.1'

    def __str__(self) -> str:
        return '\n'.join(['Cluster: {', '  Chief: %s,' % self.chief,
                          '  Servers(%d): %s,' % (self.num_servers, self.servers),
                          '  Workers(%d): %s,' % (self.num_workers, self.workers), '}'])

    def __iter__(self):
        return iter(self.settings['nodes'])

    def save(self, path: str) -> None:
        import yaml
        with open(path, 'w') as fw:
            yaml.dump(self.settings, fw)

    def make_ps_config(self) -> Dict[str, Union[str, int]]:
        single = len(self.hosts) == 1
        addr = '127.0.0.1' if single else self.chief_address
        return {
            'DMLC_PS_ROOT_URI': addr,
            'DMLC_PS_ROOT_PORT': self.get_available_port(addr),
            'DMLC_NUM_WORKER': self.num_workers,
            'DMLC_NUM_SERVER': self.num_servers,
            'DMLC_PS_VAN_TYPE': 'tcp',
        }

    @staticmethod
    def get_available_port(localhost: str = '127.0.0.1') -> int:
        for p in range(13100, 13200):
            with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
                s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
                try:
                    s.bind(('', p))
                    return p
                except OSError:
                    continue
        raise RuntimeError('no free port in 13100-13199')
