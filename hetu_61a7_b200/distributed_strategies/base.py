"""Strategy base.  reference: distributed_strategies/base.py:11-21 (reads the
cluster description that ``heturun`` wrote to /tmp/hetu_config.yml).  Without a
launcher file the cluster is derived from WORLD_SIZE (torchrun) or is a single
local worker."""
from __future__ import annotations

import os

from ..context import DeviceGroup, DistConfig

CONFIG_PATH = os.environ.get('HETU_CONFIG_PATH', '/tmp/hetu_config.yml')


class Strategy(object):
    def __init__(self, save_path=None):
        world = int(os.environ.get('WORLD_SIZE', '1'))
        if os.path.exists(CONFIG_PATH) and os.environ.get('HETU_LAUNCHED', '0') == '1':
            self.settings = DistConfig(CONFIG_PATH)
        else:
            self.settings = DistConfig(num_local_workers=world,
                                       num_local_servers=int(os.environ.get('DMLC_NUM_SERVER', '0')))
        self.save_path = save_path
        self.overlap = True
        self.use_nccl_collectives = None

    def worker_devices(self):
        devs = []
        for host, num in self.settings.workers.items():
            devs.extend('%s:gpu:%d' % (host, i) for i in range(num))
        return devs

    def set_raw_ctxs_n_states(self, node_list, memory_pool=None):
        raise NotImplementedError

    def set_overlap(self, overlap):
        self.overlap = overlap
