import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: test needs a real B200 (run with -m gpu on the GPU box)')
    config.addinivalue_line('markers', 'multigpu: test needs >1 GPU')


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason='no GPU in this container')
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


def free_port_block(span=8):
    """First port of ``span + 1`` consecutive free loopback ports, taken below the kernel's ephemeral range so that no
    outgoing connection of another process can land on one of them between this check and the bind of the process
    under test (a base port from bind(0) says nothing about base + 1 ...)."""
    import random
    import socket
    rng = random.Random(os.getpid() * 7919 + int.from_bytes(os.urandom(4), 'little'))
    for _ in range(500):
        p = rng.randint(20000, 31000 - span)
        socks = []
        try:
            for q in range(p, p + span + 1):
                s = socket.socket()
                socks.append(s)
                s.bind(('127.0.0.1', q))
            return p
        except OSError:
            continue
        finally:
            for s in socks:
                s.close()
    raise RuntimeError('no block of %d free ports found' % (span + 1))
