"""Streams and events.  reference: python/hetu/stream.py:16-105.

Cross-stream ordering is stream-ordered (``cudaStreamWaitEvent``) rather than
the reference's host-blocking ``event.sync()`` before every node
(executor.py:1034-1036); ``Event.sync`` still exists for API parity."""
from __future__ import annotations

import time

import torch


class Stream(object):
    def __init__(self, ctx=None, priority=0):
        self.ctx = ctx
        self.device = ctx.torch_device if ctx is not None else torch.device('cpu')
        self.handle = (torch.cuda.Stream(device=self.device, priority=priority)
                       if self.device.type == 'cuda' else None)

    @property
    def cuda_stream(self) -> int:
        return self.handle.cuda_stream if self.handle is not None else 0

    def sync(self):
        if self.handle is not None:
            self.handle.synchronize()

    def wait_event(self, event: 'Event'):
        if self.handle is not None and event.handle is not None:
            self.handle.wait_event(event.handle)

    def wait_stream(self, other: 'Stream'):
        if self.handle is not None and other.handle is not None:
            self.handle.wait_stream(other.handle)

    def __enter__(self):
        self._guard = torch.cuda.stream(self.handle) if self.handle is not None else None
        if self._guard is not None:
            self._guard.__enter__()
        return self

    def __exit__(self, *exc):
        if self._guard is not None:
            self._guard.__exit__(*exc)


def create_stream_handle(ctx):
    return Stream(ctx)


class Event(object):
    def __init__(self, ctx=None, timing=True):
        self.ctx = ctx
        dev = ctx.torch_device if ctx is not None else torch.device('cpu')
        self.handle = torch.cuda.Event(enable_timing=timing) if dev.type == 'cuda' else None
        self._t = None

    def record(self, stream_handle: Stream = None):
        if self.handle is not None:
            self.handle.record(stream_handle.handle if stream_handle is not None and
                               stream_handle.handle is not None else None)
        else:
            self._t = time.perf_counter()

    def sync(self):
        if self.handle is not None:
            self.handle.synchronize()

    def query(self):
        return self.handle.query() if self.handle is not None else True

    def time_since(self, event: 'Event') -> float:
        """Milliseconds from ``event`` to ``self``."""
        if self.handle is not None:
            return event.handle.elapsed_time(self.handle)
        return (self._t - event._t) * 1000.0

    def update_ts(self, ts):
        pass


def create_event_handle(ctx):
    return Event(ctx)


class PSEvent(object):
    """Wait handle on an outstanding parameter-server request (timestamp list)."""

    def __init__(self, comm, nid):
        self.comm = comm
        self.nid = nid
        self._pending = []

    def update(self, fut=None):
        if fut is not None:
            self._pending.append(fut)

    def sync(self):
        for f in self._pending:
            f.result() if hasattr(f, 'result') else f()
        self._pending.clear()
        if self.comm is not None and hasattr(self.comm, 'Wait'):
            self.comm.Wait(self.nid)

    def record(self, stream_handle=None):
        pass

    def update_ts(self, ts):
        pass


class CSEvent(object):
    """Wait handle on HET-cache futures (embedding_lookup / update / push_pull)."""

    def __init__(self, comm=None, nid=0):
        self.comm, self.nid = comm, nid
        self._ts = None

    def update_ts(self, ts):
        self._ts = ts

    def sync(self):
        if self._ts is not None:
            self._ts.wait() if hasattr(self._ts, 'wait') else self._ts.result()
            self._ts = None

    def record(self, stream_handle=None):
        pass

    def update(self):
        pass


#: reference stream.py ctypes handles; here streams / events wrap torch.cuda objects
DLStream = Stream
DLEvent = Event
