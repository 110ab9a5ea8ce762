"""Python side of the parameter server (ctypes over lib/libhetu_ps.so).

reference: the ``libps.so`` C API listed in ps-lite/src/python_binding.cc:8-149 and its use in
python/hetu/gpu_ops/executor.py:100-130 (worker_init / server_init / scheduler_init).
Roles are selected with the same ``DMLC_*`` environment variables as ps-lite
(``DMLC_ROLE``, ``DMLC_PS_ROOT_URI``, ``DMLC_PS_ROOT_PORT``, ``DMLC_NUM_WORKER``,
``DMLC_NUM_SERVER``)."""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import numpy as np

from .. import _lib as L

_I64P = ctypes.POINTER(ctypes.c_int64)
_F32P = ctypes.POINTER(ctypes.c_float)


def _lib():
    lib = L.ps_lib()
    if lib is None:
        raise RuntimeError('libhetu_ps.so is not built; run `python __graft_entry__.py`')
    lib.hps_last_error.restype = ctypes.c_char_p
    return lib


def _chk(rc, what=''):
    if rc != 0:
        raise RuntimeError('PS call %s failed: %s' % (what, _lib().hps_last_error().decode()))


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_F32P)


def _i64(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(_I64P)


def _env():
    return (os.environ.get('DMLC_PS_ROOT_URI', '127.0.0.1'), int(os.environ.get('DMLC_PS_ROOT_PORT', '13100')),
            int(os.environ.get('DMLC_NUM_WORKER', '1')), int(os.environ.get('DMLC_NUM_SERVER', '1')))


def start_scheduler_from_env(block: bool = False):
    host, port, nw, ns = _env()
    _chk(_lib().hps_start_scheduler(port, nw, ns), 'start_scheduler')
    if block:
        _lib().hps_wait_scheduler()


def stop_scheduler():
    _lib().hps_stop_scheduler()


def start_server_from_env(block: bool = False):
    host, port, nw, ns = _env()
    my_host = os.environ.get('DMLC_NODE_HOST', '127.0.0.1')
    _chk(_lib().hps_start_server(host.encode(), port, my_host.encode(), ns), 'start_server')
    if block:
        _lib().hps_wait_servers()


def stop_local_servers():
    _lib().hps_stop_servers()


def get_dead_nodes(timeout: float = 60.0, host: Optional[str] = None, port: Optional[int] = None):
    """Ids of registered nodes (server r -> 8+2r, worker r -> 9+2r) whose last heartbeat is older
    than ``timeout`` seconds.  Needs ``PS_HEARTBEAT_INTERVAL`` > 0 in the node processes.
    reference: Postoffice::GetDeadNodes (ps-lite/src/postoffice.cc:201-221)."""
    h, p, _, _ = _env()
    out = (ctypes.c_int * 1024)()
    n = ctypes.c_int(0)
    _chk(_lib().hps_query_dead_nodes((host or h).encode(), int(port or p), ctypes.c_double(timeout), out, 1024,
                                     ctypes.byref(n)), 'get_dead_nodes')
    return [out[i] for i in range(min(n.value, 1024))]


class PSWorker(object):
    """Worker-side agent.  Method names follow the reference C API."""

    def __init__(self, host: str, port: int):
        self.lib = _lib()
        _chk(self.lib.hps_worker_init(host.encode(), port), 'worker_init')
        self._keep = {}
        self._meta = {}

    @staticmethod
    def from_env() -> 'PSWorker':
        host, port, _, _ = _env()
        return PSWorker(host, port)

    def finalize(self):
        self.lib.hps_worker_finalize()

    Finalize = finalize
    finish = finalize

    def Init(self):
        """reference ``ps_comm.Init()``: the connection is made in the constructor; kept for scripts."""
        return self

    def StartServer(self):
        """reference server-role entry (executor.py:113-121): run the local server shard until shutdown."""
        start_server_from_env(block=True)

    def PushData(self, node_id, rows, data, lengths=None):
        """Row-addressed raw write (reference PushData, used by the graph/GNN stores): rows of
        ``data`` replace nothing on their own - they are added, so callers push deltas."""
        self.SparsePush(node_id, rows, data)

    def PullData(self, node_id, rows, out=None, lengths=None):
        return self.SparsePull(node_id, rows, out=out)

    def rank(self) -> int:
        return int(self.lib.hps_rank())

    def is_recovery(self) -> bool:
        """True when this worker registered after the cluster was complete and took over the rank
        of a dead worker (reference ``is_recovery``, ps-lite/src/van.cc:132-162)."""
        return bool(self.lib.hps_is_recovery())

    def num_retransmits(self) -> int:
        """Requests sent again after ``PS_RESEND_TIMEOUT`` ms without a reply (``PS_RESEND=1``)."""
        self.lib.hps_num_retransmits.restype = ctypes.c_uint64
        return int(self.lib.hps_num_retransmits())

    def get_dead_nodes(self, timeout: float = 60.0):
        return get_dead_nodes(timeout)

    def nrank(self) -> int:
        return int(self.lib.hps_nrank())

    # -- tensors ------------------------------------------------------------------------------
    def InitTensor(self, node_id, param_type, length, width, init_type, a, b, seed, opt=None):
        self._meta[int(node_id)] = (int(param_type), int(length), int(width))
        if opt is not None:
            # server-side optimizer: (kind, [lr, h0, h1, eps]); pushes to this tensor are then gradients
            # (reference ps-lite include/ps/server/optimizer.h; the stock path keeps "server does +=")
            kind, args = opt
            kinds = {'sgd': 1, 'momentum': 2, 'nesterov': 3, 'adagrad': 4, 'adam': 5}
            code = kinds[kind.lower()] if isinstance(kind, str) else int(kind)
            arr = (ctypes.c_double * len(args))(*[float(v) for v in args])
            _chk(self.lib.hps_init_tensor_opt(ctypes.c_uint64(node_id), int(param_type), ctypes.c_uint64(int(length)),
                                              ctypes.c_uint64(int(width)), int(init_type), ctypes.c_double(a), ctypes.c_double(b),
                                              ctypes.c_uint64(int(seed) & 0xFFFFFFFFFFFFFFFF), code, arr, len(args)), 'InitTensor')
            return
        _chk(self.lib.hps_init_tensor(ctypes.c_uint64(node_id), int(param_type), ctypes.c_uint64(int(length)),
                                      ctypes.c_uint64(int(width)), int(init_type), ctypes.c_double(a),
                                      ctypes.c_double(b), ctypes.c_uint64(int(seed) & 0xFFFFFFFFFFFFFFFF)), 'InitTensor')

    def _numel(self, node_id):
        t, length, width = self._meta[int(node_id)]
        return length if t == 0 else length * width

    def Pull(self, node_id, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty(self._numel(node_id), dtype=np.float32)
        _chk(self.lib.hps_pull(ctypes.c_uint64(node_id), out.ctypes.data_as(_F32P)), 'Pull')
        return out

    def Push(self, node_id, grad):
        g, gp = _f32(grad)
        _chk(self.lib.hps_push(ctypes.c_uint64(node_id), gp), 'Push')

    def DDPushPull(self, node_id, grad, out: Optional[np.ndarray] = None) -> np.ndarray:
        g, gp = _f32(grad)
        if out is None:
            out = np.empty(self._numel(node_id), dtype=np.float32)
        _chk(self.lib.hps_dd_pushpull(ctypes.c_uint64(node_id), gp, out.ctypes.data_as(_F32P)), 'DDPushPull')
        return out

    def SparsePull(self, node_id, rows, width=None, out: Optional[np.ndarray] = None) -> np.ndarray:
        r, rp = _i64(np.asarray(rows).reshape(-1))
        width = width or self._meta[int(node_id)][2]
        if out is None:
            out = np.zeros((r.size, width), dtype=np.float32)
        _chk(self.lib.hps_sparse_pull(ctypes.c_uint64(node_id), rp, ctypes.c_uint64(r.size),
                                      out.ctypes.data_as(_F32P)), 'SparsePull')
        return out

    def _check_rows(self, node_id, r, g, what):
        # the native client reads rows x width floats from `grads`: a buffer of another size would be read out of bounds
        width = self._meta[int(node_id)][2] if int(node_id) in self._meta else None
        if width and g.size != r.size * width:
            raise ValueError('%s: %d values for %d rows of width %d' % (what, g.size, r.size, width))

    def SparsePush(self, node_id, rows, grads):
        r, rp = _i64(np.asarray(rows).reshape(-1))
        g, gp = _f32(grads)
        self._check_rows(node_id, r, g, 'SparsePush')
        _chk(self.lib.hps_sparse_push(ctypes.c_uint64(node_id), rp, ctypes.c_uint64(r.size), gp), 'SparsePush')

    def SDPushPull(self, node_id, rows, grads, out: Optional[np.ndarray] = None) -> np.ndarray:
        r, rp = _i64(np.asarray(rows).reshape(-1))
        g, gp = _f32(grads)
        self._check_rows(node_id, r, g, 'SDPushPull')
        if out is None:
            out = np.empty(self._numel(node_id), dtype=np.float32)
        _chk(self.lib.hps_sd_pushpull(ctypes.c_uint64(node_id), rp, ctypes.c_uint64(r.size), gp,
                                      out.ctypes.data_as(_F32P)), 'SDPushPull')
        return out

    def SSPushPull(self, node_id, rows, grads, pull_rows, out: Optional[np.ndarray] = None) -> np.ndarray:
        r, rp = _i64(np.asarray(rows).reshape(-1))
        g, gp = _f32(grads)
        self._check_rows(node_id, r, g, 'SSPushPull')
        pr, prp = _i64(np.asarray(pull_rows).reshape(-1))
        width = self._meta[int(node_id)][2]
        if out is None:
            out = np.zeros((pr.size, width), dtype=np.float32)
        _chk(self.lib.hps_ss_pushpull(ctypes.c_uint64(node_id), rp, ctypes.c_uint64(r.size), gp, prp,
                                      ctypes.c_uint64(pr.size), out.ctypes.data_as(_F32P)), 'SSPushPull')
        return out

    # ---- requests of a cache that lives outside this library (parallel/gpu_cache.py) -----------------------------------
    def sync_embedding(self, node_id, rows, versions, bound, width):
        """rows + the versions the caller holds (-1 = not cached) -> (positions, server versions, rows) of the entries
        that are missing or more than ``bound`` versions behind."""
        r, rp = _i64(np.asarray(rows).reshape(-1))
        v, vp = _i64(np.asarray(versions).reshape(-1))
        n = r.size
        out_idx = np.empty(n, dtype=np.int64)
        out_ver = np.empty(n, dtype=np.int64)
        out_rows = np.empty((n, width), dtype=np.float32)
        k = ctypes.c_uint64(0)
        _chk(self.lib.hps_sync_embedding(ctypes.c_uint64(node_id), rp, vp, ctypes.c_uint64(n), ctypes.c_int64(int(bound)),
                                         out_idx.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                         out_ver.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                         out_rows.ctypes.data_as(_F32P), ctypes.byref(k)), 'SyncEmbedding')
        k = int(k.value)
        return out_idx[:k], out_ver[:k], out_rows[:k]

    def push_embedding(self, node_id, rows, grads, updates):
        r, rp = _i64(np.asarray(rows).reshape(-1))
        g, gp = _f32(grads)
        u, up = _i64(np.asarray(updates).reshape(-1))
        self._check_rows(node_id, r, g, 'PushEmbedding')
        _chk(self.lib.hps_push_embedding(ctypes.c_uint64(node_id), rp, gp, up, ctypes.c_uint64(r.size)), 'PushEmbedding')

    # async variants keep their numpy buffers alive until Wait()
    def PushAsync(self, node_id, grad):
        g, gp = _f32(grad)
        self._keep.setdefault(int(node_id), []).append(g)
        _chk(self.lib.hps_push_async(ctypes.c_uint64(node_id), gp), 'PushAsync')

    def PullAsync(self, node_id, out: np.ndarray):
        self._keep.setdefault(int(node_id), []).append(out)
        _chk(self.lib.hps_pull_async(ctypes.c_uint64(node_id), out.ctypes.data_as(_F32P)), 'PullAsync')

    def SparsePullAsync(self, node_id, rows, out: np.ndarray):
        """Queue a row pull on the worker's I/O thread for ``node_id`` (ordered with the other async
        requests of that key); ``out`` is filled once ``Wait(node_id)`` returns."""
        r, rp = _i64(np.asarray(rows).reshape(-1))
        self._keep.setdefault(int(node_id), []).extend([r, out])
        _chk(self.lib.hps_sparse_pull_async(ctypes.c_uint64(node_id), rp, ctypes.c_uint64(r.size),
                                            out.ctypes.data_as(_F32P)), 'SparsePullAsync')

    def Wait(self, node_id):
        _chk(self.lib.hps_wait(ctypes.c_uint64(node_id)), 'Wait')
        self._keep.pop(int(node_id), None)

    WaitData = Wait

    def SetPriority(self, node_id, priority: int):
        """Async requests of higher-priority tensors are served first (P3 van behaviour, ps-lite/src/p3_van.h:12-68);
        requests of one tensor always run in submission order."""
        _chk(self.lib.hps_set_priority(ctypes.c_uint64(node_id), int(priority)), 'SetPriority')

    def pull_full(self, node):
        return self.Pull(node.id).reshape(node.shape)

    def Clear(self, node_id):
        _chk(self.lib.hps_clear(ctypes.c_uint64(node_id)), 'Clear')

    ClearOnServer = Clear

    def SaveParam(self, node_id, path: str):
        _chk(self.lib.hps_save(ctypes.c_uint64(node_id), str(path).encode()), 'SaveParam')

    def LoadParam(self, node_id, path: str):
        _chk(self.lib.hps_load(ctypes.c_uint64(node_id), str(path).encode()), 'LoadParam')

    def BarrierWorker(self):
        _chk(self.lib.hps_barrier(0), 'BarrierWorker')

    def ssp_init(self, rank, group_size, tolerance, key=0):
        _chk(self.lib.hps_ssp_init(ctypes.c_uint64(key), int(group_size), ctypes.c_int64(int(tolerance))), 'ssp_init')

    def ssp_sync(self, rank, version, key=0):
        _chk(self.lib.hps_ssp_sync(ctypes.c_uint64(key), ctypes.c_int64(int(version))), 'ssp_sync')

    def preduce_get_partner(self, key, rank, max_worker, wait_time):
        out = (ctypes.c_int * 256)()
        n = ctypes.c_int(0)
        _chk(self.lib.hps_preduce_get_partner(ctypes.c_uint64(key), int(max_worker), ctypes.c_float(wait_time),
                                              out, ctypes.byref(n)), 'preduce_get_partner')
        return [out[i] for i in range(n.value)]

    def startRecord(self, path=None):
        self._record_path = path
        self.lib.hps_start_record()

    PSF_NAMES = ('DensePull', 'DensePush', 'DDPushPull', 'SparsePull', 'SparsePush', 'SDPushPull', 'SSPushPull', 'ParamInit',
                 'ParamClear', 'ParamSave', 'ParamLoad', 'SyncEmbedding', 'PushEmbedding', 'PushSyncEmbedding', 'SSPInit',
                 'SSPSync', 'PReduceGetPartner')

    def getLoadsByFunction(self):
        """{PS function: (bytes sent, bytes received, requests)} since ``startRecord`` (reference kvworker.h:39-51)."""
        arr = [(ctypes.c_uint64 * 17)() for _ in range(3)]
        self.lib.hps_get_loads_by_function(*arr)
        return {n: (int(arr[0][i]), int(arr[1][i]), int(arr[2][i])) for i, n in enumerate(self.PSF_NAMES) if arr[2][i]}

    def getLoads(self):
        s, r = ctypes.c_uint64(0), ctypes.c_uint64(0)
        self.lib.hps_get_loads(ctypes.byref(s), ctypes.byref(r))
        path = getattr(self, '_record_path', None)
        if path:
            os.makedirs(path, exist_ok=True)
            with open(os.path.join(path, 'loads_%d.txt' % self.rank()), 'w') as fw:
                fw.write('sent %d\nreceived %d\n' % (s.value, r.value))
                for name, (bs, br, nc) in self.getLoadsByFunction().items():
                    fw.write('%s: sent %d received %d requests %d\n' % (name, bs, br, nc))
        return s.value, r.value
