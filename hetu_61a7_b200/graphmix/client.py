"""Worker-side client of the graph servers.

reference: third_party/GraphMix/python/graphmix/client.py:17-101 (``Client.meta``, ``rank``,
``num_worker``, ``num_server``, ``barrier``, ``pull_graph(*samplers)`` / ``pull_node(ids)``
returning wait objects, ``wait(obj)``) and the ``Graph`` / ``NodePack`` classes bound from
graphmix/src/graph/python_binding.cc."""
from __future__ import annotations

import ctypes
import json
import os

import numpy as np

from . import _core as C

default_server_port = 27777


class Graph(object):
    """A sampled mini-batch sub-graph.  ``f_feat`` [n, F] float32; ``i_feat`` [n, I+1] int32 (stored int
    features, then the node's global id); ``edge_index`` [2, E] local indices; ``extra`` [n, X] sampler
    specific (GraphSage: is_seed, hop); ``type`` the sampler kind."""

    def __init__(self, type_, f_feat, i_feat, gid, edge_index, extra):
        self.type, self.f_feat, self.i_feat, self.global_id, self.edge_index, self.extra = type_, f_feat, i_feat, gid, edge_index, extra
        self.tag = C.sampler.names.get(type_, str(type_))
        self.format = 'coo'

    @property
    def num_nodes(self):
        return int(self.f_feat.shape[0])

    @property
    def num_edges(self):
        return int(self.edge_index.shape[1])

    def convert2coo(self):
        self.format = 'coo'
        return self

    def convert2csr(self):
        """(indptr, indices) by source node; edge_index stays available."""
        order = np.argsort(self.edge_index[0], kind='stable')
        indptr = np.zeros(self.num_nodes + 1, dtype=np.int64)
        np.add.at(indptr, self.edge_index[0] + 1, 1)
        self.csr = (np.cumsum(indptr), self.edge_index[1][order])
        self.format = 'csr'
        return self

    def gcn_norm(self, use_original_gcn_norm=False):
        """Edge weights of the symmetrically normalised adjacency (with self loops when
        ``use_original_gcn_norm``), as used by the GCN examples."""
        src, dst = self.edge_index
        n = self.num_nodes
        if use_original_gcn_norm:
            loops = np.arange(n, dtype=np.int64)
            src, dst = np.concatenate([src, loops]), np.concatenate([dst, loops])
        deg = np.bincount(dst, minlength=n).astype(np.float32)
        w = 1.0 / np.sqrt(np.maximum(deg[src], 1) * np.maximum(deg[dst], 1))
        return np.stack([src, dst]), w.astype(np.float32)

    def __repr__(self):
        return 'Graph(%s, nodes=%d, edges=%d)' % (self.tag, self.num_nodes, self.num_edges)


class NodePack(object):
    def __init__(self, f_feat, i_feat, ptr, nbr):
        self.f_feat, self.i_feat, self.edge_ptr, self.edges = f_feat, i_feat, ptr, nbr

    def neighbors(self, k):
        return self.edges[self.edge_ptr[k]:self.edge_ptr[k + 1]]


class _WaitObject(object):
    def __init__(self, query, is_graph_query=True, n=0):
        self.query, self.is_graph_query, self.n = query, is_graph_query, n


class Client(object):
    """``Client(port)`` in stand-alone mode (one local server), ``Client()`` under the launcher
    (server address and worker rank come from GRAPHMIX_* variables)."""

    def __init__(self, port=None, host=None):
        nw = os.environ.get('GRAPHMIX_NUM_WORKER', '0')
        self.stand_alone = nw in ('', '0')
        if self.stand_alone:
            if port is None:
                raise RuntimeError("A port should be specified in standalone mode, try 'graphmix.default_server_port'")
            self.host, self.port = host or '127.0.0.1', int(port)
            self._rank, self._nworker = 0, 1
            self.host0, self.port0 = self.host, self.port
        else:
            addrs = [a.rsplit(':', 1) for a in os.environ['GRAPHMIX_SERVERS'].split(',')]
            self._rank = int(os.environ.get('GRAPHMIX_WORKER_RANK', '0'))
            self._nworker = int(nw)
            h, p = addrs[self._rank % len(addrs)]
            self.host, self.port = h, int(p)
            self.host0, self.port0 = addrs[0][0], int(addrs[0][1])
        self._h = C.lib().hgm_client_create(self.host.encode(), self.port)
        if not self._h:
            raise RuntimeError('graph client: ' + C.lib().hgm_last_error().decode())
        n = C.lib().hgm_client_meta(ctypes.c_void_p(self._h), None, 0)
        buf = ctypes.create_string_buffer(n + 1)
        C.lib().hgm_client_meta(ctypes.c_void_p(self._h), buf, n + 1)
        self._meta = json.loads(buf.value.decode())

    @property
    def meta(self):
        return self._meta

    def rank(self):
        if self.stand_alone:
            self._handle_error('rank')
        return self._rank

    def num_server(self):
        return self._meta['num_part']

    def num_worker(self):
        if self.stand_alone:
            self._handle_error('num_worker')
        return self._nworker

    def barrier(self):
        if self.stand_alone:
            self._handle_error('barrier')
        C.check(C.lib().hgm_client_barrier(ctypes.c_void_p(self._h), ctypes.c_int64(self._nworker), self.host0.encode(),
                                           self.port0), 'barrier')

    barrier_all = barrier

    def _handle_error(self, method):
        raise RuntimeError("method Client.'%s' is not supported in standalone client mode" % method)

    def pull_graph(self, *samplers):
        if len(samplers) == 1 and type(samplers[0]) in (list, tuple):
            samplers = samplers[0]
        t = np.ascontiguousarray(list(samplers), dtype=np.int32)
        q = C.lib().hgm_client_pull_graph(ctypes.c_void_p(self._h), t.ctypes.data_as(C.I32P), int(t.size))
        return _WaitObject(q)

    def pull_node(self, node_ids):
        ids = np.ascontiguousarray(node_ids, dtype=np.int64).reshape(-1)
        q = C.lib().hgm_client_pull_node(ctypes.c_void_p(self._h), ids.ctypes.data_as(C.I64P), ctypes.c_int64(ids.size))
        return _WaitObject(q, False, int(ids.size))

    def wait(self, waitobj):
        assert type(waitobj) is _WaitObject
        h = ctypes.c_void_p(self._h)
        if waitobj.is_graph_query:
            dims = (ctypes.c_int64 * 6)()
            C.check(C.lib().hgm_client_resolve_graph(h, ctypes.c_int64(waitobj.query), dims), 'pull_graph')
            type_, n, f_len, i_len, x_len, ne = (int(d) for d in dims)
            f = np.empty((n, f_len), dtype=np.float32)
            i = np.empty((n, i_len), dtype=np.int32)
            gid = np.empty(n, dtype=np.int64)
            e = np.empty((2, ne), dtype=np.int64)
            x = np.empty((n, x_len), dtype=np.int32)
            C.check(C.lib().hgm_client_fetch_graph(h, ctypes.c_int64(waitobj.query), f.ctypes.data_as(C.F32P),
                                                   i.ctypes.data_as(C.I32P), gid.ctypes.data_as(C.I64P),
                                                   e.ctypes.data_as(C.I64P), x.ctypes.data_as(C.I32P)), 'pull_graph')
            return Graph(type_, f, i, gid, e, x)
        dims = (ctypes.c_int64 * 4)()
        C.check(C.lib().hgm_client_resolve_node(h, ctypes.c_int64(waitobj.query), dims), 'pull_node')
        n = max(waitobj.n, 1)
        f = np.empty(int(dims[0]), dtype=np.float32)
        i = np.empty(int(dims[1]), dtype=np.int32)
        ptr = np.empty(int(dims[2]), dtype=np.int64)
        nbr = np.empty(int(dims[3]), dtype=np.int64)
        C.check(C.lib().hgm_client_fetch_node(h, ctypes.c_int64(waitobj.query), f.ctypes.data_as(C.F32P), i.ctypes.data_as(C.I32P),
                                              ptr.ctypes.data_as(C.I64P), nbr.ctypes.data_as(C.I64P)), 'pull_node')
        return NodePack(f.reshape(waitobj.n, -1) if waitobj.n else f.reshape(0, 0), i.reshape(waitobj.n, -1) if waitobj.n else i.reshape(0, 0), ptr, nbr)

    def server_stats(self):
        """Counters of the graph server this client is attached to: remote-node cache lookups / hits / size and the
        number of sub-graphs it has handed out."""
        out = (ctypes.c_int64 * 4)()
        C.check(C.lib().hgm_client_server_stats(ctypes.c_void_p(self._h), out), 'server_stats')
        return {'cache_lookups': out[0], 'cache_hits': out[1], 'cache_size': out[2], 'graphs_served': out[3]}

    def stop_server(self):
        C.lib().hgm_client_stop_server(ctypes.c_void_p(self._h))

    def close(self):
        if self._h:
            C.lib().hgm_client_free(ctypes.c_void_p(self._h))
            self._h = None
