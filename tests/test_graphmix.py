"""Graph partitioning, graph servers and samplers (pattern: reference
third_party/GraphMix/tests/{test_standalone,test_samplers,test_pull_node,test_remote_sampling}.py -
every sampled sub-graph is checked against the full dataset: features and labels belong to the
node, every sampled edge exists in the original graph)."""
import os
import socket

import numpy as np
import pytest
import yaml

import graphmix
from graphmix.dataset import synthetic_dataset


def _free_port(span=8):
    from conftest import free_port_block
    return free_port_block(span)


def _check_graph(ds, edge_set, g):
    orig = g.i_feat[:, -2]                                   # original node id rides along as an int feature
    np.testing.assert_array_equal(g.f_feat, ds.x[orig])
    np.testing.assert_array_equal(g.i_feat[:, 0], ds.y[orig])
    assert len(set(g.global_id.tolist())) == g.num_nodes
    np.testing.assert_array_equal(g.i_feat[:, -1], g.global_id)
    for u, v in zip(g.edge_index[0], g.edge_index[1]):
        assert (int(orig[u]), int(orig[v])) in edge_set
    # induced: every original edge between two sampled nodes is present
    pos = {int(o): k for k, o in enumerate(orig)}
    have = set(zip(g.edge_index[0].tolist(), g.edge_index[1].tolist()))
    sampled = set(pos)
    expect = sum(1 for (a, b) in edge_set if a in sampled and b in sampled)
    assert expect == len(have)


def test_partition_is_balanced_and_beats_random(tmp_path):
    ds = synthetic_dataset('toy', 1200, 8, 4, avg_degree=8, p_in=0.95)
    meta = graphmix.part_graph(ds, 4, str(tmp_path))
    assert meta['num_part'] == 4 and sum(meta['partition']['nodes']) == 1200
    assert max(meta['partition']['nodes']) <= 1.1 * 300 + 2
    assert sum(meta['partition']['edges']) == meta['edge']
    rnd = np.random.RandomState(12345).randint(0, 4, 1200)
    e = ds.graph.edge_index
    random_cut = int((rnd[e[0]] != rnd[e[1]]).sum())
    assert meta['cut_edge'] < 0.6 * random_cut, (meta['cut_edge'], random_cut)
    # shards: contiguous ownership, edges stored with their source's owner
    sh = graphmix.Shard(os.path.join(str(tmp_path), 'toy')).load_graph_shard(2)
    lo, hi = meta['partition']['offset'][2], meta['partition']['offset'][3]
    assert sh.edges[0].min() >= lo and sh.edges[0].max() < hi and sh.f_feat.shape == (hi - lo, 8)


def test_standalone_server_all_samplers(tmp_path):
    ds = synthetic_dataset('solo', 600, 6, 3, avg_degree=6)
    graphmix.part_graph(ds, 1, str(tmp_path))
    edge_set = set(zip(ds.graph.edge_index[0].tolist(), ds.graph.edge_index[1].tolist()))
    port = _free_port()

    def server_init(server):
        server.add_sampler(graphmix.sampler.GlobalNode, batch_size=64)
        server.add_sampler(graphmix.sampler.LocalNode, batch_size=48)
        server.add_sampler(graphmix.sampler.RandomWalk, rw_head=20, rw_length=3)
        server.add_sampler(graphmix.sampler.GraphSage, batch_size=10, depth=2, width=3)
        server.is_ready()
    server = graphmix.start_standalone_server(os.path.join(str(tmp_path), 'solo'), port, server_init)
    try:
        os.environ.pop('GRAPHMIX_NUM_WORKER', None)
        comm = graphmix.Client(port)
        assert comm.meta['node'] == 600 and comm.num_server() == 1
        with pytest.raises(RuntimeError):
            comm.rank()
        seen = set()
        for kind in (graphmix.sampler.GlobalNode, graphmix.sampler.LocalNode, graphmix.sampler.RandomWalk,
                     graphmix.sampler.GraphSage) * 2:
            g = comm.wait(comm.pull_graph(kind))
            g.convert2coo()
            assert g.type == kind
            seen.add(g.tag)
            _check_graph(ds, edge_set, g)
            if kind == graphmix.sampler.GlobalNode:
                assert g.num_nodes == 64
            if kind == graphmix.sampler.LocalNode:
                assert g.num_nodes == 48
            if kind == graphmix.sampler.RandomWalk:
                assert 20 <= g.num_nodes <= 20 * 4
            if kind == graphmix.sampler.GraphSage:
                assert g.extra.shape == (g.num_nodes, 2) and g.extra[:, 0].sum() == 10
                assert g.num_nodes <= 10 * (1 + 3 + 9) and g.extra[:, 1].max() <= 2
                # every non-seed node hangs off a node one hop closer
                src, dst = g.edge_index
                for v in np.nonzero(g.extra[:, 0] == 0)[0]:
                    assert np.any(g.extra[src[dst == v], 1] == g.extra[v, 1] - 1)
        assert seen == {'GlobalNode', 'LocalNode', 'RandomWalk', 'GraphSage'}
        # several queries in flight, any sampler
        qs = [comm.pull_graph() for _ in range(6)]
        assert all(comm.wait(q).num_nodes > 0 for q in qs)
        pack = comm.wait(comm.pull_node([5, 17, 599]))
        np.testing.assert_array_equal(pack.f_feat, ds.x[pack.i_feat[:, -1]])
        gnorm_e, gnorm_w = g.gcn_norm(True)
        assert gnorm_e.shape[1] == g.num_edges + g.num_nodes and np.all(gnorm_w > 0)
        comm.close()
    finally:
        server.stop()
        server.close()


def _worker(args):
    ds = synthetic_dataset('dist', 900, 5, 3, avg_degree=6)
    edge_set = set(zip(ds.graph.edge_index[0].tolist(), ds.graph.edge_index[1].tolist()))
    comm = graphmix.Client()
    rank, nrank = comm.rank(), comm.num_worker()
    assert nrank == 2 and comm.num_server() == 3
    comm.barrier()
    off = comm.meta['partition']['offset']
    for it in range(12):
        kind = [graphmix.sampler.GraphSage, graphmix.sampler.GlobalNode, graphmix.sampler.RandomWalk][it % 3]
        g = comm.wait(comm.pull_graph(kind))
        _check_graph(ds, edge_set, g)
        if kind == graphmix.sampler.GlobalNode:                          # spans several servers' ranges
            owners = np.searchsorted(off, g.global_id, side='right') - 1
            assert len(set(owners.tolist())) > 1
    ids = np.array([0, off[1], off[2], off[3] - 1])                       # one node of every server, through one server
    pack = comm.wait(comm.pull_node(ids))
    np.testing.assert_array_equal(pack.f_feat, ds.x[pack.i_feat[:, -1]])
    for k in range(len(ids)):
        orig_u = int(pack.i_feat[k, -1])
        assert len(pack.neighbors(k)) == sum(1 for (a, b) in edge_set if a == orig_u)
    st = comm.server_stats()
    assert st['graphs_served'] >= 1 and st['cache_lookups'] > 0 and 0 < st['cache_hits'] <= st['cache_lookups'], st
    assert 0 < st['cache_size'] <= 0.3 * 900 + 1
    comm.barrier()
    with open(os.path.join(args.out, 'ok_%d' % rank), 'w') as f:
        f.write('ok')


def _server_init(server):
    server.init_cache(0.3, [graphmix.cache.LFUOpt, graphmix.cache.LFU, graphmix.cache.LRU][server.rank() % 3])
    server.add_sampler(graphmix.sampler.GlobalNode, batch_size=96)
    server.add_sampler(graphmix.sampler.RandomWalk, rw_head=32, rw_length=2)
    server.add_sampler(graphmix.sampler.GraphSage, batch_size=16, depth=2, width=2, thread=2)
    server.is_ready()


def test_three_servers_two_workers_remote_sampling(tmp_path):
    ds = synthetic_dataset('dist', 900, 5, 3, avg_degree=6)
    graphmix.part_graph(ds, 3, str(tmp_path))
    port = _free_port()
    cfg = {'env': {'GRAPHMIX_ROOT_URI': '127.0.0.1', 'GRAPHMIX_ROOT_PORT': port, 'GRAPHMIX_NUM_WORKER': 2},
           'launch': {'worker': 2, 'server': 3, 'scheduler': True, 'data': os.path.join(str(tmp_path), 'dist')}}
    cfg_path = os.path.join(str(tmp_path), 'config.yml')
    with open(cfg_path, 'w') as f:
        yaml.safe_dump(cfg, f)

    class Args(object):
        config = cfg_path
        out = str(tmp_path)
    graphmix.launcher(_worker, Args(), server_init=_server_init)
    assert os.path.exists(os.path.join(str(tmp_path), 'ok_0')) and os.path.exists(os.path.join(str(tmp_path), 'ok_1'))
