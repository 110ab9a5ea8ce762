#!/usr/bin/env python
"""Featureless-graph GNN with a trainable node-embedding table, sub-graphs from graph servers
(reference examples/gnn/run_dist_hybrid.py: GraphMix sampling + Hetu Hybrid mode - the embedding
table goes through the parameter server / HET cache, the GNN weights are all-reduced).

    python -m hetu_61a7_b200.graphmix.partition -d Cora -n 2 -p /tmp/gm --scale 0.5
    python examples/gnn/run_dist_hybrid.py --data /tmp/gm/Cora --workers 1 --gpu -1            # local
    heturun -s 1 -w 2 python examples/gnn/run_dist_hybrid.py --data /tmp/gm/Cora --hybrid      # PS + all-reduce

Under ``heturun`` every launched rank is one worker: rank 0 also starts the graph servers."""
import argparse
import os
import sys
import tempfile

import numpy as np
import yaml

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import graphmix                                                # noqa: E402
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models import gnn                          # noqa: E402


def worker(args):
    comm = graphmix.Client()
    meta = comm.meta
    rank = comm.rank()
    ctx = ht.cpu(0) if args.gpu < 0 else ht.gpu(int(os.environ.get('LOCAL_RANK', rank)) % max(1, args.gpus_per_node))
    (node_ids, y_, mp, mask), (loss, y, train) = gnn.sparse_model(meta['node'], args.embed_dim, args.hidden, meta['class'], lr=args.lr)
    kw = {}
    if args.hybrid:
        kw = dict(comm_mode='Hybrid', cstable_policy=args.cache, cache_bound=args.bound, bsp=0)
    ex = ht.Executor([loss, y, train], ctx=ctx, **kw)
    eye = np.eye(meta['class'], dtype=np.float32)
    nxt = comm.pull_graph(graphmix.sampler.GraphSage)
    for step in range(args.steps):
        g = comm.wait(nxt)
        nxt = comm.pull_graph(graphmix.sampler.GraphSage)
        (src, dst), w = g.gcn_norm(use_original_gcn_norm=True)
        adj = ht.sparse_array(w, (dst, src), shape=(g.num_nodes, g.num_nodes), ctx=ctx)
        label = g.i_feat[:, 0]
        train_mask = (g.i_feat[:, 1] * g.extra[:, 0]).astype(np.float32)
        l, p, _ = ex.run(feed_dict={node_ids: g.global_id.astype(np.float32), y_: eye[label], mp: adj, mask: train_mask},
                         convert_to_numpy_ret_vals=True)
        if step % args.log_every == 0 or step == args.steps - 1:
            held = 1 - g.i_feat[:, 1]
            acc = float(((p.argmax(1) == label) * held).sum() / max(held.sum(), 1))
            print('worker %d step %d nodes %d loss %.4f held-out acc %.3f' % (rank, step, g.num_nodes, float(l), acc), flush=True)
    comm.wait(nxt)
    comm.barrier()


def server_init(server):
    server.init_cache(0.2, graphmix.cache.LFUOpt)
    server.add_sampler(graphmix.sampler.GraphSage, batch_size=256, depth=2, width=5, thread=2)
    server.is_ready()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--config')
    ap.add_argument('--data')
    ap.add_argument('--workers', type=int, default=1)
    ap.add_argument('--port', type=int, default=graphmix.default_server_port + 100)
    ap.add_argument('--embed-dim', type=int, default=64)
    ap.add_argument('--hidden', type=int, default=64)
    ap.add_argument('--lr', type=float, default=0.3)
    ap.add_argument('--steps', type=int, default=60)
    ap.add_argument('--log-every', type=int, default=10)
    ap.add_argument('--gpu', type=int, default=0)
    ap.add_argument('--gpus-per-node', type=int, default=8)
    ap.add_argument('--hybrid', action='store_true', help='embedding table on the PS (needs heturun -s N)')
    ap.add_argument('--cache', default='LFUOpt')
    ap.add_argument('--bound', type=int, default=3)
    args = ap.parse_args()
    if not args.config:
        if not args.data:
            ap.error('give --config or --data')
        nparts = yaml.safe_load(open(os.path.join(os.path.expanduser(args.data), 'meta.yml')))['num_part']
        cfg = {'env': {'GRAPHMIX_ROOT_URI': '127.0.0.1', 'GRAPHMIX_ROOT_PORT': args.port, 'GRAPHMIX_NUM_WORKER': args.workers},
               'launch': {'worker': args.workers, 'server': nparts, 'scheduler': True, 'data': args.data}}
        fd, args.config = tempfile.mkstemp(suffix='.yml')
        with os.fdopen(fd, 'w') as f:
            yaml.safe_dump(cfg, f)
    graphmix.launcher(worker, args, server_init=server_init)


if __name__ == '__main__':
    main()
