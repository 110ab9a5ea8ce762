#!/usr/bin/env python
"""Mini-batch GCN / GraphSAGE training on sub-graphs sampled by graph servers
(reference examples/gnn/run_dist.py + third_party/GraphMix: ``graphmix.launcher(worker, args,
server_init)`` with a YAML config; workers pull GraphSage / RandomWalk batches asynchronously and
train on the induced sub-graph while the next one is in flight).

    python -m hetu_61a7_b200.graphmix.partition -d Cora -n 2 -p /tmp/gm --scale 0.5
    python examples/gnn/run_dist.py --data /tmp/gm/Cora --workers 2 --gpu -1
"""
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
    ctx = ht.cpu(0) if args.gpu < 0 else ht.gpu(rank % max(1, args.gpus_per_node))
    (feat, y_, mp, mask), (loss, y, train) = gnn.dense_model(meta['float_feature'], args.hidden, meta['class'], lr=args.lr,
                                                             layer=gnn.GCN if args.model == 'gcn' else gnn.SageConv)
    ex = ht.Executor([loss, y, train], ctx=ctx, comm_mode='AllReduce' if (args.allreduce and comm.num_worker() > 1) else None)
    eye = np.eye(meta['class'], dtype=np.float32)
    kind = graphmix.sampler.GraphSage if args.sampler == 'sage' else graphmix.sampler.RandomWalk
    nxt = comm.pull_graph(kind)
    for step in range(args.steps):
        g = comm.wait(nxt)
        nxt = comm.pull_graph(kind)                                   # overlap sampling with this step
        (src, dst), w = g.gcn_norm(use_original_gcn_norm=True)
        adj = ht.sparse_array(w, (dst, src), shape=(g.num_nodes, g.num_nodes), ctx=ctx)
        label = g.i_feat[:, 0]
        train_mask = g.i_feat[:, 1].astype(np.float32)
        if g.extra.shape[1]:
            train_mask = train_mask * g.extra[:, 0]                   # GraphSage: the loss is taken on the seeds
        l, p, _ = ex.run(feed_dict={feat: g.f_feat, y_: eye[label], mp: adj, mask: train_mask}, convert_to_numpy_ret_vals=True)
        if step % args.log_every == 0 or step == args.steps - 1:
            test = 1 - g.i_feat[:, 1]
            acc = float(((p.argmax(1) == label) * test).sum() / max(test.sum(), 1))
            print('worker %d step %d nodes %d edges %d loss %.4f held-out acc %.3f' % (rank, step, g.num_nodes, g.num_edges,
                                                                                    float(l), acc), flush=True)
    comm.wait(nxt)
    comm.barrier()


def server_init(server):
    server.init_cache(0.2, graphmix.cache.LFUOpt)
    server.add_sampler(graphmix.sampler.GraphSage, batch_size=256, depth=2, width=5, thread=2)
    server.add_sampler(graphmix.sampler.RandomWalk, rw_head=256, rw_length=2)
    server.is_ready()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--config', help='graphmix launch YAML; built from --data / --workers when omitted')
    ap.add_argument('--data', help='partitioned graph directory (holds meta.yml)')
    ap.add_argument('--workers', type=int, default=2)
    ap.add_argument('--port', type=int, default=graphmix.default_server_port)
    ap.add_argument('--model', default='gcn', choices=['gcn', 'sage'])
    ap.add_argument('--sampler', default='sage', choices=['sage', 'rw'])
    ap.add_argument('--hidden', type=int, default=64)
    ap.add_argument('--lr', type=float, default=0.3)
    ap.add_argument('--steps', type=int, default=60)
    ap.add_argument('--log-every', type=int, default=10)
    ap.add_argument('--gpu', type=int, default=0, help='-1 = CPU executor')
    ap.add_argument('--gpus-per-node', type=int, default=8)
    ap.add_argument('--allreduce', action='store_true', help='average gradients over workers (needs heturun / torchrun env)')
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
