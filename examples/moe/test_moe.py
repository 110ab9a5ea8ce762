#!/usr/bin/env python
"""MoE layer training with every gate the reference ships (examples/moe/test_moe_{top,hash,ktop1,sam,base}.py).

    python examples/moe/test_moe.py --gate top --top 2
    heturun -w 8 python examples/moe/test_moe.py --gate top --top 2        # experts sharded over 8 GPUs,
                                                                           # dispatch/combine fused with the NVLink all-to-all
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models import moe                          # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gate', default='top', choices=['top', 'hash', 'ktop1', 'sam', 'base'])
    ap.add_argument('--top', type=int, default=2)
    ap.add_argument('--batch-size', type=int, default=16)
    ap.add_argument('--num-tokens', type=int, default=1024)
    ap.add_argument('--model-dim', type=int, default=2048)
    ap.add_argument('--hidden-size', type=int, default=2048)
    ap.add_argument('--num-local-experts', type=int, default=2)
    ap.add_argument('--steps', type=int, default=20)
    args = ap.parse_args()
    world = int(os.environ.get('WORLD_SIZE', 1))
    comm = ht.wrapped_mpi_nccl_init()
    rank = comm.rank
    ctx = ht.gpu(comm.local_rank)
    x, t = ht.Variable(name='x'), ht.Variable(name='t')
    indice = ht.Variable(name='hash_indice') if args.gate == 'hash' else None
    loss, _ = moe.moe_model(x, t, batch_size=args.batch_size, num_tokens=args.num_tokens, model_dim=args.model_dim,
                            hidden_size=args.hidden_size, num_local_experts=args.num_local_experts, world=world,
                            rank=rank, top=args.top, gate=args.gate, hash_indices=indice)
    train = ht.optim.SGDOptimizer(learning_rate=0.125 / 1000).minimize(loss)
    ex = ht.Executor({'train': [loss, train]}, ctx=ctx, comm_mode='AllReduce' if world > 1 else None)
    rs = np.random.RandomState(rank)
    xv = ht.array(rs.randn(args.batch_size, args.num_tokens, args.model_dim).astype(np.float32), ctx)
    tv = ht.array(np.zeros((args.batch_size,), dtype=np.float32), ctx)
    feed = {x: xv, t: tv}
    if indice is not None:
        feed[indice] = rs.randint(0, args.num_local_experts * world,
                                  (args.batch_size * args.num_tokens,)).astype(np.float32)
    for step in range(args.steps):
        t0 = time.time()
        out = ex.run('train', feed_dict=feed)
        l = float(np.asarray(out[0].asnumpy()).reshape(-1)[0])
        if rank == 0:
            print('step %d loss %.5f  %.2f ms' % (step, l, (time.time() - t0) * 1e3), flush=True)


if __name__ == '__main__':
    main()
