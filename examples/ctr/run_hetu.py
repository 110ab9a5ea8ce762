#!/usr/bin/env python
"""CTR training: Wide&Deep / DeepFM / DCN / Deep&Cross on Criteo (reference examples/ctr/run_hetu.py).

    python examples/ctr/run_hetu.py --model wdl_criteo --batch-size 128                       # one GPU
    heturun -w 8 python examples/ctr/run_hetu.py --model wdl_criteo --comm Hybrid             # NVLink-sharded table + fused dense all-reduce
    heturun -s 1 -w 4 python examples/ctr/run_hetu.py --model wdl_criteo --comm PS --cache lfuopt --bound 10

Criteo is not shipped with the image: synthetic batches with the Criteo shape (13 dense + 26
sparse fields, zipf ids over a 33 762 577-row vocabulary) are generated when the npy files are absent.
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
import hetu as ht                                              # noqa: E402
from hetu_61a7_b200.models import ctr                          # noqa: E402


def run_adult(args, ctx, rank, n):
    """Wide&Deep on Adult (reference models/wdl_adult.py): real files through --data-dir (adult.data / adult.test),
    Adult-shaped random data otherwise."""
    from hetu_61a7_b200.utils import ctr_data
    if args.data_dir:
        deep, wide, y = ctr_data.load_adult_data(args.data_dir, return_val=False)
    else:
        deep, wide, y = ctr_data.adult_synthetic(n, seed=rank)
    n = len(y) // args.batch_size * args.batch_size
    loader = lambda a: ht.dataloader_op([ht.Dataloader(np.asarray(a[:n], np.float32), args.batch_size, 'train')])   # noqa: E731
    X_deep, X_wide, y_ = [loader(a) for a in deep], loader(wide), loader(y)
    loss, pred, label, train_op = ctr.wdl_adult(X_deep, X_wide, y_, dim_wide=wide.shape[1])
    ex = ht.Executor({'train': [loss, pred, label, train_op]}, ctx=ctx, comm_mode=args.comm, bsp=args.bsp, shard_dataloaders=False)
    for ep in range(args.nepoch):
        t0, losses, acc = time.time(), [], []
        for _ in range(ex.get_batch_num('train')):
            l, p, lab, _ = ex.run('train', convert_to_numpy_ret_vals=True)
            losses.append(float(np.mean(l)))
            acc.extend(np.argmax(p, 1) == np.argmax(lab, 1))
        print('[rank %d] epoch %d loss %.4f acc %.4f  %.0f samples/s' % (rank, ep, np.mean(losses), np.mean(acc), n / (time.time() - t0)),
              flush=True)
    if args.comm in ('PS', 'Hybrid') and os.environ.get('DMLC_PS_ROOT_URI'):
        ht.worker_finish()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='wdl_criteo', choices=['wdl_criteo', 'dfm_criteo', 'dcn_criteo', 'dc_criteo', 'wdl_adult'])
    ap.add_argument('--batch-size', type=int, default=128)
    ap.add_argument('--vocab', type=int, default=33762577)
    ap.add_argument('--comm', default=None, choices=[None, 'PS', 'Hybrid', 'AllReduce'])
    ap.add_argument('--cache', default=None, choices=[None, 'lru', 'lfu', 'lfuopt'])
    ap.add_argument('--bound', type=int, default=100)
    ap.add_argument('--bsp', type=int, default=-1, help='-1 ASP, 0 BSP, >0 SSP staleness')
    ap.add_argument('--nepoch', type=int, default=1)
    ap.add_argument('--iters', type=int, default=200, help='synthetic batches per epoch')
    ap.add_argument('--all', action='store_true', help='report AUC')
    ap.add_argument('--cpu', action='store_true')
    ap.add_argument('--data-dir', default=None, help='directory with Criteo train.csv / train.txt (or the processed .npy files); '
                    'synthetic Criteo-shaped data when omitted')
    ap.add_argument('--head', type=int, default=0, help='with --data-dir: use only the first N rows')
    args = ap.parse_args()

    comm = ht.wrapped_mpi_nccl_init() if args.comm in ('Hybrid', 'AllReduce') else None
    rank = comm.rank if comm is not None else int(os.environ.get('DMLC_WORKER_ID', os.environ.get('RANK', 0)))
    ctx = ht.cpu(0) if args.cpu else ht.gpu(int(os.environ.get('LOCAL_RANK', 0)))
    n = args.iters * args.batch_size
    if args.model == 'wdl_adult':
        return run_adult(args, ctx, rank, n)
    if args.data_dir:
        from hetu_61a7_b200.utils import ctr_data
        if args.head:
            dense_v, sparse_v, label_v = ctr_data.process_head_criteo_data(args.data_dir, nrows=args.head, return_val=False)
        else:
            dense_v, sparse_v, label_v = ctr_data.process_all_criteo_data(args.data_dir, return_val=False)
        nw = int(os.environ.get('WORLD_SIZE', os.environ.get('DMLC_NUM_WORKER', 1)))
        dense_v, sparse_v, label_v = dense_v[rank::nw], sparse_v[rank::nw].astype(np.int64), label_v[rank::nw].astype(np.float32)
        args.vocab = int(sparse_v.max()) + 1 if args.head else args.vocab
        n = len(label_v) // args.batch_size * args.batch_size
    else:
        dense_v, sparse_v, label_v = ht.data.criteo_synthetic(n, vocab=args.vocab, seed=rank)
    dense = ht.dataloader_op([ht.Dataloader(dense_v, args.batch_size, 'train')])
    sparse = ht.dataloader_op([ht.Dataloader(sparse_v, args.batch_size, 'train', dtype=np.int64)])
    y_ = ht.dataloader_op([ht.Dataloader(label_v, args.batch_size, 'train')])
    loss, pred, label, train_op = getattr(ctr, args.model)(dense, sparse, y_, feature_dimension=args.vocab, ctx=ctx)
    kw = dict(ctx=ctx, comm_mode=args.comm, bsp=args.bsp, shard_dataloaders=False)
    if args.cache:
        kw.update(cstable_policy=args.cache.upper() if args.cache != 'lfuopt' else 'LFUOpt', cache_bound=args.bound)
    ex = ht.Executor({'train': [loss, pred, label, train_op]}, **kw)
    for ep in range(args.nepoch):
        t0, losses, preds, labels = time.time(), [], [], []
        for _ in range(ex.get_batch_num('train')):
            l, p, y, _ = ex.run('train', convert_to_numpy_ret_vals=True)
            losses.append(float(np.mean(l)))
            if args.all:
                preds.append(p.reshape(-1))
                labels.append(y.reshape(-1))
        dt = time.time() - t0
        msg = '[rank %d] epoch %d loss %.4f  %.0f samples/s' % (rank, ep, np.mean(losses), n / dt)
        if args.all:
            msg += '  auc %.4f' % ht.metrics.auc(np.concatenate(labels), np.concatenate(preds)) \
                if hasattr(ht.metrics, 'auc') else ''
        print(msg, flush=True)
    if args.comm in ('PS', 'Hybrid') and os.environ.get('DMLC_PS_ROOT_URI'):
        ht.worker_finish()


if __name__ == '__main__':
    main()
