"""Data-parallel classical ML over the parameter server (reference
third_party/HetuML/examples/parallel_ml/*.py).  Start one process per role:

    python examples/hetuml/parallel_ml.py --role scheduler --scheduler 127.0.0.1:13200 -s 1 -w 2 &
    python examples/hetuml/parallel_ml.py --role server    --scheduler 127.0.0.1:13200 -s 1 -w 2 &
    python examples/hetuml/parallel_ml.py --role worker    --scheduler 127.0.0.1:13200 -s 1 -w 2 --model gbdt &
    python examples/hetuml/parallel_ml.py --role worker    --scheduler 127.0.0.1:13200 -s 1 -w 2 --model gbdt
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from hetuml.cluster import Cluster                           # noqa: E402
from hetuml.data import Dataset                              # noqa: E402
from hetuml.ensemble import GBDT                             # noqa: E402
from hetuml.linear import LogisticRegression                 # noqa: E402
from standalone_ml import synthetic                          # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--role', required=True, choices=['scheduler', 'server', 'worker'])
    ap.add_argument('--scheduler', default='127.0.0.1:13200')
    ap.add_argument('-s', '--num-servers', type=int, default=1)
    ap.add_argument('-w', '--num-workers', type=int, default=2)
    ap.add_argument('--model', default='gbdt', choices=['gbdt', 'lr'])
    ap.add_argument('--train')
    args = ap.parse_args()
    cluster = Cluster(args.scheduler, args.num_servers, args.num_workers, args.role)
    if args.role == 'worker':
        if args.train:
            data = Dataset.from_file(args.train, neg_y=args.model == 'lr', rank=cluster.rank, total_ranks=args.num_workers)
        else:
            X, y = synthetic()
            y = np.where(y > 0, 1.0, -1.0).astype(np.float32) if args.model == 'lr' else y
            data = Dataset.from_data((X[cluster.rank::args.num_workers], y[cluster.rank::args.num_workers]))
        model = GBDT(max_depth=6, parallel=True) if args.model == 'gbdt' else LogisticRegression(learning_rate=0.5, parallel=True)
        model.fit(data)
        print('worker %d: %s' % (cluster.rank, model.evaluate(data, 'log-loss,error')), flush=True)
    else:
        cluster.start_server('double')
    cluster.join()


if __name__ == '__main__':
    main()
