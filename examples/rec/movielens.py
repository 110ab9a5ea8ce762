"""reference examples/rec/movielens.py: ``getdata(dataset, data_dir)`` -> NCF training triples with sampled
negatives and leave-one-out test rows.  No downloader here: put ``ratings.csv`` under ``data_dir/<dataset>``."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
from hetu_61a7_b200.utils.ctr_data import movielens_getdata   # noqa: E402


def getdata(dataset='ml-25m', data_dir='datasets', num_negatives=4):
    return movielens_getdata(os.path.join(data_dir, dataset), num_negatives)
