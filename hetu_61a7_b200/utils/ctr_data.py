"""CTR / recommendation dataset preparation from local files.

reference: examples/ctr/models/load_data.py (Criteo: log-transformed dense features, label-encoded
sparse features with one global id space; Adult: wide one-hot + cross columns and deep
categorical ids; Avazu) and examples/rec/movielens.py (implicit-feedback pairs with sampled
negatives, leave-one-out test items).  The downloaders are left out (no network): functions take
the directory that already holds the raw files and cache the processed arrays next to them as
``.npy``, in the reference's file names."""
from __future__ import annotations

import os

import numpy as np

CRITEO_DENSE = ['I%d' % i for i in range(1, 14)]
CRITEO_SPARSE = ['C%d' % i for i in range(1, 27)]


# ----------------------------------------------------------------------------------- criteo
def process_dense_feats(data, feats):
    """log(x + 1) for x > -1, else -1; missing -> 0."""
    d = data[feats].fillna(0.0).to_numpy(dtype=np.float64)
    out = np.where(d > -1, np.log(np.maximum(d, -0.999999) + 1), -1.0)
    return out.astype(np.float32)


def process_sparse_feats(data, feats):
    """Label-encode every column (sorted unique values, missing = "-1") and shift the columns into
    one id space: column k starts where column k-1 ended."""
    cols, offset = [], 0
    for f in feats:
        vals = data[f].fillna('-1').astype(str).to_numpy()
        uniq, inv = np.unique(vals, return_inverse=True)
        cols.append(inv.astype(np.int64) + offset)
        offset += len(uniq)
    return np.stack(cols, axis=1).astype(np.int32)


def _split(dense, sparse, labels, return_val, test_num):
    if not return_val:
        return dense, sparse, labels
    return ((dense[:-test_num], dense[-test_num:]), (sparse[:-test_num], sparse[-test_num:]),
            (labels[:-test_num], labels[-test_num:]))


def read_criteo(path, nrows=None):
    """``train.csv`` with a header (label, I1..I13, C1..C26) or the raw tab-separated ``train.txt``."""
    import pandas as pd
    csv_path, txt_path = os.path.join(path, 'train.csv'), os.path.join(path, 'train.txt')
    if os.path.exists(csv_path):
        return pd.read_csv(csv_path, nrows=nrows, header=0)
    if os.path.exists(txt_path):
        return pd.read_csv(txt_path, nrows=nrows, header=None, sep='\t', names=['label'] + CRITEO_DENSE + CRITEO_SPARSE)
    raise FileNotFoundError('no train.csv / train.txt under %s (downloading is not possible here)' % path)


def process_head_criteo_data(path, nrows=20000, return_val=True):
    df = read_criteo(path, nrows)
    labels = df['label'].to_numpy().reshape(-1, 1).astype(np.float32)
    dense = process_dense_feats(df, [c for c in df.columns if c.startswith('I')])
    sparse = process_sparse_feats(df, [c for c in df.columns if c.startswith('C')])
    return _split(dense, sparse, labels, return_val, max(len(df) // 10, 1))


def process_sampled_criteo_data(path):
    return tuple(np.load(os.path.join(path, f)) for f in
                 ('sampled_dense_feats.npy', 'sampled_sparse_feats.npy', 'sampled_labels.npy'))


def process_all_criteo_data(path, return_val=True):
    names = ['train_dense_feats.npy', 'test_dense_feats.npy', 'train_sparse_feats.npy', 'test_sparse_feats.npy',
             'train_labels.npy', 'test_labels.npy']
    paths = [os.path.join(path, n) for n in names]
    if not all(os.path.exists(p) for p in paths):
        (td, vd), (ts, vs), (tl, vl) = process_head_criteo_data(path, nrows=None, return_val=True)
        for p, a in zip(paths, (td, vd, ts, vs, tl, vl)):
            np.save(p, a)
    files = [np.load(p) for p in paths]
    if return_val:
        return (files[0], files[1]), (files[2], files[3]), (files[4], files[5])
    return files[0], files[2], files[4]


# ------------------------------------------------------------------------------------ adult
ADULT_COLUMNS = ['age', 'workclass', 'fnlwgt', 'education', 'education_num', 'marital_status', 'occupation', 'relationship',
                 'race', 'gender', 'capital_gain', 'capital_loss', 'hours_per_week', 'native_country', 'income_bracket']
ADULT_CATEGORICAL = ['workclass', 'education', 'marital_status', 'occupation', 'relationship', 'race', 'gender', 'native_country']
ADULT_CONTINUOUS = ['age', 'education_num', 'capital_gain', 'capital_loss', 'hours_per_week'][:4]


def cross_columns(x_cols):
    """Names of crossed columns: ['a', 'b'] -> 'a_b' (reference load_data.py:171-178)."""
    return {'_'.join(c): list(c) for c in x_cols}


def val2idx(df, cols):
    """Per-column value -> dense index; returns the encoded frame and the vocabulary sizes."""
    sizes = {}
    out = df.copy()
    for c in cols:
        uniq, inv = np.unique(df[c].astype(str).to_numpy(), return_inverse=True)
        out[c] = inv
        sizes[c] = len(uniq)
    return out, sizes


def load_adult_data(path, return_val=True):
    """``adult.data`` / ``adult.test`` -> (deep inputs: 8 categorical id columns + 4 continuous, wide
    one-hot matrix incl. 2 crossed columns, one-hot labels).  Layout matches ``models.ctr.wdl_adult``."""
    import pandas as pd
    tr = pd.read_csv(os.path.join(path, 'adult.data'), names=ADULT_COLUMNS, skipinitialspace=True)
    te = pd.read_csv(os.path.join(path, 'adult.test'), names=ADULT_COLUMNS, skipinitialspace=True, skiprows=1)
    df = pd.concat([tr, te], ignore_index=True)
    y = df['income_bracket'].astype(str).str.contains('>50K').to_numpy().astype(np.int64)
    for name, cols in cross_columns([['education', 'occupation'], ['native_country', 'occupation']]).items():
        df[name] = df[cols[0]].astype(str) + '_' + df[cols[1]].astype(str)
    wide_cols = ADULT_CATEGORICAL + ['education_occupation', 'native_country_occupation']
    enc, sizes = val2idx(df, wide_cols)
    blocks, off = [], 0
    wide = np.zeros((len(df), sum(sizes[c] for c in wide_cols)), dtype=np.float32)
    for c in wide_cols:
        wide[np.arange(len(df)), off + enc[c].to_numpy()] = 1.0
        off += sizes[c]
    deep_cat = [np.minimum(enc[c].to_numpy(), 49).astype(np.float32) for c in ADULT_CATEGORICAL]
    cont = df[ADULT_CONTINUOUS].to_numpy(dtype=np.float32)
    cont = (cont - cont.mean(0)) / (cont.std(0) + 1e-6)
    deep = deep_cat + [cont[:, i] for i in range(cont.shape[1])]
    labels = np.eye(2, dtype=np.float32)[y]
    n_tr = len(tr)
    if not return_val:
        return [d[:n_tr] for d in deep], wide[:n_tr], labels[:n_tr]
    return ([d[:n_tr] for d in deep], [d[n_tr:] for d in deep]), (wide[:n_tr], wide[n_tr:]), (labels[:n_tr], labels[n_tr:])


def adult_synthetic(n, seed=0):
    """Adult-shaped random data (8 categorical + 4 continuous deep inputs, 809 wide columns, one-hot labels) for
    runs without the UCI files."""
    rs = np.random.RandomState(seed)
    cat = rs.randint(0, 50, (n, 8))
    cont = rs.randn(n, 4).astype(np.float32)
    wide = (rs.rand(n, 809) < 0.02).astype(np.float32)
    score = (cat[:, 0] % 2) + cont[:, 0] + wide[:, :20].sum(1) - 1.0
    y = np.eye(2, dtype=np.float32)[(score > 0).astype(np.int64)]
    deep = [cat[:, i].astype(np.float32) for i in range(8)] + [cont[:, i] for i in range(4)]
    return deep, wide, y


# ------------------------------------------------------------------------------------ avazu
def process_avazu(path, nrows=None):
    """Avazu click logs (``train.csv``): every column but id / click is categorical."""
    import pandas as pd
    df = pd.read_csv(os.path.join(path, 'train.csv'), nrows=nrows)
    feats = [c for c in df.columns if c not in ('id', 'click')]
    labels = df['click'].to_numpy().reshape(-1, 1).astype(np.float32)
    sparse = process_sparse_feats(df, feats)
    n_test = max(len(df) // 10, 1)
    return (sparse[:-n_test], sparse[-n_test:]), (labels[:-n_test], labels[-n_test:])


# -------------------------------------------------------------------------------- movielens
def movielens_getdata(data_dir, num_negatives=4, seed=0):
    """``ratings.csv`` (userId, movieId, rating, timestamp) -> implicit-feedback training triples with
    ``num_negatives`` sampled negatives per positive, and leave-one-out test rows (each user's
    latest item + 99 negatives), as in the NCF recipe (reference examples/rec/movielens.py)."""
    import pandas as pd
    cache = os.path.join(data_dir, 'ncf_train_%d.npz' % num_negatives)
    if os.path.exists(cache):
        z = np.load(cache)
        return {k: z[k] for k in z.files}
    df = pd.read_csv(os.path.join(data_dir, 'ratings.csv'))
    df.columns = ['user', 'item', 'rating', 'ts'][:len(df.columns)]
    _, df['user'] = np.unique(df['user'].to_numpy(), return_inverse=True)
    _, df['item'] = np.unique(df['item'].to_numpy(), return_inverse=True)
    nu, ni = int(df['user'].max()) + 1, int(df['item'].max()) + 1
    df = df.sort_values(['user', 'ts'])
    last = df.groupby('user').tail(1)
    train = df.drop(last.index)
    rng = np.random.RandomState(seed)
    seen = set(zip(df['user'].to_numpy().tolist(), df['item'].to_numpy().tolist()))

    def negatives(users, count):
        out = np.empty((len(users), count), dtype=np.int64)
        for r, u in enumerate(users):
            k = 0
            while k < count:
                j = int(rng.randint(ni))
                if (u, j) not in seen:
                    out[r, k] = j
                    k += 1
        return out
    tu, ti = train['user'].to_numpy(), train['item'].to_numpy()
    neg = negatives(tu.tolist(), num_negatives)
    users = np.concatenate([tu, np.repeat(tu, num_negatives)])
    items = np.concatenate([ti, neg.reshape(-1)])
    labels = np.concatenate([np.ones(len(tu)), np.zeros(len(tu) * num_negatives)]).astype(np.float32)
    perm = rng.permutation(len(users))
    lu, li = last['user'].to_numpy(), last['item'].to_numpy()
    test_items = np.concatenate([li.reshape(-1, 1), negatives(lu.tolist(), 99)], axis=1)
    out = dict(train_users=users[perm], train_items=items[perm], train_labels=labels[perm], test_users=lu,
               test_items=test_items, num_users=np.array(nu), num_items=np.array(ni))
    np.savez(cache, **out)
    return out
