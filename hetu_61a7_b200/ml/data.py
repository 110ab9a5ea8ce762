"""Input containers of the ML library: ``Dataset`` (labels + CSR features), ``CSRMatrix``,
``COOMatrix`` (ratings for matrix factorization) and ``Corpus`` (documents for LDA).

reference: third_party/HetuML/hetuml/python/hetuml/data.py:9-147 and the readers in
hetuml/cc/ml/include/data/{dataset.h,io.h}, include/model/lda/common/corpus.h:23-48.
The arrays stay numpy-owned and are handed to the native core by pointer."""
from __future__ import annotations

from typing import Optional, Sequence, Tuple, Union

import numpy as np

try:                                        # scipy is optional: only isinstance checks need it
    from scipy.sparse import coo_matrix, csc_matrix, csr_matrix, issparse
except ImportError:                         # pragma: no cover
    coo_matrix = csc_matrix = csr_matrix = ()

    def issparse(_):
        return False

from ._core import F32P, I32P, I64P


class CSRMatrix(object):
    """Row-compressed feature matrix (float32 values, int32 column ids, int64 row pointers, columns
    sorted inside each row)."""

    def __init__(self, indptr, indices, data, num_cols: Optional[int] = None):
        self.indptr = np.ascontiguousarray(indptr, dtype=np.int64)
        self.indices = np.ascontiguousarray(indices, dtype=np.int32)
        self.data = np.ascontiguousarray(data, dtype=np.float32)
        self._sort_rows()
        inferred = int(self.indices.max()) + 1 if self.indices.size else 0
        self.num_cols = max(int(num_cols or 0), inferred)

    def _sort_rows(self):
        ip, idx = self.indptr, self.indices
        if idx.size < 2:
            return
        row = np.repeat(np.arange(ip.size - 1), np.diff(ip))
        if not bool(np.any((row[1:] == row[:-1]) & (idx[1:] < idx[:-1]))):
            return
        order = np.lexsort((idx, row))
        self.indices = np.ascontiguousarray(idx[order])
        self.data = np.ascontiguousarray(self.data[order])

    @staticmethod
    def from_data(X) -> 'CSRMatrix':
        if isinstance(X, CSRMatrix):
            return X
        if issparse(X):
            X = X.tocsr()
            X.sum_duplicates()
            return CSRMatrix(X.indptr, X.indices, X.data, X.shape[1])
        X = np.asarray(X, dtype=np.float32)
        if X.ndim != 2:
            raise TypeError('Type %s cannot be converted to csr matrix' % type(X))
        mask = X != 0
        indptr = np.concatenate([[0], np.cumsum(mask.sum(1))])
        rows, cols = np.nonzero(mask)
        return CSRMatrix(indptr, cols, X[rows, cols], X.shape[1])

    def get_num_rows(self) -> int:
        return int(self.indptr.size - 1)

    get_num_instances = get_num_rows

    def get_num_cols(self) -> int:
        return self.num_cols

    get_max_dim = get_num_cols

    def ptrs(self):
        return (self.indptr.ctypes.data_as(I64P), self.indices.ctypes.data_as(I32P), self.data.ctypes.data_as(F32P))

    def rows(self, lo: int, hi: int) -> 'CSRMatrix':
        a, b = self.indptr[lo], self.indptr[hi]
        return CSRMatrix(self.indptr[lo:hi + 1] - a, self.indices[a:b], self.data[a:b], self.num_cols)


def _read_libsvm(path: str, rank: int = 0, total_ranks: int = 1):
    labels, indptr, indices, values = [], [0], [], []
    with open(path) as f:
        for ln, line in enumerate(f):
            if total_ranks > 1 and ln % total_ranks != rank:
                continue
            parts = line.split('#', 1)[0].split()
            if not parts:
                continue
            labels.append(float(parts[0]))
            for tok in parts[1:]:
                k, v = tok.split(':', 1)
                indices.append(int(k))
                values.append(float(v))
            indptr.append(len(indices))
    return np.asarray(labels, np.float32), np.asarray(indptr, np.int64), np.asarray(indices, np.int32), \
        np.asarray(values, np.float32)


class Dataset(object):
    """Labels + sparse features.  ``from_file`` reads libsvm / csv text (``rank``/``total_ranks``
    take every total_ranks-th line for data-parallel training), ``from_data`` takes ``(X, y)``."""

    def __init__(self, X: CSRMatrix, y: np.ndarray):
        self.X = X
        self.y = np.ascontiguousarray(y, dtype=np.float32).reshape(-1)
        if self.y.size != X.get_num_rows():
            raise ValueError('labels (%d) and feature rows (%d) differ' % (self.y.size, X.get_num_rows()))

    @staticmethod
    def from_file(path: str, data_format: str = 'libsvm', neg_y: bool = False, rank: int = 0,
                  total_ranks: int = 1) -> 'Dataset':
        fmt = data_format.lower()
        if fmt == 'libsvm':
            y, indptr, indices, values = _read_libsvm(path, rank, total_ranks)
            ds = Dataset(CSRMatrix(indptr, indices, values), y)
        elif fmt == 'csv':
            arr = np.loadtxt(path, delimiter=',', dtype=np.float32, ndmin=2)[rank::total_ranks]
            ds = Dataset(CSRMatrix.from_data(arr[:, 1:]), arr[:, 0])
        else:
            raise ValueError('unknown data format ' + data_format)
        if neg_y:
            ds.y = np.where(ds.y > 0, 1.0, -1.0).astype(np.float32)
        return ds

    @staticmethod
    def from_data(data: Tuple) -> 'Dataset':
        if isinstance(data, Dataset):
            return data
        if not isinstance(data, (tuple, list)) or len(data) != 2:
            raise TypeError('Expected a tuple for X and y, got %s' % type(data))
        X, y = data
        if not isinstance(y, np.ndarray):
            raise TypeError('y should be numpy array, got %s' % type(y))
        return Dataset(CSRMatrix.from_data(X), y)

    def get_num_instances(self) -> int:
        return self.X.get_num_rows()

    def get_max_dim(self) -> int:
        return self.X.get_num_cols()

    def labels01(self) -> np.ndarray:
        """Binary labels as {0, 1} whatever the file used ({-1,+1} or {0,1})."""
        return np.ascontiguousarray((self.y > 0).astype(np.float32))

    def labels_pm1(self) -> np.ndarray:
        return np.ascontiguousarray(np.where(self.y > 0, 1.0, -1.0).astype(np.float32))


class COOMatrix(object):
    """(row, col, value) triples; text files hold one ``row col value`` triple per line."""

    def __init__(self, row, col, data, shape: Optional[Tuple[int, int]] = None):
        self.row = np.ascontiguousarray(row, dtype=np.int32)
        self.col = np.ascontiguousarray(col, dtype=np.int32)
        self.data = np.ascontiguousarray(data, dtype=np.float32)
        nr = int(self.row.max()) + 1 if self.row.size else 0
        nc = int(self.col.max()) + 1 if self.col.size else 0
        self.shape = (max(nr, shape[0]), max(nc, shape[1])) if shape else (nr, nc)

    @staticmethod
    def from_file(path: str) -> 'COOMatrix':
        arr = np.loadtxt(path, dtype=np.float64, ndmin=2)
        return COOMatrix(arr[:, 0], arr[:, 1], arr[:, 2])

    @staticmethod
    def from_data(X) -> 'COOMatrix':
        if isinstance(X, COOMatrix):
            return X
        if issparse(X):
            X = X.tocoo()
            return COOMatrix(X.row, X.col, X.data, X.shape)
        if isinstance(X, (tuple, list)) and len(X) == 3:
            return COOMatrix(*X)
        raise TypeError('Type %s cannot be converted to coo matrix' % type(X))

    def get_num_rows(self) -> int:
        return self.shape[0]

    def get_num_cols(self) -> int:
        return self.shape[1]


class Corpus(object):
    """Documents as word-id sequences.  File format (reference corpus.h:23): one document per line,
    ``docId wordId wordId ...``."""

    def __init__(self, num_words: int, docs: Sequence[Sequence[int]]):
        self.num_words = int(num_words)
        lens = [len(d) for d in docs]
        self.doc_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        self.words = np.ascontiguousarray(np.concatenate([np.asarray(d, np.int32) for d in docs])
                                          if docs else np.zeros(0, np.int32), dtype=np.int32)
        if self.words.size and (self.words.min() < 0 or self.words.max() >= self.num_words):
            raise ValueError('word id out of range [0, %d)' % self.num_words)

    @staticmethod
    def from_file(num_words: int, path: str, rank: int = 0, total_ranks: int = 1) -> 'Corpus':
        docs = []
        with open(path) as f:
            for ln, line in enumerate(f):
                if total_ranks > 1 and ln % total_ranks != rank:
                    continue
                parts = line.split()
                if len(parts) > 1:
                    docs.append([int(w) for w in parts[1:]])
        return Corpus(num_words, docs)

    @staticmethod
    def from_data(num_words: int, docs: Sequence[Sequence[int]]) -> 'Corpus':
        return Corpus(num_words, docs)

    def get_num_docs(self) -> int:
        return int(self.doc_ptr.size - 1)

    def get_num_words(self) -> int:
        return self.num_words

    def get_num_tokens(self) -> int:
        return int(self.words.size)

    def get_word_size(self) -> int:
        return self.num_words

    def get_doc_size(self) -> int:
        return self.get_num_docs()
