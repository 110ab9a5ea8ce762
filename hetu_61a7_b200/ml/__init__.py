"""Classical machine-learning library (the reference's ``hetuml`` package, third_party/HetuML):
linear models, factorization machine, GBDT / random forest, kNN, naive Bayes, NMF and LDA on a
native OpenMP core (csrc_ml/ml.cc -> lib/libhetu_ml.so), optionally data-parallel over the
framework's parameter server.  Also importable as ``hetuml`` (``from hetuml.ensemble import GBDT``)."""
from . import cluster, data, decomposition, ensemble, factorization_machines, linear, model_base, naive_bayes, neighbors
from .cluster import Cluster
from .data import COOMatrix, Corpus, CSRMatrix, Dataset
from .decomposition import LDA, MF, NMF
from .ensemble import GBDT, RandomForest, RanomBoostedForest
from .factorization_machines import FM
from .linear import SVM, LinearRegression, LogisticRegression
from .naive_bayes import NaiveBayes
from .neighbors import KNN

__all__ = ['Cluster', 'Dataset', 'CSRMatrix', 'COOMatrix', 'Corpus', 'LogisticRegression', 'SVM', 'LinearRegression',
           'FM', 'GBDT', 'RandomForest', 'RanomBoostedForest', 'KNN', 'NaiveBayes', 'NMF', 'MF', 'LDA']
