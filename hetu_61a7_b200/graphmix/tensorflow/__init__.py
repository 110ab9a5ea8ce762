"""TensorFlow helper of the reference (graphmix/tensorflow/utils.py).  TensorFlow is not part of this image;
``mp_matrix`` imports it on first use."""


def mp_matrix(graph, use_original_gcn_norm=False):
    import numpy as np
    import tensorflow as tf                                     # noqa: F401  (ImportError here = TensorFlow not installed)
    graph.convert2coo()
    (src, dst), norm = graph.gcn_norm(use_original_gcn_norm)
    return tf.sparse.reorder(tf.SparseTensor(indices=np.stack([dst, src], axis=1).astype(np.int64), values=norm.astype(np.float32),
                                             dense_shape=(graph.num_nodes, graph.num_nodes)))
