"""reference graphmix/torch/utils.py: message-passing matrix of a sampled sub-graph as a torch sparse tensor."""
import numpy as np
import torch


def mp_matrix(graph, device, use_original_gcn_norm=False):
    """Symmetrically normalised adjacency (row = destination, column = source) of ``graph``;
    ``use_original_gcn_norm`` adds the self loops of the original GCN recipe."""
    graph.convert2coo()
    (src, dst), norm = graph.gcn_norm(use_original_gcn_norm)
    indices = torch.as_tensor(np.vstack((dst, src)), dtype=torch.long)
    return torch.sparse_coo_tensor(indices=indices, values=torch.as_tensor(norm, dtype=torch.float32),
                                   size=(graph.num_nodes, graph.num_nodes), device=device).coalesce()
