"""reference graphmix/torch/gcn.py: a small dense-feature GCN in plain PyTorch, for checking sampled
batches against another framework."""
import torch


class GCN(torch.nn.Module):
    """One graph convolution: ``A_hat @ (x W + b)``."""

    def __init__(self, in_features, out_features, activation=None):
        super().__init__()
        self.linear = torch.nn.Linear(in_features, out_features)
        self.activation = activation

    def forward(self, x, mp):
        x = torch.sparse.mm(mp, self.linear(x))
        return self.activation(x) if self.activation is not None else x


class PytorchGCN(torch.nn.Module):
    def __init__(self, in_features, hidden, num_classes, num_layers=2):
        super().__init__()
        dims = [in_features] + [hidden] * (num_layers - 1) + [num_classes]
        self.layers = torch.nn.ModuleList(GCN(dims[i], dims[i + 1], torch.relu if i < num_layers - 1 else None)
                                          for i in range(num_layers))

    def forward(self, x, mp):
        for layer in self.layers:
            x = layer(x, mp)
        return x
