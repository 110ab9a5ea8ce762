"""reference examples/gnn/gnn_model/{layer,model,utils}.py."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.models.gnn import GCN, SageConv, dense_model, sparse_model   # noqa: E402,F401
