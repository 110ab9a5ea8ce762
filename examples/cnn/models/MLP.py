"""reference examples/cnn/models/MLP.py."""
from hetu_61a7_b200.models.cnn import mlp   # noqa: F401
