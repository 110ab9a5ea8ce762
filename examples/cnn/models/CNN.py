"""reference examples/cnn/models/CNN.py."""
from hetu_61a7_b200.models.cnn import cnn_3_layers   # noqa: F401
