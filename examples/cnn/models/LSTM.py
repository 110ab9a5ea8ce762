"""reference examples/cnn/models/LSTM.py."""
from hetu_61a7_b200.models.cnn import lstm   # noqa: F401
