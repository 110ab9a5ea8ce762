"""reference examples/cnn/models/RNN.py."""
from hetu_61a7_b200.models.cnn import rnn   # noqa: F401
