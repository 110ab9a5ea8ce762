"""reference examples/cnn/models/LeNet.py."""
from hetu_61a7_b200.models.cnn import lenet   # noqa: F401
