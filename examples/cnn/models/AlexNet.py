"""reference examples/cnn/models/AlexNet.py."""
from hetu_61a7_b200.models.cnn import alexnet   # noqa: F401
