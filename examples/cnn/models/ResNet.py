"""reference examples/cnn/models/ResNet.py."""
from hetu_61a7_b200.models.cnn import resnet, resnet18, resnet34   # noqa: F401
