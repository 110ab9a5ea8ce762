"""reference examples/cnn/models/VGG.py."""
from hetu_61a7_b200.models.cnn import vgg, vgg16, vgg19   # noqa: F401
