"""reference examples/cnn/models/*: ``from models import MLP`` style imports; the definitions live in
hetu_61a7_b200/models/cnn.py."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.models.cnn import (alexnet, cnn_3_layers, lenet, logreg, lstm, mlp, resnet, resnet18, resnet34, rnn, vgg,   # noqa: E402,F401
                                       vgg16, vgg19)
from . import AlexNet, CNN, LSTM, LeNet, LogReg, MLP, RNN, ResNet, VGG   # noqa: E402,F401
