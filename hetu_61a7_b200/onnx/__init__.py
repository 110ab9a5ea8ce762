"""ONNX bridge (reference python/hetu/onnx): ``hetu2onnx.export`` / ``onnx2hetu.load_onnx``."""
from . import hetu2onnx, onnx2hetu, proto          # noqa: F401
from .hetu2onnx import export                      # noqa: F401
from .onnx2hetu import load_onnx                   # noqa: F401
