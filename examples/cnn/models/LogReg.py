"""reference examples/cnn/models/LogReg.py."""
from hetu_61a7_b200.models.cnn import logreg   # noqa: F401
