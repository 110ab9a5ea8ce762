"""reference spelling: ``from graphboard.graph2fig import show`` (python/graphboard/graph2fig.py)."""
from hetu_61a7_b200.graphboard import show, to_dot      # noqa: F401
