"""reference examples/rec/hetu_ncf.py: ``neural_mf(user_input, item_input, y_, num_users, num_items)``."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..'))
from hetu_61a7_b200.models.rec import neural_mf   # noqa: E402,F401
