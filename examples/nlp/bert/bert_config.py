"""reference examples/nlp/bert/bert_config.py."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.models.nlp import BertConfig   # noqa: E402,F401
