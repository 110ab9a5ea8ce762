"""reference examples/nlp/bert/tokenization.py."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.utils.tokenizers import BasicTokenizer, BertTokenizer, WordpieceTokenizer, load_vocab   # noqa: E402,F401
