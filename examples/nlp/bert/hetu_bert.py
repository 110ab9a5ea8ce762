"""reference examples/nlp/bert/hetu_bert.py: BertModel / BertForPreTraining / BertForSequenceClassification."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.models.nlp import BertForPreTraining, BertForSequenceClassification, BertModel   # noqa: E402,F401
