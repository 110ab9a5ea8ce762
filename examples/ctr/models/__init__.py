"""reference examples/ctr/models/*: model builders and data loading."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.models.ctr import dc_criteo, dcn_criteo, dfm_criteo, wdl_adult, wdl_criteo   # noqa: E402,F401
from hetu_61a7_b200.utils import ctr_data as load_data                                           # noqa: E402,F401
