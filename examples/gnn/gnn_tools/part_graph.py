#!/usr/bin/env python
"""reference examples/gnn/gnn_tools/part_graph.py: ``python part_graph.py -d Cora -n 4 -p ~/graphmix_data``."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.graphmix.partition import main          # noqa: E402

if __name__ == '__main__':
    main()
