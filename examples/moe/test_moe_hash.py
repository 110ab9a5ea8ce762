#!/usr/bin/env python
"""reference examples/moe/test_moe_hash.py: the hash-gate MoE layer benchmark (same flags as test_moe.py).
    heturun -w 8 python examples/moe/test_moe_hash.py --top 2"""
import sys

from test_moe import main

if '--gate' not in sys.argv:
    sys.argv += '--gate hash'.split()
main()
