"""reference tests/tester.py lives in the package now (hetu_61a7_b200/utils/tester.py)."""
from hetu_61a7_b200.utils.tester import HetuTester          # noqa: F401
