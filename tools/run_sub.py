"""Run one of bench.py's N=1 side records alone: python tools/run_sub.py config3|config5"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

name = sys.argv[1]
sys.argv = sys.argv[:1] + sys.argv[2:]
args = bench.parse()
print(json.dumps(getattr(bench, "sub_" + name)(args)))
