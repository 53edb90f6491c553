#!/usr/bin/env python
"""Multi-level measurement: a thin alias of `python bench.py --workload ml ...`
(the workload, its CPU-baseline leg and the JSON line live in bench.py:run_multilevel)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.execv(sys.executable, [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "ml", "--steps", "30"] + sys.argv[1:])
