"""one polish pass of bench.py's batch as a single part, for `ncu --set full` captures of individual launches
(tools/profile_mixed.py is run under ncu with --kernel-name / --launch-skip, see profiles/r1_summary.md)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from sviper_b200 import api  # noqa: E402

ctx = api.Context(0)
ctx.set_lanes(1)
res = ctx.polish_batch(bench.make_batch(100, seed=1000), with_final=True)
print(sum(r["rounds"] for r in res))
ctx.close()
