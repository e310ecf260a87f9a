# host-side breakdown of svp_polish_batch (SVP_TRACE_HOST=1): upload / rounds / final / gather / scatter per lane
import os, sys
sys.path.insert(0, '.')
os.environ["SVP_TRACE_HOST"] = "1"
import bench
from sviper_b200 import api
vs = bench.make_batch(100, seed=1000)
c = api.Context(0)
if len(sys.argv) > 1: c.set_lanes(int(sys.argv[1]))
pb = c.prepare(vs, True)
for _ in range(4):
    sys.stderr.write("---\n"); c.run_prepared(pb)
