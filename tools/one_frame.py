"""One frame launched a few times (for ncu): python tools/one_frame.py [--workload 4k]"""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import bench
from hevc_b200 import backend
wl = sys.argv[sys.argv.index("--workload") + 1] if "--workload" in sys.argv else "4k"
jobs, refs = bench.make_inputs(wl, 0, distinct=1)
ctx = backend.Context(jobs[0].sp, device=0, dpb_slots=32, jobs=2)
for s, r in enumerate(refs):
    ctx.upload(s, r)
j = ctx.begin(16); ctx.append_inplace(j, jobs[0], coeffs="sparse"); ctx.stage(j)
for _ in range(3):
    ctx.launch(j, 16); ctx.sync()
