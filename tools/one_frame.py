"""One 4K frame launched a few times (for ncu): python tools/one_frame.py [--sparse] [--workload 4k]"""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import bench
from hevc_b200 import backend
wl = sys.argv[sys.argv.index("--workload") + 1] if "--workload" in sys.argv else "4k"
jobs, refs = bench.make_inputs(wl, 0)
ctx = backend.Context(jobs[0].sp, device=0, dpb_slots=4, jobs=2)
for s in range(2): ctx.upload(s, refs[s])
j = ctx.begin(2); ctx.append(j, jobs[0], sparse="--sparse" in sys.argv); ctx.stage(j)
for _ in range(3):
    ctx.launch(j, 2); ctx.sync()
