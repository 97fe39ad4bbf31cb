import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import frame_cases
from hevc_b200 import backend
names = list(frame_cases.FRAME_CASES)[:int(sys.argv[1])]
skip = set(sys.argv[2].split(",")) if len(sys.argv) > 2 else set()
for name in names:
    if name in skip: continue
    job, refs = frame_cases.build_case(name)
    with backend.Context(job.sp) as ctx:
        for st in (1, 3, 7, 15):
            print(name, st, flush=True)
            ctx.reconstruct(job, refs, out_init=refs[-1], stages=st)
print("done")
