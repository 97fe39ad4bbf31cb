import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import frame_cases
from hevc_b200 import backend
for name in sys.argv[1:]:
    job, refs = frame_cases.build_case(name)
    print(name, "slots", job.sp.dpb_slots, "ctb", job.sp.log2_ctb_size, flush=True)
    with backend.Context(job.sp) as ctx:
        for st in (1, 3):
            ctx.reconstruct(job, refs, out_init=refs[-1], stages=st)
            print("  ok", st, flush=True)
print("done")
