import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import frame_cases, oracle_lib
from hevc_b200 import backend
name = sys.argv[1]; st = int(sys.argv[2])
job, refs = frame_cases.build_case(name)
t = job.tus
print("modes", np.unique(t["mode"], return_counts=True), "flags", np.unique(t["flags"], return_counts=True))
print("coeff_off %16:", np.unique(t["coeff_off"] % 16, return_counts=True))
with backend.Context(job.sp) as ctx:
    got = ctx.reconstruct(job, refs, out_init=refs[-1], stages=st)
want = oracle_lib.Lib("oracle").replay(job, refs, out_init=refs[-1], stages=st)
for c, (g, w) in enumerate(zip(got, want)):
    d = np.argwhere(g != w)
    print("plane", c, "diff", len(d), d[:5].tolist())
