"""Repeat one frame N times and compare every result with the oracle: finds scheduling-dependent results."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import bench, oracle_lib
from hevc_b200 import backend
wl = sys.argv[1] if len(sys.argv) > 1 else "1080p"
stages = int(sys.argv[2]) if len(sys.argv) > 2 else 0
jobs, refs = bench.make_inputs(wl, 0)
sp = jobs[0].sp
want = None
ctx = backend.Context(sp, device=0, dpb_slots=4, jobs=2)
for s in range(2): ctx.upload(s, refs[s])
sparse = ctx.sparse_lists(jobs[0])
j = ctx.begin(2); ctx.append(j, jobs[0], sparse=sparse); ctx.stage(j)
first = None
bad = 0
for i in range(40):
    ctx.launch(j, 2, stages=stages); ctx.sync(j)
    got = ctx.download(2)
    if first is None:
        first = got
        continue
    for c, (a, b) in enumerate(zip(got, first)):
        if not np.array_equal(a, b):
            ys, xs = np.nonzero(a != b)
            bad += 1
            print(f"run {i} plane {c}: {len(ys)} samples differ from run 0, rows {ys.min()}-{ys.max()} cols {xs.min()}-{xs.max()}")
if stages == 0:
    want = oracle_lib.Lib("oracle").replay(jobs[0], list(refs) + [None] * (sp.dpb_slots - 2))
    for c, (a, b) in enumerate(zip(first, want)):
        if not np.array_equal(a, b):
            ys, xs = np.nonzero(a != b)
            print(f"run 0 plane {c}: {len(ys)} samples differ from the oracle, rows {ys.min()}-{ys.max()} cols {xs.min()}-{xs.max()}")
            bad += 1
print(wl, "stages", stages, "bad", bad)
