"""mc_kernel + residual_kernel time per 4K frame with only MC, only the residual, both (CUDA events of the job's stream)."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import bench
from hevc_b200 import backend, abi
jobs, refs = bench.make_inputs("4k", 0)
ctx = backend.Context(jobs[0].sp, device=0, dpb_slots=4, jobs=2)
for s in range(2): ctx.upload(s, refs[s])
j = ctx.begin(2); ctx.append(j, jobs[0]); ctx.stage(j)
ctx.set_profiling(True)
for stages in (1, 2, 3, 15):
    acc = np.zeros(abi.NUM_STAGES)
    for i in range(12):
        ctx.launch(j, 2, stages=stages); ctx.sync(j)
        if i >= 2: acc += np.array(ctx.stage_ms(j))
    print("stages", stages, "mc %.4f ms  residual %.4f  loopfilter %.4f  pad %.4f  total %.4f" % tuple((acc / 10)[[0, 1, 2, 3, 4]]))
