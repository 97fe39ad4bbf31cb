"""recon_kernel time of one 4K Main10 frame with scaled coefficients (sparse form) and with un-scaled
levels + per-TU quantisers (levels form, dequantisation on the device).  Same lists, same pairs."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import bench
import levels_cases
from hevc_b200 import abi, backend

jobs, refs = bench.make_inputs("4k", 0)
fj = jobs[0]
levels, quant, scaling = levels_cases.frame_levels(fj, seed=5)
ctx = backend.Context(fj.sp, device=0, dpb_slots=4, jobs=2)
for s in range(2):
    ctx.upload(s, refs[s])
j0 = ctx.begin(2); ctx.append(j0, fj, sparse=True); ctx.stage(j0)
saved = fj.coeffs
fj.coeffs = levels
j1 = ctx.begin(3); ctx.append(j1, fj, levels=(quant, scaling)); ctx.stage(j1)
fj.coeffs = saved
ctx.sync()
ctx.set_profiling(True)
for name, j, slot in (("sparse (scaled coefficients)", j0, 2), ("levels (device dequantisation)", j1, 3)):
    t = np.zeros(abi.NUM_STAGES)
    for i in range(24):
        ctx.launch(j, slot); ctx.sync(j)
        if i >= 4:
            t += np.array(ctx.stage_ms(j))
    print(f"{name:34s} recon {t[abi.STAGE_RESIDUAL] / 20 + t[abi.STAGE_MC] / 20:.4f} ms")
