"""Time of one 4K Main10 frame's deblock + SAO stage with the bs planes uploaded (ohcu_set_deblock) and with
the planes derived on the device from the PU / TU lists (ohcu_set_deblock_derived: + bs_kernel)."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import bench
from hevc_b200 import abi, backend

jobs, refs = bench.make_inputs("4k", 0)
fj = jobs[0]
ctx = backend.Context(fj.sp, device=0, dpb_slots=4, jobs=2)
for s in range(2):
    ctx.upload(s, refs[s])
j0 = ctx.begin(2); ctx.append(j0, fj, sparse=True); ctx.stage(j0)
j1 = ctx.begin(3); ctx.append(j1, fj, sparse=True, derive_bs=True); ctx.stage(j1)
ctx.sync()
ctx.set_profiling(True)
for name, j, slot in (("bs planes from the host", j0, 2), ("bs derived on the device", j1, 3)):
    t = np.zeros(abi.NUM_STAGES)
    for i in range(24):
        ctx.launch(j, slot); ctx.sync(j)
        if i >= 4:
            t += np.array(ctx.stage_ms(j))
    t /= 20
    print(f"{name:28s} recon {t[0] + t[1]:.4f}  bs + deblock + SAO {t[2]:.4f}  halo {t[3]:.4f}  frame {t[4]:.4f} ms")
print("bs planes not uploaded: %d bytes per frame" % (2 * fj.vbs.nbytes))
