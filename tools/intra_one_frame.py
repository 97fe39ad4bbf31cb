"""One picture with intra blocks launched a few times (for ncu): python tools/intra_one_frame.py [--pct 100] [--free]
--free clears every availability flag: no block reads a neighbour, so no CTB waits for another (the predict time alone)."""
import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from hevc_b200 import backend, synth
pct = int(sys.argv[sys.argv.index("--pct") + 1]) if "--pct" in sys.argv else 100
w, h = (3840, 2160) if "--4k" in sys.argv else (1920, 1080)
job = synth.make_job(w, h, 8, 1, seed=77, intra_pct=pct, sp_intra=1)
if "--free" in sys.argv:
    job.ius["cand"] &= 0x20
refs = [synth.make_picture(job.sp, 770 + s) for s in range(job.sp.dpb_slots)]
ctx = backend.Context(job.sp, jobs=2)
for s, r in enumerate(refs):
    ctx.upload(s, r)
ctx.set_profiling(True)
j = ctx.begin(job.sp.dpb_slots - 1); ctx.append_inplace(j, job, coeffs="sparse"); ctx.stage(j)
for _ in range(3):
    ctx.launch(j, job.sp.dpb_slots - 1); ctx.sync(j)
    print(f"{w}x{h} intra_pct={pct} n_iu={len(job.ius)} residual+intra={ctx.stage_ms(j)[1]:.3f} ms", flush=True)
