"""Cost of the intra wavefront: one picture with a given share of intra CUs, stage times from the library's own
events (stage 1 = residual_kernel + pcm_kernel + intra_kernel).  python tools/intra_cost.py [--size 3840x2160]"""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from hevc_b200 import backend, synth

size = sys.argv[sys.argv.index("--size") + 1] if "--size" in sys.argv else "3840x2160"
w, h = (int(v) for v in size.split("x"))
for bd, cf in ((10, 1), (8, 1), (10, 3)):
    for pct in (0, 5, 25, 100):
        job = synth.make_job(w, h, bd, cf, seed=77, intra_pct=pct, sp_intra=1)
        refs = [synth.make_picture(job.sp, 770 + s) for s in range(job.sp.dpb_slots)]
        with backend.Context(job.sp, jobs=2) as ctx:
            for s, r in enumerate(refs):
                ctx.upload(s, r)
            ctx.set_profiling(True)
            j = ctx.begin(job.sp.dpb_slots - 1)
            ctx.append_inplace(j, job, coeffs="sparse")
            ctx.stage(j)
            ms = []
            for _ in range(5):
                ctx.launch(j, job.sp.dpb_slots - 1)
                ctx.sync(j)
                ms.append(ctx.stage_ms(j))
            ms = np.median(np.array(ms), axis=0)
        print(f"{w}x{h} bd={bd} cfi={cf} intra_pct={pct:3d}: n_pu={len(job.pus)} n_tu={len(job.tus)} n_iu={len(job.ius)} "
              f"mc={ms[0]:.3f} residual+intra={ms[1]:.3f} loopfilter={ms[2]:.3f} total={ms[4]:.3f} ms", flush=True)
