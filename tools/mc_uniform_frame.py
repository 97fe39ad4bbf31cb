"""mc_kernel on a picture tiled with equal PUs (for the comparison with tools/ubench_mc_mma.cu): 4K 10-bit, luma only
(chroma_format_idc 0), every PU N x N, uni-predicted from list 0, quarter-pel phases 1..3 on both axes, integer displacement
within +-20 samples.  python tools/mc_uniform_frame.py [--n 16] [--bi]"""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from hevc_b200 import backend, synth

n = int(sys.argv[sys.argv.index("--n") + 1]) if "--n" in sys.argv else 16
bi = "--bi" in sys.argv
w, h = 3840, 2160 if n < 64 else 2176
job = synth.make_job(w, h, 10, 0, seed=5)
rng = np.random.default_rng(5)
xs, ys = np.meshgrid(np.arange(0, w, n), np.arange(0, h, n))
# CTB-contiguous order
ctb = (ys // 64) * (w // 64) + xs // 64
order = np.argsort(ctb.ravel(), kind="stable")
pus = np.zeros(xs.size, synth.PU_DTYPE)
pus["x0"], pus["y0"] = xs.ravel()[order], ys.ravel()[order]
pus["w"] = pus["h"] = n
pus["pred_flag"] = 3 if bi else 1
mv = rng.integers(-20, 21, size=(len(pus), 2, 2)) * 4 + rng.integers(1, 4, size=(len(pus), 2, 2))
pus["mv"] = mv
job.pus = pus
job.tus = job.tus[:0]; job.coeffs = job.coeffs[:0]
job.slices["weight_flag"] = 0
refs = [synth.make_picture(job.sp, 50 + s) for s in range(job.sp.dpb_slots)]
with backend.Context(job.sp, jobs=2) as ctx:
    for s, r in enumerate(refs[:-1]):
        ctx.upload(s, r)
    ctx.set_profiling(True)
    j = ctx.begin(job.sp.dpb_slots - 1)
    ctx.append(j, job, use_deblock=False, use_sao=False)
    ctx.stage(j)
    ms = []
    for _ in range(20):
        ctx.launch(j, job.sp.dpb_slots - 1, stages=1)
        ctx.sync(j)
        ms.append(ctx.stage_ms(j)[0])
print(f"mc_kernel, {w}x{h} luma only, {len(pus)} PUs of {n}x{n}, {'bi' if bi else 'uni'}, hv phases: {np.median(ms):.4f} ms (min {min(ms):.4f})")
