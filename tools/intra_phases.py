"""Cycles per phase of intra_kernel (INTRA_PROFILE build): tools/build_flags_variant.sh iprof -DINTRA_PROFILE, then
OHCU_LIB=tools/variants/iprof/libohcu.so python tools/intra_phases.py [--pct 100] [--free]"""
import ctypes as C, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from hevc_b200 import abi, backend, synth
pct = int(sys.argv[sys.argv.index("--pct") + 1]) if "--pct" in sys.argv else 100
job = synth.make_job(3840, 2160, 8, 1, seed=77, intra_pct=pct, sp_intra=1)
if "--free" in sys.argv:
    job.ius["cand"] &= 0x20
refs = [synth.make_picture(job.sp, 770 + s) for s in range(job.sp.dpb_slots)]
ctx = backend.Context(job.sp, jobs=2)
lib = abi.load()
for s, r in enumerate(refs):
    ctx.upload(s, r)
j = ctx.begin(job.sp.dpb_slots - 1); ctx.append_inplace(j, job, coeffs="sparse"); ctx.stage(j)
ctx.launch(j, job.sp.dpb_slots - 1); ctx.sync(j)
out = (C.c_ulonglong * 16)()
lib.ohcu_intra_profile(out, 1)
ctx.launch(j, job.sp.dpb_slots - 1); ctx.sync(j)
lib.ohcu_intra_profile(out, 0)
n = max(1, out[7])
names = ["ticket", "records + own rows + residuals", "wait for neighbour CTBs", "left column + top row", "blocks", "fence + release"]
print(f"intra_pct={pct} free={'--free' in sys.argv}: {n} CTBs with intra blocks, cycles per CTB (thread 0):")
for k, name in enumerate(names):
    print(f"  {name:34s} {out[k] / n:10.0f}")
nb = max(1, out[14])
print(f"  4x4 / 8x8 luma blocks: {nb}, cycles per block (lane 0): sources {out[8] / nb:.0f}, fetch {out[9] / nb:.0f}, smoothing {out[10] / nb:.0f}, "
      f"predict {out[11] / nb:.0f}, residual + store {out[12] / nb:.0f}, tail {out[13] / nb:.0f}")
