"""Cycles thread 0 of every loopfilter_kernel CTA spends between phase marks (library built with -DLF_PROFILE:
cd hevc_b200/csrc && nvcc <flags of __graft_entry__> -DLF_PROFILE -o ../../tools/variants/lfprof.so *.cu;
run with OHCU_LIB=tools/variants/lfprof.so)."""
import ctypes as C, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import bench
from hevc_b200 import backend
jobs, refs = bench.make_inputs("4k", 0)
ctx = backend.Context(jobs[0].sp, device=0, dpb_slots=3, jobs=1)
for s in range(2): ctx.upload(s, refs[s])
j = ctx.begin(2); ctx.append(j, jobs[0]); ctx.stage(j)
for k in range(3): ctx.launch(0, 2); ctx.sync()
out = (C.c_ulonglong * 16)()
ctx.lib.ohcu_lf_profile(out, 1)
N = 16
for k in range(N): ctx.launch(0, 2); ctx.sync()
ctx.lib.ohcu_lf_profile(out, 1)
names = ["setup + TMA issue + barrier", "SAO records", "barrier", "wait for the tile", "vertical edges", "barrier",
         "horizontal edges", "barrier", "SAO + store", "segment lists"]
nctb = 2040 * N
tot = sum(out[:10])
for i in (0, 9, 1, 2, 3, 4, 5, 6, 7, 8):
    print(f"{names[i]:30s} {out[i] / nctb:9.0f} clk/CTB  {out[i] / tot * 100:5.1f}%")
print(f"{'total':30s} {tot / nctb:9.0f} clk/CTB")
