"""Cycles thread 0 of every recon_kernel CTA spends between phase marks (library built with -DRC_PROFILE:
cd hevc_b200/csrc && nvcc <flags of __graft_entry__> -DRC_PROFILE -o ../../tools/variants/prof.so *.cu;
run with OHCU_LIB=tools/variants/prof.so)."""
import ctypes as C, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import bench
from hevc_b200 import backend
jobs, refs = bench.make_inputs("4k", 0)
ctx = backend.Context(jobs[0].sp, device=0, dpb_slots=4, jobs=2)
for s in range(2): ctx.upload(s, refs[s])
j = ctx.begin(2); ctx.append(j, jobs[0]); ctx.stage(j)
for _ in range(3): ctx.launch(j, 2)
ctx.sync()
out = (C.c_ulonglong * 16)()
ctx.lib.ohcu_rc_profile(out, 1)
N = 10
for _ in range(N): ctx.launch(j, 2)
ctx.sync()
ctx.lib.ohcu_rc_profile(out, 1)
names = ["", "init + preload -> sync", "prepare + scans", "rounds cut + cell list", "issue + TU sort", "round loop (cells, waits)",
         "barrier after MC", "coefficient copy wait", "(residual setup)", "residual batches + barrier", "store"]
nctb = 2040 * N
tot = sum(out[1:11]) + sum(out[11:14])
names += ["  round loop: bookkeeping", "  round loop: waiting for windows", "  round loop: cells of warp 0"]
for i in range(1, 14):
    print(f"{names[i]:30s} {out[i] / nctb:9.0f} clk/CTB  {out[i] / tot * 100:5.1f}%")
print(f"{'total':30s} {tot / nctb:9.0f} clk/CTB")
