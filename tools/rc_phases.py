"""Cycles thread 0 of every recon_kernel CTA spends between phase marks (library built with -DRC_PROFILE:
cd hevc_b200/csrc && nvcc <flags of __graft_entry__> -DRC_PROFILE -o ../../tools/variants/prof.so *.cu;
run with OHCU_LIB=tools/variants/prof.so; --sparse = coefficients as (position, level) pairs)."""
import ctypes as C, sys
SPARSE = "--sparse" in sys.argv
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import bench
from hevc_b200 import backend
jobs, refs = bench.make_inputs("4k", 0)
NJ = 8 if "--cold" in sys.argv else 1          # --cold: 8 frames in turn, their lists fall out of L2 between launches
ctx = backend.Context(jobs[0].sp, device=0, dpb_slots=2 + NJ, jobs=NJ)
for s in range(2): ctx.upload(s, refs[s])
for k in range(NJ):
    j = ctx.begin(2 + k); ctx.append(j, jobs[k % len(jobs)], sparse=SPARSE); ctx.stage(j)
for k in range(3 * NJ): ctx.launch(k % NJ, 2 + k % NJ); ctx.sync()
out = (C.c_ulonglong * 16)()
ctx.lib.ohcu_rc_profile(out, 1)
N = 16
for k in range(N): ctx.launch(k % NJ, 2 + k % NJ); ctx.sync()
ctx.lib.ohcu_rc_profile(out, 1)
names = ["", "init + preload -> sync", "prepare + scans", "rounds cut + cell list", "issue + TU sort", "round loop (cells, waits)",
         "barrier after MC", "coefficient copy wait", "(residual setup)", "residual batches + barrier", "store"]
nctb = 2040 * N
tot = sum(out[1:11]) + sum(out[11:14])
names += ["  round loop: bookkeeping", "  round loop: waiting for windows", "  round loop: cells of warp 0"]
for i in range(1, 14):
    print(f"{names[i]:30s} {out[i] / nctb:9.0f} clk/CTB  {out[i] / tot * 100:5.1f}%")
print(f"{'total':30s} {tot / nctb:9.0f} clk/CTB")
