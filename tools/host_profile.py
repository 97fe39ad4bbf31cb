"""Wall-clock cost of each C-ABI call of the e2e submit path (one 4K frame), on the GPU box."""
import ctypes as C, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import bench
from hevc_b200 import abi, backend

jobs, refs = bench.make_inputs("4k", 0)
sp = jobs[0].sp
B = 8
ctx = backend.Context(sp, device=0, dpb_slots=2 + B, jobs=B)
lib, h = ctx.lib, ctx.handle
for s in range(2):
    ctx.upload(s, refs[s])
outs = [ctx.pinned_picture() for _ in range(B)]
T = {}
def timed(name, fn):
    t = time.perf_counter(); r = fn(); T[name] = T.get(name, 0.0) + time.perf_counter() - t; return r
def submit(k, copy):
    fj = jobs[k % len(jobs)]
    j = timed("begin", lambda: ctx.begin(2 + k))
    tab = abi.SliceTab.from_buffer_copy(fj.slices[0].tobytes())
    timed("set_slice", lambda: lib.ohcu_set_slice(h, j, 0, C.byref(tab)))
    timed("append_pu", lambda: lib.ohcu_append_pu(h, j, fj.pus.ctypes.data, len(fj.pus)))
    ptr = C.c_void_p(); off = C.c_uint32()
    lib.ohcu_reserve_coeffs(h, j, len(fj.coeffs), C.byref(ptr), C.byref(off))
    if copy:
        C.memmove(ptr.value, fj.coeffs.ctypes.data, fj.coeffs.nbytes)
    timed("append_tu", lambda: lib.ohcu_append_tu_inplace(h, j, fj.tus.ctypes.data, len(fj.tus)))
    d = fj.deblock_in(); s = fj.sao_in()
    timed("set_deblock", lambda: lib.ohcu_set_deblock(h, j, C.byref(d)))
    timed("set_sao", lambda: lib.ohcu_set_sao(h, j, C.byref(s)))
    timed("stage", lambda: lib.ohcu_frame_stage(h, j))
    timed("launch", lambda: lib.ohcu_frame_launch(h, j, -1))
    timed("download_async", lambda: ctx.download_async(2 + k, outs[k]))
for k in range(B):
    submit(k, True)
ctx.sync()
T.clear()
N = 5
t0 = time.perf_counter()
for _ in range(N):
    for k in range(B):
        submit(k, False)
ctx.sync()
tot = time.perf_counter() - t0
print(f"total {tot / (N * B) * 1e3:.3f} ms per frame")
for k, v in T.items():
    print(f"  {k:16s} {v / (N * B) * 1e3:.3f} ms")
