"""Host-side cost of submitting one 4K frame through the C ABI, call by call (one thread).
bench.py's e2e leg is host-bound below ~4 submitting threads; this shows which calls matter."""
import ctypes as C
import os
import sys
import time
from collections import defaultdict

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import bench
from hevc_b200 import abi, backend

workload = sys.argv[1] if len(sys.argv) > 1 else "4k"
jobs, refs = bench.make_inputs(workload, 0, 1)
sp = jobs[0].sp
E = 8
ctx = backend.Context(sp, dpb_slots=2 + E, jobs=E)
lib, h = ctx.lib, ctx.handle
for s in range(2):
    ctx.upload(s, refs[s])
sparse = [ctx.sparse_lists(fj) for fj in jobs]
outs = [ctx.pinned_picture() for _ in range(E)]
acc = defaultdict(float)


def timed(name, fn, *a):
    t = time.perf_counter()
    r = fn(*a)
    acc[name] += time.perf_counter() - t
    return r


def submit(i, copy):
    k = i % E
    fj = jobs[k % len(jobs)]
    j = timed("begin", ctx.begin, 2 + k)
    for s in range(len(fj.slices)):
        tab = abi.SliceTab.from_buffer_copy(fj.slices[s].tobytes())
        timed("set_slice", lib.ohcu_set_slice, h, j, s, C.byref(tab))
    timed("append_pu", lib.ohcu_append_pu, h, j, fj.pus.ctypes.data, len(fj.pus))
    recs, pairs = sparse[k % len(jobs)]
    ptr = C.c_void_p()
    timed("reserve_coefs", lib.ohcu_reserve_coefs, h, j, len(pairs), C.byref(ptr))
    if copy:
        C.memmove(ptr.value, pairs.ctypes.data, pairs.nbytes)
    timed("append_tu_sparse", lib.ohcu_append_tu_sparse, h, j, recs.ctypes.data, len(recs), ptr, len(pairs))
    d = fj.deblock_in(); s_ = fj.sao_in()
    timed("set_deblock", lib.ohcu_set_deblock, h, j, C.byref(d))
    timed("set_sao", lib.ohcu_set_sao, h, j, C.byref(s_))
    timed("flush", ctx.flush, j)
    timed("download_async", ctx.download_async, 2 + k, outs[k])


for i in range(E):
    ctx.sync()
    submit(i, True)
ctx.sync()
acc.clear()
n = 32
t0 = time.perf_counter()
for i in range(n):
    submit(i, False)
t1 = time.perf_counter()
ctx.sync()
print(f"{workload}: {len(jobs[0].pus)} PUs, {len(jobs[0].tus)} TUs, {len(sparse[0][1])} pairs per frame")
print(f"submit loop {1e3 * (t1 - t0) / n:.3f} ms per frame (one thread)")
for k, v in sorted(acc.items(), key=lambda kv: -kv[1]):
    print(f"  {k:18s} {1e3 * v / n:.3f} ms")
