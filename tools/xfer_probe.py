"""Which leg bounds the e2e path: list upload (ohcu_frame_stage) alone, picture download alone."""
import ctypes as C, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
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
def fill(k, copy):
    fj = jobs[k % len(jobs)]
    j = ctx.begin(2 + k)
    tab = abi.SliceTab.from_buffer_copy(fj.slices[0].tobytes())
    lib.ohcu_set_slice(h, j, 0, C.byref(tab))
    lib.ohcu_append_pu(h, j, fj.pus.ctypes.data, len(fj.pus))
    ptr = C.c_void_p(); off = C.c_uint32()
    lib.ohcu_reserve_coeffs(h, j, len(fj.coeffs), C.byref(ptr), C.byref(off))
    if copy: C.memmove(ptr.value, fj.coeffs.ctypes.data, fj.coeffs.nbytes)
    lib.ohcu_append_tu_inplace(h, j, fj.tus.ctypes.data, len(fj.tus))
    d = fj.deblock_in(); s = fj.sao_in()
    lib.ohcu_set_deblock(h, j, C.byref(d)); lib.ohcu_set_sao(h, j, C.byref(s))
    return j
for k in range(B):
    j = fill(k, True); ctx.flush(j)
ctx.sync()
N = 6
# upload only
t = time.perf_counter()
for _ in range(N):
    for k in range(B):
        j = fill(k, False); abi.check(lib.ohcu_frame_stage(h, j))
ctx.sync(); dt = time.perf_counter() - t
print("fill + stage (H2D only): %.3f ms/frame" % (dt / (N * B) * 1e3))
# download only
t = time.perf_counter()
for _ in range(N):
    for k in range(B):
        ctx.download_async(2 + k, outs[k])
for k in range(B): ctx.slot_sync(2 + k)
dt = time.perf_counter() - t
print("download only: %.3f ms/frame" % (dt / (N * B) * 1e3))
from concurrent.futures import ThreadPoolExecutor
for nthr in (1, 4):
    pool = ThreadPoolExecutor(nthr)
    for mode in ("fill only (no GPU work)", "fill+stage", "fill+flush", "fill+flush+download"):
        def one(i):
            k = i % B
            j = fill(k, False)
            if mode == "fill only (no GPU work)":
                lib.ohcu_frame_abort(h, j); return
            if mode == "fill+stage":
                abi.check(lib.ohcu_frame_stage(h, j)); return
            ctx.flush(j)
            if mode.endswith("download"):
                ctx.download_async(2 + k, outs[k])
        ctx.sync()
        t = time.perf_counter()
        list(pool.map(one, range(N * B)))
        ctx.sync()
        for k in range(B): ctx.slot_sync(2 + k)
        dt = time.perf_counter() - t
        print("%d threads, %-26s %.3f ms/frame" % (nthr, mode, dt / (N * B) * 1e3))
    pool.shutdown()
# which pair fails to overlap?
for k in range(B):
    j = fill(k, False); abi.check(lib.ohcu_frame_stage(h, j))
ctx.sync()
def timed(name, fn, n=N * B):
    ctx.sync(); t = time.perf_counter()
    for i in range(n): fn(i)
    ctx.sync()
    for k in range(B): ctx.slot_sync(2 + k)
    print("%-34s %.3f ms/frame" % (name, (time.perf_counter() - t) / n * 1e3))
timed("kernels only (resident lists)", lambda i: ctx.launch(i % B, 2 + i % B))
timed("kernels + download", lambda i: (ctx.launch(i % B, 2 + i % B), ctx.download_async(2 + i % B, outs[i % B])))
def stage_dl(i):
    j = fill(i % B, False); abi.check(lib.ohcu_frame_stage(h, j)); ctx.download_async(2 + i % B, outs[i % B])
pool = ThreadPoolExecutor(4)
ctx.sync(); t = time.perf_counter()
list(pool.map(stage_dl, range(N * B)))
ctx.sync()
for k in range(B): ctx.slot_sync(2 + k)
print("%-34s %.3f ms/frame" % ("4 thr: H2D lists + download", (time.perf_counter() - t) / (N * B) * 1e3))
