"""Stress: (a) B frames in flight on B streams, every output compared with the oracle; (b) the e2e
flow of bench.py (host threads, ring of E slots, async downloads), every downloaded frame compared."""
import ctypes as C
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from concurrent.futures import ThreadPoolExecutor
import bench, oracle_lib
from hevc_b200 import abi, backend
wl = sys.argv[1] if len(sys.argv) > 1 else "1080p"
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 30
jobs, refs = bench.make_inputs(wl, 0)
sp = jobs[0].sp
orc = oracle_lib.Lib("oracle")
want = [orc.replay(fj, list(refs) + [None] * (sp.dpb_slots - 2)) for fj in jobs]
B, E = 8, 16
ctx = backend.Context(sp, device=0, dpb_slots=2 + E, jobs=E)
lib, h = ctx.lib, ctx.handle
for s in range(2): ctx.upload(s, refs[s])
sparse = [ctx.sparse_lists(fj) for fj in jobs]


def report(tag, got, w):
    n = 0
    for c, (a, b) in enumerate(zip(got, w)):
        if not np.array_equal(a, b):
            ys, xs = np.nonzero(a != b)
            print(f"{tag} plane {c}: {len(ys)} samples differ, rows {ys.min()}-{ys.max()} cols {xs.min()}-{xs.max()}", flush=True)
            n = 1
    return n


bad = 0
for k in range(B):
    j = ctx.begin(2 + k); ctx.append(j, jobs[k % len(jobs)], sparse=sparse[k % len(jobs)]); ctx.stage(j)
ctx.sync()
for r in range(rounds):
    for rep in range(3):
        for k in range(B): ctx.launch(k, 2 + k)
    ctx.sync()
    for k in range(B):
        bad += report(f"resident round {r} frame {k}", ctx.download(2 + k), want[k % len(jobs)])
print(wl, "resident: bad", bad, flush=True)

outs = [ctx.pinned_picture() for _ in range(E)]
pending = [None] * E
bad2 = 0


def submit(i):
    global bad2
    k = i % E
    if pending[k] is not None:
        ctx.slot_sync(2 + k)
        bad2 += report(f"e2e frame {pending[k]} slot {k}", outs[k], want[pending[k] % len(jobs)])
        for p in outs[k]: p[:] = 0
    fj = jobs[k % len(jobs)]
    j = ctx.begin(2 + k)
    for s in range(len(fj.slices)):
        tab = abi.SliceTab.from_buffer_copy(fj.slices[s].tobytes())
        abi.check(lib.ohcu_set_slice(h, j, s, C.byref(tab)))
    abi.check(lib.ohcu_append_pu(h, j, fj.pus.ctypes.data, len(fj.pus)))
    recs, pairs = sparse[k % len(jobs)]
    ptr = C.c_void_p()
    abi.check(lib.ohcu_reserve_coefs(h, j, len(pairs), C.byref(ptr)))
    C.memmove(ptr.value, pairs.ctypes.data, pairs.nbytes)
    abi.check(lib.ohcu_append_tu_sparse(h, j, recs.ctypes.data, len(recs), ptr, len(pairs)))
    d = fj.deblock_in(); s_ = fj.sao_in()
    abi.check(lib.ohcu_set_deblock(h, j, C.byref(d)))
    abi.check(lib.ohcu_set_sao(h, j, C.byref(s_)))
    ctx.flush(j)
    ctx.download_async(2 + k, outs[k])
    pending[k] = i


ctx.sync()
nthreads = int(sys.argv[3]) if len(sys.argv) > 3 else 4
pool = ThreadPoolExecutor(nthreads)
list(pool.map(submit, range(rounds * E)))
ctx.sync()
for k in range(E):
    ctx.slot_sync(2 + k)
    bad2 += report(f"e2e last slot {k}", outs[k], want[pending[k] % len(jobs)])
print(wl, "e2e: bad", bad2, flush=True)
