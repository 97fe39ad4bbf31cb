"""Cost of the SEI picture hash on the device (ohcu_slot_md5) against the alternative it replaces for a conformance run:
download the picture (ohcu_slot_download) and hash it on the host (hashlib = what av_md5 computes)."""
import ctypes as C
import hashlib
import sys
import time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from hevc_b200 import abi, backend, synth

for (w, h, bd) in ((1920, 1080, 8), (3840, 2160, 10)):
    job = synth.make_job(w, h, bd, 1, seed=1)
    pic = synth.make_picture(job.sp, 9)
    with backend.Context(job.sp, dpb_slots=4, jobs=1) as ctx:
        for s in range(4):
            ctx.upload(s, pic)
        out = (C.c_uint8 * 48)()
        abi.check(ctx.lib.ohcu_slot_md5(ctx.handle, 0, out))
        t = time.perf_counter()
        abi.check(ctx.lib.ohcu_slot_md5(ctx.handle, 0, out))
        dev = time.perf_counter() - t
        ok = all(bytes(out)[16 * c:16 * c + 16] == hashlib.md5(p.tobytes()).digest() for c, p in enumerate(pic))
        t = time.perf_counter()
        got = ctx.download(0)
        d2h = time.perf_counter() - t
        t = time.perf_counter()
        for p in got:
            hashlib.md5(p.tobytes()).digest()
        host = time.perf_counter() - t
    print(f"{w}x{h} {bd}-bit: device md5 {dev * 1e3:.1f} ms per picture (3 chains, correct={ok}); download {d2h * 1e3:.2f} ms + host md5 {host * 1e3:.1f} ms (1 core)")
