"""One small synthetic frame (416x240, dense and sparse lists) against the oracle; meant for compute-sanitizer
(closed on the round-1 GPU pool, so only the comparison ran):
compute-sanitizer --tool racecheck|memcheck|synccheck python tools/sanitize_frame.py [bit depth] [chroma format]"""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import oracle_lib
from hevc_b200 import backend, synth
bd = int(sys.argv[1]) if len(sys.argv) > 1 else 10
cf = int(sys.argv[2]) if len(sys.argv) > 2 else 1
job = synth.make_job(416, 240, bd, cf, seed=5)
refs = [synth.make_picture(job.sp, 50 + s) for s in range(job.sp.dpb_slots)]
init = refs[-1]
want = oracle_lib.Lib("oracle").replay(job, refs, out_init=init)
with backend.Context(job.sp) as ctx:
    for sparse in (False, True):
        got = ctx.reconstruct(job, refs, out_init=init, sparse=sparse)
        print("sparse" if sparse else "dense", "bit-exact" if all(np.array_equal(a, b) for a, b in zip(got, want)) else "MISMATCH")
