import sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import frame_cases, oracle_lib
from hevc_b200 import backend
name = sys.argv[1]; st = int(sys.argv[2])
job, refs = frame_cases.build_case(name)
with backend.Context(job.sp) as ctx:
    got = ctx.reconstruct(job, refs, out_init=refs[-1], stages=st)
want = oracle_lib.Lib("oracle").replay(job, refs, out_init=refs[-1], stages=st)
sao = job.sao
lc = job.sp.log2_ctb_size
ctb_w = (job.sp.width + (1 << lc) - 1) >> lc
for c, (g, w) in enumerate(zip(got, want)):
    d = np.argwhere(g != w)
    print("plane", c, "diff", len(d), d[:5].tolist())
    sh = lc - (1 if c and job.sp.chroma_format_idc == 1 else 0)
    stats = {}
    for y, x in d[:20000]:
        rs = (y >> sh) * ctb_w + (x >> sh)
        key = (int(sao["type_idx"][rs][c]), int(sao["eo_class"][rs][c]), int(x) & 7)
        stats[key] = stats.get(key, 0) + 1
    print(sorted(stats.items())[:60])
    for y, x in d[:6]:
        rs = (y >> sh) * ctb_w + (x >> sh)
        print(y, x, "got", g[y, x], "want", w[y, x], "type", sao["type_idx"][rs][c], "eo", sao["eo_class"][rs][c],
              "off", sao["offset_val"][rs][c], "band", sao["band_position"][rs][c])
