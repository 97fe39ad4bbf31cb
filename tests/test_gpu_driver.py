"""The plain-C multi-threaded caller of the ABI (hevc_b200/csrc/e2e_driver.c -> libohcu_drv.so): frame threads
write lists into the jobs' pinned arenas, flush and download, with no Python in the loop.  Every downloaded
picture is compared with the oracle; the chain mode (picture t predicts from t-1 and t-2 of its stream) checks
the inter-frame ordering on DPB slots under concurrent submission."""
import numpy as np
import pytest

from hevc_b200 import backend, driver, synth

pytestmark = pytest.mark.gpu


def _diff(a, b):
    for c, (x, y) in enumerate(zip(a, b)):
        if not np.array_equal(x, y):
            d = np.argwhere(x != y)
            return f"plane {c}: {len(d)} samples differ, first {d[:6].tolist()}"
    return None


def _setup(n_jobs=4, bd=10, cf=1, w=256, h=128, **kw):
    jobs = [synth.make_job(w, h, bd, cf, seed=140 + k, **kw) for k in range(n_jobs)]
    sp = jobs[0].sp
    refs = [synth.make_picture(sp, 170 + s) for s in range(2)]
    return jobs, sp, refs


@pytest.mark.parametrize("fill,coeffs", [(driver.FILL_COPY, "sparse"), (driver.FILL_COPY, "dense"), (driver.APPEND, "sparse"),
                                         (driver.APPEND, "dense")])
def test_independent_frames(fill, coeffs, oracle):
    jobs, sp, refs = _setup()
    E = 8
    want = []
    for jb in jobs:
        jb.sp.dpb_slots = 2 + E
        want.append(oracle.replay(jb, list(refs) + [None] * E))
    with backend.Context(sp, dpb_slots=2 + E, jobs=E) as ctx:
        ctx.upload(0, refs[0]); ctx.upload(1, refs[1])
        outs = [ctx.pinned_picture() for _ in range(E)]
        frames = driver.Frames(jobs, sp, coeffs=coeffs)
        res, last = driver.run(ctx, frames, total_frames=96, threads=4, ring=E, first_slot=2, fill=fill,
                               download=driver.DOWNLOAD_PLANES, outs=outs)
        assert res.frames == 96 and res.device_ms > 0
        for k in range(E):
            assert last[k] >= 0
            assert _diff(outs[k], want[last[k]]) is None, f"ring entry {k} (frame {last[k]}): {_diff(outs[k], want[last[k]])}"
        # the big lists found in place: the same pictures with only the small tables rewritten
        if fill == driver.FILL_COPY:
            res, _ = driver.run(ctx, frames, total_frames=E, threads=1, ring=E, first_slot=2, fill=driver.FILL_COPY,
                                frame_by_job=True, timed=False)
            for o in outs:
                for p in o:
                    p[:] = 0
            res, last = driver.run(ctx, frames, total_frames=64, threads=4, ring=E, first_slot=2, fill=driver.FILL_NONE,
                                   frame_by_job=True, download=driver.DOWNLOAD_PLANES, outs=outs)
            for k in range(E):
                assert _diff(outs[k], want[last[k]]) is None, f"in place, ring entry {k}: {_diff(outs[k], want[last[k]])}"


def test_derived_bs_and_444(oracle):
    jobs, sp, refs = _setup(n_jobs=3, cf=3, w=192, h=128, rext=1)
    E = 6
    want = []
    for jb in jobs:
        jb.sp.dpb_slots = 2 + E
        jb.vbs, jb.hbs = oracle.boundary_strengths(jb)
        want.append(oracle.replay(jb, list(refs) + [None] * E))
    with backend.Context(sp, dpb_slots=2 + E, jobs=E) as ctx:
        ctx.upload(0, refs[0]); ctx.upload(1, refs[1])
        outs = [ctx.pinned_cropped((0, 0, sp.width, sp.height)) for _ in range(E)]
        frames = driver.Frames(jobs, sp, coeffs="sparse", derive_bs=True)
        res, last = driver.run(ctx, frames, total_frames=48, threads=3, ring=E, first_slot=2,
                               download=driver.DOWNLOAD_CROPPED, outs=outs)
        for k in range(E):
            assert _diff(outs[k], want[last[k]]) is None, f"ring entry {k}: {_diff(outs[k], want[last[k]])}"


@pytest.mark.parametrize("streams,threads", [(1, 1), (2, 4), (4, 6)])
def test_chained_streams(streams, threads, oracle):
    """picture t of a stream predicts from the stream's pictures t-1 and t-2: ordering by events on DPB slots while
    several threads submit"""
    jobs, sp, refs = _setup(n_jobs=5, bd=8, n_refs=2)
    R, T = 3, 11
    ring = streams * R
    nslots = 2 + ring
    # expected pictures of one stream (all streams carry the same frames)
    pics = []
    for t in range(T):
        jb = jobs[t % len(jobs)]
        a = pics[t - 1] if t >= 1 else refs[0]
        b = pics[t - 2] if t >= 2 else refs[1]
        saved = jb.slices["ref_slot"].copy()
        for l in range(2):
            for i in range(16):
                jb.slices["ref_slot"][:, l, i] = 1 if (i + l) & 1 else 0
        jb.sp.dpb_slots = nslots
        pics.append(oracle.replay(jb, [a, b] + [None] * ring))
        jb.slices["ref_slot"][:] = saved
    with backend.Context(sp, dpb_slots=nslots, jobs=max(ring, threads)) as ctx:
        ctx.upload(0, refs[0]); ctx.upload(1, refs[1])
        outs = [ctx.pinned_picture() for _ in range(ring)]
        frames = driver.Frames(jobs, sp, coeffs="sparse")
        res, last = driver.run(ctx, frames, total_frames=streams * T, threads=threads, ring=ring, first_slot=2,
                               download=driver.DOWNLOAD_PLANES, outs=outs, chain_streams=streams, ref_slot=(0, 1))
        assert res.frames == streams * T
        for s in range(streams):
            for r in range(R):
                t = max(tt for tt in range(T) if tt % R == r)
                k = s * R + r
                assert _diff(outs[k], pics[t]) is None, f"stream {s}, picture {t}: {_diff(outs[k], pics[t])}"
