"""Boundary strengths derived from the work lists (bs on the device).

CPU: the oracle's derivation from the PU list + coded luma TBs against the reference's own
ff_hevc_deblocking_boundary_strengths run over the transform-tree leaves, and against the committed
digest of the reference's planes.  GPU: the device derivation against the oracle, and a whole frame
deblocked with derived bs against the oracle replay fed the same planes."""
import hashlib
import json
import os

import numpy as np
import pytest

import bs_cases
import oracle_lib

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "bs_digests.json")


def _digest(vbs, hbs):
    return hashlib.sha256(vbs.tobytes() + hbs.tobytes()).hexdigest()[:16]


@pytest.mark.parametrize("name", list(bs_cases.BS_CASES))
def test_oracle_bs_matches_committed_reference_digest(name):
    job, disabled = bs_cases.build_case(name)
    vbs, hbs = oracle_lib.Lib("oracle").boundary_strengths(job, disabled)
    assert _digest(vbs, hbs) == json.load(open(GOLDEN))["cases"][name]


@pytest.mark.skipif(not oracle_lib.have_ref(), reason="needs the compiled reference (oracle/_ref)")
@pytest.mark.parametrize("name", list(bs_cases.BS_CASES))
def test_oracle_bs_vs_reference(name):
    job, disabled = bs_cases.build_case(name)
    got = oracle_lib.Lib("oracle").boundary_strengths(job, disabled)
    want = oracle_lib.Lib("ref").boundary_strengths(job, disabled)
    for g, w, what in zip(got, want, ("vertical_bs", "horizontal_bs")):
        bad = np.flatnonzero(g != w)
        assert len(bad) == 0, f"{name} {what}: {len(bad)} units differ, first at {bad[:5].tolist()} ours {g[bad[:5]].tolist()} ref {w[bad[:5]].tolist()}"
    if name.endswith("_nocbf"):                             # both outcomes of the motion rule between different PUs
        ones = int(np.count_nonzero(np.concatenate(want) == 1))
        far, _ = bs_cases.build_case(name)
        far.pus["mv"][:, :, 0] += (np.arange(len(far.pus)) % 7 * 64)[:, None].astype(np.int16)
        all_far = int(np.count_nonzero(np.concatenate(oracle_lib.Lib("ref").boundary_strengths(far, disabled)) == 1))
        assert 0 < ones < all_far * 0.9, (ones, all_far)


def _grid_units(sp):
    """indices of the plane entries the filters read: vertical edges x % 8 == 0, horizontal y % 8 == 0"""
    bsw = (sp.width >> 2) + 1
    ys, xs = np.meshgrid(np.arange(0, sp.height, 4), np.arange(8, sp.width, 8), indexing="ij")
    v = ((xs + ys * bsw) >> 2).ravel()
    ys, xs = np.meshgrid(np.arange(8, sp.height, 8), np.arange(0, sp.width, 4), indexing="ij")
    h = ((xs + ys * bsw) >> 2).ravel()
    return v, h


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(bs_cases.BS_CASES))
def test_device_bs_matches_oracle_and_filters_with_it(name):
    from hevc_b200 import backend
    job, disabled = bs_cases.build_case(name)
    base = bs_cases.BS_CASES[name][0]
    import frame_cases
    _, refs = frame_cases.build_case(base)
    orc = oracle_lib.Lib("oracle")
    vbs, hbs = orc.boundary_strengths(job, disabled)
    # the oracle replay filters with the derived planes ...
    job.vbs, job.hbs = vbs, hbs
    want = orc.replay(job, refs)
    # ... and so must the device, which is given none
    with backend.Context(job.sp) as ctx:
        out_slot = job.sp.dpb_slots - 1
        for i, r in enumerate(refs):
            if i != out_slot:
                ctx.upload(i, r)
        j = ctx.begin(out_slot)
        ctx.append(j, job, sparse=True, derive_bs=True if disabled is None else disabled)
        ctx.flush(j)
        ctx.sync(j)
        got = ctx.download(out_slot)
        gv, gh = np.zeros_like(vbs), np.zeros_like(hbs)
        backend.check(ctx.lib.ohcu_job_read_bs(ctx.handle, j, gv.ctypes.data, gh.ctypes.data), "ohcu_job_read_bs")
    iv, ih = _grid_units(job.sp)
    assert np.array_equal(gv[iv], vbs[iv]), f"{name}: vertical_bs differs at {np.flatnonzero(gv[iv] != vbs[iv])[:5].tolist()}"
    assert np.array_equal(gh[ih], hbs[ih]), f"{name}: horizontal_bs differs at {np.flatnonzero(gh[ih] != hbs[ih])[:5].tolist()}"
    for c, (g, w) in enumerate(zip(got, want)):
        assert np.array_equal(g, w), f"{name} plane {c}"
