"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the committed
golden fixtures.  Bit-exact: all arithmetic on this path is integer."""
import json
import os

import numpy as np
import pytest

import frame_cases
from hevc_b200 import backend, synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STAGE_SETS = {"mc": 1, "mc_res": 3, "mc_res_db": 7, "all": 15}


def _diff(a, b):
    for c, (x, y) in enumerate(zip(a, b)):
        if not np.array_equal(x, y):
            d = np.argwhere(x != y)
            return f"plane {c}: {len(d)} samples differ, first {d[:6].tolist()}"
    return None


@pytest.mark.parametrize("name", list(frame_cases.FRAME_CASES))
def test_frame_parity(name, oracle):
    gold = json.load(open(os.path.join(GOLDEN, "frames.json")))["cases"][name]
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    with backend.Context(job.sp) as ctx:
        for key, st in STAGE_SETS.items():
            got = ctx.reconstruct(job, refs, out_init=init, stages=st)
            want = oracle.replay(job, refs, out_init=init, stages=st)
            assert _diff(got, want) is None, f"{name}/{key}: {_diff(got, want)}"
            assert frame_cases.picture_digest(got) == gold[key], f"{name}/{key} differs from the reference fixture"


@pytest.mark.parametrize("name", ["main10_416x240_420_weighted", "rext_10b_422", "rext_8b_444_bypass",
                                  "slices_tiles_8b_420", "slices_10b_420_bypass", "mono_8b"])
@pytest.mark.parametrize("stage", [2, 4, 8, 12])
def test_single_stage_on_given_picture(name, stage, oracle):
    """residual / deblock / SAO alone on a supplied picture (no MC)"""
    job, refs = frame_cases.build_case(name)
    init = refs[-1]
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=init, stages=stage)
    want = oracle.replay(job, refs, out_init=init, stages=stage)
    assert _diff(got, want) is None, _diff(got, want)


def test_cfg1_golden_picture():
    z = np.load(os.path.join(GOLDEN, "cfg1_416x240.npz"))
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, out_init=refs[-1])
    for c, p in enumerate(got):
        assert np.array_equal(p, z[f"plane{c}"])


def test_unordered_pu_list_is_sorted(oracle):
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    rng = np.random.default_rng(3)
    job.pus = job.pus[rng.permutation(len(job.pus))].copy()
    with backend.Context(job.sp) as ctx:
        got = ctx.reconstruct(job, refs, stages=1)
    want = oracle.replay(job, refs, stages=1)
    assert _diff(got, want) is None


def test_frames_in_flight_and_reference_chain(oracle):
    """frame k+1 predicts from frame k's output slot: ordering by events on DPB slots"""
    sp_job = synth.make_job(256, 128, 8, 1, seed=11)
    refs = [synth.make_picture(sp_job.sp, 50 + s) for s in range(2)]
    jobs = [synth.make_job(256, 128, 8, 1, seed=20 + k, n_refs=2) for k in range(4)]
    with backend.Context(sp_job.sp, dpb_slots=6, jobs=3) as ctx:
        ctx.upload(0, refs[0]); ctx.upload(1, refs[1])
        chain = [refs[0], refs[1]]
        for k, jb in enumerate(jobs):
            # slots: references are the two most recent pictures
            for s in range(len(jb.slices)):
                jb.slices["ref_slot"][s, :, :] = np.array([[k + 1 if i % 2 == 0 else k for i in range(16)],
                                                           [k if i % 2 == 0 else k + 1 for i in range(16)]])
            j = ctx.begin(k + 2)
            ctx.append(j, jb)
            ctx.flush(j)
        ctx.sync()
        outs = [ctx.download(k + 2) for k in range(4)]
    for k, jb in enumerate(jobs):
        dpb = [None] * 6
        for i, p in enumerate(chain):
            dpb[i] = p
        jb.sp.dpb_slots = 6
        want = oracle.replay(jb, dpb)
        assert _diff(outs[k], want) is None, f"frame {k}: {_diff(outs[k], want)}"
        chain.append(want)
