"""Oracle against the committed golden fixtures (generated from the compiled
reference by tests/golden/make_golden.py).  CPU only; runs on the GPU box too,
where /root/reference does not exist."""
import hashlib
import json
import os

import numpy as np
import pytest

import frame_cases
import leaf_cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STAGE_SETS = {"mc": 1, "mc_res": 3, "mc_res_db": 7, "all": 15}


def _load(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)["cases"]


def test_leaf_goldens(oracle):
    gold = _load("leaf_digests.json")
    cases = leaf_cases.all_cases()
    assert {n for n, _ in cases} == set(gold), "case list and fixture out of sync: rerun make_golden.py"
    bad = [n for n, run in cases if hashlib.sha256(run(oracle)).hexdigest()[:16] != gold[n]]
    assert not bad, f"{len(bad)} leaf cases differ from the reference fixture: {bad[:8]}"


@pytest.mark.parametrize("name", list(frame_cases.FRAME_CASES))
def test_frame_goldens(name, oracle):
    gold = _load("frames.json")[name]
    job, refs = frame_cases.build_case(name)
    lists = frame_cases.lists_digest(job)
    assert lists == gold["lists"], "generator output changed: rerun make_golden.py"
    for key, st in STAGE_SETS.items():
        out = oracle.replay(job, refs, out_init=refs[-1], stages=st)
        assert frame_cases.picture_digest(out) == gold[key], f"{name}/{key}"


@pytest.mark.parametrize("name", list(frame_cases.FULLSIZE_CASES))
def test_fullsize_goldens(name, oracle):
    """BASELINE.json configs[1..4] at their named sizes: the restatement against the reference's digests."""
    gold = _load("frames_fullsize.json")[name]
    job, refs = frame_cases.build_case(name)
    assert (len(job.pus), len(job.tus), len(job.coeffs)) == (gold["n_pu"], gold["n_tu"], gold["ncoeff"]), \
        "generator output changed: rerun make_golden.py"
    out = oracle.replay(job, refs, out_init=refs[-1], stages=15)
    assert frame_cases.picture_digest(out) == gold["all"], name


def test_cfg1_golden_picture(oracle):
    """BASELINE.json configs[0]: the 416x240 8-bit 4:2:0 golden frame, byte for byte."""
    z = np.load(os.path.join(GOLDEN, "cfg1_416x240.npz"))
    job, refs = frame_cases.build_case("cfg1_416x240_8b_420")
    out = oracle.replay(job, refs, out_init=refs[-1])
    for c, p in enumerate(out):
        assert np.array_equal(p, z[f"plane{c}"])
