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
