"""GPU parity at the granularity of the reference's table: every HEVCDSPContext leaf the
path uses, through the per-entry shims (hevc_b200.dsp -> ohcu_dsp_*, same device code as the
frame kernels), on the seeded cases of leaf_cases.py — the cases that pin the oracle against
the compiled reference (test_oracle_vs_ref.py) and whose digests are committed in
tests/golden/leaf_digests.json.  Bit-exact."""
import hashlib
import json
import os

import pytest

import leaf_cases

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "leaf_digests.json")


def _cases():
    # sao_edge_restore has no per-entry shim (border rules live in the frame kernel)
    return [(n, f) for n, f in leaf_cases.all_cases() if "_r0_" not in n and "_r1_" not in n]


def test_every_leaf_matches_the_oracle(oracle):
    from hevc_b200 import dsp
    gpu = dsp.Dsp()
    bad = []
    cases = _cases()
    for name, run in cases:
        if run(gpu) != run(oracle):
            bad.append(name)
    assert not bad, f"{len(bad)} of {len(cases)} leaf cases differ, first: {bad[:8]}"


def test_leaves_match_the_committed_reference_digests():
    from hevc_b200 import dsp
    gpu = dsp.Dsp()
    golden = json.load(open(GOLDEN))["cases"]
    bad = []
    n = 0
    for name, run in _cases():
        if name in golden:
            n += 1
            if hashlib.sha256(run(gpu)).hexdigest()[:16] != golden[name]:
                bad.append(name)
    assert n > 1000 and not bad, f"{len(bad)} of {n} differ from the reference digests, first: {bad[:8]}"
