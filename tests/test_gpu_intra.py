"""Intra prediction on the device (ohcu_dsp_intra_pred -> intra_pred_tb, csrc/intra.cu) against the digests the
reference's compiled HEVCPredContext table produced (tests/golden/intra_digests.json) — every mode, size, availability
pattern, plane type and bit depth of tests/intra_cases.py — and block by block against the oracle on a sample."""
import json
import os

import numpy as np
import pytest

import intra_cases
from hevc_b200 import abi

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "intra_digests.json")


def _device(lib):
    def run(a, p):
        filter_ok = int(not p["smoothing_disabled"] and (p["c_idx"] == 0 or p["cfi"] == 3))
        rc = lib.ohcu_dsp_intra_pred(p["bd"], a.ctypes.data, a.strides[0], intra_cases.PW, intra_cases.PH, p["x"], p["y"], p["log2"],
                                     p["c_idx"], p["mode"], p["cand"], filter_ok, p["strong"])
        assert rc == 0, lib.ohcu_last_error()
    return run


@pytest.mark.parametrize("name", [n for n, _ in intra_cases.groups()])
def test_intra_pred_matches_reference_digests(name):
    lib = abi.load()
    gold = json.load(open(GOLDEN))["cases"]
    calls = dict(intra_cases.groups())[name]
    # the device shim costs a synchronous launch per block: every 5th call of the group, plus the oracle digest of the same subset
    sub = calls[::5]
    import oracle_lib
    orc = oracle_lib.Lib("oracle")
    assert intra_cases.run_group(lib, sub, _device(lib)) == intra_cases.run_group(orc, sub, "oracle"), name
    assert intra_cases.run_group(orc, calls, "oracle") == gold[name]          # and the oracle is the reference's (CPU, full group)


def test_intra_pred_argument_errors():
    lib = abi.load()
    a = np.zeros((64, 64), np.uint8)
    assert lib.ohcu_dsp_intra_pred(8, a.ctypes.data, 64, 64, 64, 0, 8, 3, 0, 26, abi_cand(left=1), 1, 0) == -3      # left neighbour outside
    assert lib.ohcu_dsp_intra_pred(8, a.ctypes.data, 64, 64, 64, 60, 8, 3, 0, 26, 0, 1, 0) == -3                    # block outside
    assert lib.ohcu_dsp_intra_pred(11, a.ctypes.data, 64, 64, 64, 8, 8, 3, 0, 26, 0, 1, 0) == -1


def abi_cand(bottom_left=0, left=0, up_left=0, up=0, up_right=0):
    return bottom_left | left << 1 | up_left << 2 | up << 3 | up_right << 4
