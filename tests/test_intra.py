"""Intra prediction: the oracle restatement (oracle/intra_oracle.c) pinned against the reference's
own compiled HEVCPredContext table (ff_hevc_pred_init, driven through oracle/ref_shim_pred.c), and against the
committed digests the reference produced (tests/golden/intra_digests.json).  CPU only."""
import json
import os

import intra_cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "intra_digests.json")


def test_oracle_matches_reference_intra_pred(oracle, reflib):
    bad = []
    groups = intra_cases.groups()
    for name, calls in groups:
        if intra_cases.run_group(oracle, calls, "oracle") != intra_cases.run_group(reflib, calls, "ref"):
            bad.append(name)
    assert not bad, f"{len(bad)}/{len(groups)} groups differ: {bad[:8]}"


def test_intra_goldens(oracle):
    gold = json.load(open(GOLDEN))["cases"]
    groups = intra_cases.groups()
    assert {n for n, _ in groups} == set(gold), "case list and fixture out of sync: rerun make_golden.py"
    bad = [n for n, calls in groups if intra_cases.run_group(oracle, calls, "oracle") != gold[n]]
    assert not bad, bad[:8]
