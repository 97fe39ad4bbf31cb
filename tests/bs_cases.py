"""Seeded cases for the boundary-strength derivation (bs on the device).

The frame cases' motion vectors are uniform over +-64 px, which makes nearly every PU boundary
"far"; these cases pull the vectors and reference indices of a frame together so that the motion
rule of boundary_strength (hevc_filter.c:764-880) meets its equal / near / swapped-list branches."""
import numpy as np

import frame_cases

BS_CASES = {
    # name: (frame case, seed, vector spread in quarter-pel, share of PUs that copy a neighbour's motion)
    "cfg1_near": ("cfg1_416x240_8b_420", 11, 6, 40),
    "cfg1_far": ("cfg1_416x240_8b_420", 12, 256, 0),
    "main10_near": ("main10_416x240_420_weighted", 13, 5, 50),
    "slices_tiles": ("slices_tiles_8b_420", 14, 6, 40),
    "slices_bypass": ("slices_10b_420_bypass", 15, 8, 30),
    "444_near": ("rext_10b_444", 16, 4, 60),
    "ctb32": ("ctb32_8b", 17, 6, 40),
    "ctb16_444": ("ctb16_10b_444", 18, 6, 40),
    "ragged": ("ragged_ctb_10b", 19, 6, 40),
    # no coded luma TB at all: every unit is decided by the motion rule
    "cfg1_near_nocbf": ("cfg1_416x240_8b_420", 22, 5, 50),
    "slices_tiles_nocbf": ("slices_tiles_8b_420", 23, 5, 50),
    # intra CUs between the inter ones: an edge with an intra side is 2 (hevc_filter.c:1030, :1062)
    "intra_near": ("intra_8b_420", 24, 6, 40),
    "intra_slices_tiles": ("intra_slices_tiles_8b", 25, 6, 40),
    "tiny": ("tiny_8x8", 20, 4, 50),
    "one_ctb": ("one_ctb", 21, 4, 50),
}


def build_case(name):
    """(job, deblock_disabled per CTB or None)"""
    base, seed, spread, copy_pct = BS_CASES[name]
    job, _ = frame_cases.build_case(base)
    rng = np.random.default_rng(seed)
    pus = job.pus
    n = len(pus)
    pus["mv"] = rng.integers(-spread, spread + 1, size=(n, 2, 2))
    pus["ref_idx"] = rng.integers(0, 2, size=(n, 2))
    # both lists of a slice may name the same pictures, so that the "same POC in L0 and L1" and the
    # swapped-list branches occur: list X index i -> slot (i + X * (slice parity)) % 2
    for si in range(len(job.slices)):
        for l in range(2):
            job.slices["ref_slot"][si, l, :] = (np.arange(16) + l * (si & 1) + (seed & 1)) % 2
    # a share of the PUs repeat the motion of the previous PU in the list (merge-like), some with swapped lists
    for i in range(1, n):
        r = rng.integers(0, 100)
        if r < copy_pct:
            pus["mv"][i] = pus["mv"][i - 1]
            pus["ref_idx"][i] = pus["ref_idx"][i - 1]
            if pus["w"][i] + pus["h"][i] != 12 or pus["pred_flag"][i - 1] != 3:
                pus["pred_flag"][i] = pus["pred_flag"][i - 1]
            if r < copy_pct // 4 and pus["pred_flag"][i] == 3:
                pus["mv"][i] = pus["mv"][i][::-1].copy()
                pus["ref_idx"][i] = pus["ref_idx"][i][::-1].copy()
            if r % 3 == 0:
                pus["mv"][i, rng.integers(0, 2), rng.integers(0, 2)] += rng.integers(-4, 5)
    if name.endswith("_nocbf"):
        job.tus = job.tus[job.tus["c_idx"] != 0].copy()
        job.leaves["cbf_luma"] = 0
    disabled = None
    if len(job.slices) > 1:                       # slice_deblocking_filter_disabled_flag on every third slice
        starts = np.unique(job.tab_slice_address)
        idx = np.searchsorted(starts, job.tab_slice_address)
        disabled = (idx % 3 == 2).astype(np.uint8)
    return job, disabled
