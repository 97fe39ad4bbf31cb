"""Seeded frame-level cases shared by the oracle, golden and GPU parity tests.

FRAME_CASES: name -> (width, height, bit_depth, chroma_format_idc, seed, generator overrides)
cfg1 is BASELINE.json configs[0] (416x240 8-bit 4:2:0, the golden frame).
"""
import hashlib

import numpy as np

from hevc_b200 import synth

FRAME_CASES = {
    "cfg1_416x240_8b_420": (416, 240, 8, 1, 1001, {}),
    "cfg1w_416x240_8b_420_weighted": (416, 240, 8, 1, 1002, dict(weighted=1)),
    "main10_416x240_420_weighted": (416, 240, 10, 1, 3001, dict(weighted=1)),
    "rext_10b_422": (320, 184, 10, 2, 4001, dict(rext=1, weighted=1)),
    "rext_10b_444": (320, 184, 10, 3, 4002, dict(rext=1)),
    "rext_8b_444_bypass": (256, 128, 8, 3, 4003, dict(rext=1, sp_pcm_or_bypass=1, bypass_pct=10)),
    "slices_tiles_8b_420": (448, 256, 8, 1, 5001, dict(n_slices=4, slices_no_lf=1, tile_cols=3, tile_rows=2,
                                                       sp_no_tile_filter=1)),
    "slices_10b_420_bypass": (320, 192, 10, 1, 5002, dict(n_slices=3, slices_no_lf=1, sp_pcm_or_bypass=1,
                                                         bypass_pct=12)),
    "mono_8b": (192, 128, 8, 0, 6001, {}),
    "main12_420": (256, 136, 12, 1, 6002, dict(weighted=1)),
    "bd9_420": (256, 136, 9, 1, 6003, {}),
    "main14_420": (256, 136, 14, 1, 6004, {}),
    "main14_444_weighted": (192, 128, 14, 3, 6005, dict(weighted=1, rext=1)),
    "dense_coeffs_8b": (256, 128, 8, 1, 7001, dict(dense_coeffs=1)),
    "dense_coeffs_10b_444": (192, 128, 10, 3, 7002, dict(dense_coeffs=1, rext=1)),
    "far_mv_8b": (320, 176, 8, 1, 7003, dict(far_mv_pct=60)),
    "ragged_ctb_10b": (200, 72, 10, 1, 7004, dict(weighted=1)),
    "tiny_8x8": (8, 8, 8, 1, 7005, {}),
    "one_ctb": (64, 64, 10, 1, 7006, dict(weighted=1)),
    "ctb32_8b": (288, 160, 8, 1, 7007, dict(sp_log2_ctb_size=5, split_pct=(0, 50, 40))),
    "ctb16_10b_444": (160, 96, 10, 3, 7008, dict(sp_log2_ctb_size=4, split_pct=(0, 0, 40))),
    "cfg2_small_1080like": (480, 272, 8, 1, 2001, {}),
    # sao_edge_restore[0 / 1] (hevcdsp_template.c:380-529) directed: 3x3 .. 4x3 CTBs, SAO edge in every CTB (all four
    # classes), so that every CTB meets the picture-border rule of its class and, with slices that switch
    # loop_filter_across_slices off and tiles that are not filtered across, every vert / horiz / diag restore flag
    "sao_restore_borders_8b": (192, 192, 8, 1, 8001, dict(sao_pct=(0, 100))),
    "sao_restore_slices_10b": (192, 192, 10, 1, 8002, dict(sao_pct=(0, 100), n_slices=5, slices_no_lf=1)),
    "sao_restore_tiles_8b_444": (256, 192, 8, 3, 8003, dict(sao_pct=(0, 100), tile_cols=2, tile_rows=3, sp_no_tile_filter=1)),
    "sao_restore_slices_tiles_10b_422": (256, 192, 10, 2, 8004, dict(sao_pct=(10, 90), n_slices=4, slices_no_lf=1, tile_cols=3, tile_rows=2,
                                                                     sp_no_tile_filter=1, rext=1)),
    # intra-predicted CUs between the inter ones: ohcu_iu records in hls_transform_unit order, their
    # coded TBs flagged OHCU_TUF_INTRA; prediction reads what was reconstructed before it (CTB wavefront on the device)
    "intra_8b_420": (416, 240, 8, 1, 9001, dict(intra_pct=35)),
    "intra_only_10b_420_strong": (256, 192, 10, 1, 9002, dict(intra_pct=100, sp_strong_intra_smoothing=1)),
    "intra_10b_422_rext": (320, 184, 10, 2, 9003, dict(intra_pct=40, rext=1)),
    "intra_8b_444_rext_bypass": (256, 128, 8, 3, 9004, dict(intra_pct=50, rext=1, sp_pcm_or_bypass=1, bypass_pct=10)),
    "intra_slices_tiles_8b": (448, 256, 8, 1, 9005, dict(intra_pct=40, n_slices=4, slices_no_lf=1, tile_cols=3, tile_rows=2,
                                                         sp_no_tile_filter=1, intra_drop_pct=15)),
    "intra_mono_12b_no_smoothing": (192, 128, 12, 0, 9006, dict(intra_pct=60, sp_intra_smoothing_disabled=1)),
    "intra_ctb32_8b": (288, 160, 8, 1, 9007, dict(intra_pct=50, sp_log2_ctb_size=5, split_pct=(0, 50, 40))),
    "intra_ctb16_10b_444": (160, 96, 10, 3, 9008, dict(intra_pct=50, sp_log2_ctb_size=4, split_pct=(0, 0, 40))),
    "dense_coeffs_intra_only_12b_444": (128, 128, 12, 3, 9009, dict(intra_pct=100, dense_coeffs=1, rext=1)),
}

# BASELINE.json configs[1..4] at their named sizes (one picture each): grids of 510 - 8160 CTBs,
# pitches beyond 8 KB, coordinates up to 7680, lists of 10^5 - 10^6 records.
FULLSIZE_CASES = {
    "cfg2_1920x1080_8b_420": (1920, 1080, 8, 1, 2002, {}),
    "cfg3_3840x2160_10b_420_weighted": (3840, 2160, 10, 1, 3002, dict(weighted=1)),
    "cfg4a_1920x1080_10b_422": (1920, 1080, 10, 2, 4004, dict(rext=1, weighted=1)),
    "cfg4b_1920x1080_10b_444": (1920, 1080, 10, 3, 4005, dict(rext=1)),
    "cfg5_7680x4320_10b_420_weighted": (7680, 4320, 10, 1, 5005, dict(weighted=1)),
    # a 4K P picture with a quarter of its CUs intra, and an all-intra 1080p picture (the wavefront at size)
    "cfg3i_3840x2160_10b_420_intra25": (3840, 2160, 10, 1, 3003, dict(weighted=1, intra_pct=25)),
    "cfg2i_1920x1080_8b_420_all_intra": (1920, 1080, 8, 1, 2003, dict(intra_pct=100)),
}


def build_case(name):
    w, h, bd, cf, seed, kw = (FRAME_CASES.get(name) or FULLSIZE_CASES[name])
    job = synth.make_job(w, h, bd, cf, seed=seed, **kw)
    refs = [synth.make_picture(job.sp, seed * 10 + s) for s in range(job.sp.dpb_slots)]
    return job, refs


def lists_digest(job):
    h = hashlib.sha256(job.pus.tobytes() + job.tus.tobytes() + job.coeffs.tobytes())
    if len(job.ius):
        h.update(job.ius.tobytes())
    return h.hexdigest()


def picture_digest(planes):
    h = hashlib.sha256()
    for p in planes:
        h.update(np.ascontiguousarray(p).tobytes())
    return h.hexdigest()
