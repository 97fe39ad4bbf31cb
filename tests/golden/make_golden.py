"""Regenerates tests/golden/* from the REFERENCE's own compiled C code
(oracle/_ref/liboh_ref.so, built from /root/reference by oracle/Makefile).
Run in the build container (needs /root/reference):  python tests/golden/make_golden.py

  leaf_digests.json   sha256[:16] of every output of every leaf case (tests/leaf_cases.py)
  frames.json         sha256 of the reconstructed picture of every frame case
                      (tests/frame_cases.py) after MC, MC+residual, and the full pipeline
  frames_fullsize.json  the same for BASELINE configs[1..4] at their named sizes (1080p 8-bit, 4K Main10,
                      1080p 10-bit 4:2:2 / 4:4:4, 8K Main10), MC+residual and the full pipeline
  intra_digests.json  intra prediction, every mode x size x availability pattern (tests/intra_cases.py)
  cfg1_416x240.npz    the full golden picture of BASELINE.json configs[0]
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import frame_cases  # noqa: E402
import leaf_cases  # noqa: E402
import levels_cases  # noqa: E402
import bs_cases  # noqa: E402
import intra_cases  # noqa: E402
import oracle_lib  # noqa: E402

STAGE_SETS = {"mc": 1, "mc_res": 3, "mc_res_db": 7, "all": 15}


def main():
    ref = oracle_lib.Lib("ref")
    leaves = {name: hashlib.sha256(run(ref)).hexdigest()[:16] for name, run in leaf_cases.all_cases()}
    with open(os.path.join(HERE, "leaf_digests.json"), "w") as f:
        json.dump({"source": "oracle/_ref/liboh_ref.so (reference C table)", "cases": leaves}, f,
                  indent=0, sort_keys=True)
    frames = {}
    for name in frame_cases.FRAME_CASES:
        job, refs = frame_cases.build_case(name)
        frames[name] = {
            "n_pu": int(len(job.pus)), "n_tu": int(len(job.tus)), "ncoeff": int(len(job.coeffs)),
            "lists": frame_cases.lists_digest(job),
        }
        for key, st in STAGE_SETS.items():
            out = ref.replay(job, refs, out_init=refs[-1], stages=st)
            frames[name][key] = frame_cases.picture_digest(out)
            if name.startswith("cfg1_") and key == "all":
                np.savez_compressed(os.path.join(HERE, "cfg1_416x240.npz"),
                                    **{f"plane{c}": p for c, p in enumerate(out)})
    with open(os.path.join(HERE, "frames.json"), "w") as f:
        json.dump({"source": "ohref_replay_frame (reference drivers + table)", "cases": frames}, f,
                  indent=1, sort_keys=True)
    # BASELINE configs at their named sizes: the reference's picture after MC + residual and after the full pipeline
    full = {}
    for name in frame_cases.FULLSIZE_CASES:
        job, refs = frame_cases.build_case(name)
        full[name] = {"n_pu": int(len(job.pus)), "n_tu": int(len(job.tus)), "ncoeff": int(len(job.coeffs))}
        for key in ("mc_res", "all"):
            full[name][key] = frame_cases.picture_digest(ref.replay(job, refs, out_init=refs[-1], stages=STAGE_SETS[key]))
    with open(os.path.join(HERE, "frames_fullsize.json"), "w") as f:
        json.dump({"source": "ohref_replay_frame (reference drivers + table)", "cases": full}, f, indent=1, sort_keys=True)
    # dequantisation (the levels form): the reference's own scale_and_clip_coeff on seeded blocks
    sl, sl_dc, cases = levels_cases.leaf_cases()
    outs = []
    for p, blk in cases:
        b = blk.copy()
        ref._dequant_block(p["bd"], p["cfi"], p["c_idx"], p["log2"], p["qp_y"], p["qp_offset"], p["intra"],
                           p["transform_skip"], p["sl_enabled"], sl.ctypes.data, sl_dc.ctypes.data, b.ctypes.data)
        outs.append(b)
    with open(os.path.join(HERE, "levels_digest.json"), "w") as f:
        json.dump({"source": "ohref_dequant_block (hevc_cabac.c derive_quant_parameters{,_c} + scale_and_clip_coeff)",
                   "cases": len(outs), "sha256": levels_cases.digest(outs)}, f, indent=1)
    # boundary strengths: the reference's ff_hevc_deblocking_boundary_strengths over the transform-tree leaves
    bs = {}
    for name in bs_cases.BS_CASES:
        job, disabled = bs_cases.build_case(name)
        vbs, hbs = ref.boundary_strengths(job, disabled)
        bs[name] = hashlib.sha256(vbs.tobytes() + hbs.tobytes()).hexdigest()[:16]
    with open(os.path.join(HERE, "bs_digests.json"), "w") as f:
        json.dump({"source": "ohref_boundary_strengths (hevc_filter.c ff_hevc_deblocking_boundary_strengths per leaf)",
                   "cases": bs}, f, indent=1, sort_keys=True)
    # intra prediction: the reference's HEVCPredContext table (ff_hevc_pred_init) through oracle/ref_shim_pred.c
    intra = {name: intra_cases.run_group(ref, calls, "ref") for name, calls in intra_cases.groups()}
    with open(os.path.join(HERE, "intra_digests.json"), "w") as f:
        json.dump({"source": "ohref_intra_pred (HEVCPredContext.intra_pred of ff_hevc_pred_init, hevcpred_template.c)",
                   "cases": intra}, f, indent=1, sort_keys=True)
    print(f"{len(leaves)} leaf digests, {len(frames)} frame cases, {len(outs)} dequantisation cases written")


if __name__ == "__main__":
    main()
