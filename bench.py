#!/usr/bin/env python
"""bench.py — HEVC pixel-reconstruction hot path on B200: Mpixel/s reconstructed per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload 4k|1080p|8k]

A step = one pass of the hot path (MC + residual, then deblock + SAO: two fused kernels
and the halo fill per frame) over one batch of synthetic frames.  Default workload: BASELINE.json configs[2], 3840x2160 Main10 4:2:0 full
reconstruction with weighted prediction.

  value        whole-job Mpixel/s (luma W*H*frames / s), work lists resident in HBM
  e2e          same metric through the C ABI with HOST buffers: list upload from pinned
               memory, kernels, and the reconstructed picture copied back to pinned memory
  roofline     the dominant kernel (mc_kernel: motion compensation) against the measured HBM peak
  cpu_baseline the reference's own CPU path on the host cores: its C drivers + its SSE intrinsics (oracle/_ref/
               liboh_ref_simd.so, the fastest reference path buildable here), with the pure-C arms beside it
  --impl reference   times that CPU path alone, same metric / config

One process per GPU (torchrun for N > 1), frames sharded across ranks, no collective on
the data path ("replicas only" sharding of independent frames); timing = max over ranks.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (width, height, bit_depth, chroma_format_idc, generator overrides, frames per step per GPU)
    "4k": (3840, 2160, 10, 1, dict(weighted=1), 8),
    "1080p": (1920, 1080, 8, 1, dict(), 16),
    "8k": (7680, 4320, 10, 1, dict(weighted=1), 8),
}
WORKLOAD_NAMES = {
    "4k": "3840x2160 Main10 4:2:0 full reconstruction, weighted prediction (BASELINE configs[2])",
    "1080p": "1920x1080 Main 8-bit 4:2:0 (BASELINE configs[1])",
    "8k": "7680x4320 Main10 4:2:0, 64 independent frames sharded across GPUs (BASELINE configs[4])",
}
N_DISTINCT = 4        # distinct synthetic frames; jobs reuse them round-robin
# dram__bytes_read.sum + dram__bytes_write.sum of one mc_kernel launch, from the committed
# `ncu --set full` capture of this same command (never measured inside a timed run)
NCU_TRAFFIC = {"4k": 59145984}
NCU_TRAFFIC_SOURCE = "profiles/r01_v13_ncu_kernels.txt (53.62 MB read + 5.53 MB write; the reference pictures and part of the written prediction stay in L2)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def make_inputs(workload, rank, world=1):
    """The N_DISTINCT * world synthetic frames of the job are sharded frame i -> rank i mod world."""
    from hevc_b200 import shard, synth
    w, h, bd, cf, kw, _ = WORKLOADS[workload]
    mine = shard.frames_of_rank(N_DISTINCT * world, rank, world)
    jobs = [synth.make_job(w, h, bd, cf, seed=shard.seed_of_frame(3000, i), **kw) for i in mine]
    refs = [synth.make_picture(jobs[0].sp, 7000 + s) for s in range(2)]
    return jobs, refs


def levels_inputs(jobs):
    """--coeffs levels: the frames' coefficient arenas become UN-scaled levels with a random quantiser
    per block and random scaling lists (tests/levels_cases.py); job.coeffs is replaced by the scaled
    coefficients (CPU restatement of scale_and_clip_coeff), which is what the parity check and the CPU
    arms consume - the GPU arm is handed the levels (ohcu_append_tu_levels) and scales them itself.
    Returns per job (recs, quant, pairs, (sl, sl_dc))."""
    import levels_cases
    import oracle_lib
    from hevc_b200.sparse import to_levels
    orc = oracle_lib.Lib("oracle")
    out = []
    for i, fj in enumerate(jobs):
        sp = fj.sp
        hs = 1 if sp.chroma_format_idc in (1, 2) else 0
        vs = 1 if sp.chroma_format_idc == 1 else 0
        ctb_w = (sp.width + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
        lv, quant, scaling = levels_cases.frame_levels(fj, seed=900 + i)
        fj.coeffs = orc.dequant_tus(fj.tus, quant, lv, sp.bit_depth, scaling)
        out.append(to_levels(fj.tus, quant, lv, sp.log2_ctb_size, hs, vs, ctb_w) + (scaling,))
    return out


def derived_bs_inputs(jobs):
    """--bs derived: the frames' bs planes are replaced by the ones the work lists imply (CPU restatement of
    ff_hevc_deblocking_boundary_strengths, oracle/replay.c) - what the parity check and the CPU arms filter
    with; the GPU arm is handed no planes and derives them itself (ohcu_set_deblock_derived)."""
    import oracle_lib
    orc = oracle_lib.Lib("oracle")
    for fj in jobs:
        fj.vbs, fj.hbs = orc.boundary_strengths(fj)


def algorithmic_bytes(job):
    """per kernel, bytes that must move once.
    mc          every reference sample a PU needs read once + every predicted sample written once
    residual    the int16 coefficients of the coded blocks read once + the samples of the coded
                blocks read and written once
    loopfilter  the reconstructed picture read once + the filtered picture written once
    pipeline    mc + residual + loopfilter.  pipeline_survey is SURVEY's figure for the layout it
                assumed (MC, residual RMW, deblock and SAO as four passes)."""
    sp = job.sp
    P = 2 if sp.bit_depth > 8 else 1
    hs = 1 if sp.chroma_format_idc in (1, 2) else 0
    vs = 1 if sp.chroma_format_idc == 1 else 0
    nsamp = sp.width * sp.height + (2 * (sp.width >> hs) * (sp.height >> vs) if sp.chroma_format_idc else 0)
    S = nsamp * P
    area = job.pus["w"].astype(np.int64) * job.pus["h"]
    planes_factor = nsamp / (sp.width * sp.height)
    nref = np.where(job.pus["pred_flag"] == 3, 2, 1)
    mc = int((area * (1 + nref)).sum() * planes_factor * P)        # read each ref sample once + write pred
    coded = int((1 << (2 * job.tus["log2_size"].astype(np.int64))).sum())
    residual = coded * 2 + coded * P * 2
    return {"mc": mc, "residual": residual, "loopfilter": 2 * S, "pad": 0, "pipeline": mc + residual + 2 * S,
            "pipeline_survey": mc + coded * 2 + coded * P * 2 + 4 * S, "S": S}


class ClockSampler:
    def __init__(self, device):
        self.device = device
        self.rows = []
        self.stop = threading.Event()
        self.th = threading.Thread(target=self.run, daemon=True)

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.device)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm = [int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit()]
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------ reference arm
def cpu_arm(jobs, refs, steps, warmup, which, threads):
    """One CPU arm: `threads` host threads, each replaying one frame per step through the library `which`
    (ref_simd: the reference's C drivers + its SSE intrinsics, -O3; ref: its pure C table, -O2; oracle: the port)."""
    import oracle_lib
    from concurrent.futures import ThreadPoolExecutor
    lib = oracle_lib.Lib(which)
    dpb = list(refs) + [None] * (jobs[0].sp.dpb_slots - len(refs))

    def one(i):
        lib.replay(jobs[i % len(jobs)], dpb)

    pool = ThreadPoolExecutor(threads)
    for _ in range(warmup):
        list(pool.map(one, range(threads)))
    t0 = time.perf_counter()
    for _ in range(steps):
        list(pool.map(one, range(threads)))
    dt = time.perf_counter() - t0
    pool.shutdown()
    sp = jobs[0].sp
    frames = steps * threads
    return {"value": frames * sp.width * sp.height / dt / 1e6, "frames": frames, "seconds": dt, "ms_per_step": dt / steps * 1e3}


def cpu_reference(jobs, refs, steps, warmup, max_threads=None, all_arms=False):
    """The reference's own CPU path on the host cores.  The headline arm is the FASTEST reference path this box can
    build: its C drivers with its SSE intrinsics (x86/hevc_mc_sse.c, hevc_idct_sse.c, hevc_sao_sse.c) wired into the
    table, -O3, one frame per host thread on all cores ("reference-simd").  The yasm sources of the reference are
    unavailable on this box (no yasm / nasm), so deblocking runs its C entries.  all_arms adds the pure-C table on one
    thread and on all threads."""
    import oracle_lib
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    if max_threads:
        cores = min(cores, max_threads)
    have = lambda so: os.path.exists(so) or os.path.isdir(oracle_lib.REFERENCE_ROOT)
    if have(oracle_lib.REF_SIMD_SO):
        kind, which = "reference-simd", "ref_simd"
    elif have(oracle_lib.REF_SO):
        kind, which = "reference", "ref"
    else:
        kind, which = "port", "oracle"
    main = cpu_arm(jobs, refs, steps, warmup, which, cores)
    out = {"value": main["value"], "unit": "Mpixel/s", "cores": cores, "kind": kind,
           "sample": f"{main['frames']} frames ({cores} threads x {steps} steps, one frame per thread per step)",
           "ms_per_step": main["ms_per_step"], "seconds": main["seconds"]}
    if all_arms:
        arms = {f"{kind}_{cores}t": round(main["value"], 2)}
        if which == "ref_simd":
            arms["reference-simd_1t"] = round(cpu_arm(jobs, refs, max(1, steps), 1, "ref_simd", 1)["value"], 2)
        if have(oracle_lib.REF_SO):
            arms[f"reference-c_{cores}t"] = round(cpu_arm(jobs, refs, max(1, steps // 2), 1, "ref", cores)["value"], 2)
            arms["reference-c_1t"] = round(cpu_arm(jobs, refs, 1, 1, "ref", 1)["value"], 2)
        out["arms"] = arms
    return out


# ------------------------------------------------------------------ our arm
def run_ours(args, rank, world):
    from hevc_b200 import abi, backend, shard
    jobs, refs = make_inputs(args.workload, rank, world)
    sp = jobs[0].sp
    B = WORKLOADS[args.workload][5]
    if args.frames:
        B = args.frames
    device = int(os.environ.get("LOCAL_RANK", 0))
    hbm_peak, peak_kind = peaks()
    # the e2e leg keeps 2B frames in flight (B being filled / uploaded while B run and drain)
    if B % N_DISTINCT:
        raise SystemExit(f"--frames must be a multiple of {N_DISTINCT} (the job ring is revisited with the same frame contents)")
    E = min(2 * B, 28)                       # ring depth of the e2e leg (2 reference slots + E <= 32 DPB slots)
    ctx = backend.Context(sp, device=device, dpb_slots=2 + max(B, E), jobs=max(B, E))
    lib, h = ctx.lib, ctx.handle
    for s in range(2):
        ctx.upload(s, refs[s])

    barrier = shard.barrier
    # --coeffs sparse: significant coefficients as (position, level) pairs (ohcu_append_tu_sparse)
    sparse_lists = [ctx.sparse_lists(fj) for fj in jobs] if args.coeffs == "sparse" else None
    # --coeffs levels: the pairs carry un-scaled levels, the kernel dequantises (ohcu_append_tu_levels)
    level_lists = levels_inputs(jobs) if args.coeffs == "levels" else None
    derive = args.bs == "derived"
    if derive:
        derived_bs_inputs(jobs)

    # ---------------- resident leg: lists staged once, kernels only in the timed region
    for k in range(B):
        j = ctx.begin(2 + k)
        assert j == k
        if level_lists:
            recs, q, pairs, scaling = level_lists[k % len(jobs)]
            ctx.append(j, jobs[k % len(jobs)], levels=(q, scaling, (recs, q, pairs)), derive_bs=derive)
        else:
            ctx.append(j, jobs[k % len(jobs)], sparse=sparse_lists[k % len(jobs)] if sparse_lists else False, derive_bs=derive)
        ctx.stage(j)
    ctx.sync()
    l0 = ctx.kernel_launches()
    for _ in range(args.warmup):
        for k in range(B):
            ctx.launch(k, 2 + k)
    ctx.sync()
    barrier()
    l1 = ctx.kernel_launches()
    with ClockSampler(device) as clk:
        ctx.timer_start()
        for _ in range(args.steps):
            for k in range(B):
                ctx.launch(k, 2 + k)
        ms_resident = ctx.timer_stop()
    barrier()
    launches = ctx.kernel_launches() - l1
    clocks = clk.summary()

    # ---------------- per-stage timing of single frames (events on the job's own stream)
    ctx.set_profiling(True)
    stage_ms = np.zeros(abi.NUM_STAGES)
    nprof = max(4, min(16, args.steps * B))
    for i in range(nprof):
        k = i % B
        ctx.launch(k, 2 + k)
        ctx.sync(k)
        stage_ms += np.array(ctx.stage_ms(k))
    stage_ms /= nprof
    ctx.set_profiling(False)

    # ---------------- e2e leg: host buffers -> H2D -> kernels -> D2H, through the C ABI
    full = (0, 0, sp.width, sp.height)
    if args.copy_out == "packed":
        outs = [ctx.pinned_cropped(full) for _ in range(E)]
    else:
        outs = [ctx.pinned_picture(slot_pitch=args.slot_pitch or args.copy_out == "pitch") for _ in range(E)]
    h2d = d2h = 0

    # which frame's coefficients a job arena holds (by arena address).  The submitting threads
    # take jobs from the ring in whatever order they reach ohcu_frame_begin, so a frame can land
    # on an arena that holds another frame's coefficients: then they are written again
    arena_content = {}
    rewrites = [0]

    def in_place(ptr, cidx, src, copy_coeffs):
        if copy_coeffs or arena_content.get(ptr.value) != cidx:
            C.memmove(ptr.value, src.ctypes.data, src.nbytes)
            arena_content[ptr.value] = cidx
            if not copy_coeffs:
                rewrites[0] += 1

    def submit(i, copy_coeffs=False):
        nonlocal h2d, d2h
        k = i % E                            # output slot / pinned output buffer of this frame
        cidx = k % len(jobs)
        fj = jobs[cidx]
        j = ctx.begin(2 + k)
        for i in range(len(fj.slices)):
            tab = abi.SliceTab.from_buffer_copy(fj.slices[i].tobytes())
            abi.check(lib.ohcu_set_slice(h, j, i, C.byref(tab)))
        abi.check(lib.ohcu_append_pu(h, j, fj.pus.ctypes.data, len(fj.pus)))
        # the residual parser writes coefficients straight into the job's pinned arena
        # (ohcu_reserve_coeffs, lc->tu.coeffs of hevcdec.h:333); the untimed first pass puts
        # them there, later passes find them in place like a decoder that just parsed them
        ptr = C.c_void_p(); off = C.c_uint32()
        if level_lists:
            recs, q, pairs, (sl, sl_dc) = level_lists[k % len(jobs)]
            abi.check(lib.ohcu_reserve_coefs(h, j, len(pairs), C.byref(ptr)))
            in_place(ptr, cidx, pairs, copy_coeffs)
            abi.check(lib.ohcu_set_scaling_list(h, j, sl.ctypes.data, sl_dc.ctypes.data))
            abi.check(lib.ohcu_append_tu_levels(h, j, recs.ctypes.data, q.ctypes.data, len(recs), ptr, len(pairs)))
            coef_bytes = pairs.nbytes + 8 * len(fj.sao) + q.nbytes + sl.nbytes + sl_dc.nbytes
        elif sparse_lists:
            recs, pairs = sparse_lists[k % len(jobs)]
            abi.check(lib.ohcu_reserve_coefs(h, j, len(pairs), C.byref(ptr)))
            in_place(ptr, cidx, pairs, copy_coeffs)
            abi.check(lib.ohcu_append_tu_sparse(h, j, recs.ctypes.data, len(recs), ptr, len(pairs)))
            coef_bytes = pairs.nbytes + 8 * len(fj.sao)
        else:
            abi.check(lib.ohcu_reserve_coeffs(h, j, len(fj.coeffs), C.byref(ptr), C.byref(off)))
            in_place(ptr, cidx, fj.coeffs, copy_coeffs)
            abi.check(lib.ohcu_append_tu_inplace(h, j, fj.tus.ctypes.data, len(fj.tus)))
            coef_bytes = fj.coeffs.nbytes
        s = fj.sao_in()
        if derive:
            d = abi.DeblockDeriveIn()
            d.qp_y_tab = fj.qp_y_tab.ctypes.data; d.deblock = fj.deblock.ctypes.data
            d.is_pcm = fj.is_pcm.ctypes.data if sp.pcm_or_bypass else None
            d.filter_slice_edges = fj.filter_slice_edges.ctypes.data
            d.tab_slice_address = fj.tab_slice_address.ctypes.data; d.tile_id = fj.tile_id.ctypes.data
            abi.check(lib.ohcu_set_deblock_derived(h, j, C.byref(d)))
        else:
            d = fj.deblock_in()
            abi.check(lib.ohcu_set_deblock(h, j, C.byref(d)))
        abi.check(lib.ohcu_set_sao(h, j, C.byref(s)))
        ctx.flush(j)
        if args.copy_out == "packed":
            ctx.output_cropped(2 + k, full, out=outs[k], wait=False)
        else:
            ctx.download_async(2 + k, outs[k])
        nctb = len(fj.sao)
        h2d = (fj.pus.nbytes + 8 * nctb + fj.slices.nbytes + fj.tus.nbytes + coef_bytes + 8 * nctb +
               (nctb if derive else fj.vbs.nbytes + fj.hbs.nbytes) + fj.qp_y_tab.nbytes + fj.deblock.nbytes + fj.sao.nbytes + 2 * nctb)
        d2h = sum(p.strides[0] * (p.shape[0] - 1) + p.shape[1] * p.itemsize for p in outs[k])   # bytes the copies move

    # frames are submitted by a few host threads, one frame each at a time, the way the
    # reference's frame threads (hevcdec.h:53) would drive their own jobs
    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(max(1, min(args.host_threads, B)))
    # untimed first pass over every job arena of the ring (puts the coefficients in place);
    # the resident leg left the ring at job B, so 2B submissions visit each arena once
    for i in range(max(B, E)):
        ctx.sync()
        submit(i, copy_coeffs=True)
    ctx.sync()
    list(pool.map(submit, range(args.warmup * B)))
    ctx.sync()
    for k in range(E):
        ctx.slot_sync(2 + k)
    barrier()
    ctx.timer_start()
    t_host = time.perf_counter()
    list(pool.map(submit, range(args.steps * B)))        # frames stream through, no per-step barrier
    ms_e2e = ctx.timer_stop()
    pool.shutdown()
    t_host = (time.perf_counter() - t_host) * 1e3
    barrier()

    # ---------------- correctness spot check of what the timed runs produced
    import oracle_lib
    check = "skipped"
    if rank == 0 and not args.no_check and sp.width * sp.height <= 3840 * 2160:
        want = oracle_lib.Lib("oracle").replay(jobs[0], list(refs) + [None] * (sp.dpb_slots - 2))
        got = ctx.download(2)
        check = "bit-exact" if all(np.array_equal(a, b) for a, b in zip(got, want)) and \
            all(np.array_equal(a, b) for a, b in zip(outs[0], want)) and \
            all(np.array_equal(a, b) for a, b in zip(outs[len(jobs)], want)) else "MISMATCH"

    # ---------------- reduce over ranks
    frames_rank = args.steps * B
    ms_resident, ms_e2e = shard.max_over_ranks([ms_resident, ms_e2e], device=f"cuda:{device}")
    ctx.close()
    if rank != 0:
        return
    px = sp.width * sp.height
    total_frames = frames_rank * world
    value = total_frames * px / (ms_resident * 1e-3) / 1e6
    e2e_value = total_frames * px / (ms_e2e * 1e-3) / 1e6
    ab = algorithmic_bytes(jobs[0])
    # stage events: [1] after mc_kernel, [2] after residual_kernel, [3] after loopfilter_kernel
    # (deblock + SAO fused), [4] after the halo fill
    mc_ms = float(stage_ms[abi.STAGE_MC])
    res_ms = float(stage_ms[abi.STAGE_RESIDUAL])
    lf_ms = float(stage_ms[abi.STAGE_DEBLOCK])
    pad_ms = float(stage_ms[abi.STAGE_SAO])
    achieved = ab["mc"] / (mc_ms * 1e-3) / 1e9
    kern = [("mc", mc_ms), ("residual", res_ms), ("loopfilter", lf_ms), ("pad", pad_ms)]
    line = {
        "metric": "Mpixel/s reconstructed per B200 (4K Main10); HBM GB/s vs peak", "value": round(value, 1),
        "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_resident / args.steps, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32 accumulate on u16 samples / int16 coefficients" if sp.bit_depth > 8
        else "int32 accumulate on u8 samples / int16 coefficients", "data": "synthetic",
        "config": {"workload": WORKLOAD_NAMES[args.workload], "frames_per_step_per_gpu": B,
                   "frames_in_flight": B, "frames_in_flight_e2e": E, "distinct_frames": len(jobs), "sharding": "independent frames per GPU, no collective",
                   "l2": f"inputs larger than L2: {B} frames x ~{(ab['pipeline']) / 1e6:.0f} MB algorithmic traffic per step",
                   "pus_per_frame": int(len(jobs[0].pus)), "tus_per_frame": int(len(jobs[0].tus)),
                   "coded_fraction": round(len(jobs[0].coeffs) / (ab['S'] / (2 if sp.bit_depth > 8 else 1)), 3),
                   "coeffs": args.coeffs, "bs": args.bs, "copy_out": args.copy_out,
                   "significant_fraction": round(float(np.count_nonzero(jobs[0].coeffs)) / max(1, len(jobs[0].coeffs)), 3)},
        "frames_per_s": round(total_frames / (ms_resident * 1e-3), 1),
        "gpu_launches": int(launches),
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": "Mpixel/s", "h2d_bytes_per_step": int(h2d * B),
                "d2h_bytes_per_step": int(d2h * B), "frames_per_s": round(total_frames / (ms_e2e * 1e-3), 1),
                "host_ms_per_step": round(t_host / args.steps, 3), "host_threads": max(1, min(args.host_threads, B)),
                "coefficient_rewrites": rewrites[0]},
        "roofline": {"bound": "hbm", "kernel": "mc_kernel (motion compensation: qpel / epel, uni / bi / weighted)",
                     "achieved": round(achieved, 1),
                     "peak": hbm_peak, "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})", "unit": "GB/s",
                     "frac": round(achieved / hbm_peak, 4),
                     "traffic": NCU_TRAFFIC.get(args.workload),
                     "traffic_source": NCU_TRAFFIC_SOURCE if args.workload in NCU_TRAFFIC else None,
                     "algorithmic_bytes_per_launch": int(ab["mc"]), "ms_per_launch": round(mc_ms, 4)},
        "kernels": {name: {"ms": round(ms, 4), "algorithmic_mb": round(ab[name] / 1e6, 2),
                           "gbs": round(ab[name] / (ms * 1e-3) / 1e9, 1) if ms > 0 and ab[name] else None}
                    for name, ms in kern},
        "pipeline": {"ms_per_frame_single_stream": round(float(stage_ms[abi.STAGE_TOTAL]), 4),
                     "algorithmic_mb_per_frame": round(ab["pipeline"] / 1e6, 2),
                     "algorithmic_mb_per_frame_survey_4pass": round(ab["pipeline_survey"] / 1e6, 2),
                     "gbs_at_value": round(ab["pipeline"] * total_frames / world / (ms_resident * 1e-3) / 1e9, 1),
                     "frac_of_peak_at_value": round(ab["pipeline"] * total_frames / world / (ms_resident * 1e-3) / 1e9 / hbm_peak, 4)},
        "parity_check": check,
    }
    if world == 1 and not args.no_cpu:
        cb = cpu_reference(jobs, refs, steps=args.cpu_steps, warmup=1, all_arms=True)
        line["cpu_baseline"] = {k: (round(v, 2) if isinstance(v, float) else v) for k, v in cb.items()
                                if k in ("value", "unit", "cores", "kind", "sample", "arms")}
    print(json.dumps(line))


def run_reference(args, rank, world):
    if rank != 0:
        return
    jobs, refs = make_inputs(args.workload, 0)
    if args.coeffs == "levels":
        levels_inputs(jobs)               # same coefficient values as the GPU arm reconstructs from
    if args.bs == "derived":
        derived_bs_inputs(jobs)           # same bs planes as the GPU arm derives
    cb = cpu_reference(jobs, refs, steps=args.steps, warmup=args.warmup)
    sp = jobs[0].sp
    line = {
        "impl": "reference", "metric": "Mpixel/s reconstructed per B200 (4K Main10); HBM GB/s vs peak",
        "value": round(cb["value"], 2), "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(cb["ms_per_step"], 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,
        "dtype": "int32 accumulate on u16 samples / int16 coefficients" if sp.bit_depth > 8 else
        "int32 accumulate on u8 samples / int16 coefficients",
        "data": "synthetic", "config": {"workload": WORKLOAD_NAMES[args.workload]},
        "cpu_baseline": {"value": round(cb["value"], 2), "unit": "Mpixel/s", "cores": cb["cores"], "kind": cb["kind"],
                         "sample": cb["sample"]},
        "e2e": {"value": round(cb["value"], 2), "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="4k", choices=list(WORKLOADS))
    ap.add_argument("--frames", type=int, default=0, help="frames per step per GPU (default per workload)")
    ap.add_argument("--cpu-steps", type=int, default=4, help="steps of the cpu_baseline leg")
    ap.add_argument("--host-threads", type=int, default=4, help="threads submitting frames in the e2e leg")
    ap.add_argument("--coeffs", default="sparse", choices=["dense", "sparse", "levels"],
                    help="coefficient form handed to the library: (position, level) pairs of the significant "
                         "coefficients (ohcu_append_tu_sparse, what the residual parser produces), zero-filled N x N blocks, or "
                         "un-scaled levels + a quantiser per block (ohcu_append_tu_levels: dequantisation on the device)")
    ap.add_argument("--bs", default="host", choices=["host", "derived"],
                    help="boundary strengths: planes uploaded by the host (ohcu_set_deblock) or derived on the device from the "
                         "PU / TU lists (ohcu_set_deblock_derived)")
    ap.add_argument("--copy-out", choices=["2d", "pitch", "packed"], default="2d",
                    help="e2e leg, device -> host: pitched 2-D copies into tight planes (2d), linear copies into planes with "
                         "the device pitch (pitch), or ohcu_slot_output_cropped: pack kernel + one linear copy (packed)")
    ap.add_argument("--slot-pitch", action="store_true",
                    help="e2e leg: host pictures with the device planes' row pitch (one linear copy per plane)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        local = int(os.environ.get("LOCAL_RANK", 0))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    import __graft_entry__
    if not os.path.exists(__graft_entry__.LIB):
        __graft_entry__.build()
    run_ours(args, rank, world)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
