#!/usr/bin/env python
"""bench.py — HEVC pixel-reconstruction hot path on B200: Mpixel/s reconstructed per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--workload 4k|1080p|1080p422|1080p444|8k]

A step = one pass of the hot path (mc_kernel, residual_kernel, loopfilter_kernel = deblock + SAO fused, pad_kernel)
over one batch of synthetic frames (192 4K frames per step by default, so that 20 steps time 0.55 s of device work).
Default workload: BASELINE.json configs[2], 3840x2160 Main10 4:2:0 full reconstruction with weighted prediction; the
other BASELINE configs are --workload 1080p / 1080p422 / 1080p444 / 8k (64 independent frames, strong scaling).

  value        whole-job Mpixel/s (luma W*H*frames / s), work lists resident in HBM, 8 frames in flight, 8 distinct
               frames per GPU predicting from 16 distinct reference pictures (inputs > L2)
  e2e          same metric through the C ABI with HOST buffers, driven by C frame threads (hevc_b200/csrc/e2e_driver.c):
               lists into the job's pinned arenas, H2D, prepare + kernels, D2H of every reconstructed picture
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
    # name: (width, height, bit_depth, chroma_format_idc, generator overrides,
    #        frames in flight per GPU (resident leg), distinct synthetic frames per GPU, frames per step per GPU)
    "4k": (3840, 2160, 10, 1, dict(weighted=1), 8, 8, 192),
    "1080p": (1920, 1080, 8, 1, dict(), 16, 8, 704),
    "1080p422": (1920, 1080, 10, 2, dict(rext=1, weighted=1), 16, 8, 448),
    "1080p444": (1920, 1080, 10, 3, dict(rext=1), 16, 8, 320),
    # 64 independent frames in all, sharded frame i -> rank i mod N: the job is fixed, so scaling is STRONG
    "8k": (7680, 4320, 10, 1, dict(weighted=1), 8, 64, 64),
}
WORKLOAD_NAMES = {
    "4k": "3840x2160 Main10 4:2:0 full reconstruction, weighted prediction (BASELINE configs[2])",
    "1080p": "1920x1080 Main 8-bit 4:2:0 (BASELINE configs[1])",
    "1080p422": "1920x1080 RExt 10-bit 4:2:2 (BASELINE configs[3])",
    "1080p444": "1920x1080 RExt 10-bit 4:4:4 (BASELINE configs[3])",
    "8k": "7680x4320 Main10 4:2:0, 64 independent frames sharded across GPUs (BASELINE configs[4])",
}
N_REF_PAIRS = 8       # distinct pairs of reference pictures (DPB slots 0..15); frame d predicts from pair d mod 8
FIRST_OUT = 2 * N_REF_PAIRS
RING = 16             # output slots FIRST_OUT..31 in rotation (e2e leg), pinned output pictures
# dram__bytes_read.sum + dram__bytes_write.sum of one mc_kernel launch, from the committed
# `ncu --set full` capture of this same command (never measured inside a timed run)
NCU_TRAFFIC = {"4k": 59612928}
NCU_TRAFFIC_SOURCE = "profiles/r02_final_ncu_kernels.txt (53.85 MB read + 5.76 MB write; the reference pictures and part of the written prediction stay in L2)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def frames_per_rank(workload, world):
    """(distinct frames, frames per step) of one rank.  8k: the 64 frames of the job are split over the ranks."""
    D, F = WORKLOADS[workload][6], WORKLOADS[workload][7]
    if workload == "8k":
        return max(1, D // world), max(1, F // world)
    return D, F


INTRA_PCT, INTRA_GOP = 0, 0          # --intra-pct / --gop: share of intra CUs in every frame, every GOP-th frame all intra


def make_inputs(workload, rank, world=1, distinct=None):
    """The synthetic frames of this rank (frame i of the job -> rank i mod world) and the 16 reference pictures.
    Frame d predicts from reference pair d mod 8 (DPB slots 2p, 2p + 1): 8 frames in flight read 16 different
    pictures, so reference reads are not L2 hits by construction."""
    from hevc_b200 import shard, synth
    w, h, bd, cf, kw = WORKLOADS[workload][:5]
    D = distinct or frames_per_rank(workload, world)[0]
    mine = shard.frames_of_rank(D * world, rank, world)
    jobs = []
    for d, i in enumerate(mine):
        ikw = {}
        if INTRA_PCT or INTRA_GOP:
            ikw = dict(sp_intra=1, intra_pct=100 if (INTRA_GOP and i % INTRA_GOP == 0) else INTRA_PCT)
        fj = synth.make_job(w, h, bd, cf, seed=shard.seed_of_frame(3000, i), **kw, **ikw)
        pair = d % N_REF_PAIRS
        for l in range(2):
            for r in range(16):
                fj.slices["ref_slot"][:, l, r] = 2 * pair + (r + l) % 2
        fj.sp.dpb_slots = 32
        jobs.append(fj)
    refs = [synth.make_picture(jobs[0].sp, 7000 + s) for s in range(2 * N_REF_PAIRS)]
    return jobs, refs


def levels_inputs(jobs):
    """--coeffs levels: the frames' coefficient arenas become UN-scaled levels with a random quantiser
    per block and random scaling lists (tests/levels_cases.py); job.coeffs is replaced by the scaled
    coefficients (CPU restatement of scale_and_clip_coeff), which is what the parity check and the CPU
    arms consume - the GPU arm is handed the levels (ohcu_append_tu_levels) and scales them itself.
    Returns per job (quant, levels, scaling)."""
    import levels_cases
    import oracle_lib
    orc = oracle_lib.Lib("oracle")
    out = []
    for i, fj in enumerate(jobs):
        lv, quant, scaling = levels_cases.frame_levels(fj, seed=900 + i)
        fj.coeffs = orc.dequant_tus(fj.tus, quant, lv, fj.sp.bit_depth, scaling)
        out.append((quant, lv, scaling))
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
    """SM clock and throttle reasons sampled every 20 ms DURING the timed regions (NVML; nvidia-smi as a fallback
    at whatever rate it manages)."""

    def __init__(self, device):
        self.device = device
        self.rows = []            # (sm_mhz, sm_max_mhz, [reasons])
        self.stop = threading.Event()
        self.th = threading.Thread(target=self.run, daemon=True)

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.device)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            while not self.stop.is_set():
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append((nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM), mx, [k for k, b in bits.items() if r & b]))
                self.stop.wait(0.02)
            return
        except Exception:
            pass
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.device)], capture_output=True, text=True, timeout=5).stdout
                r = [x.strip() for x in out.strip().split(",")]
                if len(r) >= 6 and r[0].isdigit():
                    self.rows.append((int(r[0]), int(r[1]), [names[i] for i in range(4) if r[2 + i].lower().startswith("active")]))
            except Exception:
                pass
            self.stop.wait(0.02)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm = [r[0] for r in self.rows]
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(r[1] for r in self.rows) if sm else None,
                "reasons": sorted({x for r in self.rows for x in r[2]}), "samples": len(sm)}


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


def bind_to_gpu_numa(device):
    """Best effort: run this rank's threads (and first-touch its pinned staging) on the NUMA node its GPU hangs off.
    Returns the node, or None when the topology cannot be read."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(device)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


# ------------------------------------------------------------------ our arm
def run_ours(args, rank, world):
    from concurrent.futures import ThreadPoolExecutor
    from hevc_b200 import abi, backend, driver, shard
    D, F = frames_per_rank(args.workload, world)
    B = args.in_flight or WORKLOADS[args.workload][5]
    if args.frames:
        F = args.frames
    jobs, refs = make_inputs(args.workload, rank, world)
    sp = jobs[0].sp
    device = int(os.environ.get("LOCAL_RANK", 0))
    hbm_peak, peak_kind = peaks()
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    # frame threads of the e2e leg: the host's cores are shared by the ranks of the node
    threads = args.host_threads or max(1, min(8, cores // max(1, world)))
    # several ranks on one node: each one's threads and pinned buffers on the NUMA node of its GPU (--no-numa: as scheduled)
    numa_node = bind_to_gpu_numa(device) if (world > 1 and not args.no_numa) else None
    # resident leg: J staged jobs (8k: every frame of the rank's share of the 64), launched round-robin
    J = min(D, 64) if args.workload == "8k" else B
    njobs = max(J, RING, threads)
    ctx = backend.Context(sp, device=device, dpb_slots=32, jobs=njobs)
    lib, h = ctx.lib, ctx.handle
    for s, r in enumerate(refs):
        ctx.upload(s, r)
    barrier = shard.barrier
    level_lists = levels_inputs(jobs) if args.coeffs == "levels" else None
    derive = args.bs == "derived"
    if derive:
        derived_bs_inputs(jobs)
    frames = driver.Frames(jobs, sp, coeffs=args.coeffs, derive_bs=derive, levels=level_lists)

    # ---------------- resident leg: lists staged once (through the in-place path), kernels only in the timed region
    out_of = lambda k: FIRST_OUT + k % RING
    res, _ = driver.run(ctx, frames, total_frames=njobs, threads=1, ring=RING, first_slot=FIRST_OUT, fill=driver.FILL_COPY,
                        frame_by_job=True, timed=False)
    ctx.sync()
    rounds = max(1, F // J)                      # launches of each staged job per step
    F = rounds * J
    for _ in range(max(1, args.warmup)):
        for r in range(rounds):
            for k in range(J):
                ctx.launch(k, out_of(k))
    ctx.sync()
    barrier()
    l1 = ctx.kernel_launches()
    with ClockSampler(device) as clk:
        ctx.timer_start()
        for _ in range(args.steps):
            for r in range(rounds):
                for k in range(J):
                    ctx.launch(k, out_of(k))
        ms_resident = ctx.timer_stop()
    barrier()
    launches = ctx.kernel_launches() - l1
    resident_pics = {k: ctx.download(out_of(k)) for k in range(max(0, J - RING), J)}    # checked below

    # ---------------- per-stage timing of single frames (events on the job's own stream)
    ctx.set_profiling(True)
    stage_ms = np.zeros(abi.NUM_STAGES)
    nprof = 16
    for i in range(nprof):
        k = i % J
        ctx.launch(k, out_of(k))
        ctx.sync(k)
        stage_ms += np.array(ctx.stage_ms(k))
    stage_ms /= nprof
    ctx.set_profiling(False)

    # ---------------- e2e leg: host lists -> pinned arenas -> H2D -> kernels -> D2H, through the C ABI, driven by the
    # plain-C frame threads of hevc_b200/csrc/e2e_driver.c (no Python inside the timed region)
    full = (0, 0, sp.width, sp.height)
    if args.copy_out == "packed":
        outs = [ctx.pinned_cropped(full) for _ in range(RING)]
        dl = driver.DOWNLOAD_CROPPED
    else:
        outs = [ctx.pinned_picture(slot_pitch=args.copy_out == "pitch") for _ in range(RING)]
        dl = driver.DOWNLOAD_PLANES
    total = args.steps * F

    def e2e_run(fill, n, timed=True):
        """inplace: every job arena is first filled with the frame it will carry (untimed, as a parser that writes its
        records where the library reads them would have), then only the small per-frame tables are written per frame"""
        by_job = fill == "inplace"
        if by_job:
            driver.run(ctx, frames, total_frames=njobs, threads=1, ring=RING, first_slot=FIRST_OUT, fill=driver.FILL_COPY,
                       frame_by_job=True, timed=False)
        mode = {"inplace": driver.FILL_NONE, "copy": driver.FILL_COPY, "append": driver.APPEND}[fill]
        return driver.run(ctx, frames, total_frames=n, threads=threads, ring=RING, first_slot=FIRST_OUT, fill=mode,
                          frame_by_job=by_job, download=dl, outs=outs, timed=timed)

    e2e_run(args.fill, max(RING, args.warmup * min(F, 4 * RING)), timed=False)
    barrier()
    with ClockSampler(device) as clk2:
        r_e2e, last = e2e_run(args.fill, total)
    barrier()
    ms_e2e, host_ms, cpu_ms = r_e2e.device_ms, r_e2e.host_ms, r_e2e.work_cpu_ms
    e2e_pics = [(int(last[k]), [p.copy() for p in outs[k]]) for k in range(RING)]
    d2h = sum(p.strides[0] * (p.shape[0] - 1) + p.shape[1] * p.itemsize for p in outs[0])
    h2d = int(np.mean(frames.h2d_bytes))
    # the other way of handing the lists over, for comparison: every list memcpy'd into the arenas per frame
    # (or, when that is the main mode, found in place)
    other = "copy" if args.fill == "inplace" else "inplace"
    r_other, _ = e2e_run(other, total)
    # leave every arena holding its own frame for the chained leg
    driver.run(ctx, frames, total_frames=njobs, threads=1, ring=RING, first_slot=FIRST_OUT, fill=driver.FILL_COPY,
               frame_by_job=True, timed=False)
    # ---------------- chained streams: picture t predicts from t-1 and t-2 of its stream (what a decoder does),
    # lists in place, no download: the kernels' throughput when frames depend on each other
    chain = {}
    if args.workload != "8k" and not args.no_chain:
        for S in (1, 2, 4):
            rc_, _ = driver.run(ctx, frames, total_frames=max(3 * S, min(total, 60 * S)), threads=min(threads, S), ring=3 * S,
                                first_slot=FIRST_OUT, fill=driver.FILL_NONE, frame_by_job=True, chain_streams=S, ref_slot=(0, 1))
            chain[f"streams_{S}"] = round(rc_.frames / (rc_.device_ms * 1e-3), 1)

    # ---------------- correctness of what the timed runs produced: EVERY distinct frame against the oracle replay
    import oracle_lib
    check, checked = "skipped", 0
    if not args.no_check:
        orc = oracle_lib.Lib("oracle")
        need = sorted({d for d, _ in e2e_pics if d >= 0} | {k % D for k in resident_pics})
        with ThreadPoolExecutor(min(cores, 16)) as pool:
            want = dict(zip(need, pool.map(lambda d: orc.replay(jobs[d], list(refs) + [None] * 16), need)))
        same = lambda a, b: all(np.array_equal(x, y) for x, y in zip(a, b))
        ok = all(same(pic, want[d]) for d, pic in e2e_pics if d >= 0) and all(same(pic, want[k % D]) for k, pic in resident_pics.items())
        checked = len(need)
        check = "bit-exact" if ok else "MISMATCH"
    bad = shard.max_over_ranks([1.0 if check == "MISMATCH" else 0.0], device=f"cuda:{device}")[0]

    # ---------------- reduce over ranks
    ms_resident, ms_e2e, ms_other, host_ms = shard.max_over_ranks([ms_resident, ms_e2e, r_other.device_ms, host_ms], device=f"cuda:{device}")
    ctx.close()
    if rank != 0:
        return
    if bad:
        check = "MISMATCH"
    px = sp.width * sp.height
    total_frames = total * world
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
    clocks = clk.summary()
    clocks["e2e"] = clk2.summary()
    line = {
        "metric": "Mpixel/s reconstructed per B200 (4K Main10); HBM GB/s vs peak", "value": round(value, 1),
        "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_resident / args.steps, 4), "higher_is_better": True,
        "scaling": "strong" if args.workload == "8k" else "weak",
        "vs_baseline": None, "dtype": "int32 accumulate on u16 samples / int16 coefficients" if sp.bit_depth > 8
        else "int32 accumulate on u8 samples / int16 coefficients", "data": "synthetic",
        "config": {"workload": WORKLOAD_NAMES[args.workload] + (f", {INTRA_PCT}% intra CUs" if INTRA_PCT else "") +
                               (f", every {INTRA_GOP}th picture all intra" if INTRA_GOP else ""),
                   "frames_per_step_per_gpu": F, "frames_in_flight": min(J, RING), "frames_in_flight_e2e": RING, "distinct_frames_per_gpu": len(jobs),
                   "reference_pictures": len(refs), "sharding": "independent frames per GPU, no collective",
                   "l2": f"inputs larger than L2: {len(jobs)} distinct list sets, {len(refs)} reference pictures + {RING} output slots "
                         f"of {ab['S'] / 1e6:.0f} MB, ~{ab['pipeline'] / 1e6:.0f} MB algorithmic traffic per frame",
                   "pus_per_frame": int(len(jobs[0].pus)), "tus_per_frame": int(len(jobs[0].tus)),
                   "coded_fraction": round(len(jobs[0].coeffs) / (ab['S'] / (2 if sp.bit_depth > 8 else 1)), 3),
                   "coeffs": args.coeffs, "bs": args.bs, "copy_out": args.copy_out, "fill": args.fill,
                   "significant_fraction": round(float(np.count_nonzero(jobs[0].coeffs)) / max(1, len(jobs[0].coeffs)), 3)},
        "frames_per_s": round(total_frames / (ms_resident * 1e-3), 1),
        "timed_region_ms": round(ms_resident, 1),
        "gpu_launches": int(launches),
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": "Mpixel/s", "h2d_bytes_per_step": int(h2d * F),
                "d2h_bytes_per_step": int(d2h * F), "frames_per_s": round(total_frames / (ms_e2e * 1e-3), 1),
                "timed_region_ms": round(ms_e2e, 1), "host_ms_per_step": round(host_ms / args.steps, 3),
                "host_cpu_ms_per_frame": round(cpu_ms / max(1, r_e2e.frames), 4), "host_threads": threads, "numa_node": numa_node,
                "driver": "hevc_b200/csrc/e2e_driver.c: C frame threads; " + {
                    "inplace": "the frame's lists sit in the job's pinned arenas where the parser wrote them (ohcu_frame_arenas_get / "
                               "ohcu_frame_commit_lists), per frame the small tables are written, everything is uploaded and validated on the device",
                    "copy": "every list memcpy'd into the job's pinned arenas per frame, validated on the device",
                    "append": "the copying, host-validated ohcu_append_* / ohcu_set_* entry points"}[args.fill],
                "lists_" + ("copied" if other == "copy" else "in_place"): {
                    "frames_per_s": round(total_frames / (ms_other * 1e-3), 1),
                    "host_cpu_ms_per_frame": round(r_other.work_cpu_ms / max(1, r_other.frames), 4)}},
        "chained_streams_frames_per_s": chain,
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
        "parity_check": check, "parity_frames_checked": checked,
    }
    if world == 1 and not args.no_cpu:
        cb = cpu_reference(jobs, refs, steps=args.cpu_steps, warmup=1, all_arms=True)
        line["cpu_baseline"] = {k: (round(v, 2) if isinstance(v, float) else v) for k, v in cb.items()
                                if k in ("value", "unit", "cores", "kind", "sample", "arms")}
    print(json.dumps(line))


def run_reference(args, rank, world):
    if rank != 0:
        return
    jobs, refs = make_inputs(args.workload, 0, distinct=min(frames_per_rank(args.workload, 1)[0], 16))
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
        "scaling": "strong" if args.workload == "8k" else "weak", "vs_baseline": None,
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
    ap.add_argument("--host-threads", type=int, default=0,
                    help="frame threads of the e2e leg (default: min(8, host cores / ranks))")
    ap.add_argument("--fill", default="inplace", choices=["inplace", "copy", "append"],
                    help="e2e leg, how a frame's lists reach the library: found in the job's pinned arenas where the parser wrote them "
                         "(inplace: ohcu_frame_arenas_get / ohcu_frame_commit_lists, validated on the device), memcpy'd into those arenas "
                         "per frame (copy), or handed to the copying, host-validated ohcu_append_* calls (append)")
    ap.add_argument("--no-chain", action="store_true", help="skip the chained-streams leg")
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
    ap.add_argument("--intra-pct", type=int, default=0, help="share of intra-predicted CUs in every frame (default 0: the headline workload)")
    ap.add_argument("--gop", type=int, default=0, help="every GOP-th frame of the job is all intra (an I picture); 0 = none")
    ap.add_argument("--in-flight", type=int, default=0, help="frames in flight of the resident leg (default per workload)")
    ap.add_argument("--no-numa", action="store_true", help="N > 1: do not bind each rank to the NUMA node of its GPU")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    global INTRA_PCT, INTRA_GOP
    INTRA_PCT, INTRA_GOP = args.intra_pct, args.gop
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
