"""ctypes binding of hevc_b200/libohcu_drv.so (csrc/e2e_driver.c): the plain-C, multi-threaded caller of the ABI
that bench.py's e2e leg and tests/test_gpu_driver.py run.  Nothing here computes; frames are prepared as C structs
that point into numpy arrays kept alive by the Frames object."""
import ctypes as C
import os

import numpy as np

from . import abi
from .sparse import ctb_index, iu_index, to_levels, to_sparse

# beside the libohcu.so in use (OHCU_LIB selects another build for A/B runs; its driver copy resolves it through $ORIGIN)
DRV_PATH = os.path.join(os.path.dirname(os.path.abspath(abi.LIB_PATH)), "libohcu_drv.so")

FILL_COPY, FILL_NONE, APPEND = 0, 1, 2
DOWNLOAD_NONE, DOWNLOAD_PLANES, DOWNLOAD_CROPPED = 0, 1, 2


class Frame(C.Structure):
    _fields_ = [
        ("pus", C.c_void_p), ("n_pu", C.c_uint32), ("n_tu", C.c_uint32), ("tus", C.c_void_p), ("quant", C.c_void_p),
        ("coefs", C.c_void_p), ("n_coef", C.c_uint64), ("coef_form", C.c_int32), ("n_slices", C.c_int32),
        ("slices", C.c_void_p), ("ctbs", C.c_void_p), ("deblock", C.c_int32), ("has_sao", C.c_int32),
        ("vbs", C.c_void_p), ("hbs", C.c_void_p), ("qp", C.c_void_p), ("db", C.c_void_p), ("is_pcm", C.c_void_p),
        ("sao", C.c_void_p), ("fse", C.c_void_p), ("tsa", C.c_void_p), ("tile", C.c_void_p),
        ("scaling_list", C.c_void_p), ("all_inter", C.c_int32), ("n_iu", C.c_uint32),
        ("ius", C.c_void_p), ("iu_ctbs", C.c_void_p),
    ]


class Config(C.Structure):
    _fields_ = [
        ("threads", C.c_int32), ("total_frames", C.c_int32), ("ring", C.c_int32), ("first_slot", C.c_int32),
        ("fill", C.c_int32), ("download", C.c_int32), ("timed", C.c_int32), ("chain_streams", C.c_int32),
        ("ref_slot", C.c_int32 * 2), ("frame_by_job", C.c_int32), ("pad", C.c_int32), ("last_frame", C.c_void_p),
        ("out_planes", C.c_void_p), ("out_strides", C.c_void_p),
    ]


class Result(C.Structure):
    _fields_ = [("device_ms", C.c_float), ("wall_ms", C.c_double), ("host_ms", C.c_double),
                ("thread_cpu_ms", C.c_double), ("work_cpu_ms", C.c_double), ("frames", C.c_int32), ("error", C.c_char * 512)]


_drv = None


def load():
    global _drv
    if _drv is None:
        abi.load()                                   # libohcu.so first (the driver's NEEDED entry resolves to it)
        if not os.path.exists(DRV_PATH):
            raise RuntimeError(f"{DRV_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
        lib = C.CDLL(DRV_PATH)
        lib.ohdrv_run.restype = C.c_int
        lib.ohdrv_run.argtypes = [C.c_void_p, C.POINTER(Frame), C.c_int, C.POINTER(Config), C.POINTER(Result)]
        _drv = lib
    return _drv


class Frames:
    """synth.FrameJob objects as an ohdrv_frame array.  coeffs: 'sparse' | 'dense' | 'levels' (levels[i] =
    (quant, lv, scaling) per job for the last)."""

    def __init__(self, jobs, sp, coeffs="sparse", derive_bs=False, levels=None, deblock=True, sao=True):
        self.keep = []
        self.n = len(jobs)
        self.arr = (Frame * self.n)()
        hs = 1 if sp.chroma_format_idc in (1, 2) else 0
        vs = 1 if sp.chroma_format_idc == 1 else 0
        ctb_w = (sp.width + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
        ctb_h = (sp.height + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
        nctb = ctb_w * ctb_h
        self.h2d_bytes = []
        for i, fj in enumerate(jobs):
            f = self.arr[i]
            quant = sl = None
            if coeffs == "levels":
                q, lv, scaling = levels[i]
                recs, quant, pairs = to_levels(fj.tus, q, lv, sp.log2_ctb_size, hs, vs, ctb_w)
                form = abi.COEF_LEVELS
                if scaling is not None:
                    sl = np.concatenate([np.ascontiguousarray(scaling[0], np.uint8).reshape(-1),
                                         np.ascontiguousarray(scaling[1], np.uint8).reshape(-1)])
            elif coeffs == "sparse":
                recs, pairs = to_sparse(fj.tus, fj.coeffs, sp.log2_ctb_size, hs, vs, ctb_w)
                form = abi.COEF_SPARSE
            else:
                recs, pairs, form = fj.tus, fj.coeffs, abi.COEF_DENSE
            idx = ctb_index(fj.pus, recs, sp.log2_ctb_size, hs, vs, ctb_w, nctb,
                            recs["pad"].astype(np.int64) if form >= abi.COEF_SPARSE else None)
            arrays = dict(pus=fj.pus, tus=recs, coefs=pairs, slices=fj.slices, ctbs=idx, vbs=fj.vbs, hbs=fj.hbs,
                          qp=fj.qp_y_tab, db=fj.deblock, sao=fj.sao, fse=fj.filter_slice_edges, tsa=fj.tab_slice_address,
                          tile=fj.tile_id)
            if quant is not None:
                arrays["quant"] = quant
            if sl is not None:
                arrays["scaling_list"] = sl
            if sp.pcm_or_bypass:
                arrays["is_pcm"] = fj.is_pcm
            ius = getattr(fj, "ius", None)
            if ius is not None and len(ius):
                arrays["ius"] = ius
                arrays["iu_ctbs"] = iu_index(ius, sp.log2_ctb_size, hs, vs, ctb_w, nctb)
                f.n_iu = len(ius)
            for name, a in arrays.items():
                a = np.ascontiguousarray(a)
                self.keep.append(a)
                setattr(f, name, a.ctypes.data)
            f.n_pu, f.n_tu, f.n_coef, f.coef_form, f.n_slices = len(fj.pus), len(recs), len(pairs), form, len(fj.slices)
            f.deblock = 0 if not deblock else (abi.DEBLOCK_DERIVED if derive_bs else abi.DEBLOCK_PLANES)
            f.has_sao = int(bool(sao))
            f.all_inter = int(int((fj.pus["w"].astype(np.int64) * fj.pus["h"]).sum()) == sp.width * sp.height)
            # bytes that cross the bus per frame (what stage uploads)
            up = fj.pus.nbytes + recs.nbytes + pairs.nbytes + fj.slices.nbytes + idx.nbytes + 3 * 8 * nctb + \
                fj.qp_y_tab.nbytes + fj.deblock.nbytes + fj.sao.nbytes + 10 * nctb + \
                (0 if derive_bs else fj.vbs.nbytes + fj.hbs.nbytes) + (quant.nbytes if quant is not None else 0) + \
                (ius.nbytes + 8 * nctb if f.n_iu else 0)
            self.h2d_bytes.append(int(up))


def run(ctx, frames, total_frames, threads, ring, first_slot, fill=FILL_COPY, download=DOWNLOAD_NONE, outs=None,
        timed=True, chain_streams=0, ref_slot=(0, 1), frame_by_job=False):
    """ohdrv_run; outs = list of `ring` pictures (lists of pinned plane arrays).  Returns (Result, last_frame array)."""
    drv = load()
    cfg = Config()
    cfg.threads, cfg.total_frames, cfg.ring, cfg.first_slot = threads, total_frames, ring, first_slot
    cfg.fill, cfg.download, cfg.timed, cfg.chain_streams = fill, download, int(timed), chain_streams
    cfg.ref_slot[0], cfg.ref_slot[1] = ref_slot
    cfg.frame_by_job = int(frame_by_job)
    last = np.full(ring, -1, np.int32)
    cfg.last_frame = last.ctypes.data
    keep = []
    if download != DOWNLOAD_NONE:
        assert outs is not None and len(outs) == ring
        planes = (C.c_void_p * (3 * ring))()
        for k, pic in enumerate(outs):
            for c, p in enumerate(pic):
                planes[3 * k + c] = p.ctypes.data
        strides = (C.c_ssize_t * 3)(*[p.strides[0] for p in outs[0]] + [0] * (3 - len(outs[0])))
        cfg.out_planes = C.cast(planes, C.c_void_p)
        cfg.out_strides = C.cast(strides, C.c_void_p)
        keep += [planes, strides]
    res = Result()
    rc = drv.ohdrv_run(ctx.handle, frames.arr, frames.n, C.byref(cfg), C.byref(res))
    if rc:
        raise abi.OhcuError(f"ohdrv_run -> {rc}: {res.error.decode(errors='replace')}")
    return res, last
