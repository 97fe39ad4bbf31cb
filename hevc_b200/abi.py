"""ctypes mirror of include/ohcu.h — the C ABI structs and the libohcu.so loader.

The product is the shared library; this module only describes its ABI to Python
(tests, bench.py, smoke).  There is no Python or CPU fallback: ``load()`` raises
if the CUDA library is missing.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# OHCU_LIB selects another build of the same library (A/B measurements of kernel variants)
LIB_PATH = os.environ.get("OHCU_LIB") or os.path.join(HERE, "libohcu.so")

OHCU_ABI_VERSION = 1
MAX_REFS = 16
MAX_DPB = 32
MAX_SLICES = 64
NOALLOC_MAX_BYTES = 65536

PF_L0, PF_L1, PF_BI = 1, 2, 3
SAO_NOT_APPLIED, SAO_BAND, SAO_EDGE = 0, 1, 2
TU_IDCT, TU_DC, TU_DST4, TU_SKIP, TU_BYPASS = 0, 1, 2, 3, 4
TUF_RDPCM, TUF_RDPCM_VERT, TUF_ROTATE, TUF_CROSS = 1, 2, 4, 8
STAGE_MC, STAGE_RESIDUAL, STAGE_DEBLOCK, STAGE_SAO, STAGE_TOTAL, NUM_STAGES = 0, 1, 2, 3, 4, 5

ERRORS = {0: "OK", -1: "EINVAL", -2: "ENOMEM", -3: "ECUDA", -4: "ESTATE", -5: "ENODEV", -6: "EOVERFLOW"}


class SeqParams(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("bit_depth", C.c_int32),
        ("chroma_format_idc", C.c_int32), ("log2_ctb_size", C.c_int32),
        ("log2_min_cb_size", C.c_int32), ("log2_min_pu_size", C.c_int32),
        ("dpb_slots", C.c_int32), ("jobs", C.c_int32), ("sao_enabled", C.c_int32),
        ("pcm_or_bypass", C.c_int32), ("cb_qp_offset", C.c_int32), ("cr_qp_offset", C.c_int32),
        ("no_tile_filter", C.c_int32), ("max_pus", C.c_int32), ("max_tus", C.c_int32),
        ("max_coeffs", C.c_int64), ("intra", C.c_int32), ("max_ius", C.c_int32),
        ("intra_smoothing_disabled", C.c_int32), ("strong_intra_smoothing", C.c_int32),
    ]


class SliceTab(C.Structure):
    _fields_ = [
        ("ref_slot", (C.c_uint8 * MAX_REFS) * 2),
        ("weight_flag", C.c_uint8), ("luma_log2_weight_denom", C.c_uint8),
        ("chroma_log2_weight_denom", C.c_uint8), ("pad", C.c_uint8),
        ("luma_weight", (C.c_int16 * MAX_REFS) * 2),
        ("luma_offset", (C.c_int16 * MAX_REFS) * 2),
        ("chroma_weight", ((C.c_int16 * 2) * MAX_REFS) * 2),
        ("chroma_offset", ((C.c_int16 * 2) * MAX_REFS) * 2),
    ]


class PU(C.Structure):
    _fields_ = [
        ("x0", C.c_uint16), ("y0", C.c_uint16), ("w", C.c_uint8), ("h", C.c_uint8),
        ("pred_flag", C.c_uint8), ("slice", C.c_uint8),
        ("mv", (C.c_int16 * 2) * 2), ("ref_idx", C.c_uint8 * 2), ("pad", C.c_uint8 * 2),
    ]


class TU(C.Structure):
    _fields_ = [
        ("x0", C.c_uint16), ("y0", C.c_uint16), ("c_idx", C.c_uint8), ("log2_size", C.c_uint8),
        ("mode", C.c_uint8), ("flags", C.c_uint8), ("col_limit", C.c_int8),
        ("res_scale_val", C.c_int8), ("pad", C.c_uint16), ("coeff_off", C.c_uint32),
        ("cross_tu", C.c_uint32),
    ]


class IU(C.Structure):
    _fields_ = [("x0", C.c_uint16), ("y0", C.c_uint16), ("c_idx", C.c_uint8), ("log2_size", C.c_uint8),
                ("mode", C.c_uint8), ("cand", C.c_uint8)]


class IuIndex(C.Structure):
    _fields_ = [("first_iu", C.c_uint32), ("n_iu", C.c_uint32)]


class DBParams(C.Structure):
    _fields_ = [("beta_offset", C.c_int8), ("tc_offset", C.c_int8)]


class SAO(C.Structure):
    _fields_ = [
        ("type_idx", C.c_uint8 * 3), ("eo_class", C.c_uint8 * 3), ("band_position", C.c_uint8 * 3),
        ("pad", C.c_uint8), ("offset_val", (C.c_int16 * 5) * 3),
    ]


class DeblockIn(C.Structure):
    _fields_ = [
        ("vertical_bs", C.c_void_p), ("horizontal_bs", C.c_void_p), ("qp_y_tab", C.c_void_p),
        ("deblock", C.c_void_p), ("is_pcm", C.c_void_p),
    ]


class DeblockDeriveIn(C.Structure):
    _fields_ = [
        ("qp_y_tab", C.c_void_p), ("deblock", C.c_void_p), ("is_pcm", C.c_void_p),
        ("filter_slice_edges", C.c_void_p), ("tab_slice_address", C.c_void_p), ("tile_id", C.c_void_p),
        ("deblock_disabled", C.c_void_p),
    ]


class Crop(C.Structure):
    _fields_ = [("left", C.c_int32), ("top", C.c_int32), ("width", C.c_int32), ("height", C.c_int32)]


class SaoIn(C.Structure):
    _fields_ = [
        ("sao", C.c_void_p), ("filter_slice_edges", C.c_void_p),
        ("tab_slice_address", C.c_void_p), ("tile_id", C.c_void_p),
    ]


class Coef(C.Structure):
    _fields_ = [("pos", C.c_uint16), ("level", C.c_int16)]


class TuQuant(C.Structure):
    _fields_ = [("qp", C.c_uint8), ("matrix_id", C.c_uint8)]


class CtbIndex(C.Structure):
    _fields_ = [("first_pu", C.c_uint32), ("n_pu", C.c_uint32), ("first_tu", C.c_uint32), ("n_tu", C.c_uint32),
                ("first_coef", C.c_uint32), ("n_coef", C.c_uint32)]


class FrameArenas(C.Structure):
    _fields_ = [
        ("pus", C.c_void_p), ("max_pus", C.c_uint32), ("max_tus", C.c_uint32), ("tus", C.c_void_p),
        ("quant", C.c_void_p), ("coefs", C.c_void_p), ("coeffs", C.c_void_p), ("max_coefs", C.c_uint64),
        ("ctbs", C.c_void_p), ("slices", C.c_void_p), ("vertical_bs", C.c_void_p), ("horizontal_bs", C.c_void_p),
        ("qp_y_tab", C.c_void_p), ("deblock", C.c_void_p), ("is_pcm", C.c_void_p), ("sao", C.c_void_p),
        ("filter_slice_edges", C.c_void_p), ("tab_slice_address", C.c_void_p), ("tile_id", C.c_void_p),
        ("deblock_disabled", C.c_void_p), ("scaling_list", C.c_void_p),
        ("ius", C.c_void_p), ("iu_ctbs", C.c_void_p), ("max_ius", C.c_uint32), ("reserved", C.c_uint32),
    ]


class FrameCounts(C.Structure):
    _fields_ = [
        ("n_pu", C.c_uint32), ("n_tu", C.c_uint32), ("n_coef", C.c_uint64), ("n_slices", C.c_int32),
        ("coef_form", C.c_int32), ("deblock", C.c_int32), ("sao", C.c_int32), ("pcm", C.c_int32),
        ("scaling_list", C.c_int32), ("all_inter", C.c_int32), ("n_iu", C.c_uint32),
    ]


class PicParams(C.Structure):
    _fields_ = [("pcm_or_bypass", C.c_int32), ("cb_qp_offset", C.c_int32), ("cr_qp_offset", C.c_int32),
                ("no_tile_filter", C.c_int32)]


NO_REF = 0xff
COEF_DENSE, COEF_SPARSE, COEF_LEVELS = 1, 2, 3
DEBLOCK_NONE, DEBLOCK_PLANES, DEBLOCK_DERIVED = 0, 1, 2
TUF_INTRA, IU_CBF = 16, 32
CAND_BOTTOM_LEFT, CAND_LEFT, CAND_UP_LEFT, CAND_UP, CAND_UP_RIGHT = 1, 2, 4, 8, 16
EINVAL, ENOMEM, ECUDA, ESTATE, ENODEV, EOVERFLOW = -1, -2, -3, -4, -5, -6

# every struct of include/ohcu.h by its C name: tests/test_abi.py compares sizeof / offsetof with a C program
STRUCTS = {
    "ohcu_seq_params": SeqParams, "ohcu_slice_tab": SliceTab, "ohcu_pu": PU, "ohcu_tu": TU, "ohcu_dbparams": DBParams,
    "ohcu_iu": IU, "ohcu_iu_index": IuIndex,
    "ohcu_sao": SAO, "ohcu_deblock_in": DeblockIn, "ohcu_deblock_derive_in": DeblockDeriveIn, "ohcu_sao_in": SaoIn,
    "ohcu_crop": Crop, "ohcu_coef": Coef, "ohcu_tu_quant": TuQuant, "ohcu_ctb_index": CtbIndex,
    "ohcu_frame_arenas": FrameArenas, "ohcu_frame_counts": FrameCounts, "ohcu_pic_params": PicParams,
}

assert C.sizeof(PU) == 20 and C.sizeof(TU) == 20 and C.sizeof(SAO) == 40 and C.sizeof(DBParams) == 2
assert C.sizeof(CtbIndex) == 24 and C.sizeof(Coef) == 4

# name -> (restype, argtypes); every symbol include/ohcu.h declares
_P = C.c_void_p
_I = C.c_int
_S = C.c_ssize_t
SYMBOLS = {
    "ohcu_abi_version": (C.c_int, []),
    "ohcu_last_error": (C.c_char_p, []),
    "ohcu_device_count": (C.c_int, []),
    "ohcu_create": (C.c_int, [C.POINTER(_P), C.c_int, C.POINTER(SeqParams)]),
    "ohcu_destroy": (C.c_int, [_P]),
    "ohcu_get_seq_params": (C.c_int, [_P, C.POINTER(SeqParams)]),
    "ohcu_slot_upload": (C.c_int, [_P, C.c_int, C.POINTER(_P * 3), C.POINTER(C.c_ssize_t * 3)]),
    "ohcu_slot_download": (C.c_int, [_P, C.c_int, C.POINTER(_P * 3), C.POINTER(C.c_ssize_t * 3)]),
    "ohcu_slot_download_async": (C.c_int, [_P, C.c_int, C.POINTER(_P * 3), C.POINTER(C.c_ssize_t * 3)]),
    "ohcu_slot_output_cropped": (C.c_int, [_P, C.c_int, C.POINTER(Crop), C.POINTER(_P * 3)]),
    "ohcu_slot_output_cropped_async": (C.c_int, [_P, C.c_int, C.POINTER(Crop), C.POINTER(_P * 3)]),
    "ohcu_slot_sync": (C.c_int, [_P, C.c_int]),
    "ohcu_slot_md5": (C.c_int, [_P, C.c_int, _P]),
    "ohcu_frame_begin": (C.c_int, [_P, C.c_int, C.c_int]),
    "ohcu_set_slice": (C.c_int, [_P, C.c_int, C.c_int, C.POINTER(SliceTab)]),
    "ohcu_append_pu": (C.c_int, [_P, C.c_int, _P, C.c_int]),
    "ohcu_append_iu": (C.c_int, [_P, C.c_int, _P, C.c_int]),
    "ohcu_append_tu": (C.c_int, [_P, C.c_int, _P, C.c_int, _P, C.c_size_t]),
    "ohcu_reserve_coeffs": (C.c_int, [_P, C.c_int, C.c_size_t, C.POINTER(_P), C.POINTER(C.c_uint32)]),
    "ohcu_append_tu_inplace": (C.c_int, [_P, C.c_int, _P, C.c_int]),
    "ohcu_reserve_coefs": (C.c_int, [_P, C.c_int, C.c_size_t, C.POINTER(_P)]),
    "ohcu_append_tu_sparse": (C.c_int, [_P, C.c_int, _P, C.c_int, _P, C.c_size_t]),
    "ohcu_append_tu_levels": (C.c_int, [_P, C.c_int, _P, _P, C.c_int, _P, C.c_size_t]),
    "ohcu_set_scaling_list": (C.c_int, [_P, C.c_int, _P, _P]),
    "ohcu_set_deblock": (C.c_int, [_P, C.c_int, C.POINTER(DeblockIn)]),
    "ohcu_set_deblock_derived": (C.c_int, [_P, C.c_int, C.POINTER(DeblockDeriveIn)]),
    "ohcu_set_sao": (C.c_int, [_P, C.c_int, C.POINTER(SaoIn)]),
    "ohcu_frame_arenas_get": (C.c_int, [_P, C.c_int, C.POINTER(FrameArenas)]),
    "ohcu_frame_commit_lists": (C.c_int, [_P, C.c_int, C.POINTER(FrameCounts)]),
    "ohcu_set_pic_params": (C.c_int, [_P, C.c_int, C.POINTER(PicParams)]),
    "ohcu_append_pcm": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "ohcu_slot_release": (C.c_int, [_P, C.c_int]),
    "ohcu_frame_flush": (C.c_int, [_P, C.c_int]),
    "ohcu_frame_abort": (C.c_int, [_P, C.c_int]),
    "ohcu_slot_pitch": (C.c_int, [_P, C.c_int, C.POINTER(C.c_ssize_t)]),
    "ohcu_frame_stage": (C.c_int, [_P, C.c_int]),
    "ohcu_frame_launch": (C.c_int, [_P, C.c_int, C.c_int]),
    "ohcu_frame_launch_stages": (C.c_int, [_P, C.c_int, C.c_int, C.c_int]),
    "ohcu_frame_sync": (C.c_int, [_P, C.c_int]),
    "ohcu_sync_all": (C.c_int, [_P]),
    "ohcu_job_read_bs": (C.c_int, [_P, C.c_int, _P, _P]),
    "ohcu_set_profiling": (C.c_int, [_P, C.c_int]),
    "ohcu_job_stage_ms": (C.c_int, [_P, C.c_int, C.POINTER(C.c_float * NUM_STAGES)]),
    "ohcu_kernel_launches": (C.c_int64, [_P]),
    "ohcu_job_stream": (_P, [_P, C.c_int]),
    "ohcu_flush_l2": (C.c_int, [_P]),
    "ohcu_timer_start": (C.c_int, [_P]),
    "ohcu_timer_stop": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "ohcu_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "ohcu_host_free": (C.c_int, [_P]),
    # per-entry test shims: same argument lists as the checkers' leaves (tests/oracle_lib.py)
    "ohcu_dsp_mc_put": (C.c_int, [_I, _I, _P, _P, _S, _I, _I, _I, _I]),
    "ohcu_dsp_mc_uni": (C.c_int, [_I, _I, _P, _S, _P, _S, _I, _I, _I, _I]),
    "ohcu_dsp_mc_uni_w": (C.c_int, [_I, _I, _P, _S, _P, _S, _I, _I, _I, _I, _I, _I, _I]),
    "ohcu_dsp_mc_bi": (C.c_int, [_I, _I, _P, _S, _P, _S, _P, _I, _I, _I, _I]),
    "ohcu_dsp_mc_bi_w": (C.c_int, [_I, _I, _P, _S, _P, _S, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I]),
    "ohcu_dsp_idct": (C.c_int, [_I, _P, _I, _I]),
    "ohcu_dsp_idct_noalloc": (C.c_int, [_I, _P, _I, _I, _I]),
    "ohcu_dsp_idct_dc": (C.c_int, [_I, _P, _I]),
    "ohcu_dsp_idct_4x4_luma": (C.c_int, [_I, _P]),
    "ohcu_dsp_transform_skip": (C.c_int, [_I, _P, _I]),
    "ohcu_dsp_transform_rdpcm": (C.c_int, [_I, _P, _I, _I]),
    "ohcu_dsp_transform_add": (C.c_int, [_I, _P, _P, _S, _I]),
    "ohcu_dsp_intra_pred": (C.c_int, [_I, _P, _S] + [_I] * 10),
    "ohcu_dsp_loop_filter_luma": (C.c_int, [_I, _I, _P, _S, _I, _P, _P, _P]),
    "ohcu_dsp_loop_filter_chroma": (C.c_int, [_I, _I, _P, _S, _P, _P, _P]),
    "ohcu_dsp_sao_band_filter": (C.c_int, [_I, _P, _P, _S, _S, _P, _I, _I, _I]),
    "ohcu_dsp_sao_edge_filter": (C.c_int, [_I, _P, _P, _S, _S, _P, _I, _I, _I]),
}

_lib = None


def load():
    """dlopen hevc_b200/libohcu.so and type every entry point.  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError = ABI mismatch, fail loudly
        fn.restype = res
        fn.argtypes = args
    if lib.ohcu_abi_version() != OHCU_ABI_VERSION:
        raise RuntimeError("libohcu.so ABI version mismatch")
    _lib = lib
    return lib


class OhcuError(RuntimeError):
    pass


def check(rc, what=""):
    if rc < 0:
        lib = load()
        msg = lib.ohcu_last_error()
        raise OhcuError(f"{what}: {ERRORS.get(rc, rc)}: {msg.decode() if msg else ''}")
    return rc
