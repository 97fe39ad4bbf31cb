"""Host-side mirror of the batched backend for Python callers (tests, bench, smoke).

``Context`` wraps one ``ohcu_ctx`` (one per GPU): a device DPB plus a ring of job
arenas.  Every call goes through the C ABI of include/ohcu.h; nothing here computes.
"""
import ctypes as C

import numpy as np

from . import abi
from .abi import check
from .sparse import ctb_index, iu_index, to_levels, to_sparse


def _plane_args(planes):
    ptrs = (C.c_void_p * 3)()
    strides = (C.c_ssize_t * 3)()
    for i, p in enumerate(planes):
        assert p.flags["C_CONTIGUOUS"] or p.strides[1] == p.itemsize
        ptrs[i] = p.ctypes.data
        strides[i] = p.strides[0]
    return ptrs, strides


class Context:
    def __init__(self, sp, device=0, dpb_slots=None, jobs=None):
        self.lib = abi.load()
        self.sp = abi.SeqParams.from_buffer_copy(sp)
        if dpb_slots is not None:
            self.sp.dpb_slots = dpb_slots
        if jobs is not None:
            self.sp.jobs = jobs
        self.handle = C.c_void_p()
        check(self.lib.ohcu_create(C.byref(self.handle), device, C.byref(self.sp)), "ohcu_create")
        self.nplanes = 3 if self.sp.chroma_format_idc else 1
        self.dtype = np.uint8 if self.sp.bit_depth == 8 else np.uint16

    def close(self):
        if self.handle:
            self.lib.ohcu_destroy(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- DPB ---------------------------------------------------------------
    def plane_shape(self, c):
        hs = 1 if (c and self.sp.chroma_format_idc in (1, 2)) else 0
        vs = 1 if (c and self.sp.chroma_format_idc == 1) else 0
        return (self.sp.height >> vs, self.sp.width >> hs)

    def upload(self, slot, planes):
        planes = [np.ascontiguousarray(p, dtype=self.dtype) for p in planes]
        ptrs, strides = _plane_args(planes)
        check(self.lib.ohcu_slot_upload(self.handle, slot, C.byref(ptrs), C.byref(strides)), "ohcu_slot_upload")

    def download(self, slot, out=None):
        if out is None:
            out = [np.empty(self.plane_shape(c), dtype=self.dtype) for c in range(self.nplanes)]
        ptrs, strides = _plane_args(out)
        check(self.lib.ohcu_slot_download(self.handle, slot, C.byref(ptrs), C.byref(strides)), "ohcu_slot_download")
        return out

    def download_async(self, slot, out):
        ptrs, strides = _plane_args(out)
        check(self.lib.ohcu_slot_download_async(self.handle, slot, C.byref(ptrs), C.byref(strides)),
              "ohcu_slot_download_async")

    def crop_shape(self, c, crop):
        hs = 1 if (c and self.sp.chroma_format_idc in (1, 2)) else 0
        vs = 1 if (c and self.sp.chroma_format_idc == 1) else 0
        return (crop[3] >> vs, crop[2] >> hs)

    def output_cropped(self, slot, crop, out=None, wait=True):
        """ohcu_slot_output_cropped: the (left, top, width, height) window of the slot's planes,
        packed on the device, into tight host planes (oh_output_cropped_cpy, openhevc.c:662-742)."""
        if out is None:
            out = [np.empty(self.crop_shape(c, crop), dtype=self.dtype) for c in range(self.nplanes)]
        ptrs = (C.c_void_p * 3)()
        for i, p in enumerate(out):
            assert p.flags["C_CONTIGUOUS"] and p.shape == self.crop_shape(i, crop) and p.dtype == self.dtype
            ptrs[i] = p.ctypes.data
        cr = abi.Crop(*[int(v) for v in crop])
        fn = self.lib.ohcu_slot_output_cropped if wait else self.lib.ohcu_slot_output_cropped_async
        check(fn(self.handle, slot, C.byref(cr), C.byref(ptrs)), "ohcu_slot_output_cropped")
        return out

    def pinned_cropped(self, crop):
        """Tight output planes for output_cropped(wait=False), following each other in one pinned
        block at 16-byte boundaries (they then leave the device as one copy)."""
        item = np.dtype(self.dtype).itemsize
        shapes = [self.crop_shape(c, crop) for c in range(self.nplanes)]
        offs, total = [], 0
        for h, w in shapes:
            offs.append(total)
            total += (h * w * item + 15) & ~15
        p = C.c_void_p()
        check(self.lib.ohcu_host_alloc(total, C.byref(p)), "ohcu_host_alloc")
        buf = (C.c_char * total).from_address(p.value)
        return [np.frombuffer(buf, dtype=self.dtype, count=h * w, offset=o).reshape(h, w) for (h, w), o in zip(shapes, offs)]

    def slot_sync(self, slot):
        check(self.lib.ohcu_slot_sync(self.handle, slot), "ohcu_slot_sync")

    # ---- frames ------------------------------------------------------------
    def begin(self, out_slot, poc=0):
        return check(self.lib.ohcu_frame_begin(self.handle, out_slot, poc), "ohcu_frame_begin")

    def sparse_lists(self, fj):
        """(recs, pairs) of a FrameJob for ohcu_append_tu_sparse (hevc_b200.sparse.to_sparse)."""
        sp = self.sp
        hs = 1 if sp.chroma_format_idc in (1, 2) else 0
        vs = 1 if sp.chroma_format_idc == 1 else 0
        ctb_w = (sp.width + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
        return to_sparse(fj.tus, fj.coeffs, sp.log2_ctb_size, hs, vs, ctb_w)

    def append(self, job_id, fj, use_deblock=True, use_sao=True, sparse=False, levels=None, derive_bs=None):
        """Append every list of a synth.FrameJob (or anything with the same arrays); sparse = the
        coefficients go as (position, level) pairs (True, or the lists sparse_lists() returned);
        levels = (quant, scaling) : fj.coeffs holds UN-scaled levels, quant one QUANT_DTYPE record per
        TU, scaling None or (sl[4][6][64], sl_dc[2][6]) uint8 arrays - the device dequantises;
        derive_bs = True or a per-CTB uint8 array of slice_deblocking_filter_disabled flags: fj.vbs / fj.hbs
        are NOT sent, the device derives the boundary strengths from the lists (ohcu_set_deblock_derived)."""
        lib, h = self.lib, self.handle
        for i in range(len(fj.slices)):
            tab = abi.SliceTab.from_buffer_copy(fj.slices[i].tobytes())
            check(lib.ohcu_set_slice(h, job_id, i, C.byref(tab)), "ohcu_set_slice")
        check(lib.ohcu_append_pu(h, job_id, fj.pus.ctypes.data, len(fj.pus)), "ohcu_append_pu")
        ius = getattr(fj, "ius", None)
        if ius is not None and len(ius):
            check(lib.ohcu_append_iu(h, job_id, ius.ctypes.data, len(ius)), "ohcu_append_iu")
        if len(fj.tus) and levels is not None:
            quant, scaling = levels[:2]
            sp = self.sp
            hs = 1 if sp.chroma_format_idc in (1, 2) else 0
            vs = 1 if sp.chroma_format_idc == 1 else 0
            ctb_w = (sp.width + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
            # a third element = the lists to_levels() already made (fj.coeffs is then not read)
            recs, q, pairs = levels[2] if len(levels) > 2 else to_levels(fj.tus, quant, fj.coeffs, sp.log2_ctb_size, hs, vs, ctb_w)
            if scaling is not None:
                sl, sl_dc = (np.ascontiguousarray(a, dtype=np.uint8) for a in scaling)
                assert sl.shape == (4, 6, 64) and sl_dc.shape == (2, 6)
                check(lib.ohcu_set_scaling_list(h, job_id, sl.ctypes.data, sl_dc.ctypes.data), "ohcu_set_scaling_list")
            check(lib.ohcu_append_tu_levels(h, job_id, recs.ctypes.data, q.ctypes.data, len(recs), pairs.ctypes.data,
                                            len(pairs)), "ohcu_append_tu_levels")
        elif len(fj.tus) and sparse:
            recs, pairs = sparse if isinstance(sparse, tuple) else self.sparse_lists(fj)
            check(lib.ohcu_append_tu_sparse(h, job_id, recs.ctypes.data, len(recs), pairs.ctypes.data, len(pairs)),
                  "ohcu_append_tu_sparse")
        elif len(fj.tus):
            check(lib.ohcu_append_tu(h, job_id, fj.tus.ctypes.data, len(fj.tus), fj.coeffs.ctypes.data,
                                     len(fj.coeffs)), "ohcu_append_tu")
        if use_deblock and derive_bs is not None and derive_bs is not False:
            d = abi.DeblockDeriveIn()
            d.qp_y_tab = fj.qp_y_tab.ctypes.data
            d.deblock = fj.deblock.ctypes.data
            d.is_pcm = fj.is_pcm.ctypes.data if self.sp.pcm_or_bypass else None
            d.filter_slice_edges = fj.filter_slice_edges.ctypes.data
            d.tab_slice_address = fj.tab_slice_address.ctypes.data
            d.tile_id = fj.tile_id.ctypes.data
            dis = None
            if derive_bs is not True:
                dis = np.ascontiguousarray(derive_bs, dtype=np.uint8)
                d.deblock_disabled = dis.ctypes.data
            check(lib.ohcu_set_deblock_derived(h, job_id, C.byref(d)), "ohcu_set_deblock_derived")
        elif use_deblock:
            d = fj.deblock_in()
            check(lib.ohcu_set_deblock(h, job_id, C.byref(d)), "ohcu_set_deblock")
        if use_sao:
            s = fj.sao_in()
            check(lib.ohcu_set_sao(h, job_id, C.byref(s)), "ohcu_set_sao")

    def append_inplace(self, job_id, fj, coeffs="sparse", levels=None, use_deblock=True, use_sao=True, derive_bs=None,
                       all_inter=None, commit=True):
        """The same lists through the ZERO-COPY path: ohcu_frame_arenas_get, records and tables written straight into
        the job's pinned arenas (what the patched parser does), ohcu_frame_commit_lists.  Nothing is validated on
        the host; prepare_kernel does it on the device.  coeffs = 'sparse' | 'dense'; levels as in append().
        commit=False returns (arenas, counts) without committing (tests tamper with them first)."""
        lib, h, sp = self.lib, self.handle, self.sp
        hs = 1 if sp.chroma_format_idc in (1, 2) else 0
        vs = 1 if sp.chroma_format_idc == 1 else 0
        ctb_w = (sp.width + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
        ctb_h = (sp.height + (1 << sp.log2_ctb_size) - 1) >> sp.log2_ctb_size
        nctb = ctb_w * ctb_h
        a = abi.FrameArenas()
        check(lib.ohcu_frame_arenas_get(h, job_id, C.byref(a)), "ohcu_frame_arenas_get")

        def put(dst, arr):
            arr = np.ascontiguousarray(arr)
            if arr.nbytes:
                C.memmove(dst, arr.ctypes.data, arr.nbytes)

        n = abi.FrameCounts()
        quant = None
        if len(fj.tus) and levels is not None:
            recs, quant, pairs = levels[2] if len(levels) > 2 else to_levels(fj.tus, levels[0], fj.coeffs, sp.log2_ctb_size, hs, vs, ctb_w)
            form = abi.COEF_LEVELS
        elif len(fj.tus) and coeffs == "sparse":
            recs, pairs = self.sparse_lists(fj)
            form = abi.COEF_SPARSE
        else:
            recs, pairs, form = fj.tus, fj.coeffs, abi.COEF_DENSE if len(fj.tus) else 0
        idx = ctb_index(fj.pus, recs, sp.log2_ctb_size, hs, vs, ctb_w, nctb,
                        recs["pad"].astype(np.int64) if form >= abi.COEF_SPARSE else None)
        assert len(fj.pus) <= a.max_pus and len(recs) <= a.max_tus
        put(a.pus, fj.pus); put(a.tus, recs); put(a.coefs, pairs); put(a.ctbs, idx); put(a.slices, fj.slices)
        if quant is not None:
            put(a.quant, quant)
        if levels is not None and levels[1] is not None:
            sl, sl_dc = (np.ascontiguousarray(x, dtype=np.uint8) for x in levels[1])
            put(a.scaling_list, sl); put(a.scaling_list + sl.nbytes, sl_dc)
            n.scaling_list = 1
        n.n_pu, n.n_tu, n.n_coef, n.n_slices, n.coef_form = len(fj.pus), len(recs), len(pairs), len(fj.slices), form
        ius = getattr(fj, "ius", None)
        if ius is not None and len(ius):
            assert a.ius and len(ius) <= a.max_ius, "the context was created without sp.intra"
            put(a.ius, ius); put(a.iu_ctbs, iu_index(ius, sp.log2_ctb_size, hs, vs, ctb_w, nctb))
            n.n_iu = len(ius)
        if use_deblock:
            put(a.qp_y_tab, fj.qp_y_tab); put(a.deblock, fj.deblock)
            if sp.pcm_or_bypass:
                put(a.is_pcm, fj.is_pcm); n.pcm = 1
            if derive_bs is not None and derive_bs is not False:
                n.deblock = abi.DEBLOCK_DERIVED
                dis = np.zeros(nctb, np.uint8) if derive_bs is True else np.ascontiguousarray(derive_bs, dtype=np.uint8)
                put(a.deblock_disabled, dis)
            else:
                n.deblock = abi.DEBLOCK_PLANES
                put(a.vertical_bs, fj.vbs); put(a.horizontal_bs, fj.hbs)
        if use_sao or n.deblock == abi.DEBLOCK_DERIVED:
            put(a.filter_slice_edges, fj.filter_slice_edges); put(a.tab_slice_address, fj.tab_slice_address)
            put(a.tile_id, fj.tile_id)
        if use_sao:
            put(a.sao, fj.sao); n.sao = 1
        if all_inter is None:
            all_inter = int((fj.pus["w"].astype(np.int64) * fj.pus["h"]).sum()) == sp.width * sp.height
        n.all_inter = int(bool(all_inter))
        if not commit:
            return a, n
        check(lib.ohcu_frame_commit_lists(h, job_id, C.byref(n)), "ohcu_frame_commit_lists")

    def stage(self, job_id):
        check(self.lib.ohcu_frame_stage(self.handle, job_id), "ohcu_frame_stage")

    def launch(self, job_id, out_slot=-1, stages=0):
        check(self.lib.ohcu_frame_launch_stages(self.handle, job_id, out_slot, stages), "ohcu_frame_launch")

    def flush(self, job_id):
        check(self.lib.ohcu_frame_flush(self.handle, job_id), "ohcu_frame_flush")

    def sync(self, job_id=None):
        if job_id is None:
            check(self.lib.ohcu_sync_all(self.handle), "ohcu_sync_all")
        else:
            check(self.lib.ohcu_frame_sync(self.handle, job_id), "ohcu_frame_sync")

    def set_profiling(self, on):
        check(self.lib.ohcu_set_profiling(self.handle, int(on)), "ohcu_set_profiling")

    def stage_ms(self, job_id):
        ms = (C.c_float * abi.NUM_STAGES)()
        check(self.lib.ohcu_job_stage_ms(self.handle, job_id, C.byref(ms)), "ohcu_job_stage_ms")
        return list(ms)

    def kernel_launches(self):
        return int(self.lib.ohcu_kernel_launches(self.handle))

    def flush_l2(self):
        check(self.lib.ohcu_flush_l2(self.handle), "ohcu_flush_l2")

    def timer_start(self):
        check(self.lib.ohcu_timer_start(self.handle), "ohcu_timer_start")

    def timer_stop(self):
        ms = C.c_float()
        check(self.lib.ohcu_timer_stop(self.handle, C.byref(ms)), "ohcu_timer_stop")
        return ms.value

    def pinned_picture(self, slot_pitch=False):
        """Output picture in pinned host memory (for download_async).  With slot_pitch the rows
        are spaced like the device planes (ohcu_slot_pitch), which makes the download one
        linear copy per plane; the returned arrays are views of the picture columns."""
        planes = []
        item = np.dtype(self.dtype).itemsize
        for c in range(self.nplanes):
            h, w = self.plane_shape(c)
            pitch = C.c_ssize_t(w * item)
            if slot_pitch:
                check(self.lib.ohcu_slot_pitch(self.handle, c, C.byref(pitch)), "ohcu_slot_pitch")
            nbytes = h * pitch.value
            p = C.c_void_p()
            check(self.lib.ohcu_host_alloc(nbytes, C.byref(p)), "ohcu_host_alloc")
            buf = (C.c_char * nbytes).from_address(p.value)
            planes.append(np.frombuffer(buf, dtype=self.dtype).reshape(h, pitch.value // item)[:, :w])
        return planes

    # ---- convenience: one frame, synchronously (tests, smoke) -----------------
    def reconstruct(self, fj, refs, out_slot=None, out_init=None, stages=0, use_deblock=True, use_sao=True, sparse=False,
                    levels=None, derive_bs=None, inplace=False):
        """Upload refs to slots 0.., run one frame job, return the reconstructed planes."""
        if out_slot is None:
            out_slot = self.sp.dpb_slots - 1
        for i, r in enumerate(refs):
            if r is not None and i != out_slot:
                self.upload(i, r)
        if out_init is not None:
            self.upload(out_slot, out_init)
        j = self.begin(out_slot)
        if inplace:
            self.append_inplace(j, fj, coeffs="sparse" if sparse else "dense", levels=levels, use_deblock=use_deblock,
                                use_sao=use_sao, derive_bs=derive_bs)
        else:
            self.append(j, fj, use_deblock=use_deblock, use_sao=use_sao, sparse=sparse, levels=levels, derive_bs=derive_bs)
        self.stage(j)
        self.launch(j, out_slot, stages)
        self.sync(j)
        return self.download(out_slot)
