/*
 * ohcu.h — C ABI of the B200-native HEVC pixel-reconstruction backend.
 *
 * This is the drop-in boundary for the reference decoder's HEVCDSPContext hot
 * path (reference: libavcodec/hevcdsp.h:51-166, filled by ff_hevc_dsp_init,
 * libavcodec/hevcdsp.c:311-587).  The reference calls the table once per block
 * (PU / TU / 8-sample edge / CTB); this ABI is one level up: the decoder's C
 * control plane appends per-frame work lists and the library runs them as
 * batched sm_100a launches over a device-resident DPB.
 *
 * Plain C, plain pointers and sizes.  No torch / C++ types cross this line.
 * Every entry point returns 0 on success or a negative OHCU_E* code; nothing
 * aborts or throws across the ABI (the reference table returns void and cannot
 * fail; CUDA errors surface at flush/sync).
 */
#ifndef OHCU_H
#define OHCU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OHCU_ABI_VERSION 1

/* ---- error codes -------------------------------------------------------- */
#define OHCU_OK          0
#define OHCU_EINVAL     -1   /* bad argument / unsupported parameter          */
#define OHCU_ENOMEM     -2   /* host or device allocation failed              */
#define OHCU_ECUDA      -3   /* CUDA runtime error (see ohcu_last_error)      */
#define OHCU_ESTATE     -4   /* call out of order (e.g. append before begin)  */
#define OHCU_ENODEV     -5   /* no usable sm_100 device                       */
#define OHCU_EOVERFLOW  -6   /* work list larger than the job's arena         */

/* ---- constants mirrored from the reference ------------------------------ */
#define OHCU_MAX_PB_SIZE   64   /* MAX_PB_SIZE, hevcdec.h:51 (int16 tmp stride) */
#define OHCU_MAX_TB_SIZE   32   /* MAX_TB_SIZE, hevcdec.h:50                    */
#define OHCU_MAX_DPB       32   /* HEVCContext.DPB[32], hevcdec.h:590           */
#define OHCU_MAX_REFS      16   /* luma_weight_l0[16], hevc_ps.h:210            */
#define OHCU_MAX_SLICES    64   /* MAX_SLICES_IN_FRAME, hevcdec.h:361           */

/* PredFlag, hevcdec.h:202-207 */
#define OHCU_PF_L0 1
#define OHCU_PF_L1 2
#define OHCU_PF_BI 3

/* SAOType / SAOEOClass, hevcdec.h:247-259 */
#define OHCU_SAO_NOT_APPLIED 0
#define OHCU_SAO_BAND        1
#define OHCU_SAO_EDGE        2
#define OHCU_SAO_EO_HORIZ 0
#define OHCU_SAO_EO_VERT  1
#define OHCU_SAO_EO_135D  2
#define OHCU_SAO_EO_45D   3

typedef struct ohcu_ctx ohcu_ctx;

/* ---- sequence-level parameters (what set_sps() derives, hevcdec.c:428-476) */
typedef struct ohcu_seq_params {
    int32_t width, height;        /* luma samples; multiples of the min CB size   */
    int32_t bit_depth;            /* 8, 9, 10, 12 or 14 (hevcdsp.c:514-545). One depth for luma+chroma:
                                     ff_hevc_dsp_init(&s->hevcdsp, bit_depth[LUMA]),
                                     hevcdec.c:447                                 */
    int32_t chroma_format_idc;    /* 0 mono, 1 4:2:0, 2 4:2:2, 3 4:4:4            */
    int32_t log2_ctb_size;        /* 4..6                                          */
    int32_t log2_min_cb_size;     /* >= 3 ; qp_y_tab granularity                   */
    int32_t log2_min_pu_size;     /* is_pcm granularity (sps->log2_min_pu_size)    */
    int32_t dpb_slots;            /* device DPB size, <= OHCU_MAX_DPB              */
    int32_t jobs;                 /* frames in flight (ring of job arenas), >= 1   */
    int32_t sao_enabled;          /* sps->sao_enabled_flag                         */
    int32_t pcm_or_bypass;        /* pcmf of hevc_filter.c:541-543:
                                     (pcm_enabled && pcm.loop_filter_disable) ||
                                     pps->transquant_bypass_enable_flag            */
    int32_t cb_qp_offset;         /* pps_cb_qp_offset (chroma_tc, hevc_filter.c:72) */
    int32_t cr_qp_offset;         /* pps_cr_qp_offset                              */
    int32_t no_tile_filter;       /* tiles_enabled && !loop_filter_across_tiles
                                     (sao_filter_CTB, hevc_filter.c:281-282)       */
    int32_t max_pus;              /* arena capacity per job; 0 = worst case        */
    int32_t max_tus;              /* arena capacity per job; 0 = worst case        */
    int64_t max_coeffs;           /* int16 arena capacity per job; 0 = worst case  */
    int32_t intra;                /* 1: pictures may carry intra-predicted blocks (ohcu_iu): the jobs
                                     get a residual plane set and an arena for the records.  0: a job
                                     with intra units is refused (OHCU_ESTATE).  bit_depth <= 12:
                                     ff_hevc_pred_init has no 14-bit predictors (hevcpred.c:62-76)  */
    int32_t max_ius;              /* arena capacity per job; 0 = worst case        */
    int32_t intra_smoothing_disabled;  /* sps->intra_smoothing_disabled_flag (hevcpred_template.c:275) */
    int32_t strong_intra_smoothing;    /* sps->sps_strong_intra_smoothing_enable_flag (:282)          */
} ohcu_seq_params;

/* ---- per-slice tables (SliceHeader weights + RefPicList, hevcdec.c:202-306) */
typedef struct ohcu_slice_tab {
    uint8_t ref_slot[2][OHCU_MAX_REFS];   /* DPB slot of refPicList[X].ref[i]     */
    uint8_t weight_flag;                  /* (P && weighted_pred) || (B && weighted_bipred),
                                             hevcdec.c:1755-1756                   */
    uint8_t luma_log2_weight_denom;
    uint8_t chroma_log2_weight_denom;
    uint8_t pad;
    int16_t luma_weight[2][OHCU_MAX_REFS];
    int16_t luma_offset[2][OHCU_MAX_REFS];
    int16_t chroma_weight[2][OHCU_MAX_REFS][2];
    int16_t chroma_offset[2][OHCU_MAX_REFS][2];
} ohcu_slice_tab;

/* ---- one inter PU: emitted where hls_prediction_unit calls luma_mc_* /
 *      chroma_mc_* (hevcdec.c:2250-2306)                                      */
typedef struct ohcu_pu {
    uint16_t x0, y0;          /* luma position                                    */
    uint8_t  w, h;            /* luma nPbW, nPbH (4..64)                          */
    uint8_t  pred_flag;       /* OHCU_PF_L0 / _L1 / _BI                           */
    uint8_t  slice;           /* index of the ohcu_slice_tab to use               */
    int16_t  mv[2][2];        /* [list][x,y] quarter-pel (Mv, hevcdec.h:297-300)  */
    uint8_t  ref_idx[2];      /* MvField.ref_idx                                  */
    uint8_t  pad[2];
} ohcu_pu;                    /* 20 bytes */

/* ---- one coded transform block: emitted where ff_hevc_hls_transform{,_c}
 *      end in transform_add (hevc_cabac.c:1818, :1905)                        */
#define OHCU_TU_IDCT    0     /* idct[log2-2](coeffs, col_limit)                  */
#define OHCU_TU_DC      1     /* idct_dc[log2-2](coeffs)                          */
#define OHCU_TU_DST4    2     /* idct_4x4_luma(coeffs)                            */
#define OHCU_TU_SKIP    3     /* transform_skip(coeffs, log2)                     */
#define OHCU_TU_BYPASS  4     /* cu_transquant_bypass: coeffs are the residual    */

#define OHCU_TUF_RDPCM      1 /* transform_rdpcm after skip / on bypass           */
#define OHCU_TUF_RDPCM_VERT 2 /* mode argument of transform_rdpcm (1 = vertical)  */
#define OHCU_TUF_ROTATE     4 /* 4x4 transform_skip rotation, hevc_cabac.c:1735   */
#define OHCU_TUF_CROSS      8 /* cross-component prediction, hevc_cabac.c:1810    */
#define OHCU_TUF_INTRA     16 /* the block sits on an intra-predicted TB: its residual is added
                                 after the prediction of the ohcu_iu at the same position  */

typedef struct ohcu_tu {
    uint16_t x0, y0;          /* position in samples of plane c_idx               */
    uint8_t  c_idx;           /* 0 Y, 1 Cb, 2 Cr                                  */
    uint8_t  log2_size;       /* 2..5                                             */
    uint8_t  mode;            /* OHCU_TU_*                                        */
    uint8_t  flags;           /* OHCU_TUF_*                                       */
    int8_t   col_limit;       /* idct only, as hevc_cabac.c:1793-1805             */
    int8_t   res_scale_val;   /* OHCU_TUF_CROSS only (lc->tu.res_scale_val)       */
    uint16_t pad;             /* sparse form: number of ohcu_coef of the block    */
    uint32_t coeff_off;       /* int16 offset of the N*N block in the arena
                                 (sparse form: inside the CTB's area)             */
    uint32_t cross_tu;        /* OHCU_TUF_CROSS: index (in this frame's TU list)
                                 of the luma TB whose residual is coeffs_y        */
} ohcu_tu;                    /* 20 bytes */

/* ---- one intra-predicted transform block: emitted where hls_transform_unit
 *      calls s->hpc.intra_pred[log2 - 2](s, x0, y0, c_idx) (hevcdec.c:1335-1526), coded or not, in
 *      that order.  A block reads the reconstructed samples left of and above it, so the list order is
 *      the reconstruction order inside a CTB; across CTBs the device runs a wavefront (left, up-left,
 *      up and up-right CTB first), which is the dependency the z-scan availability test allows.
 *      cand: the five FINAL availability flags of hevcpred_template.c:103-107 (lc->na.cand_* combined
 *      with the z-scan order test); constrained_intra_pred_flag must be 0.                          */
#define OHCU_CAND_BOTTOM_LEFT 1
#define OHCU_CAND_LEFT        2
#define OHCU_CAND_UP_LEFT     4
#define OHCU_CAND_UP          8
#define OHCU_CAND_UP_RIGHT   16
#define OHCU_IU_CBF          32   /* a coded TB (ohcu_tu with OHCU_TUF_INTRA, same c_idx, x0, y0, log2_size)
                                     carries this block's residual (transform_add after the prediction)   */
typedef struct ohcu_iu {
    uint16_t x0, y0;          /* position in samples of plane c_idx               */
    uint8_t  c_idx;           /* 0 Y, 1 Cb, 2 Cr                                  */
    uint8_t  log2_size;       /* 2..5                                             */
    uint8_t  mode;            /* 0 planar, 1 DC, 2..34 angular: lc->tu.intra_pred_mode, or
                                 intra_pred_mode_c after the 4:2:2 tab_mode_idx mapping     */
    uint8_t  cand;            /* OHCU_CAND_* | OHCU_IU_CBF                        */
} ohcu_iu;                    /* 8 bytes */

/* ---- DBParams, hevcdec.h:353-356 */
typedef struct ohcu_dbparams {
    int8_t beta_offset;
    int8_t tc_offset;
} ohcu_dbparams;

/* ---- SAOParams restricted to what the DSP reads (hevc_ps.h:338-347) */
typedef struct ohcu_sao {
    uint8_t type_idx[3];
    uint8_t eo_class[3];
    uint8_t band_position[3];
    uint8_t pad;
    int16_t offset_val[3][5];
} ohcu_sao;                   /* 40 bytes */

/* ---- deblocking side inputs: the arrays hevc_filter.c reads ------------- */
typedef struct ohcu_deblock_in {
    const uint8_t *vertical_bs;    /* bs_width*bs_height bytes, index (x + y*bs_width)>>2,
                                      bs_width=(width>>2)+1, bs_height=(height>>2)+1
                                      (hevcdec.c:127-128)                          */
    const uint8_t *horizontal_bs;
    const int8_t  *qp_y_tab;       /* min_cb_width*min_cb_height (get_qPy, hevc_filter.c:149) */
    const ohcu_dbparams *deblock;  /* ctb_width*ctb_height, raster                 */
    const uint8_t *is_pcm;         /* min_pu_width*min_pu_height, may be NULL when
                                      seq.pcm_or_bypass == 0                       */
} ohcu_deblock_in;

/* ---- deblocking with the boundary strengths DERIVED ON THE DEVICE (SURVEY section 8f
 *      rank 2).  Replaces ff_hevc_deblocking_boundary_strengths (hevc_filter.c:989-1125,
 *      called per transform-tree leaf at hevcdec.c:1677, :1707, :2603, :2687) and its
 *      boundary_strength (:764-880): vertical_bs / horizontal_bs are computed at flush from
 *      the frame's own lists -
 *        PU list + slice tables  ->  MvField of every 4x4 block and the picture each list
 *                                    refers to (the reference compares POCs; here DPB slots)
 *        coded luma TBs          ->  cbf_luma: an edge unit lies on a transform-block
 *                                    boundary with cbf_luma on a side exactly when it lies
 *                                    on the perimeter of a coded luma TB
 *      4x4 blocks no PU covers count as intra (bs 2 on their edges); the luma ohcu_iu blocks are
 *      the leaves of the intra CUs, so an edge unit inside one of them is no boundary (0).  PCM
 *      CUs carry neither PUs nor intra blocks: send the planes (ohcu_set_deblock) for pictures
 *      that use them.  The per-CTB tables say where the
 *      reference suppresses CTB-boundary edges (lc->boundary_flags, hevcdec.c:2822-2837, with
 *      slice_loop_filter_across_slices_enabled_flag of the CTB's slice and
 *      seq.no_tile_filter) and which slices have slice_deblocking_filter_disabled_flag. */
typedef struct ohcu_deblock_derive_in {
    const int8_t  *qp_y_tab;           /* as ohcu_deblock_in                          */
    const ohcu_dbparams *deblock;
    const uint8_t *is_pcm;             /* may be NULL                                 */
    const uint8_t *filter_slice_edges; /* per CTB, NULL = 1 everywhere                */
    const int32_t *tab_slice_address;  /* per CTB, NULL = one slice                   */
    const int32_t *tile_id;            /* per CTB, NULL = one tile                    */
    const uint8_t *deblock_disabled;   /* per CTB, NULL = 0 everywhere                */
} ohcu_deblock_derive_in;

/* ---- SAO side inputs (sao_filter_CTB, hevc_filter.c:264-325) ------------ */
typedef struct ohcu_sao_in {
    const ohcu_sao *sao;               /* ctb_width*ctb_height, raster             */
    const uint8_t  *filter_slice_edges;/* per CTB: slice_loop_filter_across_slices_enabled_flag */
    const int32_t  *tab_slice_address; /* per CTB                                  */
    const int32_t  *tile_id;           /* per CTB in RASTER order
                                          (= pps->tile_id[ctb_addr_rs_to_ts[rs]]);
                                          NULL = single tile                       */
} ohcu_sao_in;

/* ======================================================================== */
/* Library / context                                                        */
/* ======================================================================== */

int         ohcu_abi_version(void);
const char *ohcu_last_error(void);      /* thread-local text of the last failure   */
int         ohcu_device_count(void);

/* replaces ff_hevc_dsp_init(&s->hevcdsp, bit_depth) + pic_arrays_init +
 * the frame pool of hevc_refs.c:79-129 for this path (hevcdec.c:428-476).      */
int ohcu_create(ohcu_ctx **out, int device, const ohcu_seq_params *sp);
int ohcu_destroy(ohcu_ctx *ctx);
int ohcu_get_seq_params(const ohcu_ctx *ctx, ohcu_seq_params *out);

/* ---- device DPB (replaces AVFrame planes of HEVCFrame, hevc_refs.c:79-129) */
/* Host -> DPB slot (reference frames decoded elsewhere, test inputs).
 * strides in BYTES as everywhere in the reference; planes[1..2] ignored for mono. */
int ohcu_slot_upload(ohcu_ctx *ctx, int slot, const uint8_t *const planes[3],
                     const ptrdiff_t strides[3]);
/* DPB slot -> host (output / hash check); waits for the slot's producer. */
int ohcu_slot_download(ohcu_ctx *ctx, int slot, uint8_t *const planes[3],
                       const ptrdiff_t strides[3]);
/* Asynchronous variant into caller-pinned memory; completes at ohcu_slot_sync. */
int ohcu_slot_download_async(ohcu_ctx *ctx, int slot, uint8_t *const planes[3],
                             const ptrdiff_t strides[3]);
int ohcu_slot_sync(ohcu_ctx *ctx, int slot);   /* replaces ff_thread_await_progress */
/* verify_md5 (hevcdec.c:4035-4098) with the picture left where it is: the SEI picture hash is one MD5 (RFC 1321) per plane
 * over its rows of (width << pixel_shift) bytes, little endian; md5[c] receives plane c's digest (zeros for absent
 * planes).  Waits for the slot's producer; synchronous.  MD5 is a serial chain, one thread per plane: ~10^2 ms for a 4K
 * plane - a conformance tool, not a production path.                                          */
int ohcu_slot_md5(ohcu_ctx *ctx, int slot, uint8_t md5[3][16]);
/* Row pitch in bytes of the slots' device planes.  A host plane passed to the download
 * calls with exactly this stride (an AVFrame whose linesize the decoder's get_buffer chose
 * that way) is fetched as ONE linear copy per plane, which runs at the link rate next to
 * concurrent uploads; other strides take a pitched 2-D copy.  The bytes between the end
 * of a row and the start of the next then hold halo samples.                          */
int ohcu_slot_pitch(const ohcu_ctx *ctx, int plane, ptrdiff_t *pitch);

/* ---- output: cropped copy-out.
 * Replaces the conformance-window offset of ff_hevc_output_frame (hevc_refs.c:335-341:
 * data[i] += (left >> hshift << pixel_shift) + (top >> vshift) * linesize[i]) followed by
 * oh_output_cropped_cpy (openhevc.c:662-742: one memcpy per row into the caller's TIGHT
 * planes, row length = cropped width in bytes).  A kernel packs the window of the three
 * planes into one tight device buffer, which then crosses the bus as one linear copy per
 * plane (one in all when the caller's planes follow each other in memory) instead of a
 * pitched 2-D copy.  left/top/width/height are luma samples; chroma uses them shifted by the
 * format's subsampling, so they must be multiples of it.  planes[c] receives
 * (height >> vs) rows of (width >> hs) samples with no gap between rows.               */
typedef struct { int32_t left, top, width, height; } ohcu_crop;
int ohcu_slot_output_cropped(ohcu_ctx *ctx, int slot, const ohcu_crop *crop, uint8_t *const planes[3]);
/* Asynchronous variant into caller-pinned memory; completes at ohcu_slot_sync. */
int ohcu_slot_output_cropped_async(ohcu_ctx *ctx, int slot, const ohcu_crop *crop, uint8_t *const planes[3]);

/* ======================================================================== */
/* Per-frame work lists                                                     */
/* ======================================================================== */

/* hevc_frame_start (hevcdec.c:3575): picks the next job arena of the ring and
 * returns its id (>= 0) or a negative error.  out_slot is the DPB slot the
 * reconstructed picture is written to.                                        */
int ohcu_frame_begin(ohcu_ctx *ctx, int out_slot, int poc);

/* Threads: frame threads (the reference runs up to 16, hevcdec.h:53) may each drive
 * their own job concurrently; the lists of ONE job are filled by one thread at a time.
 * begin fails with OHCU_ESTATE when the ring's next job is still being filled (more
 * frames begun than jobs).  A stage that fails discards the job's lists;
 * ohcu_frame_abort does the same for a job whose appends failed.                  */
int ohcu_frame_abort(ohcu_ctx *ctx, int job);

int ohcu_set_slice(ohcu_ctx *ctx, int job, int slice_idx, const ohcu_slice_tab *tab);

/* replaces the immediate put_hevc_{q,e}pel* calls of hevcdec.c:1781-1786,
 * 1863-1875, 1936-1941, 2033-2048.  recs are copied before return.            */
int ohcu_append_pu(ohcu_ctx *ctx, int job, const ohcu_pu *recs, int n);

/* replaces idct / transform_skip / rdpcm + transform_add of hevc_cabac.c:1727-1818,
 * 1843-1905.  coeffs holds ncoeff int16 for the n records, each record's
 * coeff_off RELATIVE to coeffs[0]; the library rebases them into the arena.    */
int ohcu_append_tu(ohcu_ctx *ctx, int job, const ohcu_tu *recs, int n,
                   const int16_t *coeffs, size_t ncoeff);

/* replaces the s->hpc.intra_pred[] calls of hls_transform_unit (hevcdec.c:1335-1526).  recs are copied
 * before return; the blocks of one CTB must be appended contiguously and in decoding order, CTBs in any
 * order.  The residual of a block with OHCU_IU_CBF arrives through ohcu_append_tu* as a record with
 * OHCU_TUF_INTRA at the same position (either call first).  Needs ohcu_seq_params.intra.              */
int ohcu_append_iu(ohcu_ctx *ctx, int job, const ohcu_iu *recs, int n);

/* Zero-copy variant: reserve ncoeff int16 in the job's PINNED coefficient
 * arena and let the residual parser (lc->tu.coeffs, hevcdec.h:333) write there
 * directly.  *arena_off receives the absolute offset to store in coeff_off of
 * records passed to ohcu_append_tu_inplace.                                     */
int ohcu_reserve_coeffs(ohcu_ctx *ctx, int job, size_t ncoeff, int16_t **ptr,
                        uint32_t *arena_off);
int ohcu_append_tu_inplace(ohcu_ctx *ctx, int job, const ohcu_tu *recs, int n);

/* Sparse variant: only the significant coefficients cross the bus.  The residual
 * parser zeroes an N x N block and then writes one coefficient per significant
 * position (memset at hevc_cabac.c:2120, stores at :2352, :2398, :2693, :2740);
 * here it appends one ohcu_coef per such store instead and the kernel rebuilds
 * the dense blocks in shared memory.  The coefficient area of a CTB is the
 * concatenation of its blocks in append order:
 *   recs[i].coeff_off = block base inside the CTB's area = sum of N*N of the
 *                       CTB's earlier blocks (a multiple of 16, below the number
 *                       of samples of a CTB over all planes: 6144 for a 64x64
 *                       4:2:0 CTB, 12288 for 4:4:4)
 *   recs[i].pad       = number of ohcu_coef of this block
 *   coef.pos          = recs[i].coeff_off + y * N + x
 * coefs holds the pairs of the n blocks back to back, in record order.  The
 * blocks of one CTB must be appended contiguously (as hls_coding_quadtree
 * emits them), and a frame uses either the dense or the sparse form.
 * ohcu_reserve_coefs returns room for n pairs at the end of the job's pinned
 * arena (capacity max_coeffs pairs); passing that pointer as coefs appends
 * without a copy.                                                               */
typedef struct ohcu_coef {
    uint16_t pos;
    int16_t  level;
} ohcu_coef;                  /* 4 bytes */
int ohcu_reserve_coefs(ohcu_ctx *ctx, int job, size_t n, ohcu_coef **ptr);
int ohcu_append_tu_sparse(ohcu_ctx *ctx, int job, const ohcu_tu *recs, int n,
                          const ohcu_coef *coefs, size_t ncoefs);

/* Levels variant (dequantisation on the device).  The residual
 * parser scales every decoded level before it stores it: scale_and_clip_coeff
 * (hevc_cabac.c:2030-2055, called at :2349, :2395, :2690, :2737 unless
 * cu_transquant_bypass) with the per-block parameters of derive_quant_parameters /
 * derive_quant_parameters_c (:1472-1557).  Here the pairs carry the UN-scaled
 * trans_coeff_level and each record comes with those parameters; the kernel computes
 *   coeff = clip_int16((level * scale * scale_m + add) >> shift)       (64-bit product)
 *   shift = bit_depth + log2_size + 10 - 15,  add = 1 << (shift - 1)   (:1481-1482;
 *           log2_transform_range 15: no extended_precision_processing)
 *   scale = level_scale[qp % 6] << (qp / 6)                            (:1470, :1483)
 *   scale_m = 16, or the scaling-list entry of the position (dc entry for the DC
 *           of 16x16 / 32x32 blocks), :2031-2044
 * while it rebuilds the CTB's dense blocks.  OHCU_TU_BYPASS blocks are not scaled.
 * Same layout rules as ohcu_append_tu_sparse; a frame uses one of the three forms.      */
#define OHCU_QM_FLAT 0xff     /* no scaling list for this block (scale_m = 16): list
                                 disabled, or transform_skip with log2_size > 2 (:1487)  */
typedef struct ohcu_tu_quant {
    uint8_t qp;               /* quant_ctx->qp: qp_y, or the mapped chroma qp, plus
                                 qp_bd_offset (:1479, :1511-1527); < 76, the size of
                                 the reference's rem6[] / div6[] (whose last two
                                 entries are zero: qp 74, 75 scale with 40 << 0)       */
    uint8_t matrix_id;        /* 3 * (pred_mode != MODE_INTRA) + c_idx (:1491-1494,
                                 :1546-1550), or OHCU_QM_FLAT                           */
} ohcu_tu_quant;              /* 2 bytes */
int ohcu_append_tu_levels(ohcu_ctx *ctx, int job, const ohcu_tu *recs,
                          const ohcu_tu_quant *quant, int n,
                          const ohcu_coef *levels, size_t nlevels);
/* ScalingList of the active SPS / PPS (hevc_ps.h:605-610): sl[size_id][matrix_id][64],
 * sl_dc[size_id - 2][matrix_id].  Copied before return; needed only when a record's
 * matrix_id is not OHCU_QM_FLAT.                                                        */
int ohcu_set_scaling_list(ohcu_ctx *ctx, int job, const uint8_t sl[4][6][64],
                          const uint8_t sl_dc[2][6]);

/* ======================================================================== */
/* In-place work lists (zero-copy host path)                                */
/* ======================================================================== */
/* The append calls above validate and copy every record on the submitting thread (1.5 ms of
 * host time per 4K frame).  The reference's parser threads share nothing per frame and write
 * their results straight into per-thread / per-frame arrays (HEVCLocalContext, hevcdec.h:508-510;
 * s->sao, s->deblock, s->qp_y_tab, s->vertical_bs ... of pic_arrays_init, hevcdec.c:114-200).
 * The in-place path gives the patched call sites those arrays INSIDE the job's pinned staging:
 *   ohcu_frame_begin -> ohcu_frame_arenas_get (pointers) -> the parser fills records and tables in
 *   place, noting for every CTB it finishes where its records sit (ohcu_ctb_index, the CTB order
 *   of hls_decode_entry / hls_coding_quadtree, hevcdec.c:2864-2910) -> ohcu_frame_commit_lists
 *   (counts only) -> ohcu_frame_flush.
 * Nothing is copied or checked on the host.  A prepare kernel on the job's stream validates every
 * record (bounds, alignment, CTB membership, reference slots, coefficient areas), derives the
 * per-CTB ranges, the covered-CTB flags and the SAO / bs CTB flags that ohcu_set_sao and
 * ohcu_set_deblock_derived compute on the host, and on a malformed list sets the job's error word:
 * the job's kernels then do nothing and ohcu_frame_sync / ohcu_slot_sync of its output slot return
 * OHCU_EINVAL with the offending record in ohcu_last_error().                            */
typedef struct ohcu_ctb_index {      /* per CTB, raster order; zeroed by ohcu_frame_begin      */
    uint32_t first_pu, n_pu;         /* the CTB's PUs in arenas.pus (<= 128; PUs up to 64x64)   */
    uint32_t first_tu, n_tu;         /* its coded TBs in arenas.tus (<= 768)                    */
    uint32_t first_coef, n_coef;     /* sparse / levels: its ohcu_coef pairs in arenas.coefs    */
} ohcu_ctb_index;                    /* 24 bytes */
typedef struct ohcu_iu_index {       /* per CTB, raster order; zeroed by ohcu_frame_begin      */
    uint32_t first_iu, n_iu;         /* the CTB's intra blocks in arenas.ius, decoding order    */
} ohcu_iu_index;                     /* 8 bytes */

typedef struct ohcu_frame_arenas {
    ohcu_pu        *pus;             uint32_t max_pus;
    uint32_t        max_tus;
    ohcu_tu        *tus;
    ohcu_tu_quant  *quant;           /* levels form: parallel to tus                            */
    ohcu_coef      *coefs;           /* sparse / levels form: pairs, CTB after CTB              */
    int16_t        *coeffs;          /* dense form: the same memory seen as int16 (tu.coeff_off
                                        absolute, a multiple of 8)                              */
    uint64_t        max_coefs;       /* capacity in pairs (= 2 int16 each)                      */
    ohcu_ctb_index *ctbs;            /* ctb_width * ctb_height                                  */
    ohcu_slice_tab *slices;          /* OHCU_MAX_SLICES; ref_slot entries not in use = OHCU_NO_REF */
    uint8_t        *vertical_bs, *horizontal_bs;      /* as ohcu_deblock_in                     */
    int8_t         *qp_y_tab;
    ohcu_dbparams  *deblock;
    uint8_t        *is_pcm;
    ohcu_sao       *sao;             /* as ohcu_sao_in / ohcu_deblock_derive_in, per CTB:       */
    uint8_t        *filter_slice_edges;
    int32_t        *tab_slice_address;
    int32_t        *tile_id;
    uint8_t        *deblock_disabled;
    uint8_t        *scaling_list;    /* sl[4][6][64] then sl_dc[2][6] (1548 bytes)              */
    ohcu_iu        *ius;             /* intra blocks (NULL unless ohcu_seq_params.intra)        */
    ohcu_iu_index  *iu_ctbs;         /* ctb_width * ctb_height                                  */
    uint32_t        max_ius, reserved;
} ohcu_frame_arenas;

#define OHCU_NO_REF       0xff       /* ohcu_slice_tab.ref_slot entry that names no picture     */
#define OHCU_COEF_DENSE   1
#define OHCU_COEF_SPARSE  2
#define OHCU_COEF_LEVELS  3
#define OHCU_DEBLOCK_NONE    0
#define OHCU_DEBLOCK_PLANES  1       /* vertical_bs / horizontal_bs were written                */
#define OHCU_DEBLOCK_DERIVED 2       /* bs derived on the device (ohcu_set_deblock_derived)     */

typedef struct ohcu_frame_counts {
    uint32_t n_pu, n_tu;
    uint64_t n_coef;                 /* pairs (sparse / levels) or int16 (dense) written        */
    int32_t  n_slices;
    int32_t  coef_form;              /* OHCU_COEF_*; 0 = no coded block                         */
    int32_t  deblock;                /* OHCU_DEBLOCK_*                                          */
    int32_t  sao;                    /* sao / filter_slice_edges / tab_slice_address / tile_id written */
    int32_t  pcm;                    /* is_pcm written (used when the picture's pcm_or_bypass is set)  */
    int32_t  scaling_list;           /* scaling_list written                                    */
    int32_t  all_inter;              /* the caller's promise that the PUs cover every sample (a P / B
                                        picture without intra CUs): the slot's previous content is then
                                        not carried over.  Checked on the device.                */
    uint32_t n_iu;                   /* intra blocks written to arenas.ius                      */
} ohcu_frame_counts;

int ohcu_frame_arenas_get(ohcu_ctx *ctx, int job, ohcu_frame_arenas *out);
int ohcu_frame_commit_lists(ohcu_ctx *ctx, int job, const ohcu_frame_counts *counts);

/* Picture-level parameters the reference reads from the active PPS for EACH picture
 * (hevc_filter.c:72, :281-282, :541-543; hevcdec.c:2822-2837).  ohcu_create takes the stream's
 * initial values in ohcu_seq_params; a picture whose PPS differs overrides them for its job
 * (between ohcu_frame_begin and ohcu_frame_flush, before ohcu_set_sao / ohcu_set_deblock_derived).  */
typedef struct ohcu_pic_params {
    int32_t pcm_or_bypass, cb_qp_offset, cr_qp_offset, no_tile_filter;
} ohcu_pic_params;
int ohcu_set_pic_params(ohcu_ctx *ctx, int job, const ohcu_pic_params *pp);

/* PCM samples (a22): hls_pcm_sample -> put_pcm (hevcdec.c:1686-1721, hevcdsp_template.c:30-43).  The
 * bit reader stays on the host; what it produced - the N x N samples of one plane, already shifted to
 * the picture's bit depth (<< (BIT_DEPTH - pcm_bit_depth), template :39) - is placed in the picture
 * after prediction + residual and before the loop filters (whose is_pcm / pcm_loop_filter_disable
 * handling is ohcu_deblock_in.is_pcm).  samples: w*h values, uint8_t for 8-bit pictures, uint16_t above,
 * row-major; copied before return.                                                          */
int ohcu_append_pcm(ohcu_ctx *ctx, int job, int x0, int y0, int log2_size, int c_idx, const void *samples);

/* hevc_refs.c:131-166 (ff_hevc_unref_frame / ff_hevc_clear_refs): the decoder no longer needs the
 * picture in `slot`.  Reads of the slot fail with OHCU_ESTATE until it is written again; the memory
 * stays allocated (the DPB is a fixed pool).                                                  */
int ohcu_slot_release(ohcu_ctx *ctx, int slot);

/* replaces the per-edge calls of deblocking_filter_CTB (hevc_filter.c:588-599,
 * 634-654, 686-697, 735-755).  Arrays are copied before return.  Not calling it
 * for a frame = no deblocking (all bs zero).                                    */
int ohcu_set_deblock(ohcu_ctx *ctx, int job, const ohcu_deblock_in *in);
/* same, with the bs planes derived on the device from this job's PU and TU lists
 * (whatever is appended before the flush)                                          */
int ohcu_set_deblock_derived(ohcu_ctx *ctx, int job, const ohcu_deblock_derive_in *in);

/* replaces the per-CTB calls of sao_filter_CTB (hevc_filter.c:349-353, 486-494).
 * Not calling it for a frame = SAO not applied.                                 */
int ohcu_set_sao(ohcu_ctx *ctx, int job, const ohcu_sao_in *in);

/* End of frame (where ff_hevc_hls_filter ran per CTB, hevc_filter.c:1211-1245):
 * upload the job's lists (async, pinned) and enqueue the batched launches
 *   [prepare (in-place lists)] -> mc_kernel -> residual_kernel -> [pcm] -> [intra_kernel] ->
 *   loopfilter_kernel (deblock + SAO fused) -> pad_kernel (halo fill)
 * on the job's stream ([bs_kernel] beside the first two on a side stream when the boundary
 * strengths are derived).  Returns without waiting.                              */
int ohcu_frame_flush(ohcu_ctx *ctx, int job);

/* Split form of flush, used by bench.py to time the launches with the lists
 * already resident in HBM: stage = H2D only, launch = kernels only (may be
 * repeated; out_slot < 0 keeps the job's slot).                                  */
int ohcu_frame_stage(ohcu_ctx *ctx, int job);
int ohcu_frame_launch(ohcu_ctx *ctx, int job, int out_slot);

/* Test / measurement hook: run only some stages (bit mask 1 MC, 2 residual,
 * 4 deblock, 8 SAO; 0 = all).  Without MC the slot's current content is the
 * starting picture.                                                             */
int ohcu_frame_launch_stages(ohcu_ctx *ctx, int job, int out_slot, int stages);

int ohcu_frame_sync(ohcu_ctx *ctx, int job);
int ohcu_sync_all(ohcu_ctx *ctx);

/* Test hook: the boundary-strength planes the job's last launch filtered with (the
 * host's, or the ones derived on the device), bs_width*bs_height bytes each; waits for
 * the job.                                                                         */
int ohcu_job_read_bs(ohcu_ctx *ctx, int job, uint8_t *vertical_bs, uint8_t *horizontal_bs);

/* ---- measurement hooks (bench.py; CUDA events on the job's own stream) ---- */
#define OHCU_STAGE_MC       0
#define OHCU_STAGE_RESIDUAL 1
#define OHCU_STAGE_DEBLOCK  2
#define OHCU_STAGE_SAO      3
#define OHCU_STAGE_TOTAL    4
#define OHCU_NUM_STAGES     5
/* When enabled, every launch brackets each stage with cudaEvents. */
int ohcu_set_profiling(ohcu_ctx *ctx, int enable);
/* Milliseconds of the job's most recent launch, per stage; job must be synced. */
int ohcu_job_stage_ms(ohcu_ctx *ctx, int job, float ms[OHCU_NUM_STAGES]);
/* Number of kernels this context has launched so far. */
int64_t ohcu_kernel_launches(const ohcu_ctx *ctx);
/* Raw stream handle (cudaStream_t) of a job, for event timing by the caller. */
void *ohcu_job_stream(ohcu_ctx *ctx, int job);
/* Write >L2-size scratch so the next launch starts cold. */
int ohcu_flush_l2(ohcu_ctx *ctx);
/* Device-side wall clock over everything enqueued on every job stream between the
 * two calls: start synchronises the device and records an event; stop makes a
 * timing stream wait for every job's last event, records, synchronises and
 * returns the elapsed milliseconds.                                              */
int ohcu_timer_start(ohcu_ctx *ctx);
int ohcu_timer_stop(ohcu_ctx *ctx, float *ms);
/* Pinned host memory for callers without their own allocator (output pictures). */
int ohcu_host_alloc(size_t bytes, void **ptr);
int ohcu_host_free(void *ptr);

/* ======================================================================== */
/* Per-entry test shims                                                     */
/* ======================================================================== */
/* One block through the device code of the frame kernels, with the argument meaning of the
 * reference's table entries (hevcdsp.h:51-166), so that the table semantics can be tested
 * 1:1 against ff_hevc_dsp_init (tests/test_gpu_leaves.py, same cases as the oracle-vs-
 * reference test).  bd = bit depth; luma: 1 = put_hevc_qpel* (mx, my in quarter samples),
 * 0 = put_hevc_epel* (eighths).  Host pointers, synchronous, NOT a production path.
 *   put_hevc_{q,e}pel[idx][my][mx]        hevcdsp.h:97-98, :112-113   int16 dst, stride 64    */
int ohcu_dsp_mc_put(int bd, int luma, int16_t *dst, const uint8_t *src, ptrdiff_t srcstride,
                    int height, int mx, int my, int width);
/*  put_hevc_{q,e}pel_uni                  hevcdsp.h:99-100, :115-116                          */
int ohcu_dsp_mc_uni(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                    ptrdiff_t srcstride, int height, int mx, int my, int width);
/*  put_hevc_{q,e}pel_uni_w                hevcdsp.h:101-102, :117-118                         */
int ohcu_dsp_mc_uni_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                      ptrdiff_t srcstride, int height, int denom, int wx, int ox, int mx, int my,
                      int width);
/*  put_hevc_{q,e}pel_bi (src2: int16, stride 64)   hevcdsp.h:104-106, :119-121               */
int ohcu_dsp_mc_bi(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                   ptrdiff_t srcstride, const int16_t *src2, int height, int mx, int my, int width);
/*  put_hevc_{q,e}pel_bi_w, parameters in the order the implementation takes them
 *  (denom, wx0, wx1, ox0, ox1 — hevcdsp_template.c:3202-3205)   hevcdsp.h:107-110, :122-125   */
int ohcu_dsp_mc_bi_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                     ptrdiff_t srcstride, const int16_t *src2, int height, int denom, int wx0,
                     int wx1, int ox0, int ox1, int mx, int my, int width);
/*  idct[log2-2], idct_dc[log2-2], idct_4x4_luma, transform_skip, transform_rdpcm (in place),
 *  transform_add[log2-2]                                       hevcdsp.h:55-65               */
int ohcu_dsp_idct(int bd, int16_t *coeffs, int log2_size, int col_limit);
int ohcu_dsp_idct_dc(int bd, int16_t *coeffs, int log2_size);
int ohcu_dsp_idct_4x4_luma(int bd, int16_t *coeffs);
int ohcu_dsp_transform_skip(int bd, int16_t *coeffs, int log2_size);
int ohcu_dsp_transform_rdpcm(int bd, int16_t *coeffs, int log2_size, int mode);
int ohcu_dsp_transform_add(int bd, uint8_t *dst, const int16_t *coeffs, ptrdiff_t stride, int n);
/*  idct[log2-2] on nblocks consecutive N x N blocks, WITHOUT any allocation.  No counterpart in the reference:
 *  added for callers that cannot call cudaMalloc (code under CUDA graph
 *  capture, tight serving loops, harnesses that forbid it).  Same results as ohcu_dsp_idct block by block; the
 *  device staging is a static 64 KB array inside the library (held by every GPU that loads it until exit), so
 *  nblocks * N * N * 2 bytes must not exceed OHCU_NOALLOC_MAX_BYTES (OHCU_EINVAL with the numbers in
 *  ohcu_last_error() otherwise).  Runs on the CURRENT device, synchronously.  All calls share the array: they
 *  are serialised inside the library, and anything else that would overlap with them on the GPU races.       */
#define OHCU_NOALLOC_MAX_BYTES 65536
int ohcu_dsp_idct_noalloc(int bd, int16_t *coeffs, int log2_size, int col_limit, int nblocks);
/*  HEVCPredContext.intra_pred[log2_size - 2](s, x0, y0, c_idx)  hevcpred.h:30-41, hevcpred_template.c:30-345 and the
 *  predictors it dispatches to (pred_planar :359, pred_dc :388, pred_angular :419): one N x N block of a plane predicted
 *  IN PLACE from the reconstructed samples left of and above it (the block function with its table
 *  semantics; pictures go through ohcu_iu records and intra_kernel).  plane = sample (0, 0) of the plane (host), stride in
 *  bytes, pw x ph its size in samples, (x, y) the block in plane samples, mode 0 planar, 1 DC, 2..34 angular.
 *  cand = the five final availability flags the reference computes at hevcpred_template.c:103-107 (lc->na.cand_* combined
 *  with the z-scan order test), OHCU_CAND_*.  filter_ok = !sps->intra_smoothing_disabled_flag && (c_idx == 0 || chroma_format_idc == 3),
 *  strong = sps->sps_strong_intra_smoothing_enable_flag.  constrained_intra_pred_flag must be 0.                       */
int ohcu_dsp_intra_pred(int bd, uint8_t *plane, ptrdiff_t stride, int pw, int ph, int x, int y, int log2_size, int c_idx,
                        int mode, int cand, int filter_ok, int strong);
/*  hevc_{h,v}_loop_filter_{luma,chroma}: pix = first Q sample, tc / no_p / no_q per 4-line
 *  segment                                                     hevcdsp.h:128-131             */
int ohcu_dsp_loop_filter_luma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride, int beta,
                              const int *tc, const uint8_t *no_p, const uint8_t *no_q);
int ohcu_dsp_loop_filter_chroma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride,
                                const int *tc, const uint8_t *no_p, const uint8_t *no_q);
/*  sao_band_filter / sao_edge_filter with SAOParams reduced to offset_val[5] and
 *  band_position / eo_class of the plane (src carries the 1-sample halo the edge classes
 *  read).  sao_edge_restore has no per-entry shim: the frame kernel applies the border rules
 *  per CTB (tests/test_gpu_parity.py covers every flag).      hevcdsp.h:67-71               */
int ohcu_dsp_sao_band_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst,
                             ptrdiff_t stride_src, const int16_t *offset_val, int band_position,
                             int width, int height);
int ohcu_dsp_sao_edge_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst,
                             ptrdiff_t stride_src, const int16_t *offset_val, int eo, int width,
                             int height);

#ifdef __cplusplus
}
#endif
#endif /* OHCU_H */
