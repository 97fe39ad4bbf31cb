/*
 * replay.h — frame-level CPU replay of one frame's work lists.
 * TEST INFRASTRUCTURE ONLY (see hevc_oracle.h for the rule).
 *
 * A frame job is exactly what the decoder hands the C ABI (include/ohcu.h) for
 * one picture, plus host pictures for the DPB.  Two replays consume it:
 *   or_replay_frame()    oracle/replay.c   — pure restatement (hevc_oracle.c leaves)
 *   ohref_replay_frame() oracle/ref_shim.c — the reference's OWN code: its C
 *                        HEVCDSPContext table, its luma_mc_* / chroma_mc_* drivers
 *                        (hevcdec.c:1744-2049), and its ff_hevc_hls_filters driver
 *                        (hevc_filter.c:264-762, 1211-1245); only built when
 *                        /root/reference is present (oracle/Makefile -> oracle/_ref/).
 * Both process in the reference's order: all PUs (hls_prediction_unit order),
 * all TUs of inter CUs (transform_add), then the intra blocks in decoding order, each
 * predicted from the samples reconstructed so far and followed by its own transform_add
 * (hevcdec.c:1335-1526; inter reconstruction never reads the current picture, so moving
 * it ahead of the intra blocks changes nothing), then per CTB in raster order the one-CTB-delayed
 * deblock + SAO of ff_hevc_hls_filters (hevcdec.c:2898-2903).
 */
#ifndef OH_REPLAY_H
#define OH_REPLAY_H

#include "../include/ohcu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oh_picture {
    uint8_t  *data[3];
    ptrdiff_t stride[3];       /* bytes */
} oh_picture;

typedef struct oh_frame_job {
    ohcu_seq_params sp;
    int32_t n_slices;
    const ohcu_slice_tab *slices;
    int32_t n_pu;
    const ohcu_pu *pus;
    int32_t n_tu;
    const ohcu_tu *tus;
    const int16_t *coeffs;        /* arena; never modified (blocks are copied) */
    int64_t ncoeff;
    const ohcu_deblock_in *deblock;   /* NULL = no deblocking */
    const ohcu_sao_in *sao;           /* NULL = no SAO        */
    const oh_picture *dpb;            /* sp.dpb_slots reference pictures */
    oh_picture out;                   /* reconstructed picture (written) */
    int32_t stages;                   /* bit mask: 1 MC, 2 residual (+ intra blocks), 4 deblock, 8 SAO; 0 = all */
    int32_t n_iu;                     /* intra blocks (hls_transform_unit order); their coded TBs are the tus */
    const ohcu_iu *ius;               /* with OHCU_TUF_INTRA at the same position                              */
} oh_frame_job;

#define OH_STAGE_MC       1
#define OH_STAGE_RESIDUAL 2
#define OH_STAGE_DEBLOCK  4
#define OH_STAGE_SAO      8

/* ---- boundary strengths (bs derivation on the device) ----
 * ff_hevc_deblocking_boundary_strengths (hevc_filter.c:1006-1125) with boundary_strength
 * (:764-880, the TEST_MV_POC variant hevc.h:33 selects) fills vertical_bs / horizontal_bs
 * per transform-tree leaf.  or_boundary_strengths() derives the same planes from what the
 * C ABI already carries - the PU list, the coded luma TBs, the slice tables and the per-CTB
 * slice / tile tables; ohref_boundary_strengths() drives the reference's own function over
 * the transform-tree leaves (which only it reads). */
typedef struct oh_leaf {          /* one leaf of hls_transform_tree (hevcdec.c:1549), coded or not */
    uint16_t x0, y0;
    uint8_t  log2_size, cbf_luma, pad[2];
} oh_leaf;

typedef struct oh_bs_job {
    ohcu_seq_params sp;
    int32_t n_slices;  const ohcu_slice_tab *slices;
    int32_t n_pu;      const ohcu_pu *pus;
    int32_t n_tu;      const ohcu_tu *tus;        /* coded TBs; the luma ones count */
    int32_t n_leaf;    const oh_leaf *leaves;     /* decoding order */
    const uint8_t *filter_slice_edges;            /* per CTB: slice_loop_filter_across_slices_enabled_flag; NULL = 1 */
    const int32_t *tab_slice_address;             /* per CTB; NULL = one slice */
    const int32_t *tile_id;                       /* per CTB; NULL = one tile  */
    const uint8_t *deblock_disabled;              /* per CTB: slice_deblocking_filter_disabled_flag; NULL = 0 */
    uint8_t *vbs, *hbs;                           /* out: ((width >> 2) + 1) * ((height >> 2) + 1) each */
    int32_t n_iu;      const ohcu_iu *ius;        /* intra blocks; the luma ones are the leaves of the intra CUs */
} oh_bs_job;

int or_boundary_strengths(const oh_bs_job *job);
int ohref_boundary_strengths(const oh_bs_job *job);

/* Cropped copy-out: the conformance-window offset of ff_hevc_output_frame (hevc_refs.c:335-341)
 * and the row-by-row memcpy into tight planes of oh_output_cropped_cpy_from_layer
 * (openhevc.c:722-735).  Parity unpinned: openhevc.c needs a whole opened decoder to call; the
 * restatement is a pointer offset and a memcpy per row and is also checked against array
 * slicing in tests/test_output.py.  src strides in bytes; dst planes tight. */
void or_output_cropped(const uint8_t *const src[3], const ptrdiff_t stride[3], int nplanes, int pixel_shift,
                       int hshift, int vshift, int left, int top, int width, int height, uint8_t *const dst[3]);

/* returns 0, or a negative OHCU_E* code for malformed jobs */
int or_replay_frame(const oh_frame_job *job);
int ohref_replay_frame(const oh_frame_job *job);

#ifdef __cplusplus
}
#endif
#endif
