/*
 * ohsynth.h — deterministic synthetic work-list generator (test / bench input).
 *
 * Produces, for one picture, exactly what the reference decoder's control plane
 * would hand the C ABI (include/ohcu.h): PU records in hls_prediction_unit order,
 * TU records + dense int16 coefficient blocks as ff_hevc_hls_transform sees them
 * (lc->tu.coeffs, hevcdec.h:333), boundary-strength planes, qp_y_tab, DBParams,
 * is_pcm, SAOParams, slice addresses and tile ids.  Seeded (xoshiro256**), no
 * libc rand, so the same (params, seed) gives the same bytes everywhere.
 *
 * Decoder invariants respected (parity hazards):
 *   - CUs from a CTB quadtree down to 8x8, part modes incl. AMP; 8x4 / 4x8 PUs are
 *     uni-predicted only (hevc_mvs.c:549-550)
 *   - coefficients are zero after the last significant position in 4x4-CG diagonal
 *     scan order and col_limit is derived as hevc_cabac.c:1793-1805
 *     (unless dense_coeffs, the hazard-1 test: dense blocks with arbitrary col_limit)
 *   - chroma TB geometry per chroma_format_idc (4:2:0 4x4-luma groups share one 4x4
 *     chroma TB; 4:2:2 has two stacked square chroma TBs)
 *   - intra CUs (intra_pct): one ohcu_iu per predicted TB in hls_transform_unit order (luma, then the
 *     chroma blocks; the four 4x4 luma blocks of an 8x8 share the chroma of the parent), availability flags
 *     from what was reconstructed earlier in the same slice and tile; their coded TBs carry OHCU_TUF_INTRA
 *   - bs only on the 8x8 grid at CU/PU/TU boundaries; Cr SAO type/class copy Cb's
 *     (hevcdec.c:1261-1264); offset_val built as hevcdec.c:1289-1299
 */
#ifndef OHSYNTH_H
#define OHSYNTH_H
#include "../../include/ohcu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ohsynth_params {
    ohcu_seq_params sp;
    uint64_t seed;
    int32_t weighted;        /* explicit weighted prediction (uni_w / bi_w)          */
    int32_t n_slices;        /* >= 1, split on CTB rows                              */
    int32_t slices_no_lf;    /* 1: odd slices get loop_filter_across_slices = 0      */
    int32_t tile_cols;       /* >= 1 uniform tile columns (tile ids only)            */
    int32_t tile_rows;
    int32_t rext;            /* transform_skip / rdpcm / rotation / cross-component  */
    int32_t bypass_pct;      /* % of CUs cu_transquant_bypass (needs sp.pcm_or_bypass) */
    int32_t far_mv_pct;      /* % of MVs drawn from +-2048 px instead of +-64 px     */
    int32_t cbf_pct;         /* % of TBs coded                                       */
    int32_t dc_pct;          /* % of coded TBs with last = (0,0) -> idct_dc          */
    int32_t dst_pct;         /* % of 4x4 luma TBs through idct_4x4_luma              */
    int32_t dense_coeffs;    /* hazard test: dense blocks + arbitrary col_limit      */
    int32_t bi_pct;          /* % of PUs bi-predicted (rest split L0/L1)             */
    int32_t split_pct[3];    /* CU split probability at 64, 32, 16                   */
    int32_t bs_pct[2];       /* P(bs=1), P(bs=2) in % on block boundaries            */
    int32_t sao_pct[2];      /* P(band), P(edge) in % per CTB and component type     */
    int32_t n_refs;          /* reference pictures used (DPB slots 0..n_refs-1), <= 4 */
    int32_t intra_pct;       /* % of CUs intra-predicted (2Nx2N, NxN at 8x8): ohcu_iu records */
    int32_t intra_drop_pct;  /* % of intra blocks with random availability flags cleared      */
} ohsynth_params;

/* one leaf of a transform tree (hls_transform_tree, hevcdec.c:1549): what
 * ff_hevc_deblocking_boundary_strengths is called with, coded or not */
typedef struct ohsynth_leaf {
    uint16_t x0, y0;
    uint8_t  log2_size;
    uint8_t  cbf_luma;
    uint8_t  pad[2];
} ohsynth_leaf;

typedef struct ohsynth_frame {
    int32_t n_slices;  ohcu_slice_tab *slices;
    int32_t n_pu;      ohcu_pu *pus;
    int32_t n_tu;      ohcu_tu *tus;
    int64_t ncoeff;    int16_t *coeffs;
    int32_t bs_width, bs_height;           uint8_t *vbs, *hbs;
    int32_t min_cb_width, min_cb_height;   int8_t *qp_y_tab;
    int32_t ctb_width, ctb_height;         ohcu_dbparams *deblock;
    int32_t min_pu_width, min_pu_height;   uint8_t *is_pcm;
    ohcu_sao *sao;
    uint8_t *filter_slice_edges;
    int32_t *tab_slice_address;
    int32_t *tile_id;
    int32_t n_leaf;    ohsynth_leaf *leaves;    /* transform-tree leaves in decoding order */
    int32_t n_iu;      ohcu_iu *ius;            /* intra blocks in decoding order          */
} ohsynth_frame;

void ohsynth_default_params(ohsynth_params *p, int width, int height, int bit_depth,
                            int chroma_format_idc);
ohsynth_frame *ohsynth_generate(const ohsynth_params *p);
void ohsynth_free(ohsynth_frame *f);
/* reference picture plane: uniform noise blended 50% with a smooth gradient */
void ohsynth_fill_plane(uint8_t *plane, ptrdiff_t stride, int w, int h, int bit_depth,
                        uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif
