/*
 * ohcu_glue.h — decoder-side glue between the reference decoder (wei1ji/HEVC, libavcodec/hevcdec.c, hevc_cabac.c,
 * hevc_filter.c, hevc_refs.c) and the batched B200 backend (include/ohcu.h).  Compiled INTO the decoder, against the
 * decoder's own types; integration/ohcu_openhevc.patch adds the calls at the reference's call sites:
 *
 *   set_sps                      hevcdec.c:447    ff_hevc_dsp_init(...)            + ff_ohcu_set_sps
 *   hevc_frame_start             hevcdec.c:3575                                    + ff_ohcu_frame_start
 *   slice header / ref lists     hevcdec.c (after ff_hevc_slice_rpl)               + ff_ohcu_slice_start
 *   hls_prediction_unit          hevcdec.c:2250-2306  luma_mc_* / chroma_mc_*     -> ff_ohcu_emit_pu
 *   ff_hevc_hls_transform{,_c}   hevc_cabac.c:1695-1906  idct / skip / rdpcm / transform_add -> ff_ohcu_emit_tu
 *   hls_transform_unit           hevcdec.c:1430-1445, :1460-1475  transform_add of an uncoded chroma block with
 *                                cross-component prediction                         -> ff_ohcu_emit_cross
 *   intra_pred                   hevcpred_template.c:30-345 (every s->hpc.intra_pred[] call of hls_transform_unit,
 *                                hevcdec.c:1335-1526), after its availability flags (:103-107) -> ff_ohcu_emit_iu
 *   hls_pcm_sample               hevcdec.c:1686-1721  put_pcm                     -> ff_ohcu_emit_pcm
 *   ff_hevc_hls_filter           hevc_filter.c:1211-1233  deblock + SAO per CTB   -> nothing (whole picture at frame end)
 *   last CTB of the picture      hevcdec.c:2901-2903                               + ff_ohcu_frame_end
 *   ff_hevc_output_frame         hevc_refs.c:335-341                               + ff_ohcu_fetch_frame
 *   ff_hevc_unref_frame          hevc_refs.c:131                                   + ff_ohcu_release_frame
 *
 * A DPB slot of the device is the index of the HEVCFrame in s->DPB[32] (hevcdec.h:590).
 * Intra CUs are predicted on the device (intra_kernel, csrc/intra.cu) from ohcu_iu records carrying the availability flags
 * the reference's own intra_pred derived.  Two cases stay on the CPU path, picture by picture: bit depths above 12 (the
 * reference has no predictors for them, hevcpred.c:62-76) and constrained_intra_pred_flag; ff_ohcu_frame_end uploads
 * such an I picture as a reference, and a P / B picture that meets an intra CU under them fails with OHCU_EINVAL.
 */
#ifndef OHCU_GLUE_H
#define OHCU_GLUE_H

#include "ohcu.h"

struct HEVCContext;
struct HEVCLocalContext;
struct HEVCFrame;
struct HEVCSPS;
struct MvField;
struct AVFrame;

typedef struct OHCUGlue {
    ohcu_ctx *ctx;
    int device;
    int active;                    /* the current picture is reconstructed on the device                  */
    int failed;                    /* first error of the current picture (OHCU_E*), reported at frame end  */
    int job, out_slot;
    ohcu_frame_arenas a;
    ohcu_frame_counts n;
    int cur_slice;
    int cur_ctb;                   /* CTB (raster) the records being emitted belong to                     */
    uint32_t ctb_coef_base;        /* coefficients of the CTB's earlier blocks (sparse form)               */
    int64_t last_luma_tu;          /* index of the most recent luma TB (cross-component prediction)        */
    int last_luma_x, last_luma_y;
    int intra_ok;                  /* the context takes intra blocks (ohcu_seq_params.intra)               */
    int64_t last_iu[3];            /* per plane: the most recent intra block (its coded TB follows it)     */
} OHCUGlue;

int  ff_ohcu_set_sps(struct HEVCContext *s, const struct HEVCSPS *sps);
int  ff_ohcu_frame_start(struct HEVCContext *s);
void ff_ohcu_slice_start(struct HEVCContext *s);
void ff_ohcu_emit_pu(struct HEVCContext *s, int x0, int y0, int nPbW, int nPbH, const struct MvField *mv);
void ff_ohcu_emit_tu(struct HEVCContext *s, struct HEVCLocalContext *lc, int x0, int y0, int c_idx);
void ff_ohcu_emit_cross(struct HEVCContext *s, struct HEVCLocalContext *lc, int x0, int y0, int c_idx, int log2_size);
int  ff_ohcu_emit_iu(struct HEVCContext *s, int x0, int y0, int log2_size, int c_idx, int mode, int cand);
void ff_ohcu_emit_pcm(struct HEVCContext *s, int x0, int y0, int log2_size, int c_idx, const void *samples);
int  ff_ohcu_frame_end(struct HEVCContext *s);
int  ff_ohcu_fetch_frame(struct HEVCContext *s, struct HEVCFrame *frame);
void ff_ohcu_release_frame(struct HEVCContext *s, struct HEVCFrame *frame);
void ff_ohcu_close(struct HEVCContext *s);

#endif
