/*
 * ohcu_glue.c — see ohcu_glue.h.  Decoder-side code: it reads the reference decoder's structures (HEVCContext,
 * SliceHeader, MvField, HEVCTransformContext, SAOParams ... of libavcodec/hevcdec.h, hevc_ps.h) and writes the work
 * lists of include/ohcu.h straight into the job's pinned arenas (the zero-copy path: ohcu_frame_arenas_get /
 * ohcu_frame_commit_lists).  Nothing here computes samples.
 */
#include <stdlib.h>
#include <string.h>

#include "libavcodec/hevcdec.h"
#include "ohcu_glue.h"

#define G(s) ((s)->ohcu)

static void glue_fail(HEVCContext *s, int rc, const char *what)
{
    OHCUGlue *g = G(s);
    if (!g->failed) {
        g->failed = rc ? rc : OHCU_EINVAL;
        av_log(s->avctx, AV_LOG_ERROR, "ohcu: %s failed (%d): %s\n", what, rc, ohcu_last_error());
    }
}

/* set_sps (hevcdec.c:428-476): the device context follows the SPS like the DSP table does (hevcdec.c:447) */
int ff_ohcu_set_sps(HEVCContext *s, const HEVCSPS *sps)
{
    OHCUGlue *g = G(s);
    const HEVCPPS *pps = s->ps.pps;
    ohcu_seq_params sp;
    int rc;
    if (g->ctx) { ohcu_destroy(g->ctx); g->ctx = NULL; }
    memset(&sp, 0, sizeof(sp));
    sp.width = sps->width;  sp.height = sps->height;
    sp.bit_depth = sps->bit_depth[CHANNEL_TYPE_LUMA];
    sp.chroma_format_idc = sps->chroma_format_idc;
    sp.log2_ctb_size = sps->log2_ctb_size;
    sp.log2_min_cb_size = sps->log2_min_cb_size;
    sp.log2_min_pu_size = sps->log2_min_pu_size;
    sp.dpb_slots = FF_ARRAY_ELEMS(s->DPB);
    sp.jobs = 4;
    sp.sao_enabled = sps->sao_enabled_flag;
    /* intra blocks on the device: the depths ff_hevc_pred_init has predictors for (hevcpred.c:62-76) */
    sp.intra = g->intra_ok = sp.bit_depth <= 12;
    sp.intra_smoothing_disabled = sps->intra_smoothing_disabled_flag;
    sp.strong_intra_smoothing = sps->sps_strong_intra_smoothing_enable_flag;
    if (pps) {                                            /* initial values; every picture sets its own (ff_ohcu_frame_start) */
        sp.pcm_or_bypass = (sps->pcm_enabled_flag && sps->pcm.loop_filter_disable_flag) || pps->transquant_bypass_enable_flag;
        sp.cb_qp_offset = pps->pps_cb_qp_offset;
        sp.cr_qp_offset = pps->pps_cr_qp_offset;
        sp.no_tile_filter = pps->tiles_enabled_flag && !pps->loop_filter_across_tiles_enabled_flag;
    }
    rc = ohcu_create(&g->ctx, g->device, &sp);
    if (rc) glue_fail(s, rc, "ohcu_create");
    return rc;
}

/* hevc_frame_start (hevcdec.c:3575): the picture being decoded is s->ref, its slot the index in s->DPB */
int ff_ohcu_frame_start(HEVCContext *s)
{
    OHCUGlue *g = G(s);
    const HEVCSPS *sps = s->ps.sps;
    const HEVCPPS *pps = s->ps.pps;
    ohcu_pic_params pp;
    int rc;
    g->active = 0;
    g->failed = 0;
    if (!g->ctx) return 0;
    /* I pictures the device cannot predict stay on the CPU path (see ohcu_glue.h) */
    if (s->sh.slice_type == HEVC_SLICE_I && (!g->intra_ok || pps->constrained_intra_pred_flag)) return 0;
    g->out_slot = (int)(s->ref - s->DPB);
    g->job = ohcu_frame_begin(g->ctx, g->out_slot, s->poc);
    if (g->job < 0) { glue_fail(s, g->job, "ohcu_frame_begin"); return g->job; }
    pp.pcm_or_bypass = (sps->pcm_enabled_flag && sps->pcm.loop_filter_disable_flag) || pps->transquant_bypass_enable_flag;
    pp.cb_qp_offset = pps->pps_cb_qp_offset;
    pp.cr_qp_offset = pps->pps_cr_qp_offset;
    pp.no_tile_filter = pps->tiles_enabled_flag && !pps->loop_filter_across_tiles_enabled_flag;
    if ((rc = ohcu_set_pic_params(g->ctx, g->job, &pp)) || (rc = ohcu_frame_arenas_get(g->ctx, g->job, &g->a))) {
        glue_fail(s, rc, "ohcu_frame_arenas_get");
        ohcu_frame_abort(g->ctx, g->job);
        return rc;
    }
    memset(&g->n, 0, sizeof(g->n));
    g->cur_slice = -1;
    g->cur_ctb = -1;
    g->ctb_coef_base = 0;
    g->last_luma_tu = -1;
    g->last_iu[0] = g->last_iu[1] = g->last_iu[2] = -1;
    g->active = 1;
    return 0;
}

/* after the slice header is parsed and ff_hevc_slice_rpl built the lists: reference slots and prediction weights
 * of this slice (SliceHeader, hevc_ps.h:199-220; the weighted-prediction switch of hevcdec.c:1755-1756) */
void ff_ohcu_slice_start(HEVCContext *s)
{
    OHCUGlue *g = G(s);
    const SliceHeader *sh = &s->sh;
    const RefPicList *rpl = s->ref->refPicList[s->slice_idx];     /* as hls_prediction_unit takes it, hevcdec.c:2130 */
    ohcu_slice_tab *t;
    if (!g->active) return;
    if (sh->slice_type == HEVC_SLICE_I && !g->intra_ok) { glue_fail(s, OHCU_EINVAL, "intra slice inside a device picture"); return; }
    if (s->slice_idx >= OHCU_MAX_SLICES) { glue_fail(s, OHCU_EOVERFLOW, "slice table"); return; }
    g->cur_slice = s->slice_idx;
    if (g->cur_slice + 1 > g->n.n_slices) g->n.n_slices = g->cur_slice + 1;
    t = &g->a.slices[g->cur_slice];
    memset(t, 0, sizeof(*t));
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < OHCU_MAX_REFS; i++)
            t->ref_slot[l][i] = (rpl && i < rpl[l].nb_refs && rpl[l].ref[i]) ? (uint8_t)(rpl[l].ref[i] - s->DPB) : OHCU_NO_REF;
    t->weight_flag = (sh->slice_type == HEVC_SLICE_P && s->ps.pps->weighted_pred_flag) ||
                     (sh->slice_type == HEVC_SLICE_B && s->ps.pps->weighted_bipred_flag);
    t->luma_log2_weight_denom = sh->luma_log2_weight_denom;
    t->chroma_log2_weight_denom = (uint8_t)sh->chroma_log2_weight_denom;
    for (int i = 0; i < OHCU_MAX_REFS; i++) {
        t->luma_weight[0][i] = sh->luma_weight_l0[i];  t->luma_weight[1][i] = sh->luma_weight_l1[i];
        t->luma_offset[0][i] = sh->luma_offset_l0[i];  t->luma_offset[1][i] = sh->luma_offset_l1[i];
        for (int j = 0; j < 2; j++) {
            t->chroma_weight[0][i][j] = sh->chroma_weight_l0[i][j];  t->chroma_weight[1][i][j] = sh->chroma_weight_l1[i][j];
            t->chroma_offset[0][i][j] = sh->chroma_offset_l0[i][j];  t->chroma_offset[1][i][j] = sh->chroma_offset_l1[i][j];
        }
    }
}

/* records are emitted CTB after CTB (hls_decode_entry walks them); note where each CTB's records begin */
static ohcu_ctb_index *glue_ctb(HEVCContext *s, int x0, int y0)
{
    OHCUGlue *g = G(s);
    const int l2 = s->ps.sps->log2_ctb_size;
    const int ctb = (x0 >> l2) + (y0 >> l2) * s->ps.sps->ctb_width;
    ohcu_ctb_index *ix = &g->a.ctbs[ctb];
    if (ctb != g->cur_ctb) {
        g->cur_ctb = ctb;
        g->ctb_coef_base = 0;
        g->last_luma_tu = -1;
    }
    return ix;
}

/* replaces luma_mc_uni / luma_mc_bi / chroma_mc_uni / chroma_mc_bi of hls_prediction_unit (hevcdec.c:2250-2306) */
void ff_ohcu_emit_pu(HEVCContext *s, int x0, int y0, int nPbW, int nPbH, const MvField *mv)
{
    OHCUGlue *g = G(s);
    ohcu_ctb_index *ix;
    ohcu_pu *pu;
    if (!g->active || g->failed) return;
    if (g->n.n_pu >= g->a.max_pus || g->cur_slice < 0) { glue_fail(s, OHCU_EOVERFLOW, "PU list"); return; }
    ix = glue_ctb(s, x0, y0);
    if (!ix->n_pu) ix->first_pu = g->n.n_pu;
    ix->n_pu++;
    pu = &g->a.pus[g->n.n_pu++];
    pu->x0 = (uint16_t)x0;  pu->y0 = (uint16_t)y0;
    pu->w = (uint8_t)nPbW;  pu->h = (uint8_t)nPbH;
    pu->pred_flag = mv->pred_flag;                         /* PF_L0 / PF_L1 / PF_BI = OHCU_PF_* (hevcdec.h:202-207) */
    pu->slice = (uint8_t)g->cur_slice;
    for (int l = 0; l < 2; l++) {
        pu->mv[l][0] = mv->mv[l].x;  pu->mv[l][1] = mv->mv[l].y;
        pu->ref_idx[l] = mv->ref_idx[l] < OHCU_MAX_REFS ? mv->ref_idx[l] : 0;
    }
    pu->pad[0] = pu->pad[1] = 0;
}

/* replaces the tail of ff_hevc_hls_transform / ff_hevc_hls_transform_c (hevc_cabac.c:1727-1818, :1843-1905): which
 * inverse transform the block takes is decided exactly as there, the coefficients (lc->tu.coeffs, hevcdec.h:333) go
 * out as (position, level) pairs */
/* A coded block of an intra CU: intra_pred has just emitted the block's ohcu_iu (hls_transform_unit predicts, then
 * adds the residual, hevcdec.c:1335-1526).  The record learns that a residual follows, the TB that it waits for the
 * prediction. */
static void glue_mark_intra(HEVCContext *s, const HEVCLocalContext *lc, ohcu_tu *tu)
{
    OHCUGlue *g = G(s);
    ohcu_iu *u;
    if (lc->cu.pred_mode != MODE_INTRA || g->last_iu[tu->c_idx] < 0) return;
    u = &g->a.ius[g->last_iu[tu->c_idx]];
    if (u->x0 != tu->x0 || u->y0 != tu->y0 || u->log2_size != tu->log2_size) return;
    u->cand |= OHCU_IU_CBF;
    tu->flags |= OHCU_TUF_INTRA;
}

/* intra_pred (hevcpred_template.c:30-345), called by hls_transform_unit for every TB of an intra CU: after it has
 * derived cand_bottom_left .. cand_up_right (:103-107) the patched function hands the block over and returns.
 * x0, y0: luma coordinates, as intra_pred takes them.  Returns 1 when the block is the device's. */
int ff_ohcu_emit_iu(HEVCContext *s, int x0, int y0, int log2_size, int c_idx, int mode, int cand)
{
    OHCUGlue *g = G(s);
    const HEVCSPS *sps = s->ps.sps;
    ohcu_iu_index *ix;
    ohcu_iu *u;
    if (!g->active) return 0;
    if (g->failed) return 1;
    if (!g->a.ius || s->ps.pps->constrained_intra_pred_flag) { glue_fail(s, OHCU_EINVAL, "intra CU the device cannot predict"); return 1; }
    if (g->n.n_iu >= g->a.max_ius) { glue_fail(s, OHCU_EOVERFLOW, "intra block list"); return 1; }
    glue_ctb(s, x0, y0);
    ix = &g->a.iu_ctbs[g->cur_ctb];
    if (!ix->n_iu) ix->first_iu = g->n.n_iu;
    ix->n_iu++;
    u = &g->a.ius[g->n.n_iu];
    u->x0 = (uint16_t)(x0 >> sps->hshift[c_idx]);
    u->y0 = (uint16_t)(y0 >> sps->vshift[c_idx]);
    u->c_idx = (uint8_t)c_idx;
    u->log2_size = (uint8_t)log2_size;
    u->mode = (uint8_t)mode;
    u->cand = (uint8_t)(cand & 31);
    g->last_iu[c_idx] = g->n.n_iu;
    g->n.n_iu++;
    return 1;
}

/* hls_transform_unit, hevcdec.c:1430-1445 / :1460-1475: a chroma block without coded coefficients in a TU with
 * cross-component prediction still receives (res_scale_val * coeffs_y) >> 3: a record with no coefficients of its own */
void ff_ohcu_emit_cross(HEVCContext *s, HEVCLocalContext *lc, int x0, int y0, int c_idx, int log2_size)
{
    OHCUGlue *g = G(s);
    const HEVCSPS *sps = s->ps.sps;
    const int n = 1 << log2_size;
    ohcu_ctb_index *ix;
    ohcu_tu *tu;
    if (!g->active || g->failed) return;
    if (g->n.n_tu >= g->a.max_tus) { glue_fail(s, OHCU_EOVERFLOW, "TU list"); return; }
    if (g->last_luma_tu < 0 || g->last_luma_x != x0 || g->last_luma_y != y0) { glue_fail(s, OHCU_EINVAL, "cross-component block without its luma block"); return; }
    ix = glue_ctb(s, x0, y0);
    if (!ix->n_tu) { ix->first_tu = g->n.n_tu; ix->first_coef = (uint32_t)g->n.n_coef; }
    ix->n_tu++;
    tu = &g->a.tus[g->n.n_tu];
    memset(tu, 0, sizeof(*tu));
    tu->x0 = (uint16_t)(x0 >> sps->hshift[c_idx]);
    tu->y0 = (uint16_t)(y0 >> sps->vshift[c_idx]);
    tu->c_idx = (uint8_t)c_idx;
    tu->log2_size = (uint8_t)log2_size;
    tu->mode = OHCU_TU_BYPASS;                       /* zero coefficients, used as they are */
    tu->flags = OHCU_TUF_CROSS;
    tu->res_scale_val = (int8_t)lc->tu.res_scale_val;
    tu->cross_tu = (uint32_t)g->last_luma_tu;
    glue_mark_intra(s, lc, tu);
    tu->coeff_off = g->ctb_coef_base;
    g->ctb_coef_base += (uint32_t)(n * n);
    g->n.n_tu++;
}

void ff_ohcu_emit_tu(HEVCContext *s, HEVCLocalContext *lc, int x0, int y0, int c_idx)
{
    OHCUGlue *g = G(s);
    const HEVCSPS *sps = s->ps.sps;
    const HEVCTransformContext *tr = &lc->transform_ctx;
    const HEVCTransformScanContext *sc = &tr->scan_ctx;
    const int log2 = tr->log2_trafo_size, n = 1 << log2;
    const int16_t *coeffs = lc->tu.coeffs[c_idx ? 1 : 0];
    const int pred_mode_intra = c_idx ? lc->tu.intra_pred_mode_c : lc->tu.intra_pred_mode;
    ohcu_ctb_index *ix;
    ohcu_tu *tu;
    ohcu_coef *pairs;
    int np = 0;
    if (!g->active || g->failed) return;
    if (g->n.n_tu >= g->a.max_tus || g->n.n_coef + (uint64_t)n * n > g->a.max_coefs) { glue_fail(s, OHCU_EOVERFLOW, "TU list"); return; }
    ix = glue_ctb(s, x0, y0);
    if (!ix->n_tu) { ix->first_tu = g->n.n_tu; ix->first_coef = (uint32_t)g->n.n_coef; }
    ix->n_tu++;
    tu = &g->a.tus[g->n.n_tu];
    memset(tu, 0, sizeof(*tu));
    tu->x0 = (uint16_t)(x0 >> sps->hshift[c_idx]);
    tu->y0 = (uint16_t)(y0 >> sps->vshift[c_idx]);
    tu->c_idx = (uint8_t)c_idx;
    tu->log2_size = (uint8_t)log2;
    if (lc->cu.cu_transquant_bypass_flag) {
        tu->mode = OHCU_TU_BYPASS;
        if (tr->explicit_rdpcm_flag || (sps->implicit_rdpcm_enabled_flag && (pred_mode_intra == 10 || pred_mode_intra == 26))) {
            const int vert = sps->implicit_rdpcm_enabled_flag ? (pred_mode_intra == 26) : tr->explicit_rdpcm_dir_flag;
            tu->flags |= OHCU_TUF_RDPCM | (vert ? OHCU_TUF_RDPCM_VERT : 0);
        }
    } else if (tr->transform_skip_flag) {
        tu->mode = OHCU_TU_SKIP;
        if (sps->transform_skip_rotation_enabled_flag && log2 == 2 && lc->cu.pred_mode == MODE_INTRA) tu->flags |= OHCU_TUF_ROTATE;
        if (tr->explicit_rdpcm_flag)
            tu->flags |= OHCU_TUF_RDPCM | (tr->explicit_rdpcm_dir_flag ? OHCU_TUF_RDPCM_VERT : 0);
        else if (sps->implicit_rdpcm_enabled_flag && lc->cu.pred_mode == MODE_INTRA && (pred_mode_intra == 10 || pred_mode_intra == 26))
            tu->flags |= OHCU_TUF_RDPCM | (pred_mode_intra == 26 ? OHCU_TUF_RDPCM_VERT : 0);
    } else if (!c_idx && lc->cu.pred_mode == MODE_INTRA && log2 == 2) {
        tu->mode = OHCU_TU_DST4;
    } else {
        const int max_xy = FFMAX(sc->last_significant_coeff_x, sc->last_significant_coeff_y);
        if (max_xy == 0) {
            tu->mode = OHCU_TU_DC;
        } else {
            int col_limit = sc->last_significant_coeff_x + sc->last_significant_coeff_y + 4;
            if (max_xy < 4) col_limit = FFMIN(4, col_limit);
            else if (max_xy < 8) col_limit = FFMIN(8, col_limit);
            else if (max_xy < 12) col_limit = FFMIN(24, col_limit);
            tu->mode = OHCU_TU_IDCT;
            tu->col_limit = (int8_t)col_limit;
        }
    }
    if (c_idx && lc->tu.cross_pf && g->last_luma_tu >= 0 && g->last_luma_x == x0 && g->last_luma_y == y0) {
        tu->flags |= OHCU_TUF_CROSS;
        tu->res_scale_val = (int8_t)lc->tu.res_scale_val;
        tu->cross_tu = (uint32_t)g->last_luma_tu;
    }
    if (!c_idx) { g->last_luma_tu = g->n.n_tu; g->last_luma_x = x0; g->last_luma_y = y0; }
    glue_mark_intra(s, lc, tu);
    /* the significant coefficients of the block, positions inside the CTB's coefficient area */
    pairs = g->a.coefs + g->n.n_coef;
    for (int i = 0; i < n * n; i++)
        if (coeffs[i]) { pairs[np].pos = (uint16_t)(g->ctb_coef_base + i); pairs[np].level = coeffs[i]; np++; }
    tu->coeff_off = g->ctb_coef_base;
    tu->pad = (uint16_t)np;
    g->ctb_coef_base += (uint32_t)(n * n);
    g->n.n_coef += (uint64_t)np;
    ix->n_coef += (uint32_t)np;
    g->n.n_tu++;
}

/* hls_pcm_sample (hevcdec.c:1686-1721): put_pcm has read the samples with the bit reader; samples = tight N x N */
void ff_ohcu_emit_pcm(HEVCContext *s, int x0, int y0, int log2_size, int c_idx, const void *samples)
{
    OHCUGlue *g = G(s);
    int rc;
    if (!g->active || g->failed) return;
    rc = ohcu_append_pcm(g->ctx, g->job, x0 >> s->ps.sps->hshift[c_idx], y0 >> s->ps.sps->vshift[c_idx], log2_size, c_idx, samples);
    if (rc) glue_fail(s, rc, "ohcu_append_pcm");
}

/* The picture is complete (where the last ff_hevc_hls_filter ran, hevcdec.c:2901-2903): the side arrays the CPU loop
 * filters read (hevc_filter.c:264-325, :525-762) go along as they are, then one flush reconstructs the picture. */
int ff_ohcu_frame_end(HEVCContext *s)
{
    OHCUGlue *g = G(s);
    const HEVCSPS *sps = s->ps.sps;
    const HEVCPPS *pps = s->ps.pps;
    const int nctb = sps->ctb_width * sps->ctb_height;
    int rc;
    if (!g->ctx) return 0;
    if (!g->active) {
        /* a picture the CPU path decoded (intra): it becomes a device reference as it is */
        const uint8_t *planes[3] = { s->frame->data[0], s->frame->data[1], s->frame->data[2] };
        const ptrdiff_t strides[3] = { s->frame->linesize[0], s->frame->linesize[1], s->frame->linesize[2] };
        rc = ohcu_slot_upload(g->ctx, (int)(s->ref - s->DPB), planes, strides);
        if (rc) glue_fail(s, rc, "ohcu_slot_upload");
        return rc;
    }
    g->active = 0;
    if (g->failed) { ohcu_frame_abort(g->ctx, g->job); return g->failed; }
    memcpy(g->a.vertical_bs, s->vertical_bs, (size_t)s->bs_width * s->bs_height);
    memcpy(g->a.horizontal_bs, s->horizontal_bs, (size_t)s->bs_width * s->bs_height);
    memcpy(g->a.qp_y_tab, s->qp_y_tab, (size_t)sps->min_cb_width * sps->min_cb_height);
    memcpy(g->a.deblock, s->deblock, (size_t)nctb * sizeof(ohcu_dbparams));        /* DBParams, hevcdec.h:353-356: same two int8 */
    memcpy(g->a.is_pcm, s->is_pcm, (size_t)sps->min_pu_width * sps->min_pu_height);
    memcpy(g->a.filter_slice_edges, s->filter_slice_edges, (size_t)nctb);
    memcpy(g->a.tab_slice_address, s->tab_slice_address, (size_t)nctb * sizeof(int32_t));
    for (int rs = 0; rs < nctb; rs++) {
        const SAOParams *p = &s->sao[rs];
        ohcu_sao *o = &g->a.sao[rs];
        g->a.tile_id[rs] = pps->tile_id ? pps->tile_id[pps->ctb_addr_rs_to_ts[rs]] : 0;
        memset(o, 0, sizeof(*o));
        for (int c = 0; c < 3; c++) {
            o->type_idx[c] = p->type_idx[c];
            o->eo_class[c] = p->eo_class[c];
            o->band_position[c] = p->band_position[c];
            for (int k = 0; k < 5; k++) o->offset_val[c][k] = p->offset_val[c][k];
        }
    }
    g->n.coef_form = OHCU_COEF_SPARSE;
    g->n.deblock = OHCU_DEBLOCK_PLANES;
    g->n.sao = 1;
    g->n.pcm = 1;
    g->n.all_inter = 0;                     /* skipped / PCM CUs: whatever the PUs do not cover keeps the slot's content */
    if ((rc = ohcu_frame_commit_lists(g->ctx, g->job, &g->n)) || (rc = ohcu_frame_flush(g->ctx, g->job))) {
        glue_fail(s, rc, "ohcu_frame_flush");
        ohcu_frame_abort(g->ctx, g->job);
    }
    return rc;
}

/* ff_hevc_output_frame (hevc_refs.c:335-341): the picture leaves the device into the AVFrame the caller sees.
 * Also reports list errors the device found (OHCU_EINVAL from ohcu_slot_sync). */
int ff_ohcu_fetch_frame(HEVCContext *s, HEVCFrame *frame)
{
    OHCUGlue *g = G(s);
    uint8_t *planes[3] = { frame->frame->data[0], frame->frame->data[1], frame->frame->data[2] };
    const ptrdiff_t strides[3] = { frame->frame->linesize[0], frame->frame->linesize[1], frame->frame->linesize[2] };
    int rc;
    if (!g->ctx) return 0;
    rc = ohcu_slot_download(g->ctx, (int)(frame - s->DPB), planes, strides);
    if (!rc) rc = ohcu_slot_sync(g->ctx, (int)(frame - s->DPB));
    if (rc) glue_fail(s, rc, "ohcu_slot_download");
    return rc;
}

/* ff_hevc_unref_frame (hevc_refs.c:131-166) */
void ff_ohcu_release_frame(HEVCContext *s, HEVCFrame *frame)
{
    OHCUGlue *g = G(s);
    if (g && g->ctx) ohcu_slot_release(g->ctx, (int)(frame - s->DPB));
}

void ff_ohcu_close(HEVCContext *s)
{
    OHCUGlue *g = G(s);
    if (!g) return;
    if (g->ctx) ohcu_destroy(g->ctx);
    free(g);
    s->ohcu = NULL;
}
