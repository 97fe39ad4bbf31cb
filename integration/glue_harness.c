/*
 * glue_harness.c — TEST INFRASTRUCTURE.  Drives the PATCHED reference (integration/ohcu_openhevc.patch applied to a
 * scratch copy of libavcodec, compiled with CONFIG_OHCU=1) and the decoder-side glue (ohcu_glue.c) over one synthetic
 * frame, so that the work lists reach libohcu.so through the reference's own code and types:
 *   every coded block goes through the patched ff_hevc_hls_transform / ff_hevc_hls_transform_c (hevc_cabac.c), with
 *     lc->transform_ctx / lc->cu / lc->tu set the way the residual parser leaves them;
 *   every PU through ff_ohcu_emit_pu with a MvField, the call the patch puts in hls_prediction_unit (which itself
 *     reads the bitstream and cannot be driven without one);
 *   slice tables through ff_ohcu_slice_start from s->sh and s->ref->refPicList;
 *   the loop-filter inputs from s->vertical_bs / horizontal_bs / qp_y_tab / deblock / sao / ... at ff_ohcu_frame_end;
 *   the picture comes back through ff_ohcu_fetch_frame.
 * tests/test_gpu_integration.py compares it with the UNPATCHED reference's picture (oracle/_ref, ohref_replay_frame).
 */
#include <stdlib.h>
#include <string.h>

#include "libavcodec/hevcdec.h"
#include "ohcu_glue.h"
#include "replay.h"

static void *zalloc(size_t n) { return calloc(1, n ? n : 1); }

typedef struct glue_state {
    HEVCContext s;
    HEVCLocalContext lc;
    HEVCSPS sps;
    HEVCPPS pps;
    AVFrame frames[OHCU_MAX_DPB];
    RefPicList rpl[2];
} glue_state;

static void load_slice(glue_state *st, const ohcu_slice_tab *sl, int idx)
{
    HEVCContext *s = &st->s;
    s->slice_idx = idx;
    s->sh.slice_type = HEVC_SLICE_B;
    s->sh.dependent_slice_segment_flag = 0;
    st->pps.weighted_bipred_flag = sl->weight_flag;
    st->pps.weighted_pred_flag = sl->weight_flag;
    s->sh.luma_log2_weight_denom = sl->luma_log2_weight_denom;
    s->sh.chroma_log2_weight_denom = sl->chroma_log2_weight_denom;
    for (int i = 0; i < OHCU_MAX_REFS; i++) {
        s->sh.luma_weight_l0[i] = sl->luma_weight[0][i];  s->sh.luma_weight_l1[i] = sl->luma_weight[1][i];
        s->sh.luma_offset_l0[i] = sl->luma_offset[0][i];  s->sh.luma_offset_l1[i] = sl->luma_offset[1][i];
        for (int j = 0; j < 2; j++) {
            s->sh.chroma_weight_l0[i][j] = sl->chroma_weight[0][i][j];  s->sh.chroma_weight_l1[i][j] = sl->chroma_weight[1][i][j];
            s->sh.chroma_offset_l0[i][j] = sl->chroma_offset[0][i][j];  s->sh.chroma_offset_l1[i][j] = sl->chroma_offset[1][i][j];
        }
    }
    s->ref->refPicList[idx] = st->rpl;
    for (int l = 0; l < 2; l++) {                         /* RefPicList as ff_hevc_slice_rpl builds it: frames of the DPB */
        st->rpl[l].nb_refs = OHCU_MAX_REFS;
        for (int i = 0; i < OHCU_MAX_REFS; i++)
            st->rpl[l].ref[i] = sl->ref_slot[l][i] < OHCU_MAX_DPB ? &s->DPB[sl->ref_slot[l][i]] : NULL;
    }
    ff_ohcu_slice_start(s);
}

/* one coded block through the patched dispatcher */
static void emit_tu(glue_state *st, const oh_frame_job *job, const ohcu_tu *tu)
{
    HEVCContext *s = &st->s;
    HEVCLocalContext *lc = &st->lc;
    HEVCTransformContext *tr = &lc->transform_ctx;
    const int n = 1 << tu->log2_size, c = tu->c_idx;
    const int16_t *blk = job->coeffs + tu->coeff_off;
    int16_t *dst = lc->tu.coeffs[c ? 1 : 0];
    int lx = 0, ly = 0;
    memcpy(dst, blk, sizeof(int16_t) * n * n);
    for (int y = 0; y < n; y++)
        for (int x = 0; x < n; x++)
            if (blk[y * n + x]) { if (x > lx) lx = x; if (y > ly) ly = y; }
    memset(tr, 0, sizeof(*tr));
    tr->log2_trafo_size = tu->log2_size;
    tr->log2_tr_size_minus2 = tu->log2_size - 2;
    tr->transform_size = n;
    tr->scan_ctx.last_significant_coeff_x = lx;
    tr->scan_ctx.last_significant_coeff_y = ly;
    lc->cu.cu_transquant_bypass_flag = tu->mode == OHCU_TU_BYPASS;
    lc->cu.pred_mode = (tu->mode == OHCU_TU_DST4 || (tu->flags & (OHCU_TUF_ROTATE | OHCU_TUF_INTRA))) ? MODE_INTRA : MODE_INTER;
    st->sps.transform_skip_rotation_enabled_flag = !!(tu->flags & OHCU_TUF_ROTATE);
    tr->transform_skip_flag = tu->mode == OHCU_TU_SKIP;
    tr->explicit_rdpcm_flag = !!(tu->flags & OHCU_TUF_RDPCM);
    tr->explicit_rdpcm_dir_flag = !!(tu->flags & OHCU_TUF_RDPCM_VERT);
    lc->tu.intra_pred_mode = lc->tu.intra_pred_mode_c = 1;
    lc->tu.cross_pf = !!(tu->flags & OHCU_TUF_CROSS);
    lc->tu.res_scale_val = tu->res_scale_val;
    if (c) ff_hevc_hls_transform_c(s, lc, tu->x0 << st->sps.hshift[c], tu->y0 << st->sps.vshift[c], c, 3);
    else ff_hevc_hls_transform(s, lc, tu->x0, tu->y0, 3);
}

/* one intra block through the PATCHED intra_pred (hevcpred_template.c -> ff_ohcu_emit_iu).  The record carries the five
 * final availability flags; lc->na takes them and the z-scan table makes the block the last one in decoding order, so
 * that the flags the reference derives (:103-107) - the ones the patch hands to the glue - are the record's. */
static void emit_iu(glue_state *st, const ohcu_iu *iu, int *zs, int tw, int mask)
{
    HEVCContext *s = &st->s;
    HEVCLocalContext *lc = &st->lc;
    const int c = iu->c_idx, x0 = iu->x0 << st->sps.hshift[c], y0 = iu->y0 << st->sps.vshift[c];
    int *cell = &zs[tw + 1 + ((y0 >> 2) & mask) * tw + ((x0 >> 2) & mask)];
    *cell = 1 << 20;
    lc->cu.pred_mode = MODE_INTRA;
    lc->tu.intra_pred_mode = lc->tu.intra_pred_mode_c = iu->mode;
    lc->na.cand_bottom_left = !!(iu->cand & OHCU_CAND_BOTTOM_LEFT);
    lc->na.cand_left = !!(iu->cand & OHCU_CAND_LEFT);
    lc->na.cand_up_left = !!(iu->cand & OHCU_CAND_UP_LEFT);
    lc->na.cand_up = !!(iu->cand & OHCU_CAND_UP);
    lc->na.cand_up_right = !!(iu->cand & OHCU_CAND_UP_RIGHT);
    s->hpc.intra_pred[iu->log2_size - 2](s, x0, y0, c);
    *cell = -1;
}

/* The coded blocks and intra blocks of the picture in an order a decoder could produce: CTB after CTB; inside a CTB every
 * intra block followed by its own coded block (hls_transform_unit predicts, then adds the residual), then the CTB's other
 * coded blocks. */
static void emit_blocks(glue_state *st, const oh_frame_job *job)
{
    const ohcu_seq_params *sp = &job->sp;
    HEVCSPS *sps = &st->sps;
    HEVCPPS *pps = &st->pps;
    const int l2 = sp->log2_ctb_size, mask = (1 << (l2 - 2)) - 1, tw = mask + 2, nplanes = sp->chroma_format_idc ? 3 : 1;
    const int nctb = sps->ctb_width * sps->ctb_height;
    int *zs = malloc(sizeof(int) * tw * tw), cw[3];
    int32_t *map[3] = { NULL, NULL, NULL };
    int *tu_first = zalloc(sizeof(int) * (nctb + 1)), *iu_first = zalloc(sizeof(int) * (nctb + 1));
    int *tu_order = zalloc(sizeof(int) * (job->n_tu + 1)), *iu_order = zalloc(sizeof(int) * (job->n_iu + 1));
    uint8_t *tu_done = zalloc(job->n_tu + 1);
#define CTB_OF(r) (((((r)->x0 << sps->hshift[(r)->c_idx]) >> l2)) + ((((r)->y0 << sps->vshift[(r)->c_idx]) >> l2)) * sps->ctb_width)
    for (int i = 0; i < tw * tw; i++) zs[i] = -1;
    sps->log2_min_tb_size = 2;
    sps->tb_mask = mask;
    pps->min_tb_addr_zs_tab = zs;
    pps->min_tb_addr_zs = &zs[tw + 1];
    ff_hevc_pred_init(&st->s.hpc, sp->bit_depth);
    for (int c = 0; c < nplanes; c++) {
        cw[c] = ((sp->width >> sps->hshift[c]) + 3) >> 2;
        map[c] = zalloc(sizeof(int32_t) * cw[c] * (((sp->height >> sps->vshift[c]) + 3) >> 2));
    }
    /* counting sort of both lists by CTB (stable: list order inside a CTB) */
    for (int i = 0; i < job->n_tu; i++) tu_first[CTB_OF(&job->tus[i]) + 1]++;
    for (int i = 0; i < job->n_iu; i++) iu_first[CTB_OF(&job->ius[i]) + 1]++;
    for (int k = 0; k < nctb; k++) { tu_first[k + 1] += tu_first[k]; iu_first[k + 1] += iu_first[k]; }
    {
        int *tc = zalloc(sizeof(int) * nctb), *ic = zalloc(sizeof(int) * nctb);
        for (int i = 0; i < job->n_tu; i++) { const int k = CTB_OF(&job->tus[i]); tu_order[tu_first[k] + tc[k]++] = i; }
        for (int i = 0; i < job->n_iu; i++) { const int k = CTB_OF(&job->ius[i]); iu_order[iu_first[k] + ic[k]++] = i; }
        free(tc); free(ic);
    }
    for (int i = 0; i < job->n_tu; i++) {
        const ohcu_tu *tu = &job->tus[i];
        if (tu->flags & OHCU_TUF_INTRA) map[tu->c_idx][(tu->y0 >> 2) * cw[tu->c_idx] + (tu->x0 >> 2)] = i + 1;
    }
    for (int k = 0; k < nctb; k++) {
        for (int q = iu_first[k]; q < iu_first[k + 1]; q++) {
            const ohcu_iu *iu = &job->ius[iu_order[q]];
            emit_iu(st, iu, zs, tw, mask);
            if (iu->cand & OHCU_IU_CBF) {
                const int t = map[iu->c_idx][(iu->y0 >> 2) * cw[iu->c_idx] + (iu->x0 >> 2)];
                if (t) { emit_tu(st, job, &job->tus[t - 1]); tu_done[t - 1] = 1; }
            }
        }
        for (int q = tu_first[k]; q < tu_first[k + 1]; q++)
            if (!tu_done[tu_order[q]]) emit_tu(st, job, &job->tus[tu_order[q]]);
    }
#undef CTB_OF
    for (int c = 0; c < 3; c++) free(map[c]);
    free(zs); free(tu_first); free(iu_first); free(tu_order); free(iu_order); free(tu_done);
    pps->min_tb_addr_zs_tab = NULL; pps->min_tb_addr_zs = NULL;
}

int ohglue_replay_frame(const oh_frame_job *job, int device)
{
    const ohcu_seq_params *sp = &job->sp;
    glue_state *st = zalloc(sizeof(*st));
    HEVCContext *s = &st->s;
    HEVCSPS *sps = &st->sps;
    HEVCPPS *pps = &st->pps;
    const int bd = sp->bit_depth, P = bd > 8 ? 2 : 1, ctb_size = 1 << sp->log2_ctb_size;
    const int out_slot = sp->dpb_slots - 1;
    int nctb, rc = -1, last_slice = -1;
    OHCUGlue *g;

    /* --- what set_sps() / pic_arrays_init() derive (hevcdec.c:114-200, 428-476) --- */
    sps->width = sp->width;  sps->height = sp->height;
    sps->chroma_format_idc = sp->chroma_format_idc;
    sps->bit_depth[0] = sps->bit_depth[1] = bd;
    sps->pixel_shift[0] = sps->pixel_shift[1] = P - 1;
    sps->log2_ctb_size = sp->log2_ctb_size;
    sps->log2_min_cb_size = sp->log2_min_cb_size;
    sps->log2_min_pu_size = sp->log2_min_pu_size;
    sps->ctb_width = (sp->width + ctb_size - 1) >> sp->log2_ctb_size;
    sps->ctb_height = (sp->height + ctb_size - 1) >> sp->log2_ctb_size;
    sps->min_cb_width = sp->width >> sp->log2_min_cb_size;
    sps->min_cb_height = sp->height >> sp->log2_min_cb_size;
    sps->min_pu_width = sp->width >> sp->log2_min_pu_size;
    sps->min_pu_height = sp->height >> sp->log2_min_pu_size;
    sps->hshift[1] = sps->hshift[2] = (sp->chroma_format_idc == 1 || sp->chroma_format_idc == 2);
    sps->vshift[1] = sps->vshift[2] = (sp->chroma_format_idc == 1);
    sps->sao_enabled_flag = sp->sao_enabled && job->sao;
    sps->intra_smoothing_disabled_flag = sp->intra_smoothing_disabled;
    sps->sps_strong_intra_smoothing_enable_flag = sp->strong_intra_smoothing;
    pps->transquant_bypass_enable_flag = !!sp->pcm_or_bypass;
    pps->pps_cb_qp_offset = sp->cb_qp_offset;
    pps->pps_cr_qp_offset = sp->cr_qp_offset;
    pps->tiles_enabled_flag = !!sp->no_tile_filter;
    pps->loop_filter_across_tiles_enabled_flag = 0;
    nctb = sps->ctb_width * sps->ctb_height;
    pps->ctb_addr_rs_to_ts = zalloc(sizeof(int) * nctb);
    pps->tile_id = zalloc(sizeof(int) * nctb);
    for (int i = 0; i < nctb; i++) {
        pps->ctb_addr_rs_to_ts[i] = i;
        pps->tile_id[i] = (job->sao && job->sao->tile_id) ? job->sao->tile_id[i] : 0;
    }
    s->ps.sps = sps;
    s->ps.pps = pps;
    s->HEVClc = &st->lc;
    s->bs_width = (sp->width >> 2) + 1;
    s->bs_height = (sp->height >> 2) + 1;
    for (int i = 0; i < OHCU_MAX_DPB; i++) s->DPB[i].frame = &st->frames[i];
    s->ref = &s->DPB[out_slot];
    s->frame = s->ref->frame;
    s->poc = 0;

    /* --- the device backend follows the SPS (patched set_sps) --- */
    g = s->ohcu = zalloc(sizeof(*g));
    g->device = device;
    if (ff_ohcu_set_sps(s, sps) < 0) goto out;
    /* reference pictures, and the picture the out slot starts from (samples no PU predicts keep it) */
    for (int i = 0; i < sp->dpb_slots && i < OHCU_MAX_DPB; i++) {
        const oh_picture *p = i == out_slot ? &job->out : (job->dpb ? &job->dpb[i] : NULL);
        if (p && p->data[0]) {
            const uint8_t *planes[3] = { p->data[0], p->data[1], p->data[2] };
            const ptrdiff_t strides[3] = { p->stride[0], p->stride[1], p->stride[2] };
            if (ohcu_slot_upload(g->ctx, i, planes, strides)) goto out;
        }
    }

    /* --- one picture: hevc_frame_start, slices, CUs, the last CTB --- */
    s->sh.slice_type = HEVC_SLICE_B;
    if (ff_ohcu_frame_start(s) < 0 || !g->active) goto out;
    for (int i = 0; i < job->n_pu; i++) {
        const ohcu_pu *pu = &job->pus[i];
        MvField mvf;
        if (pu->slice != last_slice) { load_slice(st, &job->slices[pu->slice], pu->slice); last_slice = pu->slice; }
        memset(&mvf, 0, sizeof(mvf));
        for (int l = 0; l < 2; l++) { mvf.mv[l].x = pu->mv[l][0]; mvf.mv[l].y = pu->mv[l][1]; mvf.ref_idx[l] = pu->ref_idx[l]; }
        mvf.pred_flag = pu->pred_flag;
        ff_ohcu_emit_pu(s, pu->x0, pu->y0, pu->w, pu->h, &mvf);
    }
    emit_blocks(st, job);

    /* the arrays the CPU loop filters would read (pic_arrays_init, hevcdec.c:114-200) */
    {
        const size_t bsn = (size_t)s->bs_width * s->bs_height;
        s->vertical_bs = zalloc(bsn);  s->horizontal_bs = zalloc(bsn);
        s->qp_y_tab = zalloc((size_t)sps->min_cb_width * sps->min_cb_height);
        s->deblock = zalloc(sizeof(DBParams) * nctb);
        s->is_pcm = zalloc((size_t)(sps->min_pu_width + 1) * (sps->min_pu_height + 1));
        s->sao = zalloc(sizeof(SAOParams) * nctb);
        s->filter_slice_edges = zalloc(nctb);
        s->tab_slice_address = zalloc(sizeof(int32_t) * nctb);
        if (job->deblock) {
            memcpy(s->vertical_bs, job->deblock->vertical_bs, bsn);
            memcpy(s->horizontal_bs, job->deblock->horizontal_bs, bsn);
            memcpy(s->qp_y_tab, job->deblock->qp_y_tab, (size_t)sps->min_cb_width * sps->min_cb_height);
            for (int i = 0; i < nctb; i++) {
                s->deblock[i].beta_offset = job->deblock->deblock[i].beta_offset;
                s->deblock[i].tc_offset = job->deblock->deblock[i].tc_offset;
            }
            if (job->deblock->is_pcm && sp->pcm_or_bypass)
                memcpy(s->is_pcm, job->deblock->is_pcm, (size_t)sps->min_pu_width * sps->min_pu_height);
        }
        for (int i = 0; i < nctb; i++) {
            s->filter_slice_edges[i] = 1;
            if (job->sao) {
                const ohcu_sao *in = &job->sao->sao[i];
                for (int c = 0; c < 3; c++) {
                    s->sao[i].type_idx[c] = in->type_idx[c];
                    s->sao[i].eo_class[c] = in->eo_class[c];
                    s->sao[i].band_position[c] = in->band_position[c];
                    for (int k = 0; k < 5; k++) s->sao[i].offset_val[c][k] = in->offset_val[c][k];
                }
                if (job->sao->filter_slice_edges) s->filter_slice_edges[i] = job->sao->filter_slice_edges[i];
                if (job->sao->tab_slice_address) s->tab_slice_address[i] = job->sao->tab_slice_address[i];
            }
        }
        /* the patched filter driver is a no-op for a device picture ... */
        for (int rs = 0; rs < nctb; rs++)
            ff_hevc_hls_filter(s, (rs % sps->ctb_width) << sps->log2_ctb_size, (rs / sps->ctb_width) << sps->log2_ctb_size, ctb_size);
        /* ... and the last CTB hands the picture over */
        rc = ff_ohcu_frame_end(s);
        free(s->vertical_bs); free(s->horizontal_bs); free(s->qp_y_tab); free(s->deblock); free(s->is_pcm); free(s->sao);
        free(s->filter_slice_edges); free(s->tab_slice_address);
    }
    if (rc < 0) goto out;
    /* ff_hevc_output_frame */
    for (int c = 0; c < 3; c++) {
        s->ref->frame->data[c] = job->out.data[c];
        s->ref->frame->linesize[c] = (int)job->out.stride[c];
    }
    rc = ff_ohcu_fetch_frame(s, s->ref);
out:
    ff_ohcu_close(s);
    free(pps->ctb_addr_rs_to_ts);
    free(pps->tile_id);
    free(st);
    return rc;
}
