/*
 * ref_shim.c — thin OUR-side wrapper that exposes the reference's own compiled
 * code to the tests.  TEST INFRASTRUCTURE ONLY; built by oracle/Makefile into
 * oracle/_ref/liboh_ref.so and only when /root/reference is present.
 *
 * Nothing of the reference is copied: this translation unit #includes the
 * reference's hevcdec.c where it lies (so that its `static` MC drivers
 * luma_mc_uni / luma_mc_bi / chroma_mc_uni / chroma_mc_bi, hevcdec.c:1744-2049,
 * become callable) and links the reference's hevcdsp.c, hevc_filter.c and
 * videodsp.c objects compiled from /root/reference.
 *
 *   ohref_<leaf>()        one HEVCDSPContext entry (ff_hevc_dsp_init table, C version)
 *   ohref_replay_frame()  a frame job through the reference's drivers
 */
#include "libavcodec/hevcdec.c"

#include "replay.h"

#include <pthread.h>

/* ------------------------------------------------------------------ tables */
static HEVCDSPContext g_tab[15];
static VideoDSPContext g_vdsp[15];
static pthread_once_t g_once = PTHREAD_ONCE_INIT;

static void init_tables(void)
{
    static const int depths[] = { 8, 9, 10, 12, 14 };
    for (unsigned i = 0; i < sizeof(depths) / sizeof(depths[0]); i++) {
        ff_hevc_dsp_init(&g_tab[depths[i]], depths[i]);      /* hevcdsp.c:311 */
        ff_videodsp_init(&g_vdsp[depths[i]], depths[i]);     /* videodsp.c    */
    }
}
static HEVCDSPContext *tab(int bd)
{
    pthread_once(&g_once, init_tables);
    return &g_tab[bd];
}
static int lg2(int n) { int l = 0; while ((1 << l) < n) l++; return l; }

int ohref_abi_version(void) { return OHCU_ABI_VERSION; }
int ohref_sizeof_dsp_context(void) { return (int)sizeof(HEVCDSPContext); }

/* ------------------------------------------------------------------ residual */
void ohref_transform_add(int bd, uint8_t *dst, int16_t *coeffs, ptrdiff_t stride, int size)
{
    tab(bd)->transform_add[lg2(size) - 2](dst, coeffs, stride);
}
void ohref_transform_rdpcm(int bd, int16_t *coeffs, int log2_size, int mode)
{
    tab(bd)->transform_rdpcm(coeffs, log2_size, mode);
}
void ohref_transform_skip(int bd, int16_t *coeffs, int log2_size)
{
    tab(bd)->transform_skip(coeffs, log2_size);
}
void ohref_idct_4x4_luma(int bd, int16_t *coeffs) { tab(bd)->idct_4x4_luma(coeffs); }
void ohref_idct(int bd, int16_t *coeffs, int log2_size, int col_limit)
{
    tab(bd)->idct[log2_size - 2](coeffs, col_limit);
}
void ohref_idct_dc(int bd, int16_t *coeffs, int log2_size)
{
    tab(bd)->idct_dc[log2_size - 2](coeffs);
}

/* ------------------------------------------------------------------ SAO */
static void fill_sao(SAOParams *p, const int16_t offset_val[5], int band_position, int eo_class)
{
    memset(p, 0, sizeof(*p));
    for (int i = 0; i < 5; i++)
        p->offset_val[0][i] = offset_val[i];
    p->band_position[0] = band_position;
    p->eo_class[0] = eo_class;
}
void ohref_sao_band_filter(int bd, uint8_t *dst, uint8_t *src, ptrdiff_t stride_dst,
                           ptrdiff_t stride_src, const int16_t offset_val[5], int band_position,
                           int width, int height)
{
    SAOParams p;
    int borders[4] = { 0 };
    fill_sao(&p, offset_val, band_position, 0);
    tab(bd)->sao_band_filter(dst, src, stride_dst, stride_src, &p, borders, width, height, 0);
}
void ohref_sao_edge_filter(int bd, uint8_t *dst, uint8_t *src, ptrdiff_t stride_dst,
                           ptrdiff_t stride_src, const int16_t offset_val[5], int eo_class,
                           int width, int height)
{
    SAOParams p;
    fill_sao(&p, offset_val, 0, eo_class);
    tab(bd)->sao_edge_filter(dst, src, stride_dst, stride_src, &p, width, height, 0);
}
void ohref_sao_edge_restore(int bd, int restore_idx, uint8_t *dst, uint8_t *src,
                            ptrdiff_t stride_dst, ptrdiff_t stride_src,
                            const int16_t offset_val[5], int eo_class, const int borders_in[4],
                            int width, int height, const uint8_t vert_edge[2],
                            const uint8_t horiz_edge[2], const uint8_t diag_edge[4])
{
    SAOParams p;
    int borders[4];
    uint8_t ve[2], he[2], de[4];
    fill_sao(&p, offset_val, 0, eo_class);
    memcpy(borders, borders_in, sizeof(borders));
    memcpy(ve, vert_edge, 2); memcpy(he, horiz_edge, 2); memcpy(de, diag_edge, 4);
    tab(bd)->sao_edge_restore[restore_idx](dst, src, stride_dst, stride_src, &p, borders,
                                           width, height, 0, ve, he, de);
}

/* ------------------------------------------------------------------ MC */
#define PELIDX(w) ff_hevc_pel_weight[w]
void ohref_mc_put(int bd, int luma, int16_t *dst, uint8_t *src, ptrdiff_t srcstride,
                  int height, int mx, int my, int width)
{
    HEVCDSPContext *t = tab(bd);
    (luma ? t->put_hevc_qpel : t->put_hevc_epel)[PELIDX(width)][!!my][!!mx](
        dst, src, srcstride, height, mx, my, width);
}
void ohref_mc_uni(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, uint8_t *src,
                  ptrdiff_t srcstride, int height, int mx, int my, int width)
{
    HEVCDSPContext *t = tab(bd);
    (luma ? t->put_hevc_qpel_uni : t->put_hevc_epel_uni)[PELIDX(width)][!!my][!!mx](
        dst, dststride, src, srcstride, height, mx, my, width);
}
void ohref_mc_uni_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, uint8_t *src,
                    ptrdiff_t srcstride, int height, int denom, int wx, int ox,
                    int mx, int my, int width)
{
    HEVCDSPContext *t = tab(bd);
    (luma ? t->put_hevc_qpel_uni_w : t->put_hevc_epel_uni_w)[PELIDX(width)][!!my][!!mx](
        dst, dststride, src, srcstride, height, denom, wx, ox, mx, my, width);
}
void ohref_mc_bi(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, uint8_t *src,
                 ptrdiff_t srcstride, int16_t *src2, int height, int mx, int my, int width)
{
    HEVCDSPContext *t = tab(bd);
    (luma ? t->put_hevc_qpel_bi : t->put_hevc_epel_bi)[PELIDX(width)][!!my][!!mx](
        dst, dststride, src, srcstride, src2, height, mx, my, width);
}
void ohref_mc_bi_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, uint8_t *src,
                   ptrdiff_t srcstride, int16_t *src2, int height, int denom,
                   int wx0, int wx1, int ox0, int ox1, int mx, int my, int width)
{
    HEVCDSPContext *t = tab(bd);
    (luma ? t->put_hevc_qpel_bi_w : t->put_hevc_epel_bi_w)[PELIDX(width)][!!my][!!mx](
        dst, dststride, src, srcstride, src2, height, denom, wx0, wx1, ox0, ox1, mx, my, width);
}

/* ------------------------------------------------------------------ deblock */
void ohref_loop_filter_luma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride, int beta,
                            const int tc_in[2], const uint8_t no_p_in[2], const uint8_t no_q_in[2])
{
    HEVCDSPContext *t = tab(bd);
    int tc[2] = { tc_in[0], tc_in[1] };
    uint8_t no_p[2] = { no_p_in[0], no_p_in[1] }, no_q[2] = { no_q_in[0], no_q_in[1] };
    /* the _c entries are the same C functions (hevcdsp.c:500-507) */
    (vertical ? t->hevc_v_loop_filter_luma_c : t->hevc_h_loop_filter_luma_c)(pix, stride, beta, tc, no_p, no_q);
}
void ohref_loop_filter_chroma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride,
                              const int tc_in[2], const uint8_t no_p_in[2], const uint8_t no_q_in[2])
{
    HEVCDSPContext *t = tab(bd);
    int tc[2] = { tc_in[0], tc_in[1] };
    uint8_t no_p[2] = { no_p_in[0], no_p_in[1] }, no_q[2] = { no_q_in[0], no_q_in[1] };
    (vertical ? t->hevc_v_loop_filter_chroma_c : t->hevc_h_loop_filter_chroma_c)(pix, stride, tc, no_p, no_q);
}

/* ================================================================== frame replay */
/* apply one TU's inverse transform chain with the reference table, mirroring the
 * dispatch of ff_hevc_hls_transform (hevc_cabac.c:1727-1806): out = residual */
static void ref_tu_residual(HEVCDSPContext *t, const ohcu_tu *tu, const int16_t *arena, int16_t *out)
{
    int n2 = 1 << (2 * tu->log2_size);
    memcpy(out, arena + tu->coeff_off, sizeof(int16_t) * n2);
    switch (tu->mode) {
    case OHCU_TU_BYPASS:
        if (tu->flags & OHCU_TUF_RDPCM)
            t->transform_rdpcm(out, tu->log2_size, !!(tu->flags & OHCU_TUF_RDPCM_VERT));
        break;
    case OHCU_TU_SKIP:
        if (tu->flags & OHCU_TUF_ROTATE)
            for (int i = 0; i < 8; i++)
                FFSWAP(int16_t, out[i], out[16 - i - 1]);
        t->transform_skip(out, tu->log2_size);
        if (tu->flags & OHCU_TUF_RDPCM)
            t->transform_rdpcm(out, tu->log2_size, !!(tu->flags & OHCU_TUF_RDPCM_VERT));
        break;
    case OHCU_TU_DST4: t->idct_4x4_luma(out); break;
    case OHCU_TU_DC:   t->idct_dc[tu->log2_size - 2](out); break;
    default:           t->idct[tu->log2_size - 2](out, tu->col_limit); break;
    }
}

typedef struct ref_state {
    HEVCContext s;
    HEVCLocalContext lc;
    HEVCSPS sps;
    HEVCPPS pps;
    AVFrame cur;
    AVFrame ref[OHCU_MAX_DPB];
} ref_state;

static void ref_load_slice(ref_state *st, const ohcu_slice_tab *sl)
{
    HEVCContext *s = &st->s;
    s->sh.slice_type = HEVC_SLICE_B;
    st->pps.weighted_bipred_flag = sl->weight_flag;
    st->pps.weighted_pred_flag = sl->weight_flag;
    s->sh.luma_log2_weight_denom = sl->luma_log2_weight_denom;
    s->sh.chroma_log2_weight_denom = sl->chroma_log2_weight_denom;
    for (int i = 0; i < OHCU_MAX_REFS; i++) {
        s->sh.luma_weight_l0[i] = sl->luma_weight[0][i];
        s->sh.luma_weight_l1[i] = sl->luma_weight[1][i];
        s->sh.luma_offset_l0[i] = sl->luma_offset[0][i];
        s->sh.luma_offset_l1[i] = sl->luma_offset[1][i];
        for (int j = 0; j < 2; j++) {
            s->sh.chroma_weight_l0[i][j] = sl->chroma_weight[0][i][j];
            s->sh.chroma_weight_l1[i][j] = sl->chroma_weight[1][i][j];
            s->sh.chroma_offset_l0[i][j] = sl->chroma_offset[0][i][j];
            s->sh.chroma_offset_l1[i][j] = sl->chroma_offset[1][i][j];
        }
    }
}

/* the dispatch of hls_prediction_unit, hevcdec.c:2250-2306, calling the
 * reference's own luma_mc_* / chroma_mc_* drivers */
static void ref_predict_pu(ref_state *st, const oh_frame_job *job, const ohcu_pu *pu)
{
    HEVCContext *s = &st->s;
    const HEVCSPS *sps = &st->sps;
    const ohcu_slice_tab *sl = &job->slices[pu->slice];
    int x0 = pu->x0, y0 = pu->y0, nPbW = pu->w, nPbH = pu->h;
    MvField mvf;
    AVFrame *r0 = NULL, *r1 = NULL;
    uint8_t *dst[3];

    ref_load_slice(st, sl);
    memset(&mvf, 0, sizeof(mvf));
    for (int l = 0; l < 2; l++) {
        mvf.mv[l].x = pu->mv[l][0];
        mvf.mv[l].y = pu->mv[l][1];
        mvf.ref_idx[l] = pu->ref_idx[l];
    }
    mvf.pred_flag = pu->pred_flag;
    if (pu->pred_flag & PF_L0) r0 = &st->ref[sl->ref_slot[0][pu->ref_idx[0]]];
    if (pu->pred_flag & PF_L1) r1 = &st->ref[sl->ref_slot[1][pu->ref_idx[1]]];
    for (int c = 0; c < 3; c++)
        dst[c] = sps->chroma_format_idc || !c
                     ? &s->frame->data[c][(y0 >> sps->vshift[c]) * s->frame->linesize[c] +
                                          ((x0 >> sps->hshift[c]) << sps->pixel_shift[c ? 1 : 0])]
                     : NULL;
    {
        int x0_c = x0 >> sps->hshift[1], y0_c = y0 >> sps->vshift[1];
        int nPbW_c = nPbW >> sps->hshift[1], nPbH_c = nPbH >> sps->vshift[1];
        if (pu->pred_flag == PF_L0 || pu->pred_flag == PF_L1) {
            int l = pu->pred_flag == PF_L1;
            AVFrame *r = l ? r1 : r0;
            int ri = pu->ref_idx[l];
            luma_mc_uni(s, dst[0], s->frame->linesize[0], r, &mvf.mv[l], x0, y0, nPbW, nPbH,
                        sl->luma_weight[l][ri], sl->luma_offset[l][ri]);
            if (sps->chroma_format_idc)
                for (int c = 0; c < 2; c++)
                    chroma_mc_uni(s, dst[c + 1], s->frame->linesize[c + 1], r->data[c + 1],
                                  r->linesize[c + 1], l, x0_c, y0_c, nPbW_c, nPbH_c, &mvf,
                                  sl->chroma_weight[l][ri][c], sl->chroma_offset[l][ri][c]);
        } else if (pu->pred_flag == PF_BI) {
            luma_mc_bi(s, dst[0], s->frame->linesize[0], r0, &mvf.mv[0], x0, y0, nPbW, nPbH,
                       r1, &mvf.mv[1], &mvf);
            if (sps->chroma_format_idc)
                for (int c = 0; c < 2; c++)
                    chroma_mc_bi(s, dst[c + 1], s->frame->linesize[c + 1], r0, r1, x0_c, y0_c,
                                 nPbW_c, nPbH_c, &mvf, c);
        }
    }
}

static void *zalloc(size_t n) { void *p = calloc(1, n ? n : 1); return p; }

/* one coded TB: residual through the reference's table, then transform_add */
static void ref_residual_tu(ref_state *st, HEVCDSPContext *t, const oh_frame_job *job, const ohcu_tu *tu, int P)
{
    DECLARE_ALIGNED(32, int16_t, res)[32 * 32];
    DECLARE_ALIGNED(32, int16_t, res_y)[32 * 32];
    int c = tu->c_idx, n2 = 1 << (2 * tu->log2_size);
    ref_tu_residual(t, tu, job->coeffs, res);
    if (tu->flags & OHCU_TUF_CROSS) {
        ref_tu_residual(t, &job->tus[tu->cross_tu], job->coeffs, res_y);
        for (int k = 0; k < n2; k++)
            res[k] = res[k] + ((tu->res_scale_val * res_y[k]) >> 3);
    }
    t->transform_add[tu->log2_size - 2](
        st->cur.data[c] + tu->y0 * st->cur.linesize[c] + tu->x0 * P, res, st->cur.linesize[c]);
}

/* The intra blocks through the reference's own HEVCPredContext (ff_hevc_pred_init, hevcpred.c; intra_pred
 * hevcpred_template.c:30-345), in hls_transform_unit order with each block's transform_add behind it
 * (hevcdec.c:1335-1526).  Availability: the job carries the five final flags; the z-scan table handed to the
 * reference makes the current block the last one in order, so its own test (:103-107) leaves them as they are. */
static int ref_intra_blocks(ref_state *st, HEVCDSPContext *t, const oh_frame_job *job, int P)
{
    const ohcu_seq_params *sp = &job->sp;
    HEVCContext *s = &st->s;
    HEVCSPS *sps = &st->sps;
    HEVCPPS *pps = &st->pps;
    HEVCLocalContext *lc = &st->lc;
    const int log2_min_tb = 2, mask = (1 << (sp->log2_ctb_size - log2_min_tb)) - 1, tw = mask + 2;
    int nplanes = sp->chroma_format_idc ? 3 : 1, rc = 0, cw[3];
    int *zs = malloc(sizeof(int) * tw * tw);
    int32_t *map[3] = { NULL, NULL, NULL };

    sps->log2_min_tb_size = log2_min_tb;
    sps->tb_mask = mask;
    sps->intra_smoothing_disabled_flag = sp->intra_smoothing_disabled;
    sps->sps_strong_intra_smoothing_enable_flag = sp->strong_intra_smoothing;
    pps->constrained_intra_pred_flag = 0;
    for (int i = 0; i < tw * tw; i++) zs[i] = -1;
    pps->min_tb_addr_zs_tab = zs;
    pps->min_tb_addr_zs = &zs[tw + 1];
    ff_hevc_pred_init(&s->hpc, sp->bit_depth);
    for (int c = 0; c < nplanes; c++) {
        cw[c] = ((sp->width >> sps->hshift[c]) + 3) >> 2;
        map[c] = zalloc(sizeof(int32_t) * cw[c] * (((sp->height >> sps->vshift[c]) + 3) >> 2));
    }
    for (int i = 0; i < job->n_tu; i++) {
        const ohcu_tu *tu = &job->tus[i];
        if ((tu->flags & OHCU_TUF_INTRA) && tu->c_idx < nplanes)
            map[tu->c_idx][(tu->y0 >> 2) * cw[tu->c_idx] + (tu->x0 >> 2)] = i + 1;
    }
    for (int i = 0; i < job->n_iu; i++) {
        const ohcu_iu *iu = &job->ius[i];
        int c = iu->c_idx, x0, y0, cell;
        if (c >= nplanes || iu->log2_size < 2 || iu->log2_size > 5 || iu->mode > 34) { rc = -1; break; }
        x0 = iu->x0 << sps->hshift[c]; y0 = iu->y0 << sps->vshift[c];      /* the reference takes luma coordinates */
        cell = ((y0 >> log2_min_tb) & mask) * tw + ((x0 >> log2_min_tb) & mask);
        pps->min_tb_addr_zs[cell] = 1 << 20;
        lc->tu.intra_pred_mode = lc->tu.intra_pred_mode_c = iu->mode;
        lc->na.cand_bottom_left = !!(iu->cand & OHCU_CAND_BOTTOM_LEFT);
        lc->na.cand_left = !!(iu->cand & OHCU_CAND_LEFT);
        lc->na.cand_up_left = !!(iu->cand & OHCU_CAND_UP_LEFT);
        lc->na.cand_up = !!(iu->cand & OHCU_CAND_UP);
        lc->na.cand_up_right = !!(iu->cand & OHCU_CAND_UP_RIGHT);
        s->hpc.intra_pred[iu->log2_size - 2](s, x0, y0, c);
        pps->min_tb_addr_zs[cell] = -1;
        if (iu->cand & OHCU_IU_CBF) {
            int k = map[c][(iu->y0 >> 2) * cw[c] + (iu->x0 >> 2)];
            if (!k || job->tus[k - 1].log2_size != iu->log2_size) { rc = -1; break; }
            ref_residual_tu(st, t, job, &job->tus[k - 1], P);
        }
    }
    for (int c = 0; c < 3; c++) free(map[c]);
    free(zs);
    pps->min_tb_addr_zs_tab = NULL; pps->min_tb_addr_zs = NULL;
    return rc;
}

int ohref_replay_frame(const oh_frame_job *job)
{
    const ohcu_seq_params *sp = &job->sp;
    int bd = sp->bit_depth, P = bd > 8 ? 2 : 1;
    int stages = job->stages ? job->stages : 15;
    int nplanes = sp->chroma_format_idc ? 3 : 1;
    ref_state *st = zalloc(sizeof(*st));
    HEVCContext *s = &st->s;
    HEVCSPS *sps = &st->sps;
    HEVCPPS *pps = &st->pps;
    HEVCDSPContext *t = tab(bd);
    uint8_t *curbuf[3] = { NULL, NULL, NULL };
    int ctb_size = 1 << sp->log2_ctb_size;
    int nctb;

    /* --- what set_sps()/pic_arrays_init() derive (hevcdec.c:114-200, 428-476) --- */
    sps->width = sp->width;   sps->height = sp->height;
    sps->chroma_format_idc = sp->chroma_format_idc;
    sps->bit_depth[0] = sps->bit_depth[1] = bd;
    sps->pixel_shift[0] = sps->pixel_shift[1] = P - 1;
    sps->log2_ctb_size = sp->log2_ctb_size;
    sps->log2_min_cb_size = sp->log2_min_cb_size;
    sps->log2_min_pu_size = sp->log2_min_pu_size;
    sps->ctb_width  = (sp->width  + ctb_size - 1) >> sp->log2_ctb_size;
    sps->ctb_height = (sp->height + ctb_size - 1) >> sp->log2_ctb_size;
    sps->ctb_size = sps->ctb_width * sps->ctb_height;
    sps->min_cb_width  = sp->width  >> sp->log2_min_cb_size;
    sps->min_cb_height = sp->height >> sp->log2_min_cb_size;
    sps->min_pu_width  = sp->width  >> sp->log2_min_pu_size;
    sps->min_pu_height = sp->height >> sp->log2_min_pu_size;
    sps->hshift[0] = sps->vshift[0] = 0;
    sps->hshift[1] = sps->hshift[2] = (sp->chroma_format_idc == 1 || sp->chroma_format_idc == 2);
    sps->vshift[1] = sps->vshift[2] = (sp->chroma_format_idc == 1);
    sps->sao_enabled_flag = sp->sao_enabled && job->sao && (stages & OH_STAGE_SAO);
    pps->transquant_bypass_enable_flag = !!sp->pcm_or_bypass;
    pps->pps_cb_qp_offset = sp->cb_qp_offset;
    pps->pps_cr_qp_offset = sp->cr_qp_offset;
    pps->tiles_enabled_flag = !!sp->no_tile_filter;
    pps->loop_filter_across_tiles_enabled_flag = 0;
    nctb = sps->ctb_size;
    pps->ctb_addr_rs_to_ts = zalloc(sizeof(int) * nctb);
    pps->tile_id = zalloc(sizeof(int) * nctb);
    for (int i = 0; i < nctb; i++) {
        pps->ctb_addr_rs_to_ts[i] = i;
        pps->tile_id[i] = (job->sao && job->sao->tile_id) ? job->sao->tile_id[i] : 0;
    }
    s->ps.sps = sps;
    s->ps.pps = pps;
    s->HEVClc = &st->lc;
    s->threads_type = 0;
    s->hevcdsp = *t;
    s->vdsp = g_vdsp[bd];
    s->bs_width  = (sp->width  >> 2) + 1;
    s->bs_height = (sp->height >> 2) + 1;

    /* --- pictures: the current frame gets a padded private copy because the
     * reference's copy_CTB moves 8/16-byte groups past the picture width
     * (hevc_filter.c:157-177) like its own over-allocated AVFrames allow --- */
    for (int c = 0; c < nplanes; c++) {
        int w = sp->width >> sps->hshift[c], h = sp->height >> sps->vshift[c];
        ptrdiff_t ls = ((ptrdiff_t)w * P + 127) & ~(ptrdiff_t)63;
        curbuf[c] = zalloc((size_t)ls * (h + 4) + 128);
        st->cur.data[c] = curbuf[c] + 2 * ls + 64;
        st->cur.linesize[c] = (int)ls;
        for (int y = 0; y < h; y++)
            memcpy(st->cur.data[c] + y * ls, job->out.data[c] + y * job->out.stride[c], (size_t)w * P);
        for (int i = 0; i < sp->dpb_slots && i < OHCU_MAX_DPB; i++) {
            st->ref[i].data[c] = job->dpb ? job->dpb[i].data[c] : NULL;
            st->ref[i].linesize[c] = job->dpb ? (int)job->dpb[i].stride[c] : 0;
        }
    }
    s->frame = &st->cur;

    /* --- MC --- */
    if (stages & OH_STAGE_MC)
        for (int i = 0; i < job->n_pu; i++)
            ref_predict_pu(st, job, &job->pus[i]);

    /* --- residual: dispatcher tail of ff_hevc_hls_transform{,_c} (hevc_cabac.c:1807-1818) --- */
    if (stages & OH_STAGE_RESIDUAL) {
        for (int i = 0; i < job->n_tu; i++)
            if (!(job->tus[i].flags & OHCU_TUF_INTRA))
                ref_residual_tu(st, t, job, &job->tus[i], P);
        if (job->n_iu && ref_intra_blocks(st, t, job, P)) {
            for (int c = 0; c < nplanes; c++) free(curbuf[c]);
            free(pps->ctb_addr_rs_to_ts); free(pps->tile_id); free(st);
            return OHCU_EINVAL;
        }
    }

    /* --- in-loop filters through the reference's own per-CTB driver --- */
    if (stages & (OH_STAGE_DEBLOCK | OH_STAGE_SAO)) {
        size_t bsn = (size_t)s->bs_width * s->bs_height;
        int do_db = job->deblock && (stages & OH_STAGE_DEBLOCK);
        s->vertical_bs = zalloc(bsn);
        s->horizontal_bs = zalloc(bsn);
        s->qp_y_tab = zalloc((size_t)sps->min_cb_width * sps->min_cb_height);
        s->deblock = zalloc(sizeof(DBParams) * nctb);
        s->is_pcm = zalloc((size_t)sps->min_pu_width * sps->min_pu_height);
        s->sao = zalloc(sizeof(SAOParams) * nctb);
        s->filter_slice_edges = zalloc(nctb);
        s->tab_slice_address = zalloc(sizeof(int32_t) * nctb);
        if (do_db) {
            memcpy(s->vertical_bs, job->deblock->vertical_bs, bsn);
            memcpy(s->horizontal_bs, job->deblock->horizontal_bs, bsn);
            memcpy(s->qp_y_tab, job->deblock->qp_y_tab, (size_t)sps->min_cb_width * sps->min_cb_height);
            for (int i = 0; i < nctb; i++) {
                s->deblock[i].beta_offset = job->deblock->deblock[i].beta_offset;
                s->deblock[i].tc_offset = job->deblock->deblock[i].tc_offset;
            }
        }
        if (job->deblock && job->deblock->is_pcm && sp->pcm_or_bypass)
            memcpy(s->is_pcm, job->deblock->is_pcm, (size_t)sps->min_pu_width * sps->min_pu_height);
        if (sps->sao_enabled_flag) {
            for (int i = 0; i < nctb; i++) {
                const ohcu_sao *in = &job->sao->sao[i];
                for (int c = 0; c < 3; c++) {
                    s->sao[i].type_idx[c] = in->type_idx[c];
                    s->sao[i].eo_class[c] = in->eo_class[c];
                    s->sao[i].band_position[c] = in->band_position[c];
                    for (int k = 0; k < 5; k++)
                        s->sao[i].offset_val[c][k] = in->offset_val[c][k];
                }
                s->filter_slice_edges[i] = job->sao->filter_slice_edges ? job->sao->filter_slice_edges[i] : 1;
                s->tab_slice_address[i] = job->sao->tab_slice_address ? job->sao->tab_slice_address[i] : 0;
            }
            /* buffers of set_sps(), hevcdec.c:450-476 */
            st->lc.sao_pixel_buffer = zalloc((size_t)((FFMAX(ctb_size, 16) + 32) * (ctb_size + 2)) << (P - 1));
            for (int c = 0; c < nplanes; c++) {
                int w = sp->width >> sps->hshift[c], h = sp->height >> sps->vshift[c];
                s->sao_pixel_buffer_h[c] = zalloc((size_t)(w * 2 * sps->ctb_height) << (P - 1));
                s->sao_pixel_buffer_v[c] = zalloc((size_t)(h * 2 * sps->ctb_width) << (P - 1));
            }
        }
        /* CTB loop of hls_decode_entry, hevcdec.c:2868-2903 */
        {
            int x_ctb = 0, y_ctb = 0;
            for (int rs = 0; rs < nctb; rs++) {
                x_ctb = (rs % sps->ctb_width) << sps->log2_ctb_size;
                y_ctb = (rs / sps->ctb_width) << sps->log2_ctb_size;
                ff_hevc_hls_filters(s, x_ctb, y_ctb, ctb_size);
            }
            if (x_ctb + ctb_size >= sps->width && y_ctb + ctb_size >= sps->height)
                ff_hevc_hls_filter(s, x_ctb, y_ctb, ctb_size);
        }
        free(s->vertical_bs); free(s->horizontal_bs); free(s->qp_y_tab); free(s->deblock);
        free(s->is_pcm); free(s->sao); free(s->filter_slice_edges); free(s->tab_slice_address);
        free(st->lc.sao_pixel_buffer);
        for (int c = 0; c < 3; c++) { free(s->sao_pixel_buffer_h[c]); free(s->sao_pixel_buffer_v[c]); }
    }

    for (int c = 0; c < nplanes; c++) {
        int w = sp->width >> sps->hshift[c], h = sp->height >> sps->vshift[c];
        for (int y = 0; y < h; y++)
            memcpy(job->out.data[c] + y * job->out.stride[c], st->cur.data[c] + y * st->cur.linesize[c], (size_t)w * P);
        free(curbuf[c]);
    }
    free(pps->ctb_addr_rs_to_ts);
    free(pps->tile_id);
    free(st);
    return 0;
}

/* ------------------------------------------------------------------ boundary strengths
 * The reference's own ff_hevc_deblocking_boundary_strengths (hevc_filter.c:1006-1125), called
 * once per transform-tree leaf as hls_transform_tree / hls_coding_unit do (hevcdec.c), with
 * lc->boundary_flags from its own hls_decode_neighbour (hevcdec.c:2796-2843).  tab_mvf is
 * filled from the PU list the way hls_prediction_unit leaves it (MvField, hevcdec.h:302-311;
 * poc[] = the DPB slot of the referenced picture: one picture per slot). */
int ohref_boundary_strengths(const oh_bs_job *job)
{
    const ohcu_seq_params *sp = &job->sp;
    ref_state *st = zalloc(sizeof(*st));
    HEVCContext *s = &st->s;
    HEVCSPS *sps = &st->sps;
    HEVCPPS *pps = &st->pps;
    HEVCFrame *fr = zalloc(sizeof(*fr));
    int ctb_size = 1 << sp->log2_ctb_size, nctb, last_ctb = -1;
    size_t bsn;

    sps->width = sp->width;   sps->height = sp->height;
    sps->log2_ctb_size = sp->log2_ctb_size;
    sps->log2_min_cb_size = sp->log2_min_cb_size;
    sps->log2_min_pu_size = sp->log2_min_pu_size;
    sps->log2_min_tb_size = 2;
    sps->ctb_width  = (sp->width  + ctb_size - 1) >> sp->log2_ctb_size;
    sps->ctb_height = (sp->height + ctb_size - 1) >> sp->log2_ctb_size;
    sps->ctb_size = sps->ctb_width * sps->ctb_height;
    sps->min_pu_width  = sp->width  >> sp->log2_min_pu_size;
    sps->min_pu_height = sp->height >> sp->log2_min_pu_size;
    sps->min_tb_width  = sp->width  >> 2;
    sps->min_tb_height = sp->height >> 2;
    nctb = sps->ctb_size;
    pps->tiles_enabled_flag = !!sp->no_tile_filter;
    pps->loop_filter_across_tiles_enabled_flag = 0;
    pps->ctb_addr_rs_to_ts = zalloc(sizeof(int) * nctb);
    pps->ctb_addr_ts_to_rs = zalloc(sizeof(int) * nctb);
    pps->tile_id = zalloc(sizeof(int) * nctb);
    pps->col_idxX = zalloc(sizeof(int) * (sps->ctb_width + 1));
    pps->column_width = zalloc(sizeof(int) * 2);
    pps->column_width[0] = sps->ctb_width;
    for (int i = 0; i < nctb; i++) {
        pps->ctb_addr_rs_to_ts[i] = pps->ctb_addr_ts_to_rs[i] = i;
        pps->tile_id[i] = job->tile_id ? job->tile_id[i] : 0;
    }
    s->ps.sps = sps;
    s->ps.pps = pps;
    s->HEVClc = &st->lc;
    s->bs_width  = (sp->width  >> 2) + 1;
    s->bs_height = (sp->height >> 2) + 1;
    bsn = (size_t)s->bs_width * s->bs_height;
    s->vertical_bs = zalloc(bsn);
    s->horizontal_bs = zalloc(bsn);
    s->cbf_luma = zalloc((size_t)sps->min_tb_width * sps->min_tb_height);
    s->tab_slice_address = zalloc(sizeof(int32_t) * nctb);
    fr->tab_mvf = zalloc(sizeof(MvField) * (size_t)sps->min_pu_width * sps->min_pu_height);   /* pred_flag 0 = PF_INTRA */
    s->ref = fr;

    for (int i = 0; i < job->n_pu; i++) {
        const ohcu_pu *pu = &job->pus[i];
        const ohcu_slice_tab *sl = &job->slices[pu->slice];
        MvField mvf;
        memset(&mvf, 0, sizeof(mvf));
        for (int l = 0; l < 2; l++) {
            mvf.mv[l].x = pu->mv[l][0]; mvf.mv[l].y = pu->mv[l][1];
            mvf.ref_idx[l] = pu->ref_idx[l];
            mvf.poc[l] = sl->ref_slot[l][pu->ref_idx[l] & (OHCU_MAX_REFS - 1)];
        }
        mvf.pred_flag = pu->pred_flag;
        for (int y = pu->y0 >> sp->log2_min_pu_size; y < (pu->y0 + pu->h) >> sp->log2_min_pu_size; y++)
            for (int x = pu->x0 >> sp->log2_min_pu_size; x < (pu->x0 + pu->w) >> sp->log2_min_pu_size; x++)
                fr->tab_mvf[y * sps->min_pu_width + x] = mvf;
    }
    for (int i = 0; i < job->n_leaf; i++) {
        const oh_leaf *lf = &job->leaves[i];
        if (lf->cbf_luma)
            for (int y = lf->y0 >> 2; y < (lf->y0 >> 2) + (1 << (lf->log2_size - 2)); y++)
                memset(s->cbf_luma + y * sps->min_tb_width + (lf->x0 >> 2), 1, 1 << (lf->log2_size - 2));
    }
    for (int i = 0; i < job->n_leaf; i++) {
        const oh_leaf *lf = &job->leaves[i];
        int rs = (lf->y0 >> sp->log2_ctb_size) * sps->ctb_width + (lf->x0 >> sp->log2_ctb_size);
        if (rs != last_ctb) {                              /* CTB loop of hls_decode_entry, hevcdec.c:2868-2890 */
            s->sh.slice_addr = job->tab_slice_address ? job->tab_slice_address[rs] : 0;
            s->sh.slice_loop_filter_across_slices_enabled_flag = job->filter_slice_edges ? job->filter_slice_edges[rs] : 1;
            s->sh.disable_deblocking_filter_flag = job->deblock_disabled ? job->deblock_disabled[rs] : 0;
            hls_decode_neighbour(s, (rs % sps->ctb_width) << sp->log2_ctb_size, (rs / sps->ctb_width) << sp->log2_ctb_size, rs);
            last_ctb = rs;
        }
        if (!s->sh.disable_deblocking_filter_flag)         /* hevcdec.c: the call sites' condition */
            ff_hevc_deblocking_boundary_strengths(s, lf->x0, lf->y0, lf->log2_size);
    }
    memcpy(job->vbs, s->vertical_bs, bsn);
    memcpy(job->hbs, s->horizontal_bs, bsn);
    free(s->vertical_bs); free(s->horizontal_bs); free(s->cbf_luma); free(s->tab_slice_address);
    free(fr->tab_mvf); free(fr);
    free(pps->ctb_addr_rs_to_ts); free(pps->ctb_addr_ts_to_rs); free(pps->tile_id); free(pps->col_idxX); free(pps->column_width);
    free(st);
    return 0;
}
