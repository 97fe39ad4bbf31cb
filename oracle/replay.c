/*
 * replay.c — frame-level CPU replay of a work list on top of hevc_oracle.c.
 * TEST INFRASTRUCTURE ONLY (see hevc_oracle.h).
 *
 * Restates the reference's DRIVERS around the DSP table:
 *   MC         hls_prediction_unit dispatch hevcdec.c:2250-2306, luma_mc_uni/bi and
 *              chroma_mc_uni/bi hevcdec.c:1744-2049, ff_emulated_edge_mc
 *              videodsp_template.c:22-100 (restated as a replicate clamp of the
 *              source coordinates, applied always: it is idempotent inside the picture)
 *   residual   ff_hevc_hls_transform{,_c} hevc_cabac.c:1695-1906
 *   intra      the s->hpc.intra_pred[] calls of hls_transform_unit hevcdec.c:1335-1526 (intra_oracle.c)
 *   deblock    deblocking_filter_CTB hevc_filter.c:525-762 (per CTB, raster order)
 *   SAO        sao_filter_CTB hevc_filter.c:264-502 + restore_tqb_pixels :229-260,
 *              restated out of place from a snapshot of the deblocked picture (the
 *              reference reaches the same pre-SAO samples through its line buffers)
 * Pinned against the reference's own drivers (oracle/ref_shim.c,
 * ohref_replay_frame) in tests/test_oracle_vs_ref.py.
 */
#include "replay.h"
#include "hevc_oracle.h"

#include <stdlib.h>
#include <string.h>

typedef struct rp {
    const oh_frame_job *job;
    const ohcu_seq_params *sp;
    int bd, P;
    int hs[3], vs[3];
    int nplanes;
    int ctb_w, ctb_h, ctb_size;
    int bs_width;
    int min_cb_w, min_pu_w, min_pu_h;
    oh_picture cur;
} rp;

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline int iclip(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* ================================================================== MC */
#define EMU_STRIDE 80            /* EDGE_EMU_BUFFER_STRIDE, hevcdec.h:84 */

/* Fetch a (bw x bh) window whose top-left is (sx, sy) with every coordinate
 * clamped into the picture: what emulated_edge_mc builds near borders. */
static void fetch_clamped(const rp *r, uint8_t *buf, const uint8_t *plane, ptrdiff_t stride,
                          int sx, int sy, int bw, int bh, int pw, int ph)
{
    for (int y = 0; y < bh; y++) {
        int yy = iclip(sy + y, 0, ph - 1);
        for (int x = 0; x < bw; x++) {
            int xx = iclip(sx + x, 0, pw - 1);
            memcpy(buf + ((size_t)y * EMU_STRIDE + x) * r->P, plane + yy * stride + (size_t)xx * r->P, r->P);
        }
    }
}

/* one plane of one PU; c_idx 0 = luma_mc_*, 1/2 = chroma_mc_* */
static void predict_plane(const rp *r, const ohcu_pu *pu, const ohcu_slice_tab *sl, int c)
{
    int luma = c == 0;
    int hs = r->hs[c], vs = r->vs[c];
    int pw = r->sp->width >> hs, ph = r->sp->height >> vs;
    int bw = pu->w >> hs, bh = pu->h >> vs;
    int xo = pu->x0 >> hs, yo = pu->y0 >> vs;
    int before = luma ? 3 : 1, extra = luma ? 7 : 3;
    uint8_t *dst = r->cur.data[c] + (size_t)yo * r->cur.stride[c] + (size_t)xo * r->P;
    ptrdiff_t dstride = r->cur.stride[c];
    ptrdiff_t estride = EMU_STRIDE * r->P;
    uint8_t *emu[2];
    int16_t *tmp;
    int mx[2], my[2];
    const uint8_t *src[2] = { NULL, NULL };
    int denom = luma ? sl->luma_log2_weight_denom : sl->chroma_log2_weight_denom;
    int wx[2], ox[2];

    emu[0] = malloc((size_t)(64 + 7) * estride);
    emu[1] = malloc((size_t)(64 + 7) * estride);
    tmp = malloc(sizeof(int16_t) * 64 * 64);

    for (int l = 0; l < 2; l++) {
        int mvx = pu->mv[l][0], mvy = pu->mv[l][1];
        int ri = pu->ref_idx[l];
        const oh_picture *ref;
        int sx, sy;
        if (!(pu->pred_flag & (1 << l)))
            continue;
        ref = &r->job->dpb[sl->ref_slot[l][ri]];
        if (luma) {
            mx[l] = mvx & 3;  my[l] = mvy & 3;                 /* hevcdec.c:1753-1754 */
            sx = xo + (mvx >> 2); sy = yo + (mvy >> 2);
            wx[l] = sl->luma_weight[l][ri]; ox[l] = sl->luma_offset[l][ri];
        } else {
            /* hevcdec.c:1909-1915: phase in 1/8 (4:2:0) scaled up when not subsampled */
            int fx = mvx & ((1 << (2 + hs)) - 1), fy = mvy & ((1 << (2 + vs)) - 1);
            mx[l] = fx << (1 - hs); my[l] = fy << (1 - vs);
            sx = xo + (mvx >> (2 + hs)); sy = yo + (mvy >> (2 + vs));
            wx[l] = sl->chroma_weight[l][ri][c - 1]; ox[l] = sl->chroma_offset[l][ri][c - 1];
        }
        fetch_clamped(r, emu[l], ref->data[c], ref->stride[c], sx - before, sy - before,
                      bw + extra, bh + extra, pw, ph);
        src[l] = emu[l] + before * estride + before * r->P;
    }

    if (pu->pred_flag == OHCU_PF_BI) {
        /* luma_mc_bi / chroma_mc_bi: list 0 -> int16 tmp, list 1 combined with it */
        or_mc_put(r->bd, luma, tmp, src[0], estride, bh, mx[0], my[0], bw);
        if (!sl->weight_flag)
            or_mc_bi(r->bd, luma, dst, dstride, src[1], estride, tmp, bh, mx[1], my[1], bw);
        else
            or_mc_bi_w(r->bd, luma, dst, dstride, src[1], estride, tmp, bh, denom,
                       wx[0], wx[1], ox[0], ox[1], mx[1], my[1], bw);
    } else {
        int l = pu->pred_flag == OHCU_PF_L1;
        if (!sl->weight_flag)
            or_mc_uni(r->bd, luma, dst, dstride, src[l], estride, bh, mx[l], my[l], bw);
        else
            or_mc_uni_w(r->bd, luma, dst, dstride, src[l], estride, bh, denom, wx[l], ox[l],
                        mx[l], my[l], bw);
    }
    free(emu[0]); free(emu[1]); free(tmp);
}

/* ================================================================== residual */
static void tu_residual(const rp *r, const ohcu_tu *tu, int16_t *out)
{
    int n2 = 1 << (2 * tu->log2_size);
    memcpy(out, r->job->coeffs + tu->coeff_off, sizeof(int16_t) * n2);
    switch (tu->mode) {
    case OHCU_TU_BYPASS:                                       /* hevc_cabac.c:1727-1734 */
        if (tu->flags & OHCU_TUF_RDPCM)
            or_transform_rdpcm(out, tu->log2_size, !!(tu->flags & OHCU_TUF_RDPCM_VERT));
        break;
    case OHCU_TU_SKIP:                                         /* hevc_cabac.c:1736-1757 */
        if (tu->flags & OHCU_TUF_ROTATE)
            for (int i = 0; i < 8; i++) {
                int16_t t = out[i]; out[i] = out[15 - i]; out[15 - i] = t;
            }
        or_transform_skip(r->bd, out, tu->log2_size);
        if (tu->flags & OHCU_TUF_RDPCM)
            or_transform_rdpcm(out, tu->log2_size, !!(tu->flags & OHCU_TUF_RDPCM_VERT));
        break;
    case OHCU_TU_DST4: or_idct_4x4_luma(r->bd, out); break;       /* :1790-1791 */
    case OHCU_TU_DC:   or_idct_dc(r->bd, out, tu->log2_size); break; /* :1794-1796 */
    default:           or_idct(r->bd, out, tu->log2_size, tu->col_limit); break; /* :1805 */
    }
}

/* the intra blocks in decoding order: prediction from the picture as reconstructed so far, then the residual of
 * the coded TB at the same position (hls_transform_unit, hevcdec.c:1335-1526) */
static void residual_tu(const rp *r, const ohcu_tu *tu);
static int intra_blocks(const rp *r)
{
    const oh_frame_job *job = r->job;
    const ohcu_seq_params *sp = r->sp;
    int32_t *map[3] = { NULL, NULL, NULL };
    int cw[3], rc = 0;
    for (int c = 0; c < r->nplanes; c++) {
        cw[c] = ((sp->width >> r->hs[c]) + 3) >> 2;
        map[c] = calloc((size_t)cw[c] * (((sp->height >> r->vs[c]) + 3) >> 2), sizeof(int32_t));
    }
    for (int i = 0; i < job->n_tu; i++) {
        const ohcu_tu *tu = &job->tus[i];
        if ((tu->flags & OHCU_TUF_INTRA) && tu->c_idx < r->nplanes)
            map[tu->c_idx][(tu->y0 >> 2) * cw[tu->c_idx] + (tu->x0 >> 2)] = i + 1;
    }
    for (int i = 0; i < job->n_iu && !rc; i++) {
        const ohcu_iu *iu = &job->ius[i];
        int c = iu->c_idx, pw, ph;
        if (c >= r->nplanes || iu->log2_size < 2 || iu->log2_size > 5 || iu->mode > 34) { rc = OHCU_EINVAL; break; }
        pw = sp->width >> r->hs[c]; ph = sp->height >> r->vs[c];
        or_intra_pred(r->bd, r->cur.data[c], r->cur.stride[c], pw, ph, iu->x0, iu->y0, iu->log2_size, c, iu->mode,
                      iu->cand & 31, !sp->intra_smoothing_disabled && (c == 0 || sp->chroma_format_idc == 3),
                      sp->strong_intra_smoothing);
        if (iu->cand & OHCU_IU_CBF) {
            int t = map[c][(iu->y0 >> 2) * cw[c] + (iu->x0 >> 2)];
            if (!t || job->tus[t - 1].log2_size != iu->log2_size) { rc = OHCU_EINVAL; break; }
            residual_tu(r, &job->tus[t - 1]);
        }
    }
    for (int c = 0; c < 3; c++) free(map[c]);
    return rc;
}

static void residual_tu(const rp *r, const ohcu_tu *tu)
{
    int16_t res[32 * 32], res_y[32 * 32];
    int n = 1 << tu->log2_size;
    int c = tu->c_idx;
    tu_residual(r, tu, res);
    if (tu->flags & OHCU_TUF_CROSS) {                          /* hevc_cabac.c:1810-1817 */
        tu_residual(r, &r->job->tus[tu->cross_tu], res_y);
        for (int i = 0; i < n * n; i++)
            res[i] = (int16_t)(res[i] + ((tu->res_scale_val * res_y[i]) >> 3));
    }
    or_transform_add(r->bd, r->cur.data[c] + (size_t)tu->y0 * r->cur.stride[c] + (size_t)tu->x0 * r->P,
                     res, r->cur.stride[c], n);
}

/* ================================================================== deblock */
static const uint8_t tc_tab[54] = {     /* tctable, hevc_filter.c:53-57 */
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3,
    3, 3, 3, 4, 4, 4, 5, 5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 22, 24
};
static const uint8_t beta_tab[52] = {   /* betatable, hevc_filter.c:59-63 */
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18,
    20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64
};

static int qpy_at(const rp *r, int x, int y)                   /* get_qPy, hevc_filter.c:149-155 */
{
    int s = r->sp->log2_min_cb_size;
    return r->job->deblock->qp_y_tab[(x >> s) + (y >> s) * r->min_cb_w];
}
static int pcm_at(const rp *r, int x, int y)                   /* get_pcm, hevc_filter.c:504-518 */
{
    int s = r->sp->log2_min_pu_size;
    if (x < 0 || y < 0) return 2;
    if ((x >> s) >= r->min_pu_w || (y >> s) >= r->min_pu_h) return 2;
    return r->job->deblock->is_pcm ? r->job->deblock->is_pcm[(y >> s) * r->min_pu_w + (x >> s)] : 0;
}
static int luma_tc(int qp, int bs, int tc_offset)              /* TC_CALC, hevc_filter.c:520-523 */
{
    return tc_tab[iclip(qp + 2 * (bs - 1) + (tc_offset >> 1 << 1), 0, 53)];
}
static int chroma_tc(const rp *r, int qp_y, int c, int tc_offset)  /* hevc_filter.c:65-92 */
{
    static const uint8_t map420[14] = { 29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37 };
    int qi = iclip(qp_y + (c == 1 ? r->sp->cb_qp_offset : r->sp->cr_qp_offset), 0, 57);
    int qp;
    if (r->sp->chroma_format_idc == 1)
        qp = qi < 30 ? qi : (qi > 43 ? qi - 6 : map420[qi - 30]);
    else
        qp = iclip(qi, 0, 51);
    return tc_tab[iclip(qp + 2 + tc_offset, 0, 53)];
}
static uint8_t *pix_at(const rp *r, int c, int x, int y)       /* luma coordinates */
{
    return r->cur.data[c] + (size_t)(y >> r->vs[c]) * r->cur.stride[c] + (size_t)(x >> r->hs[c]) * r->P;
}

/* deblocking_filter_CTB, hevc_filter.c:525-762 */
static void deblock_ctb(const rp *r, int x0, int y0)
{
    const ohcu_deblock_in *d = r->job->deblock;
    int W = r->sp->width, H = r->sp->height;
    int ctb = (x0 >> r->sp->log2_ctb_size) + (y0 >> r->sp->log2_ctb_size) * r->ctb_w;
    int cur_tc = d->deblock[ctb].tc_offset, cur_beta = d->deblock[ctb].beta_offset;
    int left_tc = x0 ? d->deblock[ctb - 1].tc_offset : 0;
    int left_beta = x0 ? d->deblock[ctb - 1].beta_offset : 0;
    int pcmf = r->sp->pcm_or_bypass;
    int x_end = imin(x0 + r->ctb_size, W), y_end = imin(y0 + r->ctb_size, H);
    int x_end2 = x_end;
    int h = 1 << r->hs[1], v = 1 << r->vs[1];
    int tc[2], beta;
    uint8_t no_p[2] = { 0, 0 }, no_q[2] = { 0, 0 };
    const uint8_t *vbs = d->vertical_bs, *hbs = d->horizontal_bs;
    int bw = r->bs_width;

    /* vertical edges, luma (:565-602) */
    for (int y = y0; y < y_end; y += 8)
        for (int x = x0 ? x0 : 8; x < x_end; x += 8) {
            int bs0 = vbs[(x + y * bw) >> 2], bs1 = vbs[(x + (y + 4) * bw) >> 2];
            if (!bs0 && !bs1) continue;
            int qp = (qpy_at(r, x - 1, y) + qpy_at(r, x, y) + 1) >> 1;
            beta = beta_tab[iclip(qp + cur_beta, 0, 51)];
            tc[0] = bs0 ? luma_tc(qp, bs0, cur_tc) : 0;
            tc[1] = bs1 ? luma_tc(qp, bs1, cur_tc) : 0;
            if (pcmf) {
                no_p[0] = pcm_at(r, x - 1, y); no_p[1] = pcm_at(r, x - 1, y + 4);
                no_q[0] = pcm_at(r, x, y);     no_q[1] = pcm_at(r, x, y + 4);
            }
            or_loop_filter_luma(r->bd, 1, pix_at(r, 0, x, y), r->cur.stride[0], beta, tc, no_p, no_q);
        }
    /* vertical edges, chroma (:605-657): only bs == 2, grid 8 chroma samples */
    if (r->sp->chroma_format_idc)
        for (int y = y0; y < y_end; y += 8 * v)
            for (int x = x0 ? x0 : 8 * h; x < x_end; x += 8 * h) {
                int bs0 = vbs[(x + y * bw) >> 2], bs1 = vbs[(x + (y + 4 * v) * bw) >> 2];
                if (bs0 != 2 && bs1 != 2) continue;
                int qp0 = (qpy_at(r, x - 1, y) + qpy_at(r, x, y) + 1) >> 1;
                int qp1 = (qpy_at(r, x - 1, y + 4 * v) + qpy_at(r, x, y + 4 * v) + 1) >> 1;
                if (pcmf) {
                    no_p[0] = pcm_at(r, x - 1, y); no_p[1] = pcm_at(r, x - 1, y + 4 * v);
                    no_q[0] = pcm_at(r, x, y);     no_q[1] = pcm_at(r, x, y + 4 * v);
                }
                for (int c = 1; c <= 2; c++) {
                    tc[0] = bs0 == 2 ? chroma_tc(r, qp0, c, cur_tc) : 0;
                    tc[1] = bs1 == 2 ? chroma_tc(r, qp1, c, cur_tc) : 0;
                    or_loop_filter_chroma(r->bd, 1, pix_at(r, c, x, y), r->cur.stride[c], tc, no_p, no_q);
                }
            }

    /* horizontal edges, luma (:660-703): shifted 8 left so that the columns next to
     * the CTB's right edge wait for the neighbour's vertical pass.  The first 8
     * columns (left CTB's) take left_beta but, as in the reference, the tc offset
     * of THIS CTB. */
    if (x_end != W)
        x_end -= 8;
    for (int y = y0 ? y0 : 8; y < y_end; y += 8) {
        int boff = x0 ? left_beta : cur_beta;
        for (int x = x0 ? x0 - 8 : 0; x < x_end; x += 8) {
            int bs0 = hbs[(x + y * bw) >> 2], bs1 = hbs[((x + 4) + y * bw) >> 2];
            if (bs0 || bs1) {
                int qp = (qpy_at(r, x, y - 1) + qpy_at(r, x, y) + 1) >> 1;
                beta = beta_tab[iclip(qp + boff, 0, 51)];
                tc[0] = bs0 ? luma_tc(qp, bs0, cur_tc) : 0;
                tc[1] = bs1 ? luma_tc(qp, bs1, cur_tc) : 0;
                if (pcmf) {
                    no_p[0] = pcm_at(r, x, y - 1); no_p[1] = pcm_at(r, x + 4, y - 1);
                    no_q[0] = pcm_at(r, x, y);     no_q[1] = pcm_at(r, x + 4, y);
                }
                or_loop_filter_luma(r->bd, 0, pix_at(r, 0, x, y), r->cur.stride[0], beta, tc, no_p, no_q);
            }
            boff = cur_beta;
        }
    }
    /* horizontal edges, chroma (:706-761): first segment of the first (left CTB's)
     * group takes left_tc, its second segment and everything after cur_tc (:721) */
    if (r->sp->chroma_format_idc) {
        if (x_end2 != W)
            x_end = x_end2 - 8 * h;
        for (int y = y0 ? y0 : 8 * v; y < y_end; y += 8 * v) {
            int toff = x0 ? left_tc : cur_tc;
            for (int x = x0 ? x0 - 8 * h : 0; x < x_end; x += 8 * h) {
                int bs0 = hbs[(x + y * bw) >> 2], bs1 = hbs[((x + 4 * h) + y * bw) >> 2];
                if (bs0 == 2 || bs1 == 2) {
                    int qp0 = bs0 == 2 ? (qpy_at(r, x, y - 1) + qpy_at(r, x, y) + 1) >> 1 : 0;
                    int qp1 = bs1 == 2 ? (qpy_at(r, x + 4 * h, y - 1) + qpy_at(r, x + 4 * h, y) + 1) >> 1 : 0;
                    if (pcmf) {
                        no_p[0] = pcm_at(r, x, y - 1); no_p[1] = pcm_at(r, x + 4 * h, y - 1);
                        no_q[0] = pcm_at(r, x, y);     no_q[1] = pcm_at(r, x + 4 * h, y);
                    }
                    for (int c = 1; c <= 2; c++) {
                        tc[0] = bs0 == 2 ? chroma_tc(r, qp0, c, toff) : 0;
                        tc[1] = bs1 == 2 ? chroma_tc(r, qp1, c, cur_tc) : 0;
                        or_loop_filter_chroma(r->bd, 0, pix_at(r, c, x, y), r->cur.stride[c], tc, no_p, no_q);
                    }
                }
                toff = cur_tc;
            }
        }
    }
}

/* ================================================================== SAO */
typedef struct snap { uint8_t *buf[3]; uint8_t *org[3]; ptrdiff_t stride[3]; } snap;

/* sao_filter_CTB, hevc_filter.c:264-502.  `pre` is the deblocked picture before
 * any SAO, with a 1-sample margin so that the out-of-picture halo reads of
 * sao_edge_filter at picture borders stay in bounds (their results are replaced
 * by sao_edge_restore, exactly as in the reference). */
static void sao_ctb(const rp *r, const snap *pre, int x, int y)
{
    const ohcu_sao_in *si = r->job->sao;
    int x_ctb = x >> r->sp->log2_ctb_size, y_ctb = y >> r->sp->log2_ctb_size;
    int rs = y_ctb * r->ctb_w + x_ctb;
    const ohcu_sao *sao = &si->sao[rs];
    int lfase = si->filter_slice_edges ? si->filter_slice_edges[rs] : 1;
    int no_tile = r->sp->no_tile_filter;
    int restore = no_tile || !lfase;
    int edges[4] = { x_ctb == 0, y_ctb == 0, x_ctb == r->ctb_w - 1, y_ctb == r->ctb_h - 1 };
    uint8_t vert[2] = { 0, 0 }, horiz[2] = { 0, 0 }, diag[4] = { 0, 0, 0, 0 };
#define SLICE_AT(dx, dy) (si->tab_slice_address ? si->tab_slice_address[rs + (dy) * r->ctb_w + (dx)] : 0)
#define TILE_AT(dx, dy)  (si->tile_id ? si->tile_id[rs + (dy) * r->ctb_w + (dx)] : 0)
#define DIFF_SLICE(dx, dy) (!lfase && SLICE_AT(0, 0) != SLICE_AT(dx, dy))
#define DIFF_TILE(dx, dy)  (no_tile && TILE_AT(0, 0) != TILE_AT(dx, dy))
    if (restore) {                                             /* :292-325 */
        int lt = 0, rt = 0, ut = 0, bt = 0;
        if (!edges[0]) { lt = DIFF_TILE(-1, 0); vert[0]  = DIFF_SLICE(-1, 0) || lt; }
        if (!edges[2]) { rt = DIFF_TILE(1, 0);  vert[1]  = DIFF_SLICE(1, 0)  || rt; }
        if (!edges[1]) { ut = DIFF_TILE(0, -1); horiz[0] = DIFF_SLICE(0, -1) || ut; }
        if (!edges[3]) { bt = DIFF_TILE(0, 1);  horiz[1] = DIFF_SLICE(0, 1)  || bt; }
        if (!edges[0] && !edges[1]) diag[0] = DIFF_SLICE(-1, -1) || lt || ut;
        if (!edges[1] && !edges[2]) diag[1] = DIFF_SLICE(1, -1)  || rt || ut;
        if (!edges[2] && !edges[3]) diag[2] = DIFF_SLICE(1, 1)   || rt || bt;
        if (!edges[0] && !edges[3]) diag[3] = DIFF_SLICE(-1, 1)  || lt || bt;
    }
    for (int c = 0; c < r->nplanes; c++) {
        int x0 = x >> r->hs[c], y0 = y >> r->vs[c];
        int w = imin(r->ctb_size >> r->hs[c], (r->sp->width >> r->hs[c]) - x0);
        int h = imin(r->ctb_size >> r->vs[c], (r->sp->height >> r->vs[c]) - y0);
        uint8_t *frame = r->cur.data[c] + (size_t)y0 * r->cur.stride[c] + (size_t)x0 * r->P;
        const uint8_t *copy = pre->org[c] + (size_t)y0 * pre->stride[c] + (size_t)x0 * r->P;
        if (sao->type_idx[c] == OHCU_SAO_BAND)
            or_sao_band_filter(r->bd, frame, copy, r->cur.stride[c], pre->stride[c],
                               sao->offset_val[c], sao->band_position[c], w, h);
        else if (sao->type_idx[c] == OHCU_SAO_EDGE) {
            or_sao_edge_filter(r->bd, frame, copy, r->cur.stride[c], pre->stride[c],
                               sao->offset_val[c], sao->eo_class[c], w, h);
            or_sao_edge_restore(r->bd, restore, frame, copy, r->cur.stride[c], pre->stride[c],
                                sao->offset_val[c], sao->eo_class[c], edges, w, h, vert, horiz, diag);
        } else
            continue;
        /* restore_tqb_pixels, :229-260.  As in the reference the min-PU range is
         * computed from the LUMA origin plus the PLANE's width/height. */
        if (r->sp->pcm_or_bypass && r->job->deblock && r->job->deblock->is_pcm) {
            int l2 = r->sp->log2_min_pu_size, mps = 1 << l2;
            for (int py = y >> l2; py < ((y + h) >> l2); py++)
                for (int px = x >> l2; px < ((x + w) >> l2); px++) {
                    if (!r->job->deblock->is_pcm[py * r->min_pu_w + px]) continue;
                    int ox = ((px << l2) - x) >> r->hs[c], oy = ((py << l2) - y) >> r->vs[c];
                    for (int n = 0; n < (mps >> r->vs[c]); n++)
                        memcpy(frame + (size_t)(oy + n) * r->cur.stride[c] + (size_t)ox * r->P,
                               copy + (size_t)(oy + n) * pre->stride[c] + (size_t)ox * r->P,
                               (size_t)(mps >> r->hs[c]) * r->P);
                }
        }
    }
#undef SLICE_AT
#undef TILE_AT
#undef DIFF_SLICE
#undef DIFF_TILE
}

/* ================================================================== frame */
int or_replay_frame(const oh_frame_job *job)
{
    rp r;
    const ohcu_seq_params *sp = &job->sp;
    int stages = job->stages ? job->stages : 15;
    memset(&r, 0, sizeof(r));
    r.job = job; r.sp = sp;
    r.bd = sp->bit_depth; r.P = sp->bit_depth > 8 ? 2 : 1;
    r.nplanes = sp->chroma_format_idc ? 3 : 1;
    r.hs[1] = r.hs[2] = (sp->chroma_format_idc == 1 || sp->chroma_format_idc == 2);
    r.vs[1] = r.vs[2] = (sp->chroma_format_idc == 1);
    r.ctb_size = 1 << sp->log2_ctb_size;
    r.ctb_w = (sp->width + r.ctb_size - 1) >> sp->log2_ctb_size;
    r.ctb_h = (sp->height + r.ctb_size - 1) >> sp->log2_ctb_size;
    r.bs_width = (sp->width >> 2) + 1;
    r.min_cb_w = sp->width >> sp->log2_min_cb_size;
    r.min_pu_w = sp->width >> sp->log2_min_pu_size;
    r.min_pu_h = sp->height >> sp->log2_min_pu_size;
    r.cur = job->out;
    if (sp->bit_depth < 8 || sp->bit_depth > 14 || sp->chroma_format_idc < 0 || sp->chroma_format_idc > 3)
        return OHCU_EINVAL;
    /* Not supported (documented in DESIGN.md, parity hazards): with 16x16 CTBs and
     * horizontally subsampled chroma the reference defers a whole CTB of horizontal
     * chroma edges (x_end = x_end2 - 8*h, hevc_filter.c:715) past the point where its
     * CTB-delayed SAO reads them, so its output is not the picture-level filter order. */
    if (sp->log2_ctb_size < 4 || sp->log2_ctb_size > 6 ||
        (sp->log2_ctb_size == 4 && (sp->chroma_format_idc == 1 || sp->chroma_format_idc == 2)))
        return OHCU_EINVAL;

    if (stages & OH_STAGE_MC)
        for (int i = 0; i < job->n_pu; i++) {
            const ohcu_pu *pu = &job->pus[i];
            if (pu->slice >= job->n_slices || pu->pred_flag < 1 || pu->pred_flag > 3)
                return OHCU_EINVAL;
            for (int c = 0; c < r.nplanes; c++)
                predict_plane(&r, pu, &job->slices[pu->slice], c);
        }
    if (stages & OH_STAGE_RESIDUAL) {
        for (int i = 0; i < job->n_tu; i++)
            if (!(job->tus[i].flags & OHCU_TUF_INTRA))
                residual_tu(&r, &job->tus[i]);
        if (job->n_iu) {
            int rc;
            /* ff_hevc_pred_init (hevcpred.c:62-76) instantiates 8, 9, 10 and 12 bits only; any other depth gets the
             * 8-bit predictors, which read 16-bit planes as bytes: no defined behaviour to restate */
            if (sp->bit_depth > 12 || sp->bit_depth == 11) return OHCU_EINVAL;
            rc = intra_blocks(&r);
            if (rc) return rc;
        }
    }
    if ((stages & OH_STAGE_DEBLOCK) && job->deblock)
        for (int y = 0; y < sp->height; y += r.ctb_size)
            for (int x = 0; x < sp->width; x += r.ctb_size)
                deblock_ctb(&r, x, y);
    if ((stages & OH_STAGE_SAO) && job->sao && sp->sao_enabled) {
        snap pre;
        for (int c = 0; c < r.nplanes; c++) {
            int w = sp->width >> r.hs[c], h = sp->height >> r.vs[c];
            pre.stride[c] = (ptrdiff_t)(w + 2) * r.P;
            pre.buf[c] = calloc((size_t)(h + 2), (size_t)pre.stride[c]);
            pre.org[c] = pre.buf[c] + pre.stride[c] + r.P;
            for (int yy = 0; yy < h; yy++)
                memcpy(pre.org[c] + (size_t)yy * pre.stride[c], r.cur.data[c] + (size_t)yy * r.cur.stride[c], (size_t)w * r.P);
        }
        for (int y = 0; y < sp->height; y += r.ctb_size)
            for (int x = 0; x < sp->width; x += r.ctb_size)
                sao_ctb(&r, &pre, x, y);
        for (int c = 0; c < r.nplanes; c++)
            free(pre.buf[c]);
    }
    return 0;
}

/* ================================================================== boundary strengths
 * Restatement of ff_hevc_deblocking_boundary_strengths (hevc_filter.c:1006-1125) +
 * boundary_strength (:764-880, TEST_MV_POC variant) in terms of the lists the C ABI carries.
 * The reference visits every edge unit (4 samples on the 8x8 grid) exactly once, from the
 * leaf on its Q side (the current block):
 *   - on the leaf's own left / top boundary (:1020-1086): 2 if either side is intra, else 1 if
 *     the transform block on either side has cbf_luma, else the motion rule;
 *   - inside the leaf (:1088-1124): the motion rule only.
 * An edge unit is "a transform-block boundary with cbf_luma on a side" exactly when it lies on
 * the perimeter of a CODED luma TB, so the coded TBs the ABI already has are enough: the edge
 * separates two 4x4 cells that do not belong to the same coded TB and one of them belongs to
 * one.  Cells no PU covers count as intra: an edge with such a cell on a side is 2 - unless both cells
 * lie in the same intra leaf, which is no transform-tree boundary at all (the luma ohcu_iu blocks ARE the
 * leaves of the intra CUs).  The motion rule compares reference PICTURES (the
 * reference compares POCs; here DPB slots, one picture per slot). */
typedef struct bs_cell { int32_t pu, tb, iu; } bs_cell;

static int bs_far(const int16_t a[2], const int16_t b[2])
{
    return abs(a[0] - b[0]) >= 4 || abs(a[1] - b[1]) >= 4;
}

static int bs_motion(const oh_bs_job *job, const ohcu_pu *c, const ohcu_pu *n)
{
    const ohcu_slice_tab *sc = &job->slices[c->slice], *sn = &job->slices[n->slice];
    int pc[2], pn[2];
    for (int l = 0; l < 2; l++) {
        pc[l] = sc->ref_slot[l][c->ref_idx[l] & (OHCU_MAX_REFS - 1)];
        pn[l] = sn->ref_slot[l][n->ref_idx[l] & (OHCU_MAX_REFS - 1)];
    }
    if (c->pred_flag == OHCU_PF_BI && n->pred_flag == OHCU_PF_BI) {
        if (pc[0] == pn[0] && pc[0] == pc[1] && pn[0] == pn[1])
            return (bs_far(n->mv[0], c->mv[0]) || bs_far(n->mv[1], c->mv[1])) &&
                   (bs_far(n->mv[1], c->mv[0]) || bs_far(n->mv[0], c->mv[1]));
        if (pn[0] == pc[0] && pn[1] == pc[1])
            return bs_far(n->mv[0], c->mv[0]) || bs_far(n->mv[1], c->mv[1]);
        if (pn[1] == pc[0] && pn[0] == pc[1])
            return bs_far(n->mv[1], c->mv[0]) || bs_far(n->mv[0], c->mv[1]);
        return 1;
    }
    if (c->pred_flag != OHCU_PF_BI && n->pred_flag != OHCU_PF_BI) {
        const int lc = (c->pred_flag & 1) ? 0 : 1, ln = (n->pred_flag & 1) ? 0 : 1;
        return pc[lc] == pn[ln] ? bs_far(c->mv[lc], n->mv[ln]) : 1;
    }
    return 1;
}

static int bs_unit(const oh_bs_job *job, const bs_cell *q, const bs_cell *p)
{
    if (q->pu < 0 || p->pu < 0) return (q->iu >= 0 && q->iu == p->iu) ? 0 : 2;
    if (q->tb != p->tb && (q->tb >= 0 || p->tb >= 0)) return 1;
    if (q->pu == p->pu) return 0;
    return bs_motion(job, &job->pus[q->pu], &job->pus[p->pu]);
}

int or_boundary_strengths(const oh_bs_job *job)
{
    const ohcu_seq_params *sp = &job->sp;
    const int W = sp->width, H = sp->height, gw = W >> 2, gh = H >> 2, bsw = gw + 1, bsh = gh + 1;
    const int l2 = sp->log2_ctb_size, ctb_w = (W + (1 << l2) - 1) >> l2, cmask = (1 << l2) - 1;
    bs_cell *g = malloc(sizeof(bs_cell) * (size_t)gw * gh);
    if (!g) return OHCU_ENOMEM;
    for (int i = 0; i < gw * gh; i++) g[i].pu = g[i].tb = g[i].iu = -1;
    for (int i = 0; i < job->n_iu; i++) {
        const ohcu_iu *iu = &job->ius[i];
        const int n4 = 1 << (iu->log2_size - 2);
        if (iu->c_idx) continue;
        for (int y = iu->y0 >> 2; y < (iu->y0 >> 2) + n4 && y < gh; y++)
            for (int x = iu->x0 >> 2; x < (iu->x0 >> 2) + n4 && x < gw; x++) g[y * gw + x].iu = i;
    }
    for (int i = 0; i < job->n_pu; i++) {
        const ohcu_pu *pu = &job->pus[i];
        for (int y = pu->y0 >> 2; y < (pu->y0 + pu->h) >> 2 && y < gh; y++)
            for (int x = pu->x0 >> 2; x < (pu->x0 + pu->w) >> 2 && x < gw; x++) g[y * gw + x].pu = i;
    }
    for (int i = 0; i < job->n_tu; i++) {
        const ohcu_tu *tu = &job->tus[i];
        const int n4 = 1 << (tu->log2_size - 2);
        if (tu->c_idx) continue;
        for (int y = tu->y0 >> 2; y < (tu->y0 >> 2) + n4 && y < gh; y++)
            for (int x = tu->x0 >> 2; x < (tu->x0 >> 2) + n4 && x < gw; x++) g[y * gw + x].tb = i;
    }
    memset(job->vbs, 0, (size_t)bsw * bsh);
    memset(job->hbs, 0, (size_t)bsw * bsh);
    for (int dir = 0; dir < 2; dir++)                      /* 0: vertical edges (left neighbour), 1: horizontal (upper) */
        for (int y = dir ? 8 : 0; y < H; y += dir ? 8 : 4)
            for (int x = dir ? 0 : 8; x < W; x += dir ? 4 : 8) {
                const int cq = (y >> l2) * ctb_w + (x >> l2);
                const int cp = dir ? cq - ctb_w : cq - 1;
                const bs_cell *q = &g[(y >> 2) * gw + (x >> 2)];
                const bs_cell *p = dir ? q - gw : q - 1;
                if (job->deblock_disabled && job->deblock_disabled[cq]) continue;      /* the leaf's call is not made */
                if (!((dir ? y : x) & cmask)) {            /* a CTB boundary: slice / tile conditions, :1010-1016, :1052-1058 */
                    const int across = job->filter_slice_edges ? job->filter_slice_edges[cq] : 1;
                    if (job->tab_slice_address && job->tab_slice_address[cq] != job->tab_slice_address[cp] && !across) continue;
                    if (sp->no_tile_filter && job->tile_id && job->tile_id[cq] != job->tile_id[cp]) continue;
                }
                (dir ? job->hbs : job->vbs)[(x + y * bsw) >> 2] = (uint8_t)bs_unit(job, q, p);
            }
    free(g);
    return 0;
}

/* ---------------------------------------------------------------- cropped copy-out
 * hevc_refs.c:335-341: data[i] += (left >> hshift << pixel_shift) + (top >> vshift) * linesize[i];
 * openhevc.c:722-735: for every row, memcpy(dst + y * dst_stride, src + y * src_stride, dst_stride)
 * with dst_stride = the cropped width in bytes (set_cropped_dim_from_avframe, openhevc.c:463). */
void or_output_cropped(const uint8_t *const src[3], const ptrdiff_t stride[3], int nplanes, int pixel_shift,
                       int hshift, int vshift, int left, int top, int width, int height, uint8_t *const dst[3])
{
    for (int i = 0; i < nplanes; i++) {
        const int hs = i ? hshift : 0, vs = i ? vshift : 0;
        const uint8_t *s = src[i] + ((ptrdiff_t)(left >> hs) << pixel_shift) + (ptrdiff_t)(top >> vs) * stride[i];
        const size_t row = (size_t)(width >> hs) << pixel_shift;
        for (int y = 0; y < height >> vs; y++)
            memcpy(dst[i] + y * row, s + y * stride[i], row);
    }
}
