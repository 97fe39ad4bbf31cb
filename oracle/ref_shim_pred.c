/*
 * ref_shim_pred.c — TEST INFRASTRUCTURE ONLY.  Makes the reference's intra prediction callable on a bare plane:
 * HEVCPredContext.intra_pred[log2 - 2](s, x0, y0, c_idx) (hevcpred.h:30-41, filled by ff_hevc_pred_init, hevcpred.c,
 * compiled from /root/reference where it lies) reads its inputs from an HEVCContext; this file builds the few fields
 * it touches (hevcpred_template.c:30-345).  Built by `make ref` into oracle/_ref/liboh_ref.so.
 *
 * Availability: the reference combines lc->na.cand_* with a z-scan order test on pps->min_tb_addr_zs (:103-107).  The
 * shim hands it a table in which the current block is the last in z-scan order, so that its final flags ARE the five
 * flags passed in — the same input the oracle and the device function take.
 */
#include <stdlib.h>
#include <string.h>

#include "libavcodec/hevcdec.h"
#include "libavcodec/hevcpred.h"

int ohref_intra_pred(int bd, uint8_t *plane, ptrdiff_t stride, int pw, int ph, int x, int y, int log2_size, int c_idx,
                     int mode, int cand, int chroma_format_idc, int smoothing_disabled, int strong)
{
    HEVCContext *s = calloc(1, sizeof(*s));
    HEVCLocalContext *lc = calloc(1, sizeof(*lc));
    HEVCSPS *sps = calloc(1, sizeof(*sps));
    HEVCPPS *pps = calloc(1, sizeof(*pps));
    AVFrame *frame = calloc(1, sizeof(*frame));
    const int hs = c_idx && (chroma_format_idc == 1 || chroma_format_idc == 2), vs = c_idx && chroma_format_idc == 1;
    const int log2_min_tb = 2, log2_ctb = 6, mask = (1 << (log2_ctb - log2_min_tb)) - 1, tw = mask + 2;
    int *zs = malloc(sizeof(int) * tw * tw);
    int rc = -1;
    if (!s || !lc || !sps || !pps || !frame || !zs || log2_size < 2 || log2_size > 5) goto out;

    sps->chroma_format_idc = chroma_format_idc;
    sps->hshift[0] = sps->vshift[0] = 0;
    sps->hshift[1] = sps->hshift[2] = chroma_format_idc == 1 || chroma_format_idc == 2;
    sps->vshift[1] = sps->vshift[2] = chroma_format_idc == 1;
    sps->width = pw << hs;  sps->height = ph << vs;       /* luma dimensions */
    sps->log2_min_tb_size = log2_min_tb;
    sps->log2_ctb_size = log2_ctb;
    sps->tb_mask = mask;
    sps->log2_min_pu_size = 2;
    sps->min_pu_width = sps->width >> 2;
    sps->min_pu_height = sps->height >> 2;
    sps->intra_smoothing_disabled_flag = smoothing_disabled;
    sps->sps_strong_intra_smoothing_enable_flag = strong;
    sps->bit_depth[0] = sps->bit_depth[1] = bd;
    pps->constrained_intra_pred_flag = 0;
    for (int i = 0; i < tw * tw; i++) zs[i] = -1;          /* every other block comes earlier ... */
    pps->min_tb_addr_zs_tab = zs;
    pps->min_tb_addr_zs = &zs[tw + 1];
    {
        const int x0 = x << hs, y0 = y << vs;
        pps->min_tb_addr_zs[((y0 >> log2_min_tb) & mask) * tw + ((x0 >> log2_min_tb) & mask)] = 1 << 20;   /* ... than this one */
        frame->data[c_idx] = plane;
        frame->linesize[c_idx] = (int)stride;
        s->frame = frame;
        s->ps.sps = sps;
        s->ps.pps = pps;
        s->HEVClc = lc;
        lc->tu.intra_pred_mode = mode;
        lc->tu.intra_pred_mode_c = mode;
        lc->na.cand_bottom_left = !!(cand & 1);
        lc->na.cand_left = !!(cand & 2);
        lc->na.cand_up_left = !!(cand & 4);
        lc->na.cand_up = !!(cand & 8);
        lc->na.cand_up_right = !!(cand & 16);
        ff_hevc_pred_init(&s->hpc, bd);
        s->hpc.intra_pred[log2_size - 2](s, x0, y0, c_idx);
    }
    rc = 0;
out:
    free(zs); free(frame); free(pps); free(sps); free(lc); free(s);
    return rc;
}
