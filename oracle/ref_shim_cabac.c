/*
 * ref_shim_cabac.c — TEST INFRASTRUCTURE ONLY.  Makes the reference's dequantisation
 * callable: hevc_cabac.c keeps derive_quant_parameters{,_c} (:1472-1557) and
 * scale_and_clip_coeff (:2030-2055) `static`, so this file #includes the reference source
 * where it lies (nothing is copied) and drives exactly those functions with a minimal
 * HEVCContext.  Built by `make ref` into oracle/_ref/liboh_ref.so.
 */
#include "libavcodec/hevc_cabac.c"

#include <stdlib.h>
#include <string.h>

/* One N x N block of un-scaled levels -> coefficients, through the reference's own
 * derive_quant_parameters / _c and scale_and_clip_coeff (which it calls once per decoded,
 * i.e. non-zero, level: hevc_cabac.c:2349, :2395).  Returns quant_ctx->qp, the value the
 * C ABI takes as ohcu_tu_quant.qp.  sl / sl_dc: ScalingList.sl / .sl_dc, used when
 * sl_enabled (sps->scaling_list_enabled_flag, list taken from the SPS). */
int ohref_dequant_block(int bd, int chroma_format_idc, int c_idx, int log2_size, int qp_y, int qp_offset,
                        int intra, int transform_skip, int sl_enabled, const uint8_t *sl, const uint8_t *sl_dc,
                        int16_t *blk)
{
    HEVCContext *s = calloc(1, sizeof(*s));
    HEVCLocalContext *lc = calloc(1, sizeof(*lc));
    HEVCSPS *sps = calloc(1, sizeof(*sps));
    HEVCPPS *pps = calloc(1, sizeof(*pps));
    HEVCTransformContext tr;
    HEVCQuantContext *q = &tr.quant_ctx;
    const int n = 1 << log2_size;
    int qp;

    memset(&tr, 0, sizeof(tr));
    sps->bit_depth[CHANNEL_TYPE_LUMA] = sps->bit_depth[CHANNEL_TYPE_CHROMA] = bd;
    sps->qp_bd_offset = 6 * (bd - 8);
    sps->chroma_format_idc = chroma_format_idc;
    sps->scaling_list_enabled_flag = sl_enabled;
    if (sl_enabled) {
        memcpy(sps->scaling_list.sl, sl, sizeof(sps->scaling_list.sl));
        memcpy(sps->scaling_list.sl_dc, sl_dc, sizeof(sps->scaling_list.sl_dc));
    }
    s->ps.sps = sps;
    s->ps.pps = pps;
    s->HEVClc = lc;
    lc->qp_y = qp_y;
    lc->cu.pred_mode = intra ? MODE_INTRA : MODE_INTER;
    if (c_idx == 1) pps->pps_cb_qp_offset = qp_offset;
    if (c_idx == 2) pps->pps_cr_qp_offset = qp_offset;
    tr.log2_trafo_size = log2_size;
    tr.log2_tr_size_minus2 = log2_size - 2;
    tr.transform_size = n;
    tr.transform_skip_flag = transform_skip;
    tr.log2_transform_range = 15;                       /* hevc_cabac.c:2112, :2451 without extended precision */
    if (c_idx) derive_quant_parameters_c(s, lc, &tr, q, c_idx);
    else derive_quant_parameters(s, lc, &tr, q);
    for (int y = 0; y < n; y++)
        for (int x = 0; x < n; x++)
            if (blk[y * n + x])
                blk[y * n + x] = (int16_t)scale_and_clip_coeff(s, &tr, q, blk[y * n + x], x, y);
    qp = q->qp;
    free(pps); free(sps); free(lc); free(s);
    return qp;
}
