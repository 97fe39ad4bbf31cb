/*
 * ref_simd_init.c — wiring of the reference's OWN SSE intrinsics into its HEVCDSPContext for the
 * SIMD CPU baseline (oracle/_ref/liboh_ref_simd.so).  TEST / BENCH INFRASTRUCTURE, never the product.
 *
 * The reference's arch hook is ff_hevc_dsp_init_x86 (libavcodec/x86/hevcdsp_init.c:764-1277), called
 * at the end of ff_hevc_dsp_init (hevcdsp.c:583-586).  That file mixes two families of functions:
 *   ff_hevc_*      yasm sources (hevc_mc.asm, hevc_deblock.asm, hevc_idct.asm, hevc_res_add.asm):
 *                  there is no yasm / nasm in this image -> "unavailable on this box"
 *   ohhevc_*       C intrinsics (x86/hevc_mc_sse.c, hevc_idct_sse.c, hevc_sao_sse.c), which gcc builds
 * This init assigns exactly the ohhevc_* entries the reference's init assigns — its OH_EPEL_LINKS /
 * OH_QPEL_LINKS blocks (hevcdsp_init.c:646-668, :829-841; CONFIG_OH_OPTIM), the SSE2 transforms
 * (:801-805, :1009-1013, :1235-1239), the 8-bit SSE2 transform_add (:781-784), the SSE SAO (:840-841, :1030-1031,
 * :1245-1246) — for 8, 10 and 12 bit.
 * Deblocking stays the C table (its SIMD exists only as yasm).  Nothing of the reference is copied:
 * the functions are compiled from the sources where they lie (oracle/Makefile, target ref_simd).
 */
#include <stddef.h>
#include <stdint.h>

#include "libavcodec/hevcdsp.h"

#define MC_PROTOS(name, D)                                                                                          \
    void ohhevc_put_hevc_##name##_##D##_sse(int16_t *dst, uint8_t *src, ptrdiff_t srcstride, int height,            \
                                            intptr_t mx, intptr_t my, int width);                                   \
    void ohhevc_put_hevc_bi_##name##_##D##_sse(uint8_t *dst, ptrdiff_t dststride, uint8_t *src, ptrdiff_t srcstride, \
                                               int16_t *src2, int height, intptr_t mx, intptr_t my, int width);     \
    void ohhevc_put_hevc_uni_##name##_##D##_sse(uint8_t *dst, ptrdiff_t dststride, uint8_t *src,                    \
                                                ptrdiff_t srcstride, int height, intptr_t mx, intptr_t my,          \
                                                int width);                                                         \
    void ohhevc_put_hevc_uni_w_##name##_##D##_sse(uint8_t *dst, ptrdiff_t dststride, uint8_t *src,                  \
                                                  ptrdiff_t srcstride, int height, int denom, int wx, int ox,       \
                                                  intptr_t mx, intptr_t my, int width);                             \
    void ohhevc_put_hevc_bi_w_##name##_##D##_sse(uint8_t *dst, ptrdiff_t dststride, uint8_t *src,                   \
                                                 ptrdiff_t srcstride, int16_t *src2, int height, int denom,         \
                                                 int wx0, int wx1, int ox0, int ox1, intptr_t mx, intptr_t my,      \
                                                 int width);

#define MC_PROTOS_E(f, D) MC_PROTOS(f##4, D) MC_PROTOS(f##6, D) MC_PROTOS(f##8, D) MC_PROTOS(f##12, D) \
    MC_PROTOS(f##16, D) MC_PROTOS(f##24, D) MC_PROTOS(f##32, D) MC_PROTOS(f##48, D) MC_PROTOS(f##64, D)
#define MC_PROTOS_Q(f, D) MC_PROTOS(f##4, D) MC_PROTOS(f##8, D) MC_PROTOS(f##12, D) \
    MC_PROTOS(f##16, D) MC_PROTOS(f##24, D) MC_PROTOS(f##32, D) MC_PROTOS(f##48, D) MC_PROTOS(f##64, D)
#define MC_PROTOS_D(D) MC_PROTOS_E(pel_pixels, D) MC_PROTOS_E(epel_h, D) MC_PROTOS_E(epel_v, D) MC_PROTOS_E(epel_hv, D) \
    MC_PROTOS_Q(qpel_h, D) MC_PROTOS_Q(qpel_v, D) MC_PROTOS_Q(qpel_hv, D)
MC_PROTOS_D(8)
MC_PROTOS_D(10)
MC_PROTOS_D(12)

#define TR_PROTOS(D)                                                                    \
    void ohhevc_transform_4x4_luma_##D##_sse2(int16_t *coeffs);                          \
    void ohhevc_transform_4x4_##D##_sse2(int16_t *coeffs, int col_limit);                \
    void ohhevc_transform_8x8_##D##_sse2(int16_t *coeffs, int col_limit);                \
    void ohhevc_transform_16x16_##D##_sse2(int16_t *coeffs, int col_limit);              \
    void ohhevc_transform_32x32_##D##_sse2(int16_t *coeffs, int col_limit);              \
    void ohhevc_transform_4x4_add_##D##_sse2(uint8_t *dst, int16_t *c, ptrdiff_t stride);   \
    void ohhevc_transform_8x8_add_##D##_sse2(uint8_t *dst, int16_t *c, ptrdiff_t stride);   \
    void ohhevc_transform_16x16_add_##D##_sse2(uint8_t *dst, int16_t *c, ptrdiff_t stride); \
    void ohhevc_transform_32x32_add_##D##_sse2(uint8_t *dst, int16_t *c, ptrdiff_t stride); \
    void ohhevc_sao_band_filter_0_##D##_sse(uint8_t *dst, uint8_t *src, ptrdiff_t stride_dst, ptrdiff_t stride_src, \
                                            struct SAOParams *sao, int *borders, int width, int height, int c_idx); \
    void ohhevc_sao_edge_filter_##D##_sse(uint8_t *dst, uint8_t *src, ptrdiff_t stride_dst, ptrdiff_t stride_src,   \
                                          struct SAOParams *sao, int width, int height, int c_idx);
TR_PROTOS(8)
TR_PROTOS(10)
TR_PROTOS(12)

/* same shape as the reference's OH_PEL_LINK (x86/hevcdsp.h:42-47) */
#define LINK(dst, idx, my, mx, name, D)                                   \
    c->dst[idx][my][mx]         = ohhevc_put_hevc_##name##_##D##_sse;     \
    c->dst##_bi[idx][my][mx]    = ohhevc_put_hevc_bi_##name##_##D##_sse;  \
    c->dst##_uni[idx][my][mx]   = ohhevc_put_hevc_uni_##name##_##D##_sse; \
    c->dst##_uni_w[idx][my][mx] = ohhevc_put_hevc_uni_w_##name##_##D##_sse; \
    c->dst##_bi_w[idx][my][mx]  = ohhevc_put_hevc_bi_w_##name##_##D##_sse;
/* index = ff_hevc_pel_weight[width] (hevcdec.c:49); OH_EPEL_LINKS / OH_QPEL_LINKS, hevcdsp_init.c:646-665 */
#define LINKS_E(dst, my, mx, f, D)                                                                               \
    LINK(dst, 1, my, mx, f##4, D) LINK(dst, 2, my, mx, f##6, D) LINK(dst, 3, my, mx, f##8, D)                     \
    LINK(dst, 4, my, mx, f##12, D) LINK(dst, 5, my, mx, f##16, D) LINK(dst, 6, my, mx, f##24, D)                  \
    LINK(dst, 7, my, mx, f##32, D) LINK(dst, 8, my, mx, f##48, D) LINK(dst, 9, my, mx, f##64, D)
#define LINKS_Q(dst, my, mx, f, D)                                                                               \
    LINK(dst, 1, my, mx, f##4, D) LINK(dst, 3, my, mx, f##8, D)                                                   \
    LINK(dst, 4, my, mx, f##12, D) LINK(dst, 5, my, mx, f##16, D) LINK(dst, 6, my, mx, f##24, D)                  \
    LINK(dst, 7, my, mx, f##32, D) LINK(dst, 8, my, mx, f##48, D) LINK(dst, 9, my, mx, f##64, D)

#define INIT_DEPTH(D)                                                   \
    do {                                                                \
        LINKS_E(put_hevc_epel, 0, 0, pel_pixels, D)                     \
        LINKS_E(put_hevc_epel, 0, 1, epel_h, D)                         \
        LINKS_E(put_hevc_epel, 1, 0, epel_v, D)                         \
        LINKS_E(put_hevc_epel, 1, 1, epel_hv, D)                        \
        LINKS_Q(put_hevc_qpel, 0, 0, pel_pixels, D)                     \
        LINKS_Q(put_hevc_qpel, 0, 1, qpel_h, D)                         \
        LINKS_Q(put_hevc_qpel, 1, 0, qpel_v, D)                         \
        LINKS_Q(put_hevc_qpel, 1, 1, qpel_hv, D)                        \
        c->idct_4x4_luma = ohhevc_transform_4x4_luma_##D##_sse2;        \
        c->idct[0] = ohhevc_transform_4x4_##D##_sse2;                   \
        c->idct[1] = ohhevc_transform_8x8_##D##_sse2;                   \
        c->idct[2] = ohhevc_transform_16x16_##D##_sse2;                 \
        c->idct[3] = ohhevc_transform_32x32_##D##_sse2;                 \
        c->sao_band_filter = (void *)ohhevc_sao_band_filter_0_##D##_sse; \
        c->sao_edge_filter = (void *)ohhevc_sao_edge_filter_##D##_sse;  \
    } while (0)

/* what each group does is switchable so that the pinning test can bisect a mismatch:
 * bit 0 MC, bit 1 transforms, bit 2 transform_add, bit 3 SAO (default all) */
int ohref_simd_mask = 15;

void ff_hevc_dsp_init_x86(HEVCDSPContext *c, const int bit_depth)
{
    HEVCDSPContext keep = *c;
    switch (bit_depth) {
    case 8:
        INIT_DEPTH(8);
        /* the intrinsics transform_add is wired at 8 bit only (hevcdsp_init.c:781-784); at 10 / 12 bit the reference
         * assigns yasm functions (:1005-1007), unavailable here, so those depths keep the C transform_add.  (The
         * unwired ohhevc_transform_*_add_10_sse2 are not bit-exact with the C table on RExt residuals.) */
        c->transform_add[0] = ohhevc_transform_4x4_add_8_sse2;
        c->transform_add[1] = ohhevc_transform_8x8_add_8_sse2;
        c->transform_add[2] = ohhevc_transform_16x16_add_8_sse2;
        c->transform_add[3] = ohhevc_transform_32x32_add_8_sse2;
        break;
    case 10: INIT_DEPTH(10); break;
    case 12: INIT_DEPTH(12); break;
    default: return;                       /* 9 and 14 bit: C table (no intrinsics in the reference) */
    }
    if (!(ohref_simd_mask & 1)) {
        memcpy(c->put_hevc_epel, keep.put_hevc_epel, sizeof(keep.put_hevc_epel));
        memcpy(c->put_hevc_epel_uni, keep.put_hevc_epel_uni, sizeof(keep.put_hevc_epel_uni));
        memcpy(c->put_hevc_epel_uni_w, keep.put_hevc_epel_uni_w, sizeof(keep.put_hevc_epel_uni_w));
        memcpy(c->put_hevc_epel_bi, keep.put_hevc_epel_bi, sizeof(keep.put_hevc_epel_bi));
        memcpy(c->put_hevc_epel_bi_w, keep.put_hevc_epel_bi_w, sizeof(keep.put_hevc_epel_bi_w));
        memcpy(c->put_hevc_qpel, keep.put_hevc_qpel, sizeof(keep.put_hevc_qpel));
        memcpy(c->put_hevc_qpel_uni, keep.put_hevc_qpel_uni, sizeof(keep.put_hevc_qpel_uni));
        memcpy(c->put_hevc_qpel_uni_w, keep.put_hevc_qpel_uni_w, sizeof(keep.put_hevc_qpel_uni_w));
        memcpy(c->put_hevc_qpel_bi, keep.put_hevc_qpel_bi, sizeof(keep.put_hevc_qpel_bi));
        memcpy(c->put_hevc_qpel_bi_w, keep.put_hevc_qpel_bi_w, sizeof(keep.put_hevc_qpel_bi_w));
    }
    if (!(ohref_simd_mask & 2)) {
        c->idct_4x4_luma = keep.idct_4x4_luma;
        memcpy(c->idct, keep.idct, sizeof(keep.idct));
    }
    if (!(ohref_simd_mask & 4)) memcpy(c->transform_add, keep.transform_add, sizeof(keep.transform_add));
    if (!(ohref_simd_mask & 8)) { c->sao_band_filter = keep.sao_band_filter; c->sao_edge_filter = keep.sao_edge_filter; }
}
