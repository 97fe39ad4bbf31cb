/*
 * hevc_oracle.c — CPU restatement of the reference's HEVCDSPContext leaves.
 * TEST INFRASTRUCTURE ONLY (see hevc_oracle.h).  Each function cites the
 * reference lines it restates; all paths are libavcodec/ in /root/reference.
 *
 * Written with the bit depth as a run-time argument and plain accessor
 * helpers instead of the reference's macro template, and the inverse DCT as a
 * masked matrix product instead of the even/odd butterfly macros (the two are
 * the same exact-integer sums; tests/test_oracle_vs_ref.py proves it against
 * the compiled reference).
 */
#include "hevc_oracle.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ helpers */
static inline int px_ld(const uint8_t *base, ptrdiff_t idx, int bd)
{
    return bd == 8 ? base[idx] : ((const uint16_t *)base)[idx];
}
static inline void px_st(uint8_t *base, ptrdiff_t idx, int bd, int v)
{
    if (bd == 8) base[idx] = (uint8_t)v;
    else ((uint16_t *)base)[idx] = (uint16_t)v;
}
static inline int px_size(int bd) { return bd == 8 ? 1 : 2; }
static inline int clip3(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline int clip_px(int v, int bd) { return clip3(v, 0, (1 << bd) - 1); }
static inline int clip_s16(int v) { return clip3(v, -32768, 32767); }
/* left shift that is well defined for negative values (the reference relies on
 * gcc's arithmetic behaviour, e.g. (ox0+ox1+1) << log2Wd, template :2453) */
static inline int shl(int v, int s) { return (int)((unsigned)v << s); }

/* ------------------------------------------------------------------ tables */
/* ff_hevc_qpel_filters (hevcdsp.c:191-195), indexed by the phase directly */
const int8_t or_qpel_taps[4][8] = {
    { 0, 0,   0, 64,  0,   0, 0,  0 },
    { -1, 4, -10, 58, 17,  -5, 1,  0 },
    { -1, 4, -11, 40, 40, -11, 4, -1 },
    { 0, 1,  -5, 17, 58, -10, 4, -1 },
};
/* ff_hevc_epel_filters (hevcdsp.c:181-189), indexed by the phase directly */
const int8_t or_epel_taps[8][4] = {
    { 0, 64, 0, 0 },
    { -2, 58, 10, -2 }, { -4, 54, 16, -2 }, { -6, 46, 28, -4 }, { -4, 36, 36, -4 },
    { -4, 28, 46, -6 }, { -2, 16, 54, -4 }, { -2, 10, 58, -2 },
};

/* The 32x32 core transform (hevcdsp.c:32-97) follows one rule: entry (r,c) is
 * the tuned value of cos(r*(2c+1)*pi/64), row 0 being the constant 64.  The 33
 * magnitudes for angles m*pi/64, m = 0..32: */
static const int8_t dct_mag[33] = {
    64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64,
    61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9, 4, 0
};
int or_dct_coef(int row, int col)
{
    int m;
    if (row == 0)
        return 64;
    m = (row * (2 * col + 1)) & 127;
    if (m <= 32) return dct_mag[m];
    if (m <= 64) return -dct_mag[64 - m];
    if (m <= 96) return -dct_mag[m - 64];
    return dct_mag[128 - m];
}

/* ------------------------------------------------------------------ residual */
/* transquant_bypass / transform_add{4,8,16,32} — template :45-84 */
void or_transform_add(int bd, uint8_t *dst, const int16_t *coeffs, ptrdiff_t stride, int size)
{
    ptrdiff_t ps = stride / px_size(bd);
    for (int y = 0; y < size; y++)
        for (int x = 0; x < size; x++)
            px_st(dst, y * ps + x, bd, clip_px(px_ld(dst, y * ps + x, bd) + coeffs[y * size + x], bd));
}

/* transform_rdpcm — template :87-107.  Running sums wrap in int16. */
void or_transform_rdpcm(int16_t *c, int log2_size, int mode)
{
    int n = 1 << log2_size;
    if (mode) {            /* vertical: each row accumulates the row above */
        for (int y = 1; y < n; y++)
            for (int x = 0; x < n; x++)
                c[y * n + x] = (int16_t)(c[y * n + x] + c[(y - 1) * n + x]);
    } else {               /* horizontal prefix sum */
        for (int y = 0; y < n; y++)
            for (int x = 1; x < n; x++)
                c[y * n + x] = (int16_t)(c[y * n + x] + c[y * n + x - 1]);
    }
}

/* transform_skip — template :109-133 */
void or_transform_skip(int bd, int16_t *c, int log2_size)
{
    int shift = 15 - bd - log2_size;
    int n2 = 1 << (2 * log2_size);
    if (shift > 0) {
        int off = 1 << (shift - 1);
        for (int i = 0; i < n2; i++)
            c[i] = (int16_t)((c[i] + off) >> shift);
    } else {
        for (int i = 0; i < n2; i++)
            c[i] = (int16_t)shl(c[i], -shift);
    }
}

/* one 4-point DST-VII pass — TR_4x4_LUMA, template :140-155 */
static void dst4_pass(int16_t *v, int step, int shift)
{
    int add = 1 << (shift - 1);
    int s0 = v[0], s1 = v[step], s2 = v[2 * step], s3 = v[3 * step];
    int a = s0 + s2, b = s2 + s3, d = s0 - s3, e = 74 * s1;
    v[0]        = (int16_t)clip_s16((29 * a + 55 * b + e + add) >> shift);
    v[step]     = (int16_t)clip_s16((55 * d - 29 * b + e + add) >> shift);
    v[2 * step] = (int16_t)clip_s16((74 * (s0 - s2 + s3) + add) >> shift);
    v[3 * step] = (int16_t)clip_s16((55 * a + 29 * d - e + add) >> shift);
}
/* transform_4x4_luma — template :157-177 */
void or_idct_4x4_luma(int bd, int16_t *c)
{
    for (int i = 0; i < 4; i++) dst4_pass(c + i, 4, 7);
    for (int i = 0; i < 4; i++) dst4_pass(c + 4 * i, 1, 20 - bd);
}

/* Which input index j of an N-point pass does the reference butterfly read when
 * called with zero-skip bound `end`?  TR_32 sums odd j < end and recurses with
 * end/2 into TR_16 on the even inputs; TR_16/TR_8 called from a larger size get
 * the full bound; TR_4 ignores it (template :180-239). */
static int idct_reads(int n, int j, int end)
{
    if (n == 4) return 1;
    if (j & 1) return j < end;
    if (n == 32 && (j & 3) == 2) return (j >> 1) < (end >> 1);
    return 1;
}
/* one N-point inverse pass over v[0], v[step], ... as a masked matrix product */
static void idct_pass(int16_t *v, int step, int n, int end, int shift)
{
    int add = 1 << (shift - 1);
    int stride32 = 32 / n;          /* N-point matrix = rows 0, 32/N, ... of the 32x32 */
    int in[32], out[32];
    for (int j = 0; j < n; j++)
        in[j] = idct_reads(n, j, end) ? v[j * step] : 0;
    for (int k = 0; k < n; k++) {
        int acc = 0;
        for (int j = 0; j < n; j++)
            acc += or_dct_coef(j * stride32, k) * in[j];
        out[k] = clip_s16((acc + add) >> shift);
    }
    for (int k = 0; k < n; k++)
        v[k * step] = (int16_t)out[k];
}
/* idct_NxN — IDCT(H) macro, template :249-271, with IDCT_VAR* :241-247 */
void or_idct(int bd, int16_t *c, int log2_size, int col_limit)
{
    int n = 1 << log2_size;
    int limit = col_limit < n ? col_limit : n;
    int limit2 = col_limit + 4 < n ? col_limit + 4 : n;
    for (int i = 0; i < n; i++) {          /* columns, shift 7 */
        idct_pass(c + i, n, n, limit2, 7);
        if (limit2 < n && i % 4 == 0 && i != 0)
            limit2 -= 4;
    }
    for (int i = 0; i < n; i++)            /* rows, shift 20 - bd */
        idct_pass(c + i * n, 1, n, limit, 20 - bd);
}

/* idct_NxN_dc — IDCT_DC(H) macro, template :273-286 */
void or_idct_dc(int bd, int16_t *c, int log2_size)
{
    int shift = 14 - bd;
    int add = 1 << (shift - 1);
    int v = (((c[0] + 1) >> 1) + add) >> shift;
    int n2 = 1 << (2 * log2_size);
    for (int i = 0; i < n2; i++)
        c[i] = (int16_t)v;
}

/* ------------------------------------------------------------------ SAO */
/* sao_band_filter_0 — template :310-335 */
void or_sao_band_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst,
                        ptrdiff_t stride_src, const int16_t offset_val[5], int band_position,
                        int width, int height)
{
    int lut[32] = { 0 };
    int shift = bd - 5;
    ptrdiff_t ds = stride_dst / px_size(bd), ss = stride_src / px_size(bd);
    for (int k = 0; k < 4; k++)
        lut[(k + band_position) & 31] = offset_val[k + 1];
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++) {
            int s = px_ld(src, y * ss + x, bd);
            px_st(dst, y * ds + x, bd, clip_px(s + lut[s >> shift], bd));
        }
}

static inline int sgn3(int a, int b) { return (a > b) - (a < b); }

/* sao_edge_filter — template :342-378.  Neighbour pair per class: horizontal,
 * vertical, 135 degree (up-left / down-right), 45 degree (up-right / down-left). */
void or_sao_edge_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst,
                        ptrdiff_t stride_src, const int16_t offset_val[5], int eo_class,
                        int width, int height)
{
    static const int8_t dxy[4][4] = { { -1, 0, 1, 0 }, { 0, -1, 0, 1 }, { -1, -1, 1, 1 }, { 1, -1, -1, 1 } };
    static const uint8_t cat[5] = { 1, 2, 0, 3, 4 };
    ptrdiff_t ds = stride_dst / px_size(bd), ss = stride_src / px_size(bd);
    ptrdiff_t a = dxy[eo_class][0] + dxy[eo_class][1] * ss;
    ptrdiff_t b = dxy[eo_class][2] + dxy[eo_class][3] * ss;
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++) {
            ptrdiff_t p = y * ss + x;
            int c = px_ld(src, p, bd);
            int k = cat[2 + sgn3(c, px_ld(src, p + a, bd)) + sgn3(c, px_ld(src, p + b, bd))];
            px_st(dst, y * ds + x, bd, clip_px(c + offset_val[k], bd));
        }
}

/* sao_edge_restore_0 / _1 — template :380-436 and :438-529.
 * restore_idx selects [0] (picture borders only) or [1] (+ slice/tile rules). */
void or_sao_edge_restore(int bd, int restore_idx, uint8_t *dst, const uint8_t *src,
                         ptrdiff_t stride_dst, ptrdiff_t stride_src,
                         const int16_t offset_val[5], int eo_class, const int borders[4],
                         int width, int height, const uint8_t vert_edge[2],
                         const uint8_t horiz_edge[2], const uint8_t diag_edge[4])
{
    ptrdiff_t ds = stride_dst / px_size(bd), ss = stride_src / px_size(bd);
    int first_x = 0;
    int o0 = offset_val[0];

    /* picture-border columns / rows keep src + offset_val[0] (= src) */
    if (eo_class != 1 /* VERT */) {
        if (borders[0]) {
            for (int y = 0; y < height; y++)
                px_st(dst, y * ds, bd, clip_px(px_ld(src, y * ss, bd) + o0, bd));
            first_x = 1;
        }
        if (borders[2]) {
            for (int y = 0; y < height; y++)
                px_st(dst, y * ds + width - 1, bd, clip_px(px_ld(src, y * ss + width - 1, bd) + o0, bd));
            width--;
        }
    }
    if (eo_class != 0 /* HORIZ */) {
        if (borders[1])
            for (int x = first_x; x < width; x++)
                px_st(dst, x, bd, clip_px(px_ld(src, x, bd) + o0, bd));
        if (borders[3]) {
            for (int x = first_x; x < width; x++)
                px_st(dst, (height - 1) * ds + x, bd,
                      clip_px(px_ld(src, (height - 1) * ss + x, bd) + o0, bd));
            height--;
        }
    }
    if (!restore_idx)
        return;

    /* slice / tile boundaries that must not be filtered across — template :495-528.
     * Note width/height here are the values already shrunk above. */
    {
        int keep_ul = !diag_edge[0] && eo_class == 2 && !borders[0] && !borders[1];
        int keep_ur = !diag_edge[1] && eo_class == 3 && !borders[1] && !borders[2];
        int keep_lr = !diag_edge[2] && eo_class == 2 && !borders[2] && !borders[3];
        int keep_ll = !diag_edge[3] && eo_class == 3 && !borders[0] && !borders[3];

        if (vert_edge[0] && eo_class != 1)
            for (int y = keep_ul; y < height - keep_ll; y++)
                px_st(dst, y * ds, bd, px_ld(src, y * ss, bd));
        if (vert_edge[1] && eo_class != 1)
            for (int y = keep_ur; y < height - keep_lr; y++)
                px_st(dst, y * ds + width - 1, bd, px_ld(src, y * ss + width - 1, bd));
        if (horiz_edge[0] && eo_class != 0)
            for (int x = first_x + keep_ul; x < width - keep_ur; x++)
                px_st(dst, x, bd, px_ld(src, x, bd));
        if (horiz_edge[1] && eo_class != 0)
            for (int x = first_x + keep_ll; x < width - keep_lr; x++)
                px_st(dst, (height - 1) * ds + x, bd, px_ld(src, (height - 1) * ss + x, bd));
        if (diag_edge[0] && eo_class == 2)
            px_st(dst, 0, bd, px_ld(src, 0, bd));
        if (diag_edge[1] && eo_class == 3)
            px_st(dst, width - 1, bd, px_ld(src, width - 1, bd));
        if (diag_edge[2] && eo_class == 2)
            px_st(dst, (height - 1) * ds + width - 1, bd, px_ld(src, (height - 1) * ss + width - 1, bd));
        if (diag_edge[3] && eo_class == 3)
            px_st(dst, (height - 1) * ds, bd, px_ld(src, (height - 1) * ss, bd));
    }
}

/* ------------------------------------------------------------------ MC */
/* The 14-bit intermediate of one sample: put_hevc_pel_pixels / put_hevc_{q,e}pel_{h,v,hv}
 * (template :2359-2373, :2480-2542, :2935-2999).  `row` is the int16 tmp line
 * store of the hv case; everything is truncated to int16 exactly where the
 * reference stores into int16_t. */
typedef struct {
    int bd, ntaps, before;          /* 8/3 for luma, 4/1 for chroma */
    const int8_t *fx, *fy;          /* NULL when that phase is zero  */
} mc_kernel;

static mc_kernel mc_setup(int bd, int luma, int mx, int my)
{
    mc_kernel k;
    k.bd = bd;
    k.ntaps = luma ? 8 : 4;
    k.before = luma ? 3 : 1;
    k.fx = mx ? (luma ? or_qpel_taps[mx] : or_epel_taps[mx]) : NULL;
    k.fy = my ? (luma ? or_qpel_taps[my] : or_epel_taps[my]) : NULL;
    return k;
}

static int mc_hsum(const mc_kernel *k, const uint8_t *src, ptrdiff_t pos)
{
    int acc = 0;
    for (int t = 0; t < k->ntaps; t++)
        acc += k->fx[t] * px_ld(src, pos + t - k->before, k->bd);
    return acc >> (k->bd - 8);
}

/* intermediate block: out[y*64 + x], int16 */
static void mc_intermediate(const mc_kernel *k, int16_t *out, const uint8_t *src,
                            ptrdiff_t ss /* in pixels */, int w, int h)
{
    if (!k->fx && !k->fy) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                out[y * 64 + x] = (int16_t)shl(px_ld(src, y * ss + x, k->bd), 14 - k->bd);
    } else if (k->fx && !k->fy) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                out[y * 64 + x] = (int16_t)mc_hsum(k, src, y * ss + x);
    } else if (!k->fx) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                int acc = 0;
                for (int t = 0; t < k->ntaps; t++)
                    acc += k->fy[t] * px_ld(src, (y + t - k->before) * ss + x, k->bd);
                out[y * 64 + x] = (int16_t)(acc >> (k->bd - 8));
            }
    } else {
        /* horizontal pass over h + ntaps - 1 rows into an int16 scratch, then the
         * vertical pass with >> 6 (template :2512-2542) */
        int16_t *tmp = (int16_t *)malloc(sizeof(int16_t) * (64 + 7) * 64);
        int rows = h + k->ntaps - 1;
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < w; x++)
                tmp[y * 64 + x] = (int16_t)mc_hsum(k, src, (y - k->before) * ss + x);
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                int acc = 0;
                for (int t = 0; t < k->ntaps; t++)
                    acc += k->fy[t] * tmp[(y + t) * 64 + x];
                out[y * 64 + x] = (int16_t)(acc >> 6);
            }
        free(tmp);
    }
}

/* put_hevc_qpel / put_hevc_epel [idx][!!my][!!mx] */
void or_mc_put(int bd, int luma, int16_t *dst, const uint8_t *src, ptrdiff_t srcstride,
               int height, int mx, int my, int width)
{
    mc_kernel k = mc_setup(bd, luma, mx, my);
    mc_intermediate(&k, dst, src, srcstride / px_size(bd), width, height);
}

/* put_hevc_*_uni — template :2375-2389 (full-pel = memcpy), :2544-2691, :3001-3095 */
void or_mc_uni(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
               ptrdiff_t srcstride, int height, int mx, int my, int width)
{
    ptrdiff_t ds = dststride / px_size(bd), ss = srcstride / px_size(bd);
    if (!mx && !my) {
        for (int y = 0; y < height; y++)
            memcpy(dst + y * dststride, src + y * srcstride, (size_t)width * px_size(bd));
        return;
    }
    mc_kernel k = mc_setup(bd, luma, mx, my);
    int16_t *im = (int16_t *)malloc(sizeof(int16_t) * 64 * 64);
    int shift = 14 - bd;
    int offset = bd < 14 ? 1 << (shift - 1) : 0;
    mc_intermediate(&k, im, src, ss, width, height);
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++)
            px_st(dst, y * ds + x, bd, clip_px((im[y * 64 + x] + offset) >> shift, bd));
    free(im);
}

/* put_hevc_*_uni_w — template :2417-2439, :2733-2882, :3097-3200 */
void or_mc_uni_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                 ptrdiff_t srcstride, int height, int denom, int wx, int ox,
                 int mx, int my, int width)
{
    ptrdiff_t ds = dststride / px_size(bd), ss = srcstride / px_size(bd);
    mc_kernel k = mc_setup(bd, luma, mx, my);
    int16_t *im = (int16_t *)malloc(sizeof(int16_t) * 64 * 64);
    int shift = denom + 14 - bd;
    int offset = bd < 14 ? 1 << (shift - 1) : 0;
    ox = ox * (1 << (bd - 8));
    mc_intermediate(&k, im, src, ss, width, height);
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++)
            px_st(dst, y * ds + x, bd, clip_px(((im[y * 64 + x] * wx + offset) >> shift) + ox, bd));
    free(im);
}

/* put_hevc_*_bi — template :2391-2415, :2570-2731, :3003-3200 */
void or_mc_bi(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
              ptrdiff_t srcstride, const int16_t *src2, int height, int mx, int my, int width)
{
    ptrdiff_t ds = dststride / px_size(bd), ss = srcstride / px_size(bd);
    mc_kernel k = mc_setup(bd, luma, mx, my);
    int16_t *im = (int16_t *)malloc(sizeof(int16_t) * 64 * 64);
    int shift = 14 + 1 - bd;
    int offset = bd < 14 ? 1 << (shift - 1) : 0;
    mc_intermediate(&k, im, src, ss, width, height);
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++)
            px_st(dst, y * ds + x, bd,
                  clip_px((im[y * 64 + x] + src2[y * 64 + x] + offset) >> shift, bd));
    free(im);
}

/* put_hevc_*_bi_w — template :2441-2465, :2760-2922, :3202-3357.
 * wx1 weighs the directly filtered reference, wx0 the int16 src2. */
void or_mc_bi_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                ptrdiff_t srcstride, const int16_t *src2, int height, int denom,
                int wx0, int wx1, int ox0, int ox1, int mx, int my, int width)
{
    ptrdiff_t ds = dststride / px_size(bd), ss = srcstride / px_size(bd);
    mc_kernel k = mc_setup(bd, luma, mx, my);
    int16_t *im = (int16_t *)malloc(sizeof(int16_t) * 64 * 64);
    int log2Wd = denom + 14 - bd;
    ox0 = ox0 * (1 << (bd - 8));
    ox1 = ox1 * (1 << (bd - 8));
    mc_intermediate(&k, im, src, ss, width, height);
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++)
            px_st(dst, y * ds + x, bd,
                  clip_px((im[y * 64 + x] * wx1 + src2[y * 64 + x] * wx0 +
                           shl(ox0 + ox1 + 1, log2Wd)) >> (log2Wd + 1), bd));
    free(im);
}

/* ------------------------------------------------------------------ deblock */
/* hevc_loop_filter_luma — template :3377-3471.  One call = 8 samples along the
 * edge = two 4-line segments.  `across` steps from P to Q, `along` steps to the
 * next line (both in pixels). */
void or_loop_filter_luma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride, int beta_in,
                         const int tc_in[2], const uint8_t no_p_in[2], const uint8_t no_q_in[2])
{
    ptrdiff_t ps = stride / px_size(bd);
    ptrdiff_t across = vertical ? 1 : ps, along = vertical ? ps : 1;
    int beta = beta_in << (bd - 8);

    for (int seg = 0; seg < 2; seg++) {
        ptrdiff_t base = (ptrdiff_t)seg * 4 * along;
        int tc = tc_in[seg] << (bd - 8);
        int no_p = no_p_in[seg], no_q = no_q_in[seg];
        int s[2][8];                /* lines 0 and 3: P3..P0 Q0..Q3 */
        for (int l = 0; l < 2; l++)
            for (int i = 0; i < 8; i++)
                s[l][i] = px_ld(pix, base + (l ? 3 : 0) * along + (i - 4) * across, bd);
        int dp0 = abs(s[0][1] - 2 * s[0][2] + s[0][3]);
        int dq0 = abs(s[0][6] - 2 * s[0][5] + s[0][4]);
        int dp3 = abs(s[1][1] - 2 * s[1][2] + s[1][3]);
        int dq3 = abs(s[1][6] - 2 * s[1][5] + s[1][4]);
        int d0 = dp0 + dq0, d3 = dp3 + dq3;
        if (d0 + d3 >= beta)
            continue;

        int tc25 = (tc * 5 + 1) >> 1;
        int strong =
            abs(s[0][0] - s[0][3]) + abs(s[0][7] - s[0][4]) < (beta >> 3) && abs(s[0][3] - s[0][4]) < tc25 &&
            abs(s[1][0] - s[1][3]) + abs(s[1][7] - s[1][4]) < (beta >> 3) && abs(s[1][3] - s[1][4]) < tc25 &&
            (d0 << 1) < (beta >> 2) && (d3 << 1) < (beta >> 2);
        int side = (beta + (beta >> 1)) >> 3;
        int nd_p = (dp0 + dp3 < side) ? 2 : 1;
        int nd_q = (dq0 + dq3 < side) ? 2 : 1;

        for (int d = 0; d < 4; d++) {
            ptrdiff_t o = base + d * along;
            int p3 = px_ld(pix, o - 4 * across, bd), p2 = px_ld(pix, o - 3 * across, bd);
            int p1 = px_ld(pix, o - 2 * across, bd), p0 = px_ld(pix, o - across, bd);
            int q0 = px_ld(pix, o, bd), q1 = px_ld(pix, o + across, bd);
            int q2 = px_ld(pix, o + 2 * across, bd), q3 = px_ld(pix, o + 3 * across, bd);
            if (strong) {
                int tc2 = tc << 1;
                if (!no_p) {
                    px_st(pix, o - across, bd,     p0 + clip3(((p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3) - p0, -tc2, tc2));
                    px_st(pix, o - 2 * across, bd, p1 + clip3(((p2 + p1 + p0 + q0 + 2) >> 2) - p1, -tc2, tc2));
                    px_st(pix, o - 3 * across, bd, p2 + clip3(((2 * p3 + 3 * p2 + p1 + p0 + q0 + 4) >> 3) - p2, -tc2, tc2));
                }
                if (!no_q) {
                    px_st(pix, o, bd,              q0 + clip3(((p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3) - q0, -tc2, tc2));
                    px_st(pix, o + across, bd,     q1 + clip3(((p0 + q0 + q1 + q2 + 2) >> 2) - q1, -tc2, tc2));
                    px_st(pix, o + 2 * across, bd, q2 + clip3(((2 * q3 + 3 * q2 + q1 + q0 + p0 + 4) >> 3) - q2, -tc2, tc2));
                }
            } else {
                int tc_2 = tc >> 1;
                int delta0 = (9 * (q0 - p0) - 3 * (q1 - p1) + 8) >> 4;
                if (abs(delta0) < 10 * tc) {
                    delta0 = clip3(delta0, -tc, tc);
                    if (!no_p) px_st(pix, o - across, bd, clip_px(p0 + delta0, bd));
                    if (!no_q) px_st(pix, o, bd, clip_px(q0 - delta0, bd));
                    if (!no_p && nd_p > 1) {
                        int dp = clip3((((p2 + p0 + 1) >> 1) - p1 + delta0) >> 1, -tc_2, tc_2);
                        px_st(pix, o - 2 * across, bd, clip_px(p1 + dp, bd));
                    }
                    if (!no_q && nd_q > 1) {
                        int dq = clip3((((q2 + q0 + 1) >> 1) - q1 - delta0) >> 1, -tc_2, tc_2);
                        px_st(pix, o + across, bd, clip_px(q1 + dq, bd));
                    }
                }
            }
        }
    }
}

/* hevc_loop_filter_chroma — template :3473-3505 */
void or_loop_filter_chroma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride,
                           const int tc_in[2], const uint8_t no_p_in[2], const uint8_t no_q_in[2])
{
    ptrdiff_t ps = stride / px_size(bd);
    ptrdiff_t across = vertical ? 1 : ps, along = vertical ? ps : 1;
    for (int seg = 0; seg < 2; seg++) {
        int tc = tc_in[seg] << (bd - 8);
        if (tc <= 0)
            continue;
        for (int d = 0; d < 4; d++) {
            ptrdiff_t o = (ptrdiff_t)(seg * 4 + d) * along;
            int p1 = px_ld(pix, o - 2 * across, bd), p0 = px_ld(pix, o - across, bd);
            int q0 = px_ld(pix, o, bd), q1 = px_ld(pix, o + across, bd);
            int delta0 = clip3((((q0 - p0) * 4) + p1 - q1 + 4) >> 3, -tc, tc);
            if (!no_p_in[seg]) px_st(pix, o - across, bd, clip_px(p0 + delta0, bd));
            if (!no_q_in[seg]) px_st(pix, o, bd, clip_px(q0 - delta0, bd));
        }
    }
}

/* ---------------------------------------------------------------- dequantisation
 * derive_quant_parameters{,_c} (hevc_cabac.c:1472-1557): shift = bit_depth + log2 + 10 -
 * log2_transform_range (15 without extended_precision_processing), add = 1 << (shift - 1),
 * scale = level_scale[qp % 6] << (qp / 6); scale_and_clip_coeff (:2030-2055): scale_m from
 * the scaling list by position (8x8 lists sub-sampled for 16x16 / 32x32, dc_scale for their
 * DC), 64-bit product, clip to int16. */
#include "../include/ohcu.h"

void or_dequant_block(int16_t *blk, int log2_size, int bd, int qp, const uint8_t *matrix, int dc_scale)
{
    static const int level_scale[6] = {40, 45, 51, 57, 64, 72};
    const int n = 1 << log2_size, shift = bd + log2_size + 10 - 15;
    /* rem6[] / div6[] (hevc_cabac.c:1456-1468) are declared with 76 entries but initialised with
     * 74: for qp 74 and 75 (12-bit video, qp_y >= 50) the reference reads zeros, i.e. scales with
     * level_scale[0] << 0.  Reproduced: parity is against the reference, not the standard. */
    const int rem = qp < 74 ? qp % 6 : 0, div = qp < 74 ? qp / 6 : 0;
    const int64_t add = (int64_t)1 << (shift - 1), scale = (int64_t)level_scale[rem] << div;
    for (int y = 0; y < n; y++)
        for (int x = 0; x < n; x++) {
            const int64_t level = blk[y * n + x];
            int64_t m = 16, t;
            if (!level) continue;
            if (matrix) {
                if (x || y || log2_size < 4) {
                    int pos;
                    switch (log2_size) {
                    case 3: pos = (y << 3) + x; break;
                    case 4: pos = ((y >> 1) << 3) + (x >> 1); break;
                    case 5: pos = ((y >> 2) << 3) + (x >> 2); break;
                    default: pos = (y << 2) + x; break;
                    }
                    m = matrix[pos];
                } else {
                    m = dc_scale;
                }
            }
            t = (level * scale * m + add) >> shift;
            blk[y * n + x] = (int16_t)(t < -32768 ? -32768 : t > 32767 ? 32767 : t);
        }
}

void or_dequant_tus(const struct ohcu_tu *tus, const struct ohcu_tu_quant *quant, int n, int16_t *arena,
                    int bd, const uint8_t *sl, const uint8_t *sl_dc)
{
    for (int i = 0; i < n; i++) {
        const int l2 = tus[i].log2_size, id = quant[i].matrix_id;
        const uint8_t *m = (id != OHCU_QM_FLAT && sl) ? sl + ((l2 - 2) * 6 + id) * 64 : 0;
        if (tus[i].mode == OHCU_TU_BYPASS) continue;          /* cu_transquant_bypass: not scaled */
        or_dequant_block(arena + tus[i].coeff_off, l2, bd, quant[i].qp, m, (m && l2 >= 4) ? sl_dc[(l2 - 4) * 6 + id] : 16);
    }
}
