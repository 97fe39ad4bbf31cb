/*
 * intra_oracle.c — CPU restatement of the reference's intra prediction.
 * TEST INFRASTRUCTURE: only tests/ and the checkers link it; the product never does.
 *
 * Follows libavcodec/hevcpred_template.c:
 *   reference samples + substitution of unavailable neighbours      intra_pred :30-272
 *   [1 2 1] / strong (bilinear) smoothing of the reference samples  :274-326
 *   pred_planar :359-372, pred_dc :388-417, pred_angular :419-505
 * with constrained_intra_pred_flag == 0 (the CIP substitution of :112-150, :172-243 is not restated; the
 * reference's own function is the checker there).  Availability arrives as the five final cand_* flags the
 * reference computes at :103-107 (lc->na.* combined with the z-scan order test).
 * Pinned against the reference's compiled ff_hevc_pred_init table through oracle/ref_shim_pred.c
 * (tests/test_intra.py), every mode x size x availability pattern x bit depth.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "hevc_oracle.h"

static inline int ipx_ld(const uint8_t *base, ptrdiff_t idx, int bd) { return bd == 8 ? base[idx] : ((const uint16_t *)base)[idx]; }
static inline void ipx_st(uint8_t *base, ptrdiff_t idx, int bd, int v)
{
    if (bd == 8) base[idx] = (uint8_t)v;
    else ((uint16_t *)base)[idx] = (uint16_t)v;
}
static inline int iclipp(int v, int bd) { return v < 0 ? 0 : (v > (1 << bd) - 1 ? (1 << bd) - 1 : v); }
static inline int iabs(int v) { return v < 0 ? -v : v; }

/* intra_pred_angle[] / inv_angle[], template :428-435 */
static const int8_t k_angle[33] = { 32, 26, 21, 17, 13, 9, 5, 2, 0, -2, -5, -9, -13, -17, -21, -26, -32,
                                    -26, -21, -17, -13, -9, -5, -2, 0, 2, 5, 9, 13, 17, 21, 26, 32 };
static const int16_t k_inv_angle[15] = { -4096, -1638, -910, -630, -482, -390, -315, -256, -315, -390, -482, -630, -910, -1638, -4096 };

/* One N x N block of plane c_idx predicted in place.
 *   plane / stride (bytes) / pw, ph : the plane being reconstructed (neighbours are read from it)
 *   x, y          : block position in plane samples
 *   cand          : OR_CAND_* bits, the final availability of the five neighbour groups
 *   filter_ok     : !sps->intra_smoothing_disabled_flag && (c_idx == 0 || chroma_format_idc == 3)   (:275)
 *   strong        : sps->sps_strong_intra_smoothing_enable_flag                                     (:282) */
void or_intra_pred(int bd, uint8_t *plane, ptrdiff_t stride, int pw, int ph, int x, int y, int log2_size, int c_idx,
                   int mode, int cand, int filter_ok, int strong)
{
    const int n = 1 << log2_size, ps = (int)(stride / (bd == 8 ? 1 : 2));
    int L[2 * 32 + 1], T[2 * 32 + 1], FL[2 * 32 + 1], FT[2 * 32 + 1];
    int *left = L + 1, *top = T + 1;                       /* left[-1] == top[-1] == the corner */
    int bl = cand & OR_CAND_BOTTOM_LEFT, lf = cand & OR_CAND_LEFT, ul = cand & OR_CAND_UP_LEFT, up = cand & OR_CAND_UP,
        ur = cand & OR_CAND_UP_RIGHT;
    /* rows / columns of the second half that exist in the picture (:109-112) */
    int bl_size = (y + 2 * n < ph ? y + 2 * n : ph) - (y + n);
    int tr_size = (x + 2 * n < pw ? x + 2 * n : pw) - (x + n);
    const ptrdiff_t o = (ptrdiff_t)y * ps + x;
#define PX(dx, dy) ipx_ld(plane, o + (ptrdiff_t)(dy) * ps + (dx), bd)
    if (bl_size < 0) bl_size = 0;
    if (tr_size < 0) tr_size = 0;
    for (int i = 0; i <= 2 * n; i++) L[i] = T[i] = 0;

    /* ---- what exists (:152-171) */
    if (ul) left[-1] = top[-1] = PX(-1, -1);
    if (up) for (int i = 0; i < n; i++) top[i] = PX(i, -1);
    if (ur) {
        for (int i = 0; i < tr_size; i++) top[n + i] = PX(n + i, -1);
        for (int i = tr_size; i < n; i++) top[n + i] = PX(n + tr_size - 1, -1);
    }
    if (lf) for (int i = 0; i < n; i++) left[i] = PX(-1, i);
    if (bl) {
        for (int i = 0; i < bl_size; i++) left[n + i] = PX(-1, n + i);
        for (int i = bl_size; i < n; i++) left[n + i] = PX(-1, n + bl_size - 1);
    }
    /* ---- substitution: walk from the bottom-left end towards the top-right end, every missing group takes the
     *      nearest sample already known (:245-272) */
    if (!bl) {
        if (lf) {
            for (int i = n; i < 2 * n; i++) left[i] = left[n - 1];
        } else {
            int v;
            if (ul) v = left[-1];
            else if (up) { v = top[0]; ul = 1; }
            else if (ur) { v = top[n]; for (int i = 0; i < n; i++) top[i] = v; up = ul = 1; }
            else { v = 1 << (bd - 1); for (int i = 0; i < 2 * n; i++) top[i] = v; }
            left[-1] = v;
            for (int i = 0; i < 2 * n; i++) left[i] = v;
            if (cand & (OR_CAND_UP_LEFT | OR_CAND_UP | OR_CAND_UP_RIGHT)) lf = 1;
        }
    }
    if (!lf) for (int i = 0; i < n; i++) left[i] = left[n];
    if (!ul) left[-1] = left[0];
    if (!up) for (int i = 0; i < n; i++) top[i] = left[-1];
    if (!ur) for (int i = n; i < 2 * n; i++) top[i] = top[n - 1];
    top[-1] = left[-1];

    /* ---- smoothing of the reference samples (:274-326) */
    if (filter_ok && mode != 1 && n != 4) {
        static const int thresh[3] = { 7, 1, 0 };
        const int d26 = iabs(mode - 26), d10 = iabs(mode - 10);
        if ((d26 < d10 ? d26 : d10) > thresh[log2_size - 3]) {
            int *fl = FL + 1, *ft = FT + 1;
            if (strong && c_idx == 0 && log2_size == 5 && iabs(top[-1] + top[63] - 2 * top[31]) < (1 << (bd - 5)) &&
                iabs(left[-1] + left[63] - 2 * left[31]) < (1 << (bd - 5))) {
                fl[-1] = ft[-1] = top[-1];
                ft[63] = top[63]; fl[63] = left[63];
                for (int i = 0; i < 63; i++) {
                    ft[i] = ((63 - i) * top[-1] + (i + 1) * top[63] + 32) >> 6;
                    fl[i] = ((63 - i) * left[-1] + (i + 1) * left[63] + 32) >> 6;
                }
            } else {
                fl[2 * n - 1] = left[2 * n - 1];
                ft[2 * n - 1] = top[2 * n - 1];
                for (int i = 2 * n - 2; i >= 0; i--) {
                    fl[i] = (left[i + 1] + 2 * left[i] + left[i - 1] + 2) >> 2;
                    ft[i] = (top[i + 1] + 2 * top[i] + top[i - 1] + 2) >> 2;
                }
                fl[-1] = ft[-1] = (left[0] + 2 * left[-1] + top[0] + 2) >> 2;
            }
            left = fl; top = ft;
        }
    }

    /* ---- the three predictors */
#define OUT(xx, yy, v) ipx_st(plane, o + (ptrdiff_t)(yy) * ps + (xx), bd, (v))
    if (mode == 0) {                                       /* planar, :359-372 */
        for (int yy = 0; yy < n; yy++)
            for (int xx = 0; xx < n; xx++)
                OUT(xx, yy, ((n - 1 - xx) * left[yy] + (xx + 1) * top[n] + (n - 1 - yy) * top[xx] + (yy + 1) * left[n] + n) >> (log2_size + 1));
    } else if (mode == 1) {                                /* DC, :388-417 */
        int dc = n;
        for (int i = 0; i < n; i++) dc += left[i] + top[i];
        dc >>= log2_size + 1;
        for (int yy = 0; yy < n; yy++)
            for (int xx = 0; xx < n; xx++) OUT(xx, yy, dc);
        if (c_idx == 0 && n < 32) {
            OUT(0, 0, (left[0] + 2 * dc + top[0] + 2) >> 2);
            for (int i = 1; i < n; i++) { OUT(i, 0, (top[i] + 3 * dc + 2) >> 2); OUT(0, i, (left[i] + 3 * dc + 2) >> 2); }
        }
    } else {                                               /* angular 2..34, :419-505 */
        const int angle = k_angle[mode - 2], vertical = mode >= 18;
        const int *main_ = vertical ? top : left, *side = vertical ? left : top;
        int R[3 * 32 + 4], *ref = R + 32;                 /* ref[k] = main[k - 1], extended to negative k from the side */
        const int last = (n * angle) >> 5;
        for (int k = 0; k <= 2 * n; k++) ref[k] = main_[k - 1];
        if (angle < 0 && last < -1)
            for (int k = last; k <= -1; k++) ref[k] = side[-1 + ((k * k_inv_angle[mode - 11] + 128) >> 8)];
        for (int a = 0; a < n; a++) {                      /* a runs along the prediction direction's own axis */
            const int idx = ((a + 1) * angle) >> 5, fact = ((a + 1) * angle) & 31;
            for (int b = 0; b < n; b++) {
                const int v = fact ? ((32 - fact) * ref[b + idx + 1] + fact * ref[b + idx + 2] + 16) >> 5 : ref[b + idx + 1];
                if (vertical) OUT(b, a, v); else OUT(a, b, v);
            }
        }
        if (c_idx == 0 && n < 32) {                        /* edge filters of the pure vertical / horizontal modes */
            if (mode == 26) for (int i = 0; i < n; i++) OUT(0, i, iclipp(top[0] + ((left[i] - left[-1]) >> 1), bd));
            if (mode == 10) for (int i = 0; i < n; i++) OUT(i, 0, iclipp(left[0] + ((top[i] - top[-1]) >> 1), bd));
        }
    }
#undef OUT
#undef PX
}
