/*
 * hevc_oracle.h — CPU restatement of the reference's HEVCDSPContext leaf
 * arithmetic (libavcodec/hevcdsp_template.c) with the bit depth as a run-time
 * argument instead of a template include.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it, and only as the checker / baseline.  The shipped path is
 * hevc_b200/libohcu.so (CUDA) and never calls into this directory.
 *
 * Parity pinning: every function here is checked bit-for-bit against the
 * reference's own compiled C table (oracle/_ref/liboh_ref.so, built from
 * /root/reference by oracle/Makefile) in tests/test_oracle_vs_ref.py, and
 * against the committed fixtures in tests/golden/ that were generated from
 * that compiled reference (tests/golden/make_golden.py).
 *
 * Conventions follow the reference table: strides in BYTES, `pixel` is uint8_t
 * when bd == 8 else uint16_t, int16 intermediates have row stride 64
 * (MAX_PB_SIZE, hevcdec.h:51), coefficient blocks are contiguous N*N int16.
 */
#ifndef HEVC_ORACLE_H
#define HEVC_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* constant tables (values of hevcdsp.c:181-195 and the 32x32 DCT matrix :32-97,
 * regenerated from the cosine index rule, see hevc_oracle.c) */
extern const int8_t or_qpel_taps[4][8];   /* [mx] , row 0 = full-pel (unused) */
extern const int8_t or_epel_taps[8][4];   /* [mx] , row 0 = full-pel (unused) */
int  or_dct_coef(int row, int col);       /* transform[row][col]              */

/* ---- residual (hevcdsp_template.c:45-296) ---- */
void or_transform_add(int bd, uint8_t *dst, const int16_t *coeffs, ptrdiff_t stride, int size);
void or_transform_rdpcm(int16_t *coeffs, int log2_size, int mode);
void or_transform_skip(int bd, int16_t *coeffs, int log2_size);
void or_idct_4x4_luma(int bd, int16_t *coeffs);
void or_idct(int bd, int16_t *coeffs, int log2_size, int col_limit);
void or_idct_dc(int bd, int16_t *coeffs, int log2_size);

/* ---- SAO (hevcdsp_template.c:310-531); argument order as the table:
 *      dst first (the frame), src second (the pre-SAO copy) ---- */
void or_sao_band_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst,
                        ptrdiff_t stride_src, const int16_t offset_val[5], int band_position,
                        int width, int height);
void or_sao_edge_filter(int bd, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst,
                        ptrdiff_t stride_src, const int16_t offset_val[5], int eo_class,
                        int width, int height);
void or_sao_edge_restore(int bd, int restore_idx, uint8_t *dst, const uint8_t *src,
                         ptrdiff_t stride_dst, ptrdiff_t stride_src,
                         const int16_t offset_val[5], int eo_class, const int borders[4],
                         int width, int height, const uint8_t vert_edge[2],
                         const uint8_t horiz_edge[2], const uint8_t diag_edge[4]);

/* ---- motion compensation (hevcdsp_template.c:2359-3357).
 *      luma != 0 -> 8-tap qpel (mx,my in 0..3), else 4-tap epel (mx,my in 0..7).
 *      The [idx][!!my][!!mx] table dispatch is folded into mx/my being zero. ---- */
void or_mc_put(int bd, int luma, int16_t *dst, const uint8_t *src, ptrdiff_t srcstride,
               int height, int mx, int my, int width);
void or_mc_uni(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
               ptrdiff_t srcstride, int height, int mx, int my, int width);
void or_mc_uni_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                 ptrdiff_t srcstride, int height, int denom, int wx, int ox,
                 int mx, int my, int width);
void or_mc_bi(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
              ptrdiff_t srcstride, const int16_t *src2, int height, int mx, int my, int width);
void or_mc_bi_w(int bd, int luma, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src,
                ptrdiff_t srcstride, const int16_t *src2, int height, int denom,
                int wx0, int wx1, int ox0, int ox1, int mx, int my, int width);

/* ---- deblocking (hevcdsp_template.c:3377-3553).  vertical != 0 filters a
 *      vertical edge (hevc_v_loop_filter_*), else a horizontal one. ---- */
void or_loop_filter_luma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride, int beta,
                         const int tc[2], const uint8_t no_p[2], const uint8_t no_q[2]);
void or_loop_filter_chroma(int bd, int vertical, uint8_t *pix, ptrdiff_t stride,
                           const int tc[2], const uint8_t no_p[2], const uint8_t no_q[2]);

/* ---- dequantisation (hevc_cabac.c:1470-1557, :2030-2055): the device-side step of
 *      ohcu_append_tu_levels.  blk: N x N un-scaled levels, scaled in place (zeros stay
 *      zero, as the reference only scales the levels it decodes).  matrix: the 64-entry
 *      scaling list of the block's size and matrix id, or NULL for the flat 16; dc_scale
 *      is used for the DC of 16x16 / 32x32 blocks when matrix != NULL. ---- */
void or_dequant_block(int16_t *blk, int log2_size, int bd, int qp, const uint8_t *matrix, int dc_scale);
/* every block of a TU list whose records carry arena offsets (dense layout): what the
 * device does for a levels frame.  sl = ScalingList.sl[4][6][64], sl_dc = .sl_dc[2][6]
 * (hevc_ps.h:605-610), both may be NULL when every matrix_id is OHCU_QM_FLAT. */
struct ohcu_tu; struct ohcu_tu_quant;
void or_dequant_tus(const struct ohcu_tu *tus, const struct ohcu_tu_quant *quant, int n, int16_t *arena,
                    int bd, const uint8_t *sl, const uint8_t *sl_dc);

#ifdef __cplusplus
}
#endif
/* ---- intra prediction (intra_oracle.c; hevcpred_template.c) */
#define OR_CAND_BOTTOM_LEFT 1
#define OR_CAND_LEFT        2
#define OR_CAND_UP_LEFT     4
#define OR_CAND_UP          8
#define OR_CAND_UP_RIGHT   16
void or_intra_pred(int bd, uint8_t *plane, ptrdiff_t stride, int pw, int ph, int x, int y, int log2_size, int c_idx,
                   int mode, int cand, int filter_ok, int strong);

#endif
