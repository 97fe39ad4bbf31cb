// deblock.cu — picture-level deblocking, one pass, edge-parallel and wavefront-free.
//
// Replaces deblocking_filter_CTB (hevc_filter.c:525-762) and the table entries
// hevc_{h,v}_loop_filter_{luma,chroma}(_c) (hevcdsp_template.c:3377-3553).
//
// HEVC filters edges on an 8x8 sample grid, reads at most 4 and modifies at most 3
// samples on each side, and takes its decisions per 4-line segment.  Hence the 8x8
// blocks CENTRED on the grid crossings, [8i-4, 8i+4) x [8j-4, 8j+4), are independent:
// each contains one 8-sample piece of a vertical edge (two segments) and one of a
// horizontal edge, and everything those read or write.  One thread owns one such block
// of one plane: load 8x8, filter the vertical edge, then the horizontal edge on the
// result (the standard's order: all vertical edges of the picture before horizontal
// ones), store.  tc / beta / pcm flags are derived on the device from the bs planes,
// qp_y_tab and the per-CTB offsets exactly as the reference driver does, including its
// choice of which CTB's offsets apply (hevc_filter.c:666, 699, 712, 721).
#include "ohcu_internal.h"

namespace ohcu {

__constant__ uint8_t c_tctable[54] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3,
    3, 3, 3, 4, 4, 4, 5, 5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 22, 24};
__constant__ uint8_t c_betatable[52] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18,
    20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64};
__constant__ uint8_t c_qpc420[14] = {29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37};

__device__ __forceinline__ int qpy_at(const FrameDev &f, int x, int y)
{
    return f.qp_y[(x >> f.log2_min_cb) + (y >> f.log2_min_cb) * f.min_cb_w];
}
__device__ __forceinline__ int pcm_at(const FrameDev &f, int x, int y)     // get_pcm
{
    if (x < 0 || y < 0) return 2;
    const int xp = x >> f.log2_min_pu, yp = y >> f.log2_min_pu;
    if (xp >= f.min_pu_w || yp >= f.min_pu_h) return 2;
    return f.is_pcm ? f.is_pcm[yp * f.min_pu_w + xp] : 0;
}
__device__ __forceinline__ int ctb_at(const FrameDev &f, int x, int y)
{
    return (x >> f.log2_ctb) + (y >> f.log2_ctb) * f.ctb_w;
}
__device__ __forceinline__ int chroma_tc(const FrameDev &f, int qp_y, int c, int tc_offset)
{
    const int qi = clip3i(qp_y + (c == 1 ? f.cb_qp_offset : f.cr_qp_offset), 0, 57);
    int qp;
    if (f.cfi == 1) qp = qi < 30 ? qi : (qi > 43 ? qi - 6 : c_qpc420[qi - 30]);
    else qp = min(qi, 51);
    return c_tctable[clip3i(qp + 2 + tc_offset, 0, 53)];
}

// one 4-line luma segment: l[line][0..7] = P3 P2 P1 P0 Q0 Q1 Q2 Q3 (template :3390-3470)
template <int BD>
__device__ __forceinline__ void luma_segment(int (&l)[4][8], int beta, int tc, int no_p, int no_q)
{
    beta <<= BD - 8;
    tc <<= BD - 8;
    const int dp0 = abs(l[0][1] - 2 * l[0][2] + l[0][3]), dq0 = abs(l[0][6] - 2 * l[0][5] + l[0][4]);
    const int dp3 = abs(l[3][1] - 2 * l[3][2] + l[3][3]), dq3 = abs(l[3][6] - 2 * l[3][5] + l[3][4]);
    const int d0 = dp0 + dq0, d3 = dp3 + dq3;
    if (d0 + d3 >= beta) return;
    const int beta_3 = beta >> 3, beta_2 = beta >> 2, tc25 = (tc * 5 + 1) >> 1;
    const bool strong =
        abs(l[0][0] - l[0][3]) + abs(l[0][7] - l[0][4]) < beta_3 && abs(l[0][3] - l[0][4]) < tc25 &&
        abs(l[3][0] - l[3][3]) + abs(l[3][7] - l[3][4]) < beta_3 && abs(l[3][3] - l[3][4]) < tc25 &&
        (d0 << 1) < beta_2 && (d3 << 1) < beta_2;
    if (strong) {
        const int tc2 = tc << 1;
#pragma unroll
        for (int d = 0; d < 4; d++) {
            const int p3 = l[d][0], p2 = l[d][1], p1 = l[d][2], p0 = l[d][3];
            const int q0 = l[d][4], q1 = l[d][5], q2 = l[d][6], q3 = l[d][7];
            if (!no_p) {
                l[d][3] = p0 + clip3i(((p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3) - p0, -tc2, tc2);
                l[d][2] = p1 + clip3i(((p2 + p1 + p0 + q0 + 2) >> 2) - p1, -tc2, tc2);
                l[d][1] = p2 + clip3i(((2 * p3 + 3 * p2 + p1 + p0 + q0 + 4) >> 3) - p2, -tc2, tc2);
            }
            if (!no_q) {
                l[d][4] = q0 + clip3i(((p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3) - q0, -tc2, tc2);
                l[d][5] = q1 + clip3i(((p0 + q0 + q1 + q2 + 2) >> 2) - q1, -tc2, tc2);
                l[d][6] = q2 + clip3i(((2 * q3 + 3 * q2 + q1 + q0 + p0 + 4) >> 3) - q2, -tc2, tc2);
            }
        }
    } else {
        const int side = (beta + (beta >> 1)) >> 3, tc_2 = tc >> 1;
        const bool nd_p2 = dp0 + dp3 < side, nd_q2 = dq0 + dq3 < side;
#pragma unroll
        for (int d = 0; d < 4; d++) {
            const int p2 = l[d][1], p1 = l[d][2], p0 = l[d][3];
            const int q0 = l[d][4], q1 = l[d][5], q2 = l[d][6];
            int delta0 = (9 * (q0 - p0) - 3 * (q1 - p1) + 8) >> 4;
            if (abs(delta0) < 10 * tc) {
                delta0 = clip3i(delta0, -tc, tc);
                if (!no_p) l[d][3] = clip_px<BD>(p0 + delta0);
                if (!no_q) l[d][4] = clip_px<BD>(q0 - delta0);
                if (!no_p && nd_p2)
                    l[d][2] = clip_px<BD>(p1 + clip3i((((p2 + p0 + 1) >> 1) - p1 + delta0) >> 1, -tc_2, tc_2));
                if (!no_q && nd_q2)
                    l[d][5] = clip_px<BD>(q1 + clip3i((((q2 + q0 + 1) >> 1) - q1 - delta0) >> 1, -tc_2, tc_2));
            }
        }
    }
}

// one 4-line chroma segment: l[line][2..5] = P1 P0 Q0 Q1 (template :3473-3505)
template <int BD>
__device__ __forceinline__ void chroma_segment(int (&l)[4][8], int tc, int no_p, int no_q)
{
    tc <<= BD - 8;
    if (tc <= 0) return;
#pragma unroll
    for (int d = 0; d < 4; d++) {
        const int p1 = l[d][2], p0 = l[d][3], q0 = l[d][4], q1 = l[d][5];
        const int delta0 = clip3i((((q0 - p0) * 4) + p1 - q1 + 4) >> 3, -tc, tc);
        if (!no_p) l[d][3] = clip_px<BD>(p0 + delta0);
        if (!no_q) l[d][4] = clip_px<BD>(q0 - delta0);
    }
}

// Filter parameters of one 4-line segment.
// vertical edge at luma column x, segment rows [ys, ys+4): unit row yu = ys & ~(8v-1) ...
struct SegParam { int beta, tc, no_p, no_q; };

template <int BD>
__global__ void __launch_bounds__(128) deblock_kernel(const __grid_constant__ FrameDev f, int c)
{
    typedef typename Px<BD>::T pixel;
    const int pw = f.pw[c], ph = f.ph[c];
    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
    const int nbx = (pw + 11) >> 3, nby = (ph + 11) >> 3;
    const int bx = blockIdx.x * 32 + (threadIdx.x & 31);
    const int by = blockIdx.y * 4 + (threadIdx.x >> 5);
    if (bx >= nbx || by >= nby) return;
    const int X0 = 8 * bx - 4, Y0 = 8 * by - 4;         // plane coordinates of the block
    const int ex = 8 * bx, ey = 8 * by;                   // the grid crossing
    // luma-unit coordinates of the edges / segment starts
    const int lx = ex << hs, ly = ey << vs;
    const int W = f.width, H = f.height;
    const bool has_v = bx > 0 && ex < pw, has_h = by > 0 && ey < ph;

    // ---- parameters (before touching pixels, so that blocks with nothing to do exit early)
    SegParam vp[2], hp[2];
#pragma unroll
    for (int s = 0; s < 2; s++) { vp[s] = SegParam{0, 0, 0, 0}; hp[s] = SegParam{0, 0, 0, 0}; }
    bool any = false;
    if (has_v) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
            // segment rows (plane) [Y0 + 4s, Y0 + 4s + 4) -> luma row of its first line
            const int ysl = (Y0 + 4 * s) << vs;
            if (ysl < 0 || ysl >= H) continue;
            const int bs = f.vbs[(lx + ysl * f.bs_width) >> 2];
            const int cb = ctb_at(f, lx, ysl);
            if (c == 0) {
                if (!bs) continue;
                const int yu = ysl & ~7;                                   // 8-row unit, qp at its first row (:571)
                const int qp = (qpy_at(f, lx - 1, yu) + qpy_at(f, lx, yu) + 1) >> 1;
                vp[s].beta = c_betatable[clip3i(qp + f.db[cb].beta_offset, 0, 51)];
                const int tco = f.db[cb].tc_offset;
                vp[s].tc = c_tctable[clip3i(qp + 2 * (bs - 1) + (tco >> 1 << 1), 0, 53)];
            } else {
                if (bs != 2) continue;
                const int qp = (qpy_at(f, lx - 1, ysl) + qpy_at(f, lx, ysl) + 1) >> 1;   // per segment (:614-615)
                vp[s].tc = chroma_tc(f, qp, c, f.db[cb].tc_offset);
            }
            if (f.pcm_or_bypass) { vp[s].no_p = pcm_at(f, lx - 1, ysl); vp[s].no_q = pcm_at(f, lx, ysl); }
            any = true;
        }
    }
    if (has_h) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
            const int xsl = (X0 + 4 * s) << hs;             // luma column of the segment's first line
            if (xsl < 0 || xsl >= W) continue;
            const int bs = f.hbs[(xsl + ly * f.bs_width) >> 2];
            const int unit = 8 << hs;                        // edge unit length in luma columns
            const int xu = xsl & ~(unit - 1);
            // CTB that runs this unit in the reference: the one owning column xu + unit,
            // the last CTB column taking the remainder (hevc_filter.c:661-667, 715-718)
            const int xproc = (xu + unit < W) ? xu + unit : xu;
            const int cb_proc = ctb_at(f, xproc, ly), cb_here = ctb_at(f, xu, ly);
            if (c == 0) {
                if (!bs) continue;
                const int qp = (qpy_at(f, xu, ly - 1) + qpy_at(f, xu, ly) + 1) >> 1;
                hp[s].beta = c_betatable[clip3i(qp + f.db[cb_here].beta_offset, 0, 51)];
                const int tco = f.db[cb_proc].tc_offset;
                hp[s].tc = c_tctable[clip3i(qp + 2 * (bs - 1) + (tco >> 1 << 1), 0, 53)];
            } else {
                if (bs != 2) continue;
                const int qp = (qpy_at(f, xsl, ly - 1) + qpy_at(f, xsl, ly) + 1) >> 1;
                const bool first = xsl == xu;                // first segment of the unit
                hp[s].tc = chroma_tc(f, qp, c, f.db[first ? cb_here : cb_proc].tc_offset);
            }
            if (f.pcm_or_bypass) { hp[s].no_p = pcm_at(f, xsl, ly - 1); hp[s].no_q = pcm_at(f, xsl, ly); }
            any = true;
        }
    }
    if (!any) return;

    // ---- load the block (only in-picture samples exist)
    uint8_t *plane = f.work.p[c];
    const int pitch = f.work.pitch[c];
    int px[8][8];
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const int y = Y0 + r;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int x = X0 + k;
            px[r][k] = (y >= 0 && y < ph && x >= 0 && x < pw) ? ((const pixel *)(plane + (size_t)y * pitch))[x] : 0;
        }
    }
    // ---- vertical edge: lines are rows
    if (has_v) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
            if (vp[s].tc == 0 && c != 0) continue;
            int l[4][8];
#pragma unroll
            for (int d = 0; d < 4; d++)
#pragma unroll
                for (int k = 0; k < 8; k++) l[d][k] = px[4 * s + d][k];
            if (c == 0) { if (vp[s].tc || vp[s].beta) luma_segment<BD>(l, vp[s].beta, vp[s].tc, vp[s].no_p, vp[s].no_q); }
            else chroma_segment<BD>(l, vp[s].tc, vp[s].no_p, vp[s].no_q);
#pragma unroll
            for (int d = 0; d < 4; d++)
#pragma unroll
                for (int k = 0; k < 8; k++) px[4 * s + d][k] = l[d][k];
        }
    }
    // ---- horizontal edge: lines are columns
    if (has_h) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
            if (hp[s].tc == 0 && c != 0) continue;
            int l[4][8];
#pragma unroll
            for (int d = 0; d < 4; d++)
#pragma unroll
                for (int k = 0; k < 8; k++) l[d][k] = px[k][4 * s + d];
            if (c == 0) { if (hp[s].tc || hp[s].beta) luma_segment<BD>(l, hp[s].beta, hp[s].tc, hp[s].no_p, hp[s].no_q); }
            else chroma_segment<BD>(l, hp[s].tc, hp[s].no_p, hp[s].no_q);
#pragma unroll
            for (int d = 0; d < 4; d++)
#pragma unroll
                for (int k = 0; k < 8; k++) px[k][4 * s + d] = l[d][k];
        }
    }
    // ---- store
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const int y = Y0 + r;
        if (y < 0 || y >= ph) continue;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int x = X0 + k;
            if (x >= 0 && x < pw) ((pixel *)(plane + (size_t)y * pitch))[x] = (pixel)px[r][k];
        }
    }
}

int launch_deblock(const FrameDev &f, cudaStream_t st)
{
    if (!f.vbs) return 0;
    int n = 0;
    for (int c = 0; c < f.nplanes; c++) {
        const int nbx = (f.pw[c] + 11) >> 3, nby = (f.ph[c] + 11) >> 3;
        dim3 grid((nbx + 31) / 32, (nby + 3) / 4);
        OHCU_DISPATCH_BD(f.bd, (deblock_kernel<BD><<<grid, 128, 0, st>>>(f, c)));
        n++;
    }
    return n;
}

}  // namespace ohcu
