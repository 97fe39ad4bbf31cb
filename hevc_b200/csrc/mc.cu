// mc.cu — batched inter prediction: luma qpel (8-tap) and chroma epel (4-tap) motion
// compensation with uni / bi / weighted prediction for every PU of a picture.
//
// Replaces, per picture, the per-block calls of hls_prediction_unit
// (hevcdec.c:2250-2306) into luma_mc_uni / luma_mc_bi / chroma_mc_uni / chroma_mc_bi
// (hevcdec.c:1744-2049) and the table entries put_hevc_{qpel,epel}{,_uni,_uni_w,_bi,_bi_w}
// (hevcdsp_template.c:2359-3357).  ff_emulated_edge_mc (videodsp_template.c:22-100) is
// folded in as a replicate clamp of the reference coordinates while staging.
//
// One CTA per CTB; its warps take (PU, plane) tasks round-robin.  A task is processed
// in chunks of <= 16 output rows: the reference window (+ filter halo) is staged in the
// warp's shared memory, filtered horizontally then vertically (14-bit int16
// intermediates exactly as the reference's tmp arrays), combined and written.
#include "ohcu_internal.h"

namespace ohcu {

// ff_hevc_qpel_filters / ff_hevc_epel_filters (hevcdsp.c:181-195), row = phase
__constant__ int8_t c_qpel[4][8] = {
    {0, 0, 0, 64, 0, 0, 0, 0}, {-1, 4, -10, 58, 17, -5, 1, 0},
    {-1, 4, -11, 40, 40, -11, 4, -1}, {0, 1, -5, 17, 58, -10, 4, -1}};
__constant__ int8_t c_epel[8][4] = {
    {0, 64, 0, 0}, {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-6, 46, 28, -4},
    {-4, 36, 36, -4}, {-4, 28, 46, -6}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};

constexpr int MC_WARPS = 4;
constexpr int MC_CHUNK = 16;               // output rows per chunk
constexpr int MC_SRC_STRIDE = 72;          // staged row stride in samples (64 + 7, even)
constexpr int MC_SRC_ROWS = MC_CHUNK + 7;

struct McWarpSmem {
    uint16_t src[MC_SRC_ROWS * MC_SRC_STRIDE];   // reference window, any depth as u16
    int16_t hbuf[MC_SRC_ROWS * OHCU_MAX_PB_SIZE]; // horizontally filtered rows (hv case)
    int16_t tmp0[MC_CHUNK * OHCU_MAX_PB_SIZE];    // list-0 intermediate (lc->tmp, stride 64)
};

__device__ __forceinline__ int pow2_lanes(int n)   // min(32, next power of two >= n)
{
    int p = 1;
    while (p < n && p < 32) p <<= 1;
    return p;
}

template <int BD, bool LUMA>
__device__ void mc_task(const FrameDev &f, const ohcu_pu &pu, int c, McWarpSmem &sm, int lane)
{
    typedef typename Px<BD>::T pixel;
    constexpr int NT = LUMA ? 8 : 4;
    constexpr int BEFORE = LUMA ? 3 : 1;
    const ohcu_slice_tab &sl = f.slices[pu.slice];
    const int hs = LUMA ? 0 : f.hs, vs = LUMA ? 0 : f.vs;
    const int bw = pu.w >> hs, bh = pu.h >> vs;
    const int xo = pu.x0 >> hs, yo = pu.y0 >> vs;
    const int pw = f.pw[c], ph = f.ph[c];
    const bool bi = pu.pred_flag == OHCU_PF_BI;
    const bool wflag = sl.weight_flag != 0;
    const int denom = LUMA ? sl.luma_log2_weight_denom : sl.chroma_log2_weight_denom;

    int wx[2], ox[2];
#pragma unroll
    for (int l = 0; l < 2; l++) {
        int ri = pu.ref_idx[l] & (OHCU_MAX_REFS - 1);
        wx[l] = LUMA ? sl.luma_weight[l][ri] : sl.chroma_weight[l][ri][c - 1];
        ox[l] = (LUMA ? sl.luma_offset[l][ri] : sl.chroma_offset[l][ri][c - 1]) * (1 << (BD - 8));
    }

    const int cpw = pow2_lanes(bw);              // lanes per output row
    const int rpi = 32 / cpw;                    // output rows per warp iteration
    const int lcol = lane & (cpw - 1), lrow = lane / cpw;

    uint8_t *dplane = f.work.p[c];
    const int dpitch = f.work.pitch[c];

    for (int cy = 0; cy < bh; cy += MC_CHUNK) {
        const int rows = min(MC_CHUNK, bh - cy);
        for (int l = 0; l < 2; l++) {
            if (!(pu.pred_flag & (1 << l))) continue;
            const int mvx = pu.mv[l][0], mvy = pu.mv[l][1];
            int mx, my, sx, sy;
            if (LUMA) {                                        // hevcdec.c:1753-1760
                mx = mvx & 3; my = mvy & 3;
                sx = xo + (mvx >> 2); sy = yo + (mvy >> 2);
            } else {                                           // hevcdec.c:1909-1915
                mx = (mvx & ((1 << (2 + hs)) - 1)) << (1 - hs);
                my = (mvy & ((1 << (2 + vs)) - 1)) << (1 - vs);
                sx = xo + (mvx >> (2 + hs)); sy = yo + (mvy >> (2 + vs));
            }
            const int8_t *fx = LUMA ? c_qpel[mx] : c_epel[mx];
            const int8_t *fy = LUMA ? c_qpel[my] : c_epel[my];
            const PlaneSet &rp = f.dpb[sl.ref_slot[l][pu.ref_idx[l] & (OHCU_MAX_REFS - 1)] & (OHCU_MAX_DPB - 1)];
            const uint8_t *ref = rp.p[c];
            const int rpitch = rp.pitch[c];

            // ---- stage the window; coordinates clamped = emulated_edge_mc
            const int hb = mx ? BEFORE : 0, vb = my ? BEFORE : 0;
            const int sw = bw + (mx ? NT - 1 : 0), sh = rows + (my ? NT - 1 : 0);
            {
                const int spw = pow2_lanes(sw), srpi = 32 / spw;
                const int scol0 = lane & (spw - 1), srow0 = lane / spw;
                for (int r = srow0; r < sh; r += srpi) {
                    const int y = clip3i(sy + cy - vb + r, 0, ph - 1);
                    const pixel *row = (const pixel *)(ref + (size_t)y * rpitch);
                    for (int col = scol0; col < sw; col += spw) {
                        const int x = clip3i(sx - hb + col, 0, pw - 1);
                        sm.src[r * MC_SRC_STRIDE + col] = __ldg(row + x);
                    }
                }
            }
            __syncwarp();

            // ---- horizontal pass of the hv case into int16 rows (tmp_array, template :2522)
            if (mx && my) {
                for (int r = lrow; r < sh; r += rpi)
                    for (int col = lcol; col < bw; col += cpw) {
                        int acc = 0;
#pragma unroll
                        for (int t = 0; t < NT; t++) acc += fx[t] * sm.src[r * MC_SRC_STRIDE + col + t];
                        sm.hbuf[r * OHCU_MAX_PB_SIZE + col] = (int16_t)(acc >> (BD - 8));
                    }
                __syncwarp();
            }

            const bool last = !bi || l == 1;
            for (int r = lrow; r < rows; r += rpi)
                for (int col = lcol; col < bw; col += cpw) {
                    int val = 0;
                    if (mx && my) {
#pragma unroll
                        for (int t = 0; t < NT; t++) val += fy[t] * sm.hbuf[(r + t) * OHCU_MAX_PB_SIZE + col];
                        val >>= 6;
                    } else if (mx) {
#pragma unroll
                        for (int t = 0; t < NT; t++) val += fx[t] * sm.src[r * MC_SRC_STRIDE + col + t];
                        val >>= (BD - 8);
                    } else if (my) {
#pragma unroll
                        for (int t = 0; t < NT; t++) val += fy[t] * sm.src[(r + t) * MC_SRC_STRIDE + col];
                        val >>= (BD - 8);
                    } else {
                        val = sm.src[r * MC_SRC_STRIDE + col] << (14 - BD);
                    }
                    if (!last) {
                        sm.tmp0[r * OHCU_MAX_PB_SIZE + col] = (int16_t)val;
                        continue;
                    }
                    int o;
                    if (!bi) {
                        if (!wflag) {                          // *_uni, template :2544-2568
                            o = (val + (1 << (13 - BD))) >> (14 - BD);
                        } else {                               // *_uni_w, template :2733-2757
                            const int shift = denom + 14 - BD;
                            o = ((val * wx[l] + (1 << (shift - 1))) >> shift) + ox[l];
                        }
                    } else {
                        const int v0 = sm.tmp0[r * OHCU_MAX_PB_SIZE + col];
                        if (!wflag) {                          // *_bi, template :2570-2594
                            o = (val + v0 + (1 << (14 - BD))) >> (15 - BD);
                        } else {                               // *_bi_w, template :2760-2790
                            const int log2Wd = denom + 14 - BD;
                            o = (val * wx[1] + v0 * wx[0] + (ox[0] + ox[1] + 1) * (1 << log2Wd)) >> (log2Wd + 1);
                        }
                    }
                    pixel *d = (pixel *)(dplane + (size_t)(yo + cy + r) * dpitch) + xo + col;
                    *d = (pixel)clip_px<BD>(o);
                }
            __syncwarp();
        }
    }
}

template <int BD>
__global__ void __launch_bounds__(MC_WARPS * 32) mc_kernel(const __grid_constant__ FrameDev f)
{
    __shared__ McWarpSmem sm[MC_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int2 range = f.ctb_pu_range[blockIdx.x];
    const int p0 = range.x;
    const int ntask = range.y * f.nplanes;
    for (int t = warp; t < ntask; t += MC_WARPS) {
        const int pi = p0 + t / f.nplanes, c = t % f.nplanes;
        const ohcu_pu pu = f.pus[pi];
        if (pu.pred_flag < 1 || pu.pred_flag > 3) continue;
        if (c == 0) mc_task<BD, true>(f, pu, 0, sm[warp], lane);
        else        mc_task<BD, false>(f, pu, c, sm[warp], lane);
    }
}

int launch_mc(const FrameDev &f, cudaStream_t st)
{
    if (f.n_pu <= 0) return 0;
    const int nctb = f.ctb_w * f.ctb_h;
    OHCU_DISPATCH_BD(f.bd, (mc_kernel<BD><<<nctb, MC_WARPS * 32, 0, st>>>(f)));
    return 1;
}

}  // namespace ohcu
