// recon.cu — fused inter reconstruction of one picture: motion compensation + residual.
//
// One CTA owns one CTB.  It builds the CTB's prediction in a shared-memory tile (every
// inter PU: luma qpel 8-tap / chroma epel 4-tap, uni / bi / weighted), adds the residual
// of every coded TB of the CTB into the same tile (inverse DCT 4..32 / DC / 4x4 DST /
// transform_skip / rdpcm / cross-component), and writes the CTB to the picture once with
// full-line stores.  The prediction is never written to HBM and read back (the
// transform_add RMW of the reference disappears).
//
// Replaces, per picture:
//   hls_prediction_unit -> luma_mc_uni/bi, chroma_mc_uni/bi     hevcdec.c:1744-2049, 2250-2306
//   put_hevc_{qpel,epel}{,_uni,_uni_w,_bi,_bi_w}[idx][my][mx]   hevcdsp_template.c:2359-3357
//   ff_emulated_edge_mc (replicate clamp while staging)         videodsp_template.c:22-100
//   ff_hevc_hls_transform / _c + transform_add                  hevc_cabac.c:1695-1906,
//                                                               hevcdsp_template.c:45-296
//
// Arithmetic.  Samples are staged as 16-bit lanes, two per 32-bit word, so the FIR taps
// run on the integer dot-product pipe: IDP.2A (dp2a) does two 16x8-bit MACs with int32
// accumulation per instruction, at the same issue rate as IMAD on sm_100a
// (profiles/r01_ubench_int_pipes.txt).  Horizontal pairs come straight from the staged
// words (+ one funnel shift for odd phases), vertical pairs are built with PRMT.  The
// int16 intermediates, shifts and clips are exactly the reference's.
#include "ohcu_internal.h"

namespace ohcu {

int upload_dct_table();

constexpr int RC_WARPS = 8;
constexpr int RC_THREADS = RC_WARPS * 32;
constexpr int MC_CHUNK = 16;                 // output rows per task
constexpr int SRC_ROWS = MC_CHUNK + 7;       // staged rows (8-tap halo)
constexpr int SRC_PITCH = 36;                // words per staged row: 72 samples, == 4 mod 32
constexpr int HB_PITCH = 36;                 // words per row of the horizontal-pass output
constexpr int MAX_CTB_TUS = 768;             // 4x4 TBs of a 64x64 CTB, 3 full-size planes

struct WarpScratch {
    uint32_t src[SRC_ROWS * SRC_PITCH];      // reference window, 2 samples per word
    uint32_t hb[SRC_ROWS * HB_PITCH];        // int16 rows after the horizontal pass
};

// ff_hevc_qpel_filters / ff_hevc_epel_filters (hevcdsp.c:181-195) packed four int8 taps
// per word: [phase][0] = taps 0..3, [phase][1] = taps 4..7
__constant__ uint32_t c_qpel4[4][2];
__constant__ uint32_t c_epel4[8][2];

int upload_mc_taps()
{
    static const int8_t q[4][8] = {{0, 0, 0, 64, 0, 0, 0, 0}, {-1, 4, -10, 58, 17, -5, 1, 0},
                                   {-1, 4, -11, 40, 40, -11, 4, -1}, {0, 1, -5, 17, 58, -10, 4, -1}};
    static const int8_t e[8][4] = {{0, 64, 0, 0}, {-2, 58, 10, -2}, {-4, 54, 16, -2}, {-6, 46, 28, -4},
                                   {-4, 36, 36, -4}, {-4, 28, 46, -6}, {-2, 16, 54, -4}, {-2, 10, 58, -2}};
    uint32_t qp[4][2], ep[8][2];
    for (int p = 0; p < 4; p++)
        for (int h = 0; h < 2; h++) {
            uint32_t v = 0;
            for (int k = 0; k < 4; k++) v |= (uint32_t)(uint8_t)q[p][4 * h + k] << (8 * k);
            qp[p][h] = v;
        }
    for (int p = 0; p < 8; p++) {
        uint32_t v = 0;
        for (int k = 0; k < 4; k++) v |= (uint32_t)(uint8_t)e[p][k] << (8 * k);
        ep[p][0] = v; ep[p][1] = 0;
    }
    if (cudaMemcpyToSymbol(c_qpel4, qp, sizeof(qp)) != cudaSuccess) return -1;
    if (cudaMemcpyToSymbol(c_epel4, ep, sizeof(ep)) != cudaSuccess) return -1;
    return 0;
}

// ---------------------------------------------------------------- dot-product helpers
__device__ __forceinline__ int dp2lo(uint32_t a, uint32_t b, int c) { return __dp2a_lo((int)a, (int)b, c); }
__device__ __forceinline__ int dp2hi(uint32_t a, uint32_t b, int c) { return __dp2a_hi((int)a, (int)b, c); }

// NT-tap FIR over the 16 samples held in W[0..7] (sample 2j in the low half of W[j]):
// out[i] = sum_k tap[k] * sample[i + k], i = 0..7
template <int NT>
__device__ __forceinline__ void fir_row(const uint32_t (&W)[8], uint32_t F0, uint32_t F1, int (&out)[8])
{
    uint32_t S[7];
#pragma unroll
    for (int j = 0; j < 7; j++) S[j] = __funnelshift_r(W[j], W[j + 1], 16);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if (NT == 8) {
            out[2 * j] = dp2hi(W[j + 3], F1, dp2lo(W[j + 2], F1, dp2hi(W[j + 1], F0, dp2lo(W[j], F0, 0))));
            out[2 * j + 1] = dp2hi(S[j + 3], F1, dp2lo(S[j + 2], F1, dp2hi(S[j + 1], F0, dp2lo(S[j], F0, 0))));
        } else {
            out[2 * j] = dp2hi(W[j + 1], F0, dp2lo(W[j], F0, 0));
            out[2 * j + 1] = dp2hi(S[j + 1], F0, dp2lo(S[j], F0, 0));
        }
    }
}

// NT-tap FIR down two columns at once: V[i] holds the samples of rows i (two columns per
// word); outA/outB[r] = sum_k tap[k] * column{A,B}[r + k], r = 0..R-1
template <int NT, int R>
__device__ __forceinline__ void fir_col(const uint32_t (&V)[R + NT - 1], uint32_t F0, uint32_t F1,
                                        int (&outA)[R], int (&outB)[R])
{
    uint32_t PA[R + NT - 2], PB[R + NT - 2];
#pragma unroll
    for (int i = 0; i < R + NT - 2; i++) {
        PA[i] = __byte_perm(V[i], V[i + 1], 0x5410);
        PB[i] = __byte_perm(V[i], V[i + 1], 0x7632);
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
        if (NT == 8) {
            outA[r] = dp2hi(PA[r + 6], F1, dp2lo(PA[r + 4], F1, dp2hi(PA[r + 2], F0, dp2lo(PA[r], F0, 0))));
            outB[r] = dp2hi(PB[r + 6], F1, dp2lo(PB[r + 4], F1, dp2hi(PB[r + 2], F0, dp2lo(PB[r], F0, 0))));
        } else {
            outA[r] = dp2hi(PA[r + 2], F0, dp2lo(PA[r], F0, 0));
            outB[r] = dp2hi(PB[r + 2], F0, dp2lo(PB[r], F0, 0));
        }
    }
}

// ---------------------------------------------------------------- prediction epilogue
// All five table families reduce to  o = ((val*wA + tmp*wB + add) >> sh) + post  on the
// 14-bit intermediate `val` of this list and `tmp` of list 0 (bi only):
//   uni     (template :2544)  wA 1          add 1<<(13-BD)            sh 14-BD      post 0
//   uni_w   (template :2733)  wA wx         add 1<<(shift-1)          sh shift      post ox
//   bi      (template :2570)  wA 1   wB 1   add 1<<(14-BD)            sh 15-BD      post 0
//   bi_w    (template :2760)  wA wx1 wB wx0 add (ox0+ox1+1)<<log2Wd   sh log2Wd+1   post 0
// store_tmp: list 0 of a bi PU parks its intermediate in the tile (the reference's
// lc->tmp, hevcdec.c:1854), list 1 picks it up.
struct Epi {
    int wA, wB, add, sh, post, maxv;
    bool store_tmp, bi;
};

__device__ __forceinline__ void epi_pair(const Epi &e, int va, int vb, uint32_t *tw)
{
    if (e.store_tmp) {
        *tw = __byte_perm((uint32_t)va, (uint32_t)vb, 0x5410);
        return;
    }
    int ta = 0, tb = 0;
    if (e.bi) {
        const uint32_t t = *tw;
        ta = (int)(int16_t)(t & 0xffffu);
        tb = (int)t >> 16;
    }
    int oa = ((va * e.wA + ta * e.wB + e.add) >> e.sh) + e.post;
    int ob = ((vb * e.wA + tb * e.wB + e.add) >> e.sh) + e.post;
    oa = min(max(oa, 0), e.maxv);
    ob = min(max(ob, 0), e.maxv);
    *tw = (uint32_t)oa | ((uint32_t)ob << 16);
}

__device__ __forceinline__ int fastdiv(int n, int inv) { return (n * inv) >> 16; }   // n < 4096, divisor <= 32
__constant__ int c_inv[41] = {0, 65536, 32768, 21846, 16384, 13108, 10923, 9363, 8192, 7282, 6554, 5958, 5462, 5042,
                              4682, 4370, 4096, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428,
                              2341, 2260, 2185, 2115, 2048, 1986, 1928, 1873, 1821, 1772, 1725, 1681, 1639};
__device__ __forceinline__ int fastinv(int d) { return c_inv[d]; }                   // ceil(65536 / d), d <= 40

// ---------------------------------------------------------------- reference window staging
// Copies the window [x0, x0+cols) x [y0, y0+rows) of a reference plane into src[] as
// 16-bit samples, two per word, window column 0 in the low half of word 0 of each row.
// Coordinates outside the plane replicate the border sample = ff_emulated_edge_mc
// (videodsp_template.c:22-100).  Windows that do not touch the left/right border take the
// word path: one (HIGH: or two) aligned 32-bit loads per lane, realigned with a funnel shift.
template <bool HIGH>
__device__ __forceinline__ void stage_window(uint32_t *src, const uint8_t *__restrict__ ref, int rpitch, int pw,
                                             int ph, int x0, int y0, int cols, int rows, int lane)
{
    const int nw = (cols + 1) >> 1;                       // smem words per row
    if (x0 >= 0 && x0 + cols <= pw) {
        if (HIGH) {
            const int inv = fastinv(nw), items = rows * nw;
            const int wa = x0 >> 1;
            const bool odd = x0 & 1;
            for (int it = lane; it < items; it += 32) {
                const int r = fastdiv(it, inv), j = it - r * nw;
                const int y = clip3i(y0 + r, 0, ph - 1);
                const uint32_t *g = (const uint32_t *)(ref + (size_t)y * rpitch) + wa + j;
                uint32_t v = __ldg(g);
                if (odd) v = __funnelshift_r(v, __ldg(g + 1), 16);
                src[r * SRC_PITCH + j] = v;
            }
        } else {
            const int nq = (nw + 1) >> 1;                 // 4-sample units per row
            const int inv = fastinv(nq), items = rows * nq;
            const int wa = x0 >> 2, sh = (x0 & 3) * 8;
            for (int it = lane; it < items; it += 32) {
                const int r = fastdiv(it, inv), j = it - r * nq;
                const int y = clip3i(y0 + r, 0, ph - 1);
                const uint32_t *g = (const uint32_t *)(ref + (size_t)y * rpitch) + wa + j;
                uint32_t v = __ldg(g);
                if (sh) v = __funnelshift_r(v, __ldg(g + 1), sh);
                uint2 o;
                o.x = __byte_perm(v, 0, 0x4140);
                o.y = __byte_perm(v, 0, 0x4342);
                *(uint2 *)&src[r * SRC_PITCH + 2 * j] = o;
            }
        }
    } else {
        const int inv = fastinv(nw), items = rows * nw;
        for (int it = lane; it < items; it += 32) {
            const int r = fastdiv(it, inv), j = it - r * nw;
            const int y = clip3i(y0 + r, 0, ph - 1);
            const int xa = clip3i(x0 + 2 * j, 0, pw - 1), xb = clip3i(x0 + 2 * j + 1, 0, pw - 1);
            const uint8_t *row = ref + (size_t)y * rpitch;
            uint32_t a, b;
            if (HIGH) { a = __ldg((const uint16_t *)row + xa); b = __ldg((const uint16_t *)row + xb); }
            else      { a = __ldg(row + xa); b = __ldg(row + xb); }
            src[r * SRC_PITCH + j] = a | (b << 16);
        }
    }
}

// ---------------------------------------------------------------- one MC task
// One warp predicts rows [cy, cy+rows) of plane c of one PU: for each list, stage the
// window, filter (one of four cases by fractional phase), and run the epilogue into the
// CTB tile.  tile = word pointer to the PU's first row/column inside the CTB tile.
template <bool HIGH, bool LUMA>
__device__ void mc_task(const FrameDev &f, const ohcu_pu &pu, int c, int cy, int rows, uint32_t *tile, int tpitch,
                        WarpScratch &ws, int lane)
{
    constexpr int NT = LUMA ? 8 : 4;
    constexpr int BEFORE = LUMA ? 3 : 1;
    constexpr int R = 4;                                   // rows per vertical-pass item
    const int bd = f.bd;
    const ohcu_slice_tab &sl = f.slices[pu.slice];
    const int hs = LUMA ? 0 : f.hs, vs = LUMA ? 0 : f.vs;
    const int bw = pu.w >> hs;
    const int xo = pu.x0 >> hs, yo = (pu.y0 >> vs) + cy;
    const int pw = f.pw[c], ph = f.ph[c];
    const bool bi = pu.pred_flag == OHCU_PF_BI;
    const bool wflag = sl.weight_flag != 0;
    const int denom = LUMA ? sl.luma_log2_weight_denom : sl.chroma_log2_weight_denom;
    const int ncp = bw >> 1;                               // column pairs
    const int ng = (bw + 7) >> 3;                          // 8-column groups
    const int inv_cp = fastinv(ncp), inv_g = fastinv(ng);

    int wx[2], ox[2];
#pragma unroll
    for (int l = 0; l < 2; l++) {
        const int ri = pu.ref_idx[l] & (OHCU_MAX_REFS - 1);
        wx[l] = LUMA ? sl.luma_weight[l][ri] : sl.chroma_weight[l][ri][c - 1];
        ox[l] = (LUMA ? sl.luma_offset[l][ri] : sl.chroma_offset[l][ri][c - 1]) * (1 << (bd - 8));
    }

#pragma unroll 1
    for (int l = 0; l < 2; l++) {
        if (!(pu.pred_flag & (1 << l))) continue;
        // ---- epilogue of this list
        Epi e;
        e.maxv = (1 << bd) - 1;
        e.bi = bi;
        e.store_tmp = bi && l == 0;
        e.wB = 0; e.post = 0;
        if (!bi) {
            if (!wflag) { e.wA = 1; e.sh = 14 - bd; e.add = 1 << (13 - bd); }
            else { e.wA = wx[l]; e.sh = denom + 14 - bd; e.add = 1 << (e.sh - 1); e.post = ox[l]; }
        } else {
            if (!wflag) { e.wA = 1; e.wB = 1; e.sh = 15 - bd; e.add = 1 << (14 - bd); }
            else {
                const int log2Wd = denom + 14 - bd;
                e.wA = wx[1]; e.wB = wx[0]; e.sh = log2Wd + 1; e.add = (ox[0] + ox[1] + 1) * (1 << log2Wd);
            }
        }
        // ---- phases and integer position (hevcdec.c:1753-1760, 1909-1915)
        const int mvx = pu.mv[l][0], mvy = pu.mv[l][1];
        int mx, my, sx, sy;
        if (LUMA) {
            mx = mvx & 3; my = mvy & 3;
            sx = xo + (mvx >> 2); sy = yo + (mvy >> 2);
        } else {
            mx = (mvx & ((1 << (2 + hs)) - 1)) << (1 - hs);
            my = (mvy & ((1 << (2 + vs)) - 1)) << (1 - vs);
            sx = xo + (mvx >> (2 + hs)); sy = yo + (mvy >> (2 + vs));
        }
        const uint32_t FX0 = LUMA ? c_qpel4[mx][0] : c_epel4[mx][0], FX1 = LUMA ? c_qpel4[mx][1] : 0;
        const uint32_t FY0 = LUMA ? c_qpel4[my][0] : c_epel4[my][0], FY1 = LUMA ? c_qpel4[my][1] : 0;
        const PlaneSet &rp = f.dpb[sl.ref_slot[l][pu.ref_idx[l] & (OHCU_MAX_REFS - 1)] & (OHCU_MAX_DPB - 1)];
        const int hb = mx ? BEFORE : 0, vb = my ? BEFORE : 0;
        const int scols = bw + (mx ? NT - 1 : 0), srows = rows + (my ? NT - 1 : 0);

        stage_window<HIGH>(ws.src, rp.p[c], rp.pitch[c], pw, ph, sx - hb, sy - vb, scols, srows, lane);
        __syncwarp();

        const int sh1 = bd - 8;                            // first-stage shift (template :2486)
        if (mx && !my) {
            // ---- horizontal only: rows x 8-column groups
            const int items = rows * ng;
            for (int it = lane; it < items; it += 32) {
                const int r = fastdiv(it, inv_g), g = it - r * ng;
                uint32_t W[8];
                const uint4 a = *(const uint4 *)&ws.src[r * SRC_PITCH + 4 * g];
                const uint4 b = *(const uint4 *)&ws.src[r * SRC_PITCH + 4 * g + 4];
                W[0] = a.x; W[1] = a.y; W[2] = a.z; W[3] = a.w; W[4] = b.x; W[5] = b.y; W[6] = b.z; W[7] = b.w;
                int o[8];
                fir_row<NT>(W, FX0, FX1, o);
                uint32_t *tw = tile + r * tpitch + 4 * g;
#pragma unroll
                for (int k = 0; k < 4; k++)
                    if (8 * g + 2 * k < bw) epi_pair(e, o[2 * k] >> sh1, o[2 * k + 1] >> sh1, tw + k);
            }
        } else if (mx && my) {
            // ---- horizontal pass over rows + NT-1 rows into int16 rows (tmp_array, template :2522)
            const int items = srows * ng;
            for (int it = lane; it < items; it += 32) {
                const int r = fastdiv(it, inv_g), g = it - r * ng;
                uint32_t W[8];
                const uint4 a = *(const uint4 *)&ws.src[r * SRC_PITCH + 4 * g];
                const uint4 b = *(const uint4 *)&ws.src[r * SRC_PITCH + 4 * g + 4];
                W[0] = a.x; W[1] = a.y; W[2] = a.z; W[3] = a.w; W[4] = b.x; W[5] = b.y; W[6] = b.z; W[7] = b.w;
                int o[8];
                fir_row<NT>(W, FX0, FX1, o);
                uint4 p;
                p.x = __byte_perm((uint32_t)(o[0] >> sh1), (uint32_t)(o[1] >> sh1), 0x5410);
                p.y = __byte_perm((uint32_t)(o[2] >> sh1), (uint32_t)(o[3] >> sh1), 0x5410);
                p.z = __byte_perm((uint32_t)(o[4] >> sh1), (uint32_t)(o[5] >> sh1), 0x5410);
                p.w = __byte_perm((uint32_t)(o[6] >> sh1), (uint32_t)(o[7] >> sh1), 0x5410);
                *(uint4 *)&ws.hb[r * HB_PITCH + 4 * g] = p;
            }
            __syncwarp();
        }
        if (my) {
            // ---- vertical pass: column pairs x groups of R rows, from pixels (shift BD-8)
            //      or from the horizontal-pass rows (shift 6)
            const uint32_t *in = mx ? ws.hb : ws.src;
            const int ipitch = mx ? HB_PITCH : SRC_PITCH;
            const int sh2 = mx ? 6 : sh1;
            const int nrg = (rows + R - 1) / R;
            const int items = nrg * ncp;
            for (int it = lane; it < items; it += 32) {
                const int rg = fastdiv(it, inv_cp), cp = it - rg * ncp;
                uint32_t V[R + NT - 1];
#pragma unroll
                for (int i = 0; i < R + NT - 1; i++) V[i] = in[(rg * R + i) * ipitch + cp];
                int oa[R], ob[R];
                fir_col<NT, R>(V, FY0, FY1, oa, ob);
                uint32_t *tw = tile + rg * R * tpitch + cp;
#pragma unroll
                for (int i = 0; i < R; i++)
                    if (rg * R + i < rows) epi_pair(e, oa[i] >> sh2, ob[i] >> sh2, tw + i * tpitch);
            }
        } else if (!mx) {
            // ---- full-pel: val = sample << (14 - BD) (template :2359-2373)
            const int items = rows * ncp;
            const int sh = 14 - bd;
            for (int it = lane; it < items; it += 32) {
                const int r = fastdiv(it, inv_cp), cp = it - r * ncp;
                const uint32_t v = ws.src[r * SRC_PITCH + cp];
                epi_pair(e, (int)(v & 0xffffu) << sh, (int)(v >> 16) << sh, tile + r * tpitch + cp);
            }
        }
        __syncwarp();
    }
}

// ================================================================= residual
// transform[32][32] (hevcdsp.c:32-97) as 32-bit constants so that the unrolled butterflies
// take them straight from the constant bank; filled at load time from the cosine index rule.
__constant__ int c_dct[32][32];

int upload_dct_table()
{
    static const int mag[33] = {64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64,
                                61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9, 4, 0};
    int t[32][32];
    for (int r = 0; r < 32; r++)
        for (int c = 0; c < 32; c++) {
            int m = (r * (2 * c + 1)) & 127;
            t[r][c] = r == 0 ? 64 : (m <= 32 ? mag[m] : (m <= 64 ? -mag[64 - m] : (m <= 96 ? -mag[m - 64] : mag[128 - m])));
        }
    return cudaMemcpyToSymbol(c_dct, t, sizeof(t)) == cudaSuccess ? 0 : -1;
}

// N-point inverse DCT of in[0..N) into out[0..N): even/odd decomposition, exact integer
// sums (TR_4..TR_32, template :180-239).
template <int N>
__device__ __forceinline__ void idct1d(const int *in, int *out)
{
    if constexpr (N == 4) {
        const int e0 = 64 * in[0] + 64 * in[2], e1 = 64 * in[0] - 64 * in[2];
        const int o0 = 83 * in[1] + 36 * in[3], o1 = 36 * in[1] - 83 * in[3];
        out[0] = e0 + o0; out[1] = e1 + o1; out[2] = e1 - o1; out[3] = e0 - o0;
    } else {
        int ev[N / 2], e[N / 2];
#pragma unroll
        for (int j = 0; j < N / 2; j++) ev[j] = in[2 * j];
        idct1d<N / 2>(ev, e);
#pragma unroll
        for (int k = 0; k < N / 2; k++) {
            int o = 0;
#pragma unroll
            for (int j = 1; j < N; j += 2) o += c_dct[j * (32 / N)][k] * in[j];
            out[k] = e[k] + o;
            out[N - 1 - k] = e[k] - o;
        }
    }
}

// does the reference butterfly read input j when called with zero-skip bound `end`?
template <int N>
__device__ __forceinline__ bool idct_reads(int j, int end)
{
    if (N == 4) return true;
    if (j & 1) return j < end;
    if (N == 32 && (j & 3) == 2) return (j >> 1) < (end >> 1);
    return true;
}

template <int N> struct Stride { static constexpr int S = N + 2; };   // int16 row stride, S/2 odd

// one TB: coefficients -> residual, in place in blk[N][S]; N lanes (g = 0..N-1) cooperate
template <int LOG2>
__device__ void tb_residual(const FrameDev &f, const ohcu_tu &tu, int16_t *blk, int g, unsigned mask)
{
    constexpr int N = 1 << LOG2, S = Stride<N>::S;
    const int bd = f.bd;
    {
        const uint32_t *src = (const uint32_t *)(f.coeffs + tu.coeff_off);
        for (int i = g; i < N * N / 2; i += N) {
            const uint32_t v = __ldg(src + i);
            const int e = 2 * i, r = e >> LOG2, c = e & (N - 1);
            *(uint32_t *)&blk[r * S + c] = v;
        }
    }
    __syncwarp(mask);

    const int mode = tu.mode;
    if (mode == OHCU_TU_IDCT) {
        const int col_limit = tu.col_limit;
        int in[N], out[N];
        // columns (shift 7); zero-skip bound shrinks by 4 every 4 columns (template :256-262)
        {
            int limit2 = min(col_limit + 4, N);
            if (limit2 < N && g >= 1) limit2 -= 4 * ((g - 1) >> 2);
#pragma unroll
            for (int j = 0; j < N; j++) in[j] = idct_reads<N>(j, limit2) ? blk[j * S + g] : 0;
            idct1d<N>(in, out);
#pragma unroll
            for (int k = 0; k < N; k++) blk[k * S + g] = (int16_t)clip_s16((out[k] + 64) >> 7);
        }
        __syncwarp(mask);
        // rows (shift 20 - BD)
        {
            const int limit = min(col_limit, N);
            const int sh = 20 - bd, add = 1 << (sh - 1);
#pragma unroll
            for (int j = 0; j < N; j++) in[j] = idct_reads<N>(j, limit) ? blk[g * S + j] : 0;
            idct1d<N>(in, out);
#pragma unroll
            for (int k = 0; k < N; k++) blk[g * S + k] = (int16_t)clip_s16((out[k] + add) >> sh);
        }
    } else if (mode == OHCU_TU_DC) {                           // IDCT_DC, template :273-286
        const int v = (((blk[0] + 1) >> 1) + (1 << (13 - bd))) >> (14 - bd);
        __syncwarp(mask);
#pragma unroll
        for (int k = 0; k < N; k++) blk[g * S + k] = (int16_t)v;
    } else if (mode == OHCU_TU_DST4) {                         // transform_4x4_luma, template :140-177
        if constexpr (N == 4) {
#pragma unroll
            for (int pass = 0; pass < 2; pass++) {
                const int st = pass ? 1 : S, base = pass ? g * S : g;
                const int shift = pass ? 20 - bd : 7, add = 1 << (shift - 1);
                const int s0 = blk[base], s1 = blk[base + st], s2 = blk[base + 2 * st], s3 = blk[base + 3 * st];
                const int a = s0 + s2, b = s2 + s3, d = s0 - s3, e = 74 * s1;
                blk[base]          = (int16_t)clip_s16((29 * a + 55 * b + e + add) >> shift);
                blk[base + st]     = (int16_t)clip_s16((55 * d - 29 * b + e + add) >> shift);
                blk[base + 2 * st] = (int16_t)clip_s16((74 * (s0 - s2 + s3) + add) >> shift);
                blk[base + 3 * st] = (int16_t)clip_s16((55 * a + 29 * d - e + add) >> shift);
                __syncwarp(mask);
            }
        }
    } else {
        // OHCU_TU_SKIP / OHCU_TU_BYPASS (hevc_cabac.c:1727-1757)
        if (mode == OHCU_TU_SKIP) {
            if constexpr (N == 4) {
                if (tu.flags & OHCU_TUF_ROTATE) {              // coeffs[i] <-> coeffs[15 - i]
                    int16_t v[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) v[k] = blk[(3 - g) * S + (3 - k)];
                    __syncwarp(mask);
#pragma unroll
                    for (int k = 0; k < 4; k++) blk[g * S + k] = v[k];
                    __syncwarp(mask);
                }
            }
            const int shift = 15 - bd - LOG2;                  // transform_skip, template :109-133
#pragma unroll
            for (int k = 0; k < N; k++) {
                const int v = blk[g * S + k];
                blk[g * S + k] = shift > 0 ? (int16_t)((v + (1 << (shift - 1))) >> shift) : (int16_t)(v * (1 << -shift));
            }
            __syncwarp(mask);
        }
        if (tu.flags & OHCU_TUF_RDPCM) {                       // transform_rdpcm, template :87-107
            if (tu.flags & OHCU_TUF_RDPCM_VERT) {
                int16_t acc = blk[g];
                for (int y = 1; y < N; y++) { acc = (int16_t)(acc + blk[y * S + g]); blk[y * S + g] = acc; }
            } else {
                int16_t acc = blk[g * S];
                for (int x = 1; x < N; x++) { acc = (int16_t)(acc + blk[g * S + x]); blk[g * S + x] = acc; }
            }
        }
    }
    __syncwarp(mask);
}

// A warp takes 32/N TBs of one size class at a time (lane group = one TB) and adds their
// residual into the CTB tile: transform_add (template :45-84).
struct CtbTile {
    uint32_t *p[3];       // tile words of each plane
    int pitch[3];         // words per tile row
};

template <int LOG2>
__device__ void tu_batch(const FrameDev &f, const CtbTile &tile, int tu0, const uint16_t *list, int first, int count,
                         int cx0, int cy0, WarpScratch &ws, int lane)
{
    constexpr int N = 1 << LOG2, S = Stride<N>::S, GROUPS = 32 / N;
    const int grp = lane >> LOG2, g = lane & (N - 1);
    if (first + grp >= count) return;
    const unsigned mask = N == 32 ? 0xffffffffu : (((1u << N) - 1u) << (lane & ~(N - 1)));
    const ohcu_tu tu = f.tus[tu0 + list[first + grp]];
    int16_t *blk = (int16_t *)&ws + (size_t)grp * 2 * N * S;
    static_assert(GROUPS * 2 * N * S * 2 <= (int)sizeof(WarpScratch), "block buffers must fit the warp scratch");

    tb_residual<LOG2>(f, tu, blk, g, mask);
    if (tu.flags & OHCU_TUF_CROSS) {                           // hevc_cabac.c:1810-1817
        int16_t *by = blk + N * S;
        const ohcu_tu ty = f.tus[tu.cross_tu];
        if (ty.log2_size == LOG2) {
            tb_residual<LOG2>(f, ty, by, g, mask);
            const int scale = tu.res_scale_val;
#pragma unroll
            for (int k = 0; k < N; k++)
                blk[g * S + k] = (int16_t)(blk[g * S + k] + ((scale * by[g * S + k]) >> 3));
            __syncwarp(mask);
        }
    }
    const int c = tu.c_idx;
    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
    const int maxv = (1 << f.bd) - 1;
    uint16_t *t = (uint16_t *)tile.p[c] + (tu.y0 - (cy0 >> vs)) * (2 * tile.pitch[c]) + (tu.x0 - (cx0 >> hs)) + g;
#pragma unroll 4
    for (int r = 0; r < N; r++) {
        uint16_t *d = t + r * 2 * tile.pitch[c];
        *d = (uint16_t)min(max((int)*d + blk[r * S + g], 0), maxv);
    }
}

// ================================================================= the kernel
__device__ __forceinline__ ohcu_pu load_pu(const ohcu_pu *p)
{
    union { ohcu_pu pu; uint32_t w[5]; } u;
    static_assert(sizeof(ohcu_pu) == 20, "PU record is five words");
    const uint32_t *s = (const uint32_t *)p;
#pragma unroll
    for (int i = 0; i < 5; i++) u.w[i] = __ldg(s + i);
    return u.pu;
}

// flags: bit 0 run MC, bit 1 run the residual
template <bool HIGH>
__global__ void __launch_bounds__(RC_THREADS, 3) recon_kernel(const __grid_constant__ FrameDev f, int flags)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    WarpScratch *scratch = (WarpScratch *)smem_raw;
    uint32_t *tile_base = (uint32_t *)(smem_raw + sizeof(WarpScratch) * RC_WARPS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ctb = blockIdx.x;
    const int ctbx = ctb % f.ctb_w, ctby = ctb / f.ctb_w;
    const int cs = 1 << f.log2_ctb;
    const int cx0 = ctbx << f.log2_ctb, cy0 = ctby << f.log2_ctb;      // luma origin of the CTB

    CtbTile tile;
    int tw[3], th[3], vw[3], vh[3];
    {
        uint32_t *p = tile_base;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
            tw[c] = cs >> hs; th[c] = cs >> vs;
            vw[c] = min(tw[c], f.pw[c] - (cx0 >> hs));
            vh[c] = min(th[c], f.ph[c] - (cy0 >> vs));
            tile.pitch[c] = (tw[c] >> 1) + 2;
            tile.p[c] = p;
            if (c < f.nplanes) p += tile.pitch[c] * th[c];
        }
        // after the tiles: the size-sorted TB list and its counters
        tile_base = p;
    }
    uint16_t *tu_list = (uint16_t *)(tile_base + 16);
    int *tu_cnt = (int *)tile_base;                                      // [0..3] counts, [4..7] bases, [8..11] cursors

    const int2 prange = f.ctb_pu_range[ctb];
    const int p0 = prange.x, np = prange.y & 0xffffff;
    const bool covered = (prange.y >> 30) & 1;
    const bool do_mc = (flags & 1) && np > 0;
    const bool do_res = (flags & 2) && f.n_tu > 0;
    const int2 trange = do_res ? f.ctb_tu_range[ctb] : make_int2(0, 0);
    const int t0 = trange.x, nt = min(trange.y, MAX_CTB_TUS);

    // ---- phase 0: start from the current picture where this launch does not produce every sample
    if (!do_mc || !covered) {
        for (int c = 0; c < f.nplanes; c++) {
            const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
            const uint8_t *plane = f.work.p[c] + (size_t)(cy0 >> vs) * f.work.pitch[c];
            const int units = vw[c] >> 2, items = units * vh[c], inv = fastinv(units);
            for (int it = threadIdx.x; it < items; it += RC_THREADS) {
                const int r = fastdiv(it, inv), u = it - r * units;
                uint2 o;
                if (HIGH) {
                    o = *(const uint2 *)(plane + (size_t)r * f.work.pitch[c] + (size_t)((cx0 >> hs) + 4 * u) * 2);
                } else {
                    const uint32_t v = *(const uint32_t *)(plane + (size_t)r * f.work.pitch[c] + (cx0 >> hs) + 4 * u);
                    o.x = __byte_perm(v, 0, 0x4140);
                    o.y = __byte_perm(v, 0, 0x4342);
                }
                *(uint2 *)&tile.p[c][r * tile.pitch[c] + 2 * u] = o;
            }
        }
    }
    if (threadIdx.x < 12) tu_cnt[threadIdx.x] = 0;
    __syncthreads();

    // ---- phase 1: prediction.  Tasks = (PU, plane, chunk of <= 16 rows), round-robin over warps.
    if (do_mc) {
        int counter = 0;
        for (int i = 0; i < np; i++) {
            const ohcu_pu pu = load_pu(f.pus + p0 + i);
            const int chl = (pu.h + MC_CHUNK - 1) / MC_CHUNK;
            const int bhc = pu.h >> f.vs;
            const int chc = f.nplanes == 3 ? (bhc + MC_CHUNK - 1) / MC_CHUNK : 0;
            const int total = chl + 2 * chc;
            // does any of counter .. counter+total-1 belong to this warp?
            const int firstmine = (warp - counter) & (RC_WARPS - 1);
            if (firstmine < total) {
                for (int t = firstmine; t < total; t += RC_WARPS) {
                    int c, ch;
                    if (t < chl) { c = 0; ch = t; }
                    else if (t < chl + chc) { c = 1; ch = t - chl; }
                    else { c = 2; ch = t - chl - chc; }
                    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
                    const int bh = pu.h >> vs;
                    const int cy = ch * MC_CHUNK, rows = min(MC_CHUNK, bh - cy);
                    uint32_t *tp = tile.p[c] + (((pu.y0 - cy0) >> vs) + cy) * tile.pitch[c] + (((pu.x0 - cx0) >> hs) >> 1);
                    if (c == 0) mc_task<HIGH, true>(f, pu, 0, cy, rows, tp, tile.pitch[0], scratch[warp], lane);
                    else        mc_task<HIGH, false>(f, pu, c, cy, rows, tp, tile.pitch[c], scratch[warp], lane);
                }
            }
            counter = (counter + total) & (RC_WARPS - 1);
        }
    }

    // ---- phase 2: residual.  Counting sort of the CTB's TBs by size, then batches per warp.
    if (nt > 0) {
        for (int i = threadIdx.x; i < nt; i += RC_THREADS) atomicAdd(&tu_cnt[f.tus[t0 + i].log2_size - 2], 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            // largest first: 32x32 batches are the long ones
            int acc = 0;
            for (int l = 3; l >= 0; l--) { tu_cnt[4 + l] = acc; acc += tu_cnt[l]; }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < nt; i += RC_THREADS) {
            const int l = f.tus[t0 + i].log2_size - 2;
            tu_list[tu_cnt[4 + l] + atomicAdd(&tu_cnt[8 + l], 1)] = (uint16_t)i;
        }
    }
    __syncthreads();          // prediction complete, TB list sorted
    if (nt > 0) {
        int b = 0;            // running batch number over all size classes
#pragma unroll
        for (int l = 3; l >= 0; l--) {
            const int cnt = tu_cnt[l], base = tu_cnt[4 + l];
            const int per = 32 >> (l + 2);                 // TBs per batch
            const int nb = (cnt + per - 1) / per;
            for (int k = (warp - b) & (RC_WARPS - 1); k < nb; k += RC_WARPS) {
                if (l == 3) tu_batch<5>(f, tile, t0, tu_list + base, k * per, cnt, cx0, cy0, scratch[warp], lane);
                if (l == 2) tu_batch<4>(f, tile, t0, tu_list + base, k * per, cnt, cx0, cy0, scratch[warp], lane);
                if (l == 1) tu_batch<3>(f, tile, t0, tu_list + base, k * per, cnt, cx0, cy0, scratch[warp], lane);
                if (l == 0) tu_batch<2>(f, tile, t0, tu_list + base, k * per, cnt, cx0, cy0, scratch[warp], lane);
                __syncwarp();
            }
            b = (b + nb) & (RC_WARPS - 1);
        }
    }
    __syncthreads();

    // ---- phase 3: the CTB leaves the SM once, in 8-byte (HIGH) / 4-byte units of four samples
    for (int c = 0; c < f.nplanes; c++) {
        const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
        uint8_t *plane = f.work.p[c] + (size_t)(cy0 >> vs) * f.work.pitch[c];
        const int units = vw[c] >> 2, items = units * vh[c];
        const int inv = fastinv(units);
        for (int it = threadIdx.x; it < items; it += RC_THREADS) {
            const int r = fastdiv(it, inv), u = it - r * units;
            const uint2 v = *(const uint2 *)&tile.p[c][r * tile.pitch[c] + 2 * u];
            if (HIGH)
                *(uint2 *)(plane + (size_t)r * f.work.pitch[c] + (size_t)((cx0 >> hs) + 4 * u) * 2) = v;
            else
                *(uint32_t *)(plane + (size_t)r * f.work.pitch[c] + (cx0 >> hs) + 4 * u) = __byte_perm(v.x, v.y, 0x6420);
        }
    }
}

size_t recon_smem_bytes(const FrameDev &f)
{
    const int cs = 1 << f.log2_ctb;
    size_t words = 0;
    for (int c = 0; c < f.nplanes; c++) {
        const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
        words += (size_t)(((cs >> hs) >> 1) + 2) * (cs >> vs);
    }
    return sizeof(WarpScratch) * RC_WARPS + words * 4 + 16 * 4 + MAX_CTB_TUS * 2;
}

int launch_recon(const FrameDev &f, int flags, cudaStream_t st)
{
    if (!(flags & 3)) return 0;
    if (!((flags & 1) && f.n_pu > 0) && !((flags & 2) && f.n_tu > 0)) return 0;
    const size_t smem = recon_smem_bytes(f);
    static bool attr_set = false;
    if (!attr_set) {
        if (cudaFuncSetAttribute(recon_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(recon_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024) != cudaSuccess)
            return -1;
        attr_set = true;
    }
    const int nctb = f.ctb_w * f.ctb_h;
    if (f.bd > 8) recon_kernel<true><<<nctb, RC_THREADS, smem, st>>>(f, flags);
    else          recon_kernel<false><<<nctb, RC_THREADS, smem, st>>>(f, flags);
    return 1;
}

}  // namespace ohcu
