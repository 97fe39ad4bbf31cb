// recon.cu — inter reconstruction of one picture: motion compensation, then the residual.
//
// Two kernels, one CTA (256 threads) per CTB in both.  mc_kernel writes the prediction of
// every inter PU (luma qpel 8-tap / chroma epel 4-tap, uni / bi / weighted) straight to the
// job's picture.  residual_kernel then takes the CTBs that hold coded blocks: the CTB enters
// a shared-memory tile (the prediction was just written: L2 hits), the residual of every coded
// TB is added into the tile (inverse DCT 4..32 / DC / 4x4 DST / transform_skip / rdpcm /
// cross-component), and the CTB is written back once with full-line stores; the per-block
// transform_add RMW of the reference disappears.  Until r01 v11 both lived in one kernel whose
// tile never left the SM; its 170 KB of code thrashed the 32 KB instruction cache (26 % of
// the warp stalls) and its shared memory took the L1 the windows are read through, and the two
// launches are 10 % faster per picture (profiles/README.md, r01 v12).
//
// Replaces, per picture:
//   hls_prediction_unit -> luma_mc_uni/bi, chroma_mc_uni/bi     hevcdec.c:1744-2049, 2250-2306
//   put_hevc_{qpel,epel}{,_uni,_uni_w,_bi,_bi_w}[idx][my][mx]   hevcdsp_template.c:2359-3357
//   ff_emulated_edge_mc                                         videodsp_template.c:22-100
//   ff_hevc_hls_transform / _c + transform_add                  hevc_cabac.c:1695-1906,
//                                                               hevcdsp_template.c:45-296
//
// Motion compensation.
//  * Reference samples come from HALO-PADDED DPB planes (PAD_X / PAD_Y replicated samples,
//    written once per picture by pad_kernel).  Window positions outside the halo are clamped to
//    it, which is bit-identical to the reference's per-block edge emulation.
//  * Work is cut into CELLS of 4x4 luma samples (and their chroma): one thread per cell,
//    every thread running the same straight-line code, so a CTB of 64 tiny PUs keeps the
//    lanes as busy as a CTB with one 64x64 PU.  A 16x16 map of the CTB says which PU owns a
//    cell; the PUs enter themselves into it, so there is no sorting and no scan.  The
//    [my][mx] table dispatch is gone: every cell runs the separable h+v path with the phase-0
//    taps {0,0,0,64,0,0,0,0} standing in for "no filter", which gives the reference's results
//    exactly (64*x >> 6 == x, and (x << (14-BD)) is what (64*x) >> (BD-8) computes).
//  * A cell reads its window STRAIGHT FROM THE DPB PLANE through L1 (ld.global.nc, 8-byte
//    loads): neighbouring cells of a PU share their rows in L1 (mc_kernel's CTAs hold 10 KB of
//    shared memory each, so L1 keeps most of the SM's 256 KB).  The earlier form (one TMA box per PU and list into a
//    shared-memory pool, cells run in rounds as the boxes land) spent more time on window
//    geometry, scans, round bookkeeping and mbarrier waits than on filtering: 0.205 ms per 4K
//    frame against 0.137 ms for this one (profiles/README.md, r01 v9 -> v10).
//  * Samples are 16-bit lanes, two per 32-bit word, so the taps run on the integer
//    dot-product pipe: IDP.2A does two 16x8-bit MACs per instruction at IMAD issue rate
//    (profiles/r01_ubench_int_pipes.txt).  Horizontal pairs come straight from the window
//    words (+ funnel shifts for odd positions), vertical pairs are built with PRMT.
#include <mutex>

#include "ohcu_internal.h"

#ifndef MC_PREFETCH
#define MC_PREFETCH 0      // 1: prefetch.global.L1 of the window's rows ahead of the row loop (A/B builds)
#endif
#ifndef MC_PREFETCH_FROM
#define MC_PREFETCH_FROM 2
#endif

namespace ohcu {

constexpr int RC_WARPS = 8;
constexpr int RC_THREADS = RC_WARPS * 32;
constexpr int MAX_CTB_PUS = 128;             // 8x4 PUs of a 64x64 CTB
constexpr int MAX_CTB_TUS = 768;             // 4x4 TBs of a 64x64 CTB, 3 full-size planes
constexpr int POOL_UNIT = 128;               // granule of the residual area
#ifndef RC_MINB
#define RC_MINB 4                            // CTAs per SM the register budget is set for: 64 registers.  Against 3 (80 registers,
                                             // more L1 left): 4K equal, 8K -2 %, 1080p -14 % (510 CTBs fit one wave of 592)
#endif

// scratch of one warp in the residual phase: one 32x32 int16 block
struct WarpScratch { uint32_t w[32 * 40 / 2]; };   // 32 x (32 + 8) int16: idct_mma's pass-1 output
// The residual area of a CTA: per-warp scratch, then the int16 coefficients of the CTB
// (`units` x 128 B in all; a 64x64 4:2:0 CTB has 6144 coefficients).
__host__ __device__ constexpr int coef_area(int units)          // int16, multiple of 8
{
    return ((units * POOL_UNIT - RC_WARPS * (int)sizeof(WarpScratch)) / 2) & ~7;
}

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

// ---------------------------------------------------------------- mbarrier / TMA
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity)
{
    uint32_t ok = 0;
    long long t0 = 0;
    for (int spin = 0; !ok; spin++) {
        // try_wait suspends the warp until the phase completes or the hint (ns) runs out
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(smem_u32(b)), "r"(parity), "r"(20000) : "memory");
        // a transfer that never completes is a bug, not a reason to hang the GPU: fail the launch
        if (!ok && (spin & 63) == 63) {
            if (!t0) t0 = clock64();
            else if (clock64() - t0 > 4000000000ll) __trap();
        }
    }
}
// ---------------------------------------------------------------- dot-product helpers
__device__ __forceinline__ int dp2lo(uint32_t a, uint32_t b, int c) { return __dp2a_lo((int)a, (int)b, c); }
__device__ __forceinline__ int dp2hi(uint32_t a, uint32_t b, int c) { return __dp2a_hi((int)a, (int)b, c); }

// NT-tap FIR along a row: W[j] holds samples 2j (low half) and 2j+1;
// out[i] = sum_k tap[k] * sample[i + k], i = 0..OW-1
// four unsigned bytes x four signed bytes + c
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

template <int NT, int OW>
__device__ __forceinline__ void fir_row(const uint32_t (&W)[(OW + NT) / 2], uint32_t F0, uint32_t F1, int (&out)[OW])
{
    constexpr int NW = (OW + NT) / 2;
    uint32_t S[NW - 1];
#pragma unroll
    for (int j = 0; j < NW - 1; j++) S[j] = __funnelshift_r(W[j], W[j + 1], 16);
#pragma unroll
    for (int j = 0; j < OW / 2; j++) {
        if (NT == 8) {
            out[2 * j] = dp2hi(W[j + 3], F1, dp2lo(W[j + 2], F1, dp2hi(W[j + 1], F0, dp2lo(W[j], F0, 0))));
            out[2 * j + 1] = dp2hi(S[j + 3], F1, dp2lo(S[j + 2], F1, dp2hi(S[j + 1], F0, dp2lo(S[j], F0, 0))));
        } else {
            out[2 * j] = dp2hi(W[j + 1], F0, dp2lo(W[j], F0, 0));
            out[2 * j + 1] = dp2hi(S[j + 1], F0, dp2lo(S[j], F0, 0));
        }
    }
}

// NT-tap FIR down two columns at once: V[i] holds the samples of row i (two columns per
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
// 14-bit intermediate `val` of this list and `tmp` of the first list (bi only):
//   uni     (template :2544)  wA 1          add 1<<(13-BD)            sh 14-BD      post 0
//   uni_w   (template :2733)  wA wx         add 1<<(shift-1)          sh shift      post ox
//   bi      (template :2570)  wA 1   wB 1   add 1<<(14-BD)            sh 15-BD      post 0
//   bi_w    (template :2760)  wA wx1 wB wx0 add (ox0+ox1+1)<<log2Wd   sh log2Wd+1   post 0
struct Epi {
    int wA, wB, add, sh, post, maxv;
};

// val / tmp: two 14-bit intermediates each (int16 pairs); tmp = first list of a bi PU (the
// reference's lc->tmp, hevcdec.c:1854), 0 otherwise
__device__ __forceinline__ void epi_pair(const Epi &e, uint32_t val, uint32_t tmp, uint32_t *tw)
{
    const int va = (int)(int16_t)(val & 0xffffu), vb = (int)val >> 16;
    const int ta = (int)(int16_t)(tmp & 0xffffu), tb = (int)tmp >> 16;
    int oa = ((va * e.wA + ta * e.wB + e.add) >> e.sh) + e.post;
    int ob = ((vb * e.wA + tb * e.wB + e.add) >> e.sh) + e.post;
    oa = min(max(oa, 0), e.maxv);
    ob = min(max(ob, 0), e.maxv);
    *tw = (uint32_t)oa | ((uint32_t)ob << 16);
}

__device__ __forceinline__ int fastdiv(int n, int inv) { return (n * inv) >> 16; }   // n < 2048, divisor <= 32
__constant__ int c_inv[41] = {0, 65536, 32768, 21846, 16384, 13108, 10923, 9363, 8192, 7282, 6554, 5958, 5462, 5042,
                              4682, 4370, 4096, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428,
                              2341, 2260, 2185, 2115, 2048, 1986, 1928, 1873, 1821, 1772, 1725, 1681, 1639};
__device__ __forceinline__ int fastinv(int d) { return c_inv[d]; }                   // ceil(65536 / d), d <= 40

// ================================================================= residual
// transform[32][32] (hevcdsp.c:32-97) as 32-bit constants so that the unrolled butterflies
// take them straight from the constant bank; filled at load time from the cosine index rule.
__constant__ int c_dct[32][32];

// constant mma fragments of the transform matrix, see idct_mma below
__device__ uint32_t g_idct_frag[2][2][32][4];       // [N == 32][k-half][output index i][t]

static void fill_idct_frag(const int (*t)[32], uint32_t (*out)[2][32][4])
{
    for (int big = 0; big < 2; big++) {
        const int n = big ? 32 : 16, step = 32 / n;
        for (int h = 0; h < 2; h++)
            for (int i = 0; i < 32; i++)
                for (int q = 0; q < 4; q++) {
                    uint32_t w = 0;
                    for (int s = 0; s < 4; s++) {
                        const int j = 16 * h + (s < 2 ? 2 * q + s : 8 + 2 * q + s - 2);
                        const int v = (j < n && i < n) ? t[j * step][i] : 0;
                        w |= (uint32_t)(v & 0xff) << (8 * s);
                    }
                    out[big][h][i][q] = w;
                }
    }
}

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
    static uint32_t frag[2][2][32][4];
    fill_idct_frag(t, frag);
    if (cudaMemcpyToSymbol(g_idct_frag, frag, sizeof(frag)) != cudaSuccess) return -1;
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
// coef: the TB's N*N coefficients (shared-memory copy of the CTB's arena range, or global)
template <int LOG2>
__device__ __noinline__ void tb_residual(int bd, int mode, int tuflags, int col_limit, const int16_t *coef, int16_t *blk, int g,
                                         unsigned mask)
{
    constexpr int N = 1 << LOG2, S = Stride<N>::S;
    {
        const uint32_t *src = (const uint32_t *)coef;
        for (int i = g; i < N * N / 2; i += N) {
            const uint32_t v = src[i];
            const int e = 2 * i, r = e >> LOG2, c = e & (N - 1);
            *(uint32_t *)&blk[r * S + c] = v;
        }
    }
    __syncwarp(mask);

    if (mode == OHCU_TU_IDCT) {
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
                if (tuflags & OHCU_TUF_ROTATE) {              // coeffs[i] <-> coeffs[15 - i]
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
        if (tuflags & OHCU_TUF_RDPCM) {                       // transform_rdpcm, template :87-107
            if (tuflags & OHCU_TUF_RDPCM_VERT) {
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

// ---------------------------------------------------------------- IDCT 16 / 32 on the tensor cores
// The two passes of IDCT(16|32) (template :249-291) are matrix products with the int8
// transform matrix: integer mma.sync (exact s32 accumulation) on the int16 data split into a
// signed high and an unsigned low byte plane, v = 256 * hi + lo:
//   pass 1   tmp[y][x] = clip16((sum_j T[j][y] * coef[j][x] + 64) >> 7)         A = T^T (constant), B = coef
//   pass 2   res[y][m] = clip16((sum_x tmp[y][x] * T[x][m] + add) >> (20 - BD))  A = tmp, B = T (constant)
// The contraction index must be contiguous in both fragments.  ldmatrix(.trans) delivers 16-bit
// pairs, so the four bytes of a fragment register are data elements {2t, 2t+1, 8+2t, 9+2t} (+16
// for the second k-half) of a column (pass 1, .trans on the row-major coefficient block) or of a
// row (pass 2) — a fixed permutation of the contraction index that the constant fragments
// repeat.  One table serves both passes: the pass-1 A fragment of output row i is the pass-2 B
// fragment of output column i.
// Zero skipping: the reference skips single inputs, not tiles — an input j of a butterfly called
// with bound `end` is read iff idct_reads(j, end) (TR_8..TR_32, template :193-239), with
// end = col_limit + 4 shrinking by 4 every 4 columns in pass 1 (:256-262) and end = col_limit in
// pass 2.  Inputs it does not read are masked to zero here (registers of pass 1, the stored
// pass-1 output for pass 2), so the result equals the reference's for ANY block content, not
// only for blocks whose skipped coefficients are zero.
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void *p)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)) : "memory");
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], const void *p)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)) : "memory");
}
// d += A(16x16) * B(16x8); SA / SB: operand is signed
template <bool SA, bool SB>
__device__ __forceinline__ void imma16(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t b)
{
    if (SA && SB)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a0), "r"(a1), "r"(b));
    else if (SA)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a0), "r"(a1), "r"(b));
    else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3]) : "r"(a0), "r"(a1), "r"(b));
}
__device__ __forceinline__ uint32_t pack_sat16(int lo, int hi)
{
    uint32_t d;
    asm("cvt.pack.sat.s16.s32 %0, %1, %2;" : "=r"(d) : "r"(hi), "r"(lo));
    return d;
}
// tile word (two samples) += two residuals, clipped to the pixel range: transform_add (template :45-84)
__device__ __forceinline__ void tile_add_pair(uint32_t *w, int r0, int r1, int maxv)
{
    const uint32_t v = *w;
    const int lo = min(max((int)(v & 0xffffu) + r0, 0), maxv), hi = min(max((int)(v >> 16) + r1, 0), maxv);
    *w = (uint32_t)lo | ((uint32_t)hi << 16);
}

// One N x N block by the whole warp.  src: row-major coefficients in shared memory, 16-byte
// aligned; t1: N x (N + 8) int16 of scratch; tw: tile word of the block's first sample.
// raw: store the residual pairs instead of adding them (per-entry test shim ohcu_dsp_idct).
template <int LOG2>
__device__ __noinline__ void idct_mma(const int16_t *src, int16_t *t1, uint32_t *tw, int tpitch, int col_limit, int bd, int lane,
                                      bool raw = false)
{
    constexpr int N = 1 << LOG2, ST = N + 8, NQ = N / 8, NH = N / 16;
    const int g = lane >> 2, t = lane & 3;
    uint32_t F[NQ][NH];
#pragma unroll
    for (int q = 0; q < NQ; q++)
#pragma unroll
        for (int h = 0; h < NH; h++) F[q][h] = __ldg(&g_idct_frag[LOG2 - 4][h][g + 8 * q][t]);
    const int limit = min(col_limit, N), lim2 = min(col_limit + 4, N);
    // keep-masks of a 16-bit pair (even element e, odd element e + 1) under bound `end`
    auto pair_mask = [](int e, int end) -> uint32_t {
        const bool keep_even = (N == 32 && (e & 3) == 2) ? (e >> 1) < (end >> 1) : true;
        return (keep_even ? 0xffffu : 0u) | (e + 1 < end ? 0xffff0000u : 0u);
    };

    // ---- pass 1: columns.  One ldmatrix.x4.trans per step: lane i supplies row (i & 7) of matrix (i >> 3).
    //   N = 32: the four row tiles (j = 0..31) of column tile xt -> both k-halves
    //   N = 16: the two row tiles of column tiles 0 and 1 -> the whole block in one step
    constexpr int XSTEP = N == 16 ? 2 : 1;
#pragma unroll 1
    for (int xt0 = 0; xt0 < N / 8 / XSTEP; xt0++) {
        uint32_t r[4];
        if (N == 32) ldsm_x4_trans(r, src + lane * N + 8 * xt0);
        else ldsm_x4_trans(r, src + (8 * ((lane >> 3) & 1) + (lane & 7)) * N + 8 * (lane >> 4));
        if (lim2 < N) {                                    // inputs the reference's column pass skips
#pragma unroll
            for (int m = 0; m < 4; m++) {
                const int x = 8 * (N == 32 ? xt0 : (m >> 1)) + g;
                const int end = lim2 - (x ? 4 * ((x - 1) >> 2) : 0);
                r[m] &= pair_mask(8 * (N == 32 ? m : (m & 1)) + 2 * t, end);
            }
        }
#pragma unroll
        for (int u = 0; u < XSTEP; u++) {
            uint32_t bl[NH], bh[NH];
            if (N == 32) {
#pragma unroll
                for (int h = 0; h < NH; h++) {
                    bl[h] = __byte_perm(r[2 * h], r[2 * h + 1], 0x6420);
                    bh[h] = __byte_perm(r[2 * h], r[2 * h + 1], 0x7531);
                }
            } else {
                bl[0] = __byte_perm(r[2 * u], r[2 * u + 1], 0x6420);
                bh[0] = __byte_perm(r[2 * u], r[2 * u + 1], 0x7531);
            }
            // pass 2 does not read every column: mask of this thread's output pair (x, x + 1)
            const int x = 8 * (xt0 * XSTEP + u) + 2 * t;
            const uint32_t keep = pair_mask(x, limit);
#pragma unroll
            for (int mt = 0; mt < NH; mt++) {
                int acc[4] = {0, 0, 0, 0};
#pragma unroll
                for (int h = 0; h < NH; h++) imma16<true, true>(acc, F[2 * mt][h], F[2 * mt + 1][h], bh[h]);
#pragma unroll
                for (int k = 0; k < 4; k++) acc[k] = acc[k] * 256 + 64;
#pragma unroll
                for (int h = 0; h < NH; h++) imma16<true, false>(acc, F[2 * mt][h], F[2 * mt + 1][h], bl[h]);
                const int y = 16 * mt + g;
                *(uint32_t *)&t1[y * ST + x] = pack_sat16(acc[0] >> 7, acc[1] >> 7) & keep;
                *(uint32_t *)&t1[(y + 8) * ST + x] = pack_sat16(acc[2] >> 7, acc[3] >> 7) & keep;
            }
        }
    }
    __syncwarp();

    // ---- pass 2: rows, added into the tile
    const int sh = 20 - bd, add = 1 << (sh - 1), maxv = (1 << bd) - 1;
#pragma unroll 1
    for (int mt = 0; mt < NH; mt++) {
        uint32_t al[NH][2], ah[NH][2];
#pragma unroll
        for (int h = 0; h < NH; h++) {
            {
                uint32_t r[4];
                // matrices: (rows 16mt.., cols 16h), (rows 16mt.., cols 16h + 8), (rows 16mt + 8.., cols 16h), (.., 16h + 8)
                const int m = lane >> 3;
                ldsm_x4(r, t1 + (16 * mt + 8 * (m >> 1) + (lane & 7)) * ST + 16 * h + 8 * (m & 1));
                al[h][0] = __byte_perm(r[0], r[1], 0x6420); ah[h][0] = __byte_perm(r[0], r[1], 0x7531);
                al[h][1] = __byte_perm(r[2], r[3], 0x6420); ah[h][1] = __byte_perm(r[2], r[3], 0x7531);
            }
        }
#pragma unroll
        for (int nt = 0; nt < NQ; nt++) {
            int acc[4] = {0, 0, 0, 0};
#pragma unroll
            for (int h = 0; h < NH; h++)
                imma16<true, true>(acc, ah[h][0], ah[h][1], F[nt][h]);
#pragma unroll
            for (int k = 0; k < 4; k++) acc[k] = acc[k] * 256 + add;
#pragma unroll
            for (int h = 0; h < NH; h++)
                imma16<false, true>(acc, al[h][0], al[h][1], F[nt][h]);
            const int y = 16 * mt + g, xw = 4 * nt + t;
            if (raw) {
                tw[y * tpitch + xw] = pack_sat16(acc[0] >> sh, acc[1] >> sh);
                tw[(y + 8) * tpitch + xw] = pack_sat16(acc[2] >> sh, acc[3] >> sh);
            } else {
                tile_add_pair(tw + y * tpitch + xw, clip_s16(acc[0] >> sh), clip_s16(acc[1] >> sh), maxv);
                tile_add_pair(tw + (y + 8) * tpitch + xw, clip_s16(acc[2] >> sh), clip_s16(acc[3] >> sh), maxv);
            }
        }
    }
    __syncwarp();
}

// IDCT_DC (template :273-286) of one N x N block by the whole warp
__device__ __forceinline__ void dc_block(int log2, int c0, uint32_t *tw, int tpitch, int bd, int lane, bool raw = false)
{
    const int v = (((c0 + 1) >> 1) + (1 << (13 - bd))) >> (14 - bd), maxv = (1 << bd) - 1;
    const int n = 1 << log2, wl = log2 - 1;                // words per row = n / 2
    if (raw) {
        const uint32_t vv = (uint32_t)(v & 0xffff) * 0x10001u;
        for (int i = lane; i < (n << wl); i += 32) tw[(i >> wl) * tpitch + (i & ((1 << wl) - 1))] = vv;
        return;
    }
    for (int i = lane; i < (n << wl); i += 32) tile_add_pair(tw + (i >> wl) * tpitch + (i & ((1 << wl) - 1)), v, v, maxv);
}

// A warp takes 32/N TBs of one size class at a time (lane group = one TB) and adds their
// residual into the CTB tile: transform_add (template :45-84).
struct CtbTile {
    uint32_t *base;       // plane 0; planes 1 and 2 follow (same geometry)
    int size0, sizec;     // words of plane 0 / of one chroma plane
    int pitch0, pitchc;   // words per tile row
    // no arrays: a run-time plane index into a local array would go through local memory
    __device__ __forceinline__ uint32_t *p(int c) const { return base + (c > 0 ? size0 : 0) + (c > 1 ? sizec : 0); }
    __device__ __forceinline__ int pitch(int c) const { return c ? pitchc : pitch0; }
};

// Where the coefficients of this CTB are: csm != nullptr -> arena range [clo, ...) copied to
// shared memory by one bulk copy, else read from global memory.
struct CoefSrc {
    const int16_t *csm;
    uint32_t clo, chi;
    const int16_t *glob;
    __device__ __forceinline__ const int16_t *at(uint32_t off, int n) const
    {
        return (csm && off >= clo && off + n <= chi) ? csm + (off - clo) : glob + off;
    }
};

template <int LOG2>
__device__ void tu_batch(const FrameDev &f, const CtbTile &tile, int tu0, const uint16_t *list, int first, int count,
                         int cx0, int cy0, WarpScratch &ws, WarpScratch *ws2, const CoefSrc &cs, int lane)
{
    constexpr int N = 1 << LOG2, S = Stride<N>::S, GROUPS = 32 / N;
    constexpr bool MMA = LOG2 >= 4;
    const int grp = lane >> LOG2, g = lane & (N - 1);
    const bool have = first + grp < count;
    const unsigned mask = N == 32 ? 0xffffffffu : (((1u << N) - 1u) << (lane & ~(N - 1)));
    ohcu_tu tu = {};
    if (have) tu = f.tus[tu0 + list[first + grp]];
    static_assert(GROUPS * N * S * 2 <= (int)sizeof(WarpScratch), "block buffers must fit the warp scratch");
    // 16x16 / 32x32 IDCT and DC blocks go through the warp-wide paths below
    bool wide = false;
    if (MMA && have && !(tu.flags & OHCU_TUF_CROSS)) {
        const bool in_smem = cs.csm && tu.coeff_off >= cs.clo && tu.coeff_off + N * N <= cs.chi && !((tu.coeff_off - cs.clo) & 7);
        wide = tu.mode == OHCU_TU_DC || (tu.mode == OHCU_TU_IDCT && in_smem);
    }
    if (have && !wide) {
        int16_t *blk = (int16_t *)&ws + (size_t)grp * N * S;
        tb_residual<LOG2>(f.bd, tu.mode, tu.flags, tu.col_limit, cs.at(tu.coeff_off, N * N), blk, g, mask);
        if ((tu.flags & OHCU_TUF_CROSS) && ws2) {                  // hevc_cabac.c:1810-1817
            int16_t *by = (int16_t *)ws2 + (size_t)grp * N * S;
            const ohcu_tu ty = f.tus[tu.cross_tu];
            if (ty.log2_size == LOG2) {
                tb_residual<LOG2>(f.bd, ty.mode, ty.flags, ty.col_limit, cs.at(ty.coeff_off, N * N), by, g, mask);
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
        if (tu.flags & OHCU_TUF_INTRA) {
            // the block's prediction does not exist yet: the residual waits in the job's residual plane (intra.cu adds it)
            int16_t *d = (int16_t *)((uint8_t *)f.resid[c] + (size_t)tu.y0 * f.resid_pitch[c]) + tu.x0 + g;
            const int rp = f.resid_pitch[c] >> 1;
#pragma unroll 4
            for (int r = 0; r < N; r++) d[r * rp] = blk[r * S + g];
        } else {
            uint16_t *t = (uint16_t *)tile.p(c) + (tu.y0 - (cy0 >> vs)) * (2 * tile.pitch(c)) + (tu.x0 - (cx0 >> hs)) + g;
#pragma unroll 4
            for (int r = 0; r < N; r++) {
                uint16_t *d = t + r * 2 * tile.pitch(c);
                *d = (uint16_t)min(max((int)*d + blk[r * S + g], 0), maxv);
            }
        }
    }
    if constexpr (MMA) {
        __syncwarp();
#pragma unroll 1
        for (int gi = 0; gi < GROUPS; gi++) {
            const int sl = gi << LOG2;                                 // lane that holds this block's record
            if (!__shfl_sync(0xffffffffu, (int)wide, sl)) continue;
            const int c = __shfl_sync(0xffffffffu, (int)tu.c_idx, sl);
            const int x0 = __shfl_sync(0xffffffffu, (int)tu.x0, sl), y0 = __shfl_sync(0xffffffffu, (int)tu.y0, sl);
            const int mode = __shfl_sync(0xffffffffu, (int)tu.mode, sl);
            const int col_limit = __shfl_sync(0xffffffffu, (int)tu.col_limit, sl);
            const uint32_t off = __shfl_sync(0xffffffffu, tu.coeff_off, sl);
            const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
            const bool intra = __shfl_sync(0xffffffffu, (int)tu.flags, sl) & OHCU_TUF_INTRA;
            uint32_t *tw = tile.p(c) + (y0 - (cy0 >> vs)) * tile.pitch(c) + ((x0 - (cx0 >> hs)) >> 1);
            int tp = tile.pitch(c);
            if (intra) {                                               // residual plane instead of the tile, stored raw
                tw = (uint32_t *)((uint8_t *)f.resid[c] + (size_t)y0 * f.resid_pitch[c]) + (x0 >> 1);
                tp = f.resid_pitch[c] >> 2;
            }
            if (mode == OHCU_TU_DC) dc_block(LOG2, cs.at(off, 1)[0], tw, tp, f.bd, lane, intra);
            else idct_mma<LOG2>(cs.csm + (off - cs.clo), (int16_t *)&ws, tw, tp, col_limit, f.bd, lane, intra);
            __syncwarp();
        }
    }
}

// ---------------------------------------------------------------- levels -> coefficients
// The "levels" form of the sparse upload (ohcu_append_tu_levels): the pairs carry the un-scaled
// trans_coeff_level and are scaled while they are scattered into the CTB's dense blocks:
// scale_and_clip_coeff (hevc_cabac.c:2030-2055) with the parameters of
// derive_quant_parameters{,_c} (:1472-1557).  The reference scales exactly the levels it decodes,
// i.e. the pairs.  A pair only knows its position in the CTB's coefficient area, so a table with
// one word per 16 coefficients (blocks are 16-aligned multiples of 16) says which block it is in:
//   block base / 16 | (log2 - 2) << 10 | qp << 12 | matrix id << 19 (7 = flat) | bypass << 22
constexpr int QTAB_FLAT = 7;

__device__ __forceinline__ uint32_t qtab_word(const ohcu_tu &tu, const ohcu_tu_quant q)
{
    return (tu.coeff_off >> 4) | (uint32_t)(tu.log2_size - 2) << 10 | (uint32_t)q.qp << 12 |
           (uint32_t)(q.matrix_id == OHCU_QM_FLAT ? QTAB_FLAT : q.matrix_id) << 19 | (tu.mode == OHCU_TU_BYPASS ? 1u << 22 : 0u);
}

__device__ __forceinline__ int dequant_level(int v, uint32_t pos, uint32_t w, const uint8_t *scaling, int bd)
{
    if ((w >> 22) & 1) return v;                                      // cu_transquant_bypass: levels are the residual
    const int log2 = 2 + ((w >> 10) & 3), qp = (w >> 12) & 127, mid = (w >> 19) & 7;
    const int e = (int)pos - (int)((w & 1023u) << 4);
    const int shift = bd + log2 - 5;                                  // bit_depth + log2 + 10 - log2_transform_range (15)
    int sm = 16;
    if (mid != QTAB_FLAT && scaling) {
        const int x = e & ((1 << log2) - 1), y = e >> log2;
        const int ds = log2 > 3 ? log2 - 3 : 0, lw = log2 == 2 ? 2 : 3;  // 16x16 / 32x32 share an 8x8 list
        sm = (log2 >= 4 && !(x | y)) ? scaling[4 * 6 * 64 + (log2 - 4) * 6 + mid]
                                     : scaling[((log2 - 2) * 6 + mid) * 64 + ((y >> ds) << lw) + (x >> ds)];
    }
    // level_scale[] = {40, 45, 51, 57, 64, 72} (hevc_cabac.c:1470) as bytes of two words; rem6[] / div6[] of the
    // reference hold zeros at qp 74 and 75 (76 entries declared, 74 initialised, :1456-1468): reproduced
    const int rem = qp < 74 ? qp % 6 : 0, div = qp < 74 ? qp / 6 : 0;
    const int scale = (int)(((rem < 4 ? 0x39332d28u >> (8 * rem) : 0x00004840u >> (8 * (rem - 4))) & 0xffu) << div);
    const long long t = ((long long)(v * sm) * scale + (1ll << (shift - 1))) >> shift;
    return (int)(t < -32768 ? -32768 : (t > 32767 ? 32767 : t));
}

// ---------------------------------------------------------------- per-PU records
__device__ __forceinline__ ohcu_pu load_pu(const ohcu_pu *p)
{
    union { ohcu_pu pu; uint32_t w[5]; } u;
    static_assert(sizeof(ohcu_pu) == 20, "PU record is five words");
    const uint32_t *s = (const uint32_t *)p;
#pragma unroll
    for (int i = 0; i < 5; i++) u.w[i] = __ldg(s + i);
    return u.pu;
}

// What the cells of a PU need, prepared once by the four threads that own the PU.
//   rect  x/4 | y/4 << 4 | w/4 << 8 | h/4 << 13 | bi << 18                        (position inside the CTB)
//   org   [list slot A,B][luma, chroma]: window origin (first needed sample) in plane coordinates,
//         x | y << 16 (int16 each; negative = inside the replicated halo)
//   par   [..][..]: DPB slot | mx << 8 | my << 11
//   epi   per plane: (wA & 0xffff) | wB << 16,  add,  sh | post << 16
struct PuRec {
    uint32_t rect;
    uint32_t org[2][2];
    uint32_t par[2][2];
    uint32_t epi[3][3];
};
constexpr int CMAP_STRIDE = 16;              // 4x4 cells per CTB row (64x64 CTB)

// One reference window: the samples list `l` of the PU reads in plane type luma / chroma
// (both chroma planes share it).  Always the full separable window (NT-1 extra rows and
// columns).  Coordinates are clamped so that a window lying entirely outside the picture is
// fetched from the nearest part of the replicated halo — same samples as
// ff_emulated_edge_mc would build (videodsp_template.c:37-83).
struct WinGeom {
    int x0, y0;           // plane coordinates of the first needed sample
    int mx, my;           // fractional phase: qpel 0..3 / epel 0..7
};

__device__ __forceinline__ WinGeom win_geom(const FrameDev &f, const ohcu_pu &pu, int l, bool luma)
{
    WinGeom g;
    const int hs = luma ? 0 : f.hs, vs = luma ? 0 : f.vs;
    const int NT = luma ? 8 : 4, BEFORE = luma ? 3 : 1;
    const int bw = pu.w >> hs, bh = pu.h >> vs;
    const int xo = pu.x0 >> hs, yo = pu.y0 >> vs;
    const int pw = f.pw[luma ? 0 : 1], ph = f.ph[luma ? 0 : 1];
    // explicit selects: a run-time index into the record would push it to local memory
    const int mvx = l ? pu.mv[1][0] : pu.mv[0][0], mvy = l ? pu.mv[1][1] : pu.mv[0][1];
    int sx, sy;
    if (luma) {                                            // hevcdec.c:1753-1760
        g.mx = mvx & 3; g.my = mvy & 3;
        sx = xo + (mvx >> 2); sy = yo + (mvy >> 2);
    } else {                                               // hevcdec.c:1909-1915
        g.mx = (mvx & ((1 << (2 + hs)) - 1)) << (1 - hs);
        g.my = (mvy & ((1 << (2 + vs)) - 1)) << (1 - vs);
        sx = xo + (mvx >> (2 + hs)); sy = yo + (mvy >> (2 + vs));
    }
    g.x0 = clip3i(sx - BEFORE, -(bw + NT - 1), pw);
    g.y0 = clip3i(sy - BEFORE, -(bh + NT - 1), ph);
    return g;
}

// ---------------------------------------------------------------- one cell
// OW x OH output samples of one plane from the window at `win` (row pitch in bytes):
// horizontal pass over OH + NT - 1 rows into packed int16 registers (the reference's
// tmp_array, template :2522), vertical pass; res[r][k] = the 14-bit intermediates of output
// row r, columns 2k and 2k + 1 as an int16 pair.
//   col0  first output column of the cell relative to the window's first needed column
//   row0  first output row of the cell relative to the window's first row
template <bool HIGH, int NT, int OW, int OH>
__device__ __forceinline__ void mc_filter(const uint8_t *win, int pitch, int col0, int row0, uint32_t FX0, uint32_t FX1,
                                          uint32_t FY0, uint32_t FY1, int sh1, uint32_t (&res)[OH][OW / 2])
{
    constexpr int NW = (OW + NT) / 2, NR = OH + NT - 1;
    uint32_t hb[NR][OW / 2];
    const uint8_t *rowp = win + row0 * pitch;
#if MC_PREFETCH
    // every row of the cell's window requested at once: the misses of the later rows overlap with the filtering of the first
#pragma unroll
    for (int r = MC_PREFETCH_FROM; r < NR; r++)
        asm volatile("prefetch.global.L1 [%0];" ::"l"(rowp + r * pitch + col0 * (HIGH ? 2 : 1)));
#endif
#pragma unroll
    for (int r = 0; r < NR; r++) {
        uint32_t W[NW];
        if (HIGH) {
            const int sh = (col0 & 1) * 16;
            uint32_t T[NW];
            // 8-byte loads from the aligned pair of words at or left of the first one needed
            constexpr int NQ = NW / 2 + 1;
            const uint2 *q = (const uint2 *)rowp + (col0 >> 2);
            const bool odd = (col0 >> 1) & 1;
            uint32_t U[2 * NQ];
#pragma unroll
            for (int j = 0; j < NQ; j++) { const uint2 v = __ldg(q + j); U[2 * j] = v.x; U[2 * j + 1] = v.y; }
#pragma unroll
            for (int j = 0; j < NW; j++) T[j] = odd ? U[j + 1] : U[j];
#pragma unroll
            for (int j = 0; j < NW; j++) W[j] = __funnelshift_r(T[j], T[j + 1 < NW ? j + 1 : j], sh);
        }
        int o[OW];
        if (HIGH) {
            fir_row<NT, OW>(W, FX0, FX1, o);
        } else {
            // 8-bit samples stay packed four to a word: one dp4a (u8 x s8) per four taps instead of two dp2a on widened pairs
            constexpr int NB = (NW + 1) / 2 + 1;
            const uint32_t *p = (const uint32_t *)rowp + (col0 >> 2);
            const int sh = (col0 & 3) * 8;
            uint32_t T[NB], A[NB - 1];
#pragma unroll
            for (int j = 0; j < NB; j++) T[j] = __ldg(p + j);
#pragma unroll
            for (int j = 0; j < NB - 1; j++) A[j] = __funnelshift_r(T[j], T[j + 1], sh);
#pragma unroll
            for (int k = 0; k < OW; k++) {
                o[k] = dp4a_us(k ? __funnelshift_r(A[0], A[1], 8 * k) : A[0], FX0, 0);
                if (NT == 8) o[k] = dp4a_us(k ? __funnelshift_r(A[1], A[2 < NB - 1 ? 2 : 0], 8 * k) : A[1], FX1, o[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < OW / 2; k++)
            hb[r][k] = __byte_perm((uint32_t)(o[2 * k] >> sh1), (uint32_t)(o[2 * k + 1] >> sh1), 0x5410);
        rowp += pitch;
    }
#pragma unroll
    for (int k = 0; k < OW / 2; k++) {
        uint32_t V[NR];
#pragma unroll
        for (int r = 0; r < NR; r++) V[r] = hb[r][k];
        int oa[OH], ob[OH];
        fir_col<NT, OH>(V, FY0, FY1, oa, ob);
#pragma unroll
        for (int r = 0; r < OH; r++) res[r][k] = __byte_perm((uint32_t)(oa[r] >> 6), (uint32_t)(ob[r] >> 6), 0x5410);
    }
}

// One cell of one plane: list slot A, then (bi) slot B, then the epilogue into the CTB tile.
// One copy of the filter code: the lists run through the same loop body.  The windows are read
// from the DPB plane of the list's reference picture (org / par: PuRec).
template <bool HIGH, int NT, int OW, int OH>
__device__ __forceinline__ void mc_cell(const FrameDev &f, int c, uint32_t orgA, uint32_t parA, uint32_t orgB, uint32_t parB, bool bi,
                                          int ccol, int crow, int sh1, const Epi &e, uint32_t *tw, int tpitch)
{
    uint32_t cur[OH][OW / 2], prev[OH][OW / 2];
#pragma unroll
    for (int r = 0; r < OH; r++)
#pragma unroll
        for (int k = 0; k < OW / 2; k++) prev[r][k] = 0;
#pragma unroll 1
    for (int s = 0; s < (bi ? 2 : 1); s++) {
        const uint32_t org = s ? orgB : orgA, par = s ? parB : parA;
        const int slot = par & 0xff, mx = (par >> 8) & 7, my = (par >> 11) & 7;
        const int x0 = (int)(int16_t)(org & 0xffffu), y0 = (int)org >> 16;
        const uint8_t *plane = f.dpb[slot].p[c];
        const int pitch = f.dpb[slot].pitch[c];
        if (s) {
#pragma unroll
            for (int r = 0; r < OH; r++)
#pragma unroll
                for (int k = 0; k < OW / 2; k++) prev[r][k] = cur[r][k];
        }
        if (NT == 8)
            mc_filter<HIGH, NT, OW, OH>(plane, pitch, x0 + ccol, y0 + crow, c_qpel4[mx][0], c_qpel4[mx][1], c_qpel4[my][0],
                                        c_qpel4[my][1], sh1, cur);
        else
            mc_filter<HIGH, NT, OW, OH>(plane, pitch, x0 + ccol, y0 + crow, c_epel4[mx][0], 0, c_epel4[my][0], 0, sh1, cur);
    }
#pragma unroll
    for (int r = 0; r < OH; r++)
#pragma unroll
        for (int k = 0; k < OW / 2; k++) epi_pair(e, cur[r][k], prev[r][k], tw + r * tpitch + k);
}

// ---- PU records and the cell map of one CTB (the preparation of mc_kernel):
// four threads per PU: thread `sub` prepares the window of kind `sub` (list slot sub & 1, luma / chroma
// sub >> 1) and the epilogue constants of plane `sub`, and enters the PU into a quarter of its cells of
// the cell map; 64 PUs per pass.
__device__ __forceinline__ void build_pu_records(const FrameDev &f, PuRec *recs, uint8_t *cmap, int p0, int np, int cx0, int cy0,
                                                 int bd, int tid)
{
    const int sub = tid & 3;
#pragma unroll 1
    for (int base = 0; base < np; base += RC_THREADS / 4) {
        const int p = base + (tid >> 2);
        if (p < np) {
            const ohcu_pu pu = load_pu(f.pus + p0 + p);
            const ohcu_slice_tab &sl = f.slices[pu.slice];
            const bool bi = pu.pred_flag == OHCU_PF_BI;
            PuRec &rc = recs[p];
            const int lA = (pu.pred_flag & 1) ? 0 : 1;
            const int x4 = (pu.x0 - cx0) >> 2, y4 = (pu.y0 - cy0) >> 2, w4 = pu.w >> 2, h4 = pu.h >> 2;
            if (sub == 0)
                rc.rect = (uint32_t)(x4 | (y4 << 4) | (w4 << 8) | (h4 << 13) | ((bi ? 1 : 0) << 18));
            {
                const int s = sub & 1, pt = sub >> 1;
                uint32_t org = 0, par = 0;
                if ((s == 0 || bi) && (pt == 0 || f.nplanes == 3)) {
                    const int l = s ? 1 : lA;
                    const WinGeom g = win_geom(f, pu, l, pt == 0);
                    const int slot = sl.ref_slot[l][(l ? pu.ref_idx[1] : pu.ref_idx[0]) & (OHCU_MAX_REFS - 1)] & (OHCU_MAX_DPB - 1);
                    org = ((uint32_t)g.x0 & 0xffffu) | ((uint32_t)g.y0 << 16);
                    par = (uint32_t)slot | ((uint32_t)g.mx << 8) | ((uint32_t)g.my << 11);
                }
                rc.org[s][pt] = org;
                rc.par[s][pt] = par;
            }
            if (sub < 3) {                                 // epilogue constants of plane `sub` (see Epi)
                const int c = sub;
                const bool wflag = sl.weight_flag != 0;
                const int denom = c ? sl.chroma_log2_weight_denom : sl.luma_log2_weight_denom;
                int wx[2], ox[2];
#pragma unroll
                for (int l = 0; l < 2; l++) {
                    const int ri = (l ? pu.ref_idx[1] : pu.ref_idx[0]) & (OHCU_MAX_REFS - 1);
                    wx[l] = c ? sl.chroma_weight[l][ri][c - 1] : sl.luma_weight[l][ri];
                    ox[l] = (c ? sl.chroma_offset[l][ri][c - 1] : sl.luma_offset[l][ri]) * (1 << (bd - 8));
                }
                int wA, wB = 0, add, sh, post = 0;
                if (!bi) {
                    // rounding offsets are 0 at 14 bits (#if BIT_DEPTH < 14, template :2381, :2400, :2427)
                    if (!wflag) { wA = 1; sh = 14 - bd; add = bd < 14 ? 1 << (13 - bd) : 0; }
                    else { wA = wx[lA]; sh = denom + 14 - bd; add = bd < 14 ? 1 << (sh - 1) : 0; post = ox[lA]; }
                } else {
                    if (!wflag) { wA = 1; wB = 1; sh = 15 - bd; add = bd < 14 ? 1 << (14 - bd) : 0; }
                    else {
                        const int log2Wd = denom + 14 - bd;
                        wA = wx[1]; wB = wx[0]; sh = log2Wd + 1; add = (ox[0] + ox[1] + 1) * (1 << log2Wd);
                    }
                }
                rc.epi[c][0] = ((uint32_t)wA & 0xffffu) | ((uint32_t)wB << 16);
                rc.epi[c][1] = (uint32_t)add;
                rc.epi[c][2] = ((uint32_t)sh & 0xffffu) | ((uint32_t)post << 16);
            }
            const int inv = fastinv(w4);
            for (int i = sub; i < w4 * h4; i += 4) {
                const int yy = fastdiv(i, inv), xx = i - yy * w4;
                cmap[(y4 + yy) * CMAP_STRIDE + x4 + xx] = (uint8_t)p;
            }
        }
    }
}

// ================================================================= the kernel
__device__ __forceinline__ void unpack_epi(const uint32_t (&w)[3], int bd, Epi &e)
{
    e.wA = (int)(int16_t)(w[0] & 0xffffu);
    e.wB = (int)w[0] >> 16;
    e.add = (int)w[1];
    e.sh = (int)(w[2] & 0xffffu);
    e.post = (int)w[2] >> 16;
    e.maxv = (1 << bd) - 1;
}

// residual_kernel: the coded blocks of one CTB added to the prediction mc_kernel left in the job's
// picture.  CTBs without coded blocks leave at once.
// -DRC_PROFILE: thread 0 of every CTA adds the cycles between phase marks to g_rc_prof (tools/rc_phases.py)
#ifdef RC_PROFILE
__device__ unsigned long long g_rc_prof[16];
#define RC_MARK(i) do { if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_rc_prof[i], (unsigned long long)(t_ - rc_t)); rc_t = t_; } } while (0)
#else
#define RC_MARK(i) do { } while (0)
#endif

template <bool HIGH>
__global__ void __launch_bounds__(RC_THREADS, RC_MINB) residual_kernel(const __grid_constant__ FrameDev f, int area_units)
{
#ifdef RC_PROFILE
    long long rc_t = clock64();
#endif
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ctb = blockIdx.x;
    const int ctbx = ctb % f.ctb_w, ctby = ctb / f.ctb_w;
    const int cs = 1 << f.log2_ctb;
    const int cx0 = ctbx << f.log2_ctb, cy0 = ctby << f.log2_ctb;      // luma origin of the CTB
    const int bd = f.bd;

    // ---- shared memory carve-up: [residual area | tile, TB lists, barrier].  The residual area: per-warp
    //      scratch + the CTB's coefficients, `area_units` x 128 B.
    uint8_t *pool = smem_raw;
    const int pool_bytes = area_units * POOL_UNIT;
    CtbTile tile;
    uint32_t *sp = (uint32_t *)(smem_raw + pool_bytes);
    // valid samples of this CTB per plane type (the picture may end inside it)
    const int vw0 = min(cs, f.pw[0] - cx0), vh0 = min(cs, f.ph[0] - cy0);
    const int vwc = f.nplanes == 3 ? min(cs >> f.hs, f.pw[1] - (cx0 >> f.hs)) : 0;
    const int vhc = f.nplanes == 3 ? min(cs >> f.vs, f.ph[1] - (cy0 >> f.vs)) : 0;
    tile.base = sp;
    tile.pitch0 = (cs >> 1) + 2;
    tile.pitchc = ((cs >> f.hs) >> 1) + 2;
    tile.size0 = tile.pitch0 * cs;
    tile.sizec = f.nplanes == 3 ? tile.pitchc * (cs >> f.vs) : 0;
    sp += tile.size0 + 2 * tile.sizec;
    int *tu_cnt = (int *)sp;                    sp += 18;                         // [0..3] counts, [4..7] bases, [8..11] cursors, [12] coef lo, [13] coef hi, [14] batch counter, [15..16] sparse pair range
    uint16_t *tu_list = (uint16_t *)sp;         sp += MAX_CTB_TUS / 2;
    uint64_t *mbar = (uint64_t *)(((uintptr_t)sp + 7) & ~(uintptr_t)7);           sp += 4;                     // coefficient bulk copy
    WarpScratch *scratch2 = f.cfi == 3 ? (WarpScratch *)(((uintptr_t)sp + 15) & ~(uintptr_t)15) : nullptr;   // cross-component

    const int2 trange = f.ctb_tu_range[ctb];
    const int t0 = trange.x, nt = min(trange.y, MAX_CTB_TUS);
    if (nt == 0) return;                      // nothing to add to this CTB
    if (f.err && *f.err) return;              // malformed lists (prepare.cu)

    // ---- phase 0: the CTB's samples (the prediction, or whatever the picture holds) enter the tile
    {
        for (int c = 0; c < f.nplanes; c++) {
            const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
            const uint8_t *plane = f.work.p[c] + (size_t)(cy0 >> vs) * f.work.pitch[c];
            const int units = (c ? vwc : vw0) >> 2, items = units * (c ? vhc : vh0), inv = fastinv(units);
            for (int it = tid; it < items; it += RC_THREADS) {
                const int r = fastdiv(it, inv), u = it - r * units;
                uint2 o;
                if (HIGH) {
#ifndef RC_NO_LDGSTS
                    // same format in the picture and in the tile: asynchronous copies, awaited ahead of the
                    // batches (the TB sort and the coefficient scatter run meanwhile)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(&tile.p(c)[r * tile.pitch(c) + 2 * u])),
                                 "l"(plane + (size_t)r * f.work.pitch[c] + (size_t)((cx0 >> hs) + 4 * u) * 2) : "memory");
                    continue;
#endif
                    o = *(const uint2 *)(plane + (size_t)r * f.work.pitch[c] + (size_t)((cx0 >> hs) + 4 * u) * 2);
                } else {
                    const uint32_t v = *(const uint32_t *)(plane + (size_t)r * f.work.pitch[c] + (cx0 >> hs) + 4 * u);
                    o.x = __byte_perm(v, 0, 0x4140);
                    o.y = __byte_perm(v, 0, 0x4342);
                }
                *(uint2 *)&tile.p(c)[r * tile.pitch(c) + 2 * u] = o;
            }
        }
    }
    if (tid < 12) tu_cnt[tid] = 0;
    if (tid == 12) { tu_cnt[12] = 0x7fffffff; tu_cnt[13] = 0; tu_cnt[14] = 0; }
    if (tid == 0) mbar_init(mbar, 1);
    __syncthreads();
    RC_MARK(1);

    // ---- TBs of the CTB sorted by size (counting sort) and the arena range their coefficients
    //      occupy
    auto sort_tus = [&]() {
        if (nt > 0) {
            if (tid == RC_THREADS - 1 && f.ctb_coef_range) { const int2 cr = f.ctb_coef_range[ctb]; tu_cnt[15] = cr.x; tu_cnt[16] = cr.y; }
            for (int i = tid; i < nt; i += RC_THREADS) {
                const ohcu_tu *tu = f.tus + t0 + i;
                const int l2 = tu->log2_size;
                const uint32_t off = tu->coeff_off;
                atomicAdd(&tu_cnt[l2 - 2], 1);
                atomicMin(&tu_cnt[12], (int)off);
                atomicMax(&tu_cnt[13], (int)(off + (1u << (2 * l2))));
            }
            __syncthreads();
            if (tid == 0) {
                int acc = 0;                               // largest first: 32x32 batches are the long ones
                for (int l = 3; l >= 0; l--) { tu_cnt[4 + l] = acc; acc += tu_cnt[l]; }
            }
            __syncthreads();
            for (int i = tid; i < nt; i += RC_THREADS) {
                const int l = f.tus[t0 + i].log2_size - 2;
                tu_list[tu_cnt[4 + l] + atomicAdd(&tu_cnt[8 + l], 1)] = (uint16_t)i;
            }
        }
    };

    sort_tus();

    RC_MARK(5);
    // ================================================================ phase 2: residual
    // sparse coefficients: the first (position, level) pairs of the CTB are fetched ahead of the barrier
    const bool sparse = nt > 0 && f.ctb_coef_range != nullptr;
    uint32_t pr[4] = {0xffffu, 0xffffu, 0xffffu, 0xffffu};
    if (sparse) {
        const uint32_t *src = (const uint32_t *)f.coeffs + tu_cnt[15];
        const int npair = tu_cnt[16];
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (k * RC_THREADS + tid < npair) pr[k] = __ldg(src + k * RC_THREADS + tid);
    }
    __syncthreads();
    RC_MARK(6);
    if (nt > 0) {
        // the CTB's coefficients sit in one arena range (TBs are appended CTB by CTB): one bulk
        // copy brings them next to the per-warp block buffers; ranges that do not fit are read
        // from global memory block by block
        int16_t *coef_sm = (int16_t *)(pool + RC_WARPS * sizeof(WarpScratch));
        const uint32_t clo = sparse ? 0u : (uint32_t)tu_cnt[12], chi = (uint32_t)tu_cnt[13];
        const uint32_t cbytes = (chi - clo) * 2;
        const bool staged = sparse || (chi > clo && cbytes <= (uint32_t)coef_area(area_units) * 2 && !(cbytes & 15) && !(clo & 7));
        if (sparse) {
            // pairs -> dense blocks: clear the CTB's coefficient area, then scatter
            const uint32_t *src = (const uint32_t *)f.coeffs + tu_cnt[15];
            const int npair = tu_cnt[16];
            const uint32_t area = min(chi, (uint32_t)coef_area(area_units));
            for (uint32_t o = tid * 8; o < area; o += RC_THREADS * 8) *(uint4 *)(coef_sm + o) = make_uint4(0, 0, 0, 0);
            // levels form: which block a position belongs to (the per-warp scratch is free until the batches)
            uint32_t *qtab = (uint32_t *)pool;
            const bool levels = f.tu_quant != nullptr;
            if (levels) {
                for (int i = tid; i < nt; i += RC_THREADS) {
                    const ohcu_tu tu = f.tus[t0 + i];
                    const uint32_t w = qtab_word(tu, f.tu_quant[t0 + i]), n16 = 1u << (2 * tu.log2_size - 4);
                    if (tu.coeff_off + 16 * n16 <= area)
                        for (uint32_t k = 0; k < n16; k++) qtab[(tu.coeff_off >> 4) + k] = w;
                }
            }
            __syncthreads();
            for (int base = 0;;) {
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t pos = pr[k] & 0xffffu;
                    if (pos < area) {
                        int v = (int)(int16_t)(pr[k] >> 16);
                        if (levels) v = dequant_level(v, pos, qtab[pos >> 4], f.scaling, bd);
                        coef_sm[pos] = (int16_t)v;
                    }
                }
                base += 4 * RC_THREADS;
                if (base >= npair) break;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int q = base + k * RC_THREADS + tid;
                    pr[k] = q < npair ? __ldg(src + q) : 0xffffu;
                }
            }
            asm volatile("cp.async.wait_all;" ::: "memory");         // the tile (phase 0)
            __syncthreads();
            RC_MARK(7);
        } else if (staged) {
            if (tid == 0) {
                mbar_expect_tx(mbar, cbytes);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(coef_sm)), "l"(f.coeffs + clo), "r"(cbytes), "r"(smem_u32(mbar)) : "memory");
            }
            mbar_wait(mbar, 0);
            RC_MARK(7);
        }
        if (!sparse) {
            asm volatile("cp.async.wait_all;" ::: "memory");         // the tile (phase 0)
            __syncthreads();
        }
        CoefSrc cs;
        cs.csm = staged ? coef_sm : nullptr; cs.clo = clo; cs.chi = chi; cs.glob = f.coeffs;
        WarpScratch &ws = ((WarpScratch *)pool)[warp];
        WarpScratch *ws2 = scratch2 ? scratch2 + warp : nullptr;
        // batches (32/N TBs of one size) are handed out dynamically, largest blocks first
        const int nb3 = tu_cnt[3], nb2 = (tu_cnt[2] + 1) >> 1, nb1 = (tu_cnt[1] + 3) >> 2, nb0 = (tu_cnt[0] + 7) >> 3;
        const int nbatch = nb3 + nb2 + nb1 + nb0;
        RC_MARK(8);
        for (;;) {
            int b = 0;
            if (lane == 0) b = atomicAdd(&tu_cnt[14], 1);
            b = __shfl_sync(0xffffffffu, b, 0);
            if (b >= nbatch) break;
            if (b < nb3) tu_batch<5>(f, tile, t0, tu_list + tu_cnt[4 + 3], b, tu_cnt[3], cx0, cy0, ws, ws2, cs, lane);
            else if ((b -= nb3) < nb2) tu_batch<4>(f, tile, t0, tu_list + tu_cnt[4 + 2], b * 2, tu_cnt[2], cx0, cy0, ws, ws2, cs, lane);
            else if ((b -= nb2) < nb1) tu_batch<3>(f, tile, t0, tu_list + tu_cnt[4 + 1], b * 4, tu_cnt[1], cx0, cy0, ws, ws2, cs, lane);
            else tu_batch<2>(f, tile, t0, tu_list + tu_cnt[4 + 0], (b - nb1) * 8, tu_cnt[0], cx0, cy0, ws, ws2, cs, lane);
            __syncwarp();
        }
    }
    __syncthreads();
    RC_MARK(9);

    // ================================================================ phase 3: the CTB leaves the SM once
    for (int c = 0; c < f.nplanes; c++) {
        const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
        uint8_t *plane = f.work.p[c] + (size_t)(cy0 >> vs) * f.work.pitch[c];
        const int units = (c ? vwc : vw0) >> 2, items = units * (c ? vhc : vh0);
        const int inv = fastinv(units);
        for (int it = tid; it < items; it += RC_THREADS) {
            const int r = fastdiv(it, inv), u = it - r * units;
            const uint2 v = *(const uint2 *)&tile.p(c)[r * tile.pitch(c) + 2 * u];
            if (HIGH)
                *(uint2 *)(plane + (size_t)r * f.work.pitch[c] + (size_t)((cx0 >> hs) + 4 * u) * 2) = v;
            else
                *(uint32_t *)(plane + (size_t)r * f.work.pitch[c] + (cx0 >> hs) + 4 * u) = __byte_perm(v.x, v.y, 0x6420);
        }
    }
    RC_MARK(10);
}

#ifdef RC_PROFILE
int rc_profile_read(unsigned long long *out, int reset)
{
    if (cudaMemcpyFromSymbol(out, g_rc_prof, sizeof(g_rc_prof)) != cudaSuccess) return -1;
    if (reset) { unsigned long long z[16] = {}; cudaMemcpyToSymbol(g_rc_prof, z, sizeof(z)); }
    return 0;
}
extern "C" int ohcu_rc_profile(unsigned long long *out, int reset) { return rc_profile_read(out, reset); }
#endif

// ================================================================= prediction alone
// mc_kernel: the prediction samples of one CTB written straight to the job's picture.  PU records
// and the cell map (build_pu_records), then one thread per 4x4 luma cell and its chroma; no tile, no
// TB lists: 10 KB of static shared memory per CTA, so the L1 the windows are read through keeps
// most of the SM's 256 KB and only the filter code is resident in the instruction caches (fused
// with the residual in one kernel, 26 % of the warp stalls were instruction fetches).
// launch_recon runs it ahead of residual_kernel, which starts from the picture (its phase 0) in
// the CTBs that hold coded blocks.
#ifndef MC_MINB
#define MC_MINB 4
#endif
template <bool HIGH, int OW, int OH>
__device__ __forceinline__ void store_cell(uint8_t *dst, int pitch, const uint32_t (&o)[OH][OW / 2])
{
#pragma unroll
    for (int r = 0; r < OH; r++) {
        uint8_t *row = dst + (size_t)r * pitch;
        if (HIGH) {
            if (OW == 4) *(uint2 *)row = make_uint2(o[r][0], o[r][1]);
            else *(uint32_t *)row = o[r][0];
        } else {
            if (OW == 4) *(uint32_t *)row = __byte_perm(o[r][0], o[r][1], 0x6420);
            else *(uint16_t *)row = (uint16_t)__byte_perm(o[r][0], 0, 0x4420);
        }
    }
}

template <bool HIGH, int NT, int OW, int OH>
__device__ __forceinline__ void mc_cell_out(const FrameDev &f, int c, const PuRec &rc, int pt, bool bi, int ccol, int crow, int bd,
                                            uint8_t *dst, int pitch)
{
    Epi e;
    const uint32_t ew[3] = {rc.epi[c][0], rc.epi[c][1], rc.epi[c][2]};
    unpack_epi(ew, bd, e);
    uint32_t o[OH][OW / 2];
    mc_cell<HIGH, NT, OW, OH>(f, c, rc.org[0][pt], rc.par[0][pt], rc.org[1][pt], rc.par[1][pt], bi, ccol, crow, bd - 8, e, &o[0][0],
                              OW / 2);
    store_cell<HIGH, OW, OH>(dst, pitch, o);
}

template <bool HIGH>
__global__ void __launch_bounds__(RC_THREADS, MC_MINB) mc_kernel(const __grid_constant__ FrameDev f)
{
    __shared__ __align__(16) PuRec recs[MAX_CTB_PUS];
    __shared__ __align__(16) uint8_t cmap[CMAP_STRIDE * CMAP_STRIDE];
    constexpr int P = HIGH ? 2 : 1;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ctb = blockIdx.x;
    const int ctbx = ctb % f.ctb_w, ctby = ctb / f.ctb_w;
    const int cs = 1 << f.log2_ctb;
    const int cx0 = ctbx << f.log2_ctb, cy0 = ctby << f.log2_ctb;
    const int bd = f.bd;
    const int2 prange = f.ctb_pu_range[ctb];
    const int p0 = prange.x, np = min(prange.y & 0xffffff, MAX_CTB_PUS);
    if (np == 0) return;
    if (f.err && *f.err) return;              // malformed lists (prepare.cu)
    if (tid < CMAP_STRIDE * CMAP_STRIDE / 4) ((uint32_t *)cmap)[tid] = 0xffffffffu;
    __syncthreads();
    build_pu_records(f, recs, cmap, p0, np, cx0, cy0, bd, tid);
    __syncthreads();
    // one thread per 4x4 cell (and its chroma); a warp covers 8 x 4 cells of a 64x64 CTB
    int cx, cy;
    if (f.log2_ctb == 6) { cx = (warp & 1) * 8 + (lane & 7); cy = (warp >> 1) * 4 + (lane >> 3); }
    else { const int n4 = cs >> 2; cy = tid >> (f.log2_ctb - 2); cx = tid & (n4 - 1); }
    const int p = (cy << 2) < cs ? cmap[cy * CMAP_STRIDE + cx] : 0xff;
    if (p == 0xff) return;
    const PuRec &rc = recs[p];
    const uint32_t rect = rc.rect;
    const int ccx = cx - (int)(rect & 15), ccy = cy - (int)((rect >> 4) & 15);
    const bool pbi = (rect >> 18) & 1;
    const int x = cx0 + 4 * cx, y = cy0 + 4 * cy;                                   // luma position of the cell
    mc_cell_out<HIGH, 8, 4, 4>(f, 0, rc, 0, pbi, ccx * 4, ccy * 4, bd, f.work.p[0] + (size_t)y * f.work.pitch[0] + (size_t)x * P,
                               f.work.pitch[0]);
    if (f.nplanes == 3) {
        const int hs = f.hs, vs = f.vs;
#pragma unroll 1
        for (int c = 1; c <= 2; c++) {
            uint8_t *dst = f.work.p[c] + (size_t)(y >> vs) * f.work.pitch[c] + (size_t)(x >> hs) * P;
            if (hs && vs) mc_cell_out<HIGH, 4, 2, 2>(f, c, rc, 1, pbi, ccx * 2, ccy * 2, bd, dst, f.work.pitch[c]);
            else if (hs) mc_cell_out<HIGH, 4, 2, 4>(f, c, rc, 1, pbi, ccx * 2, ccy * 4, bd, dst, f.work.pitch[c]);
            else mc_cell_out<HIGH, 4, 4, 4>(f, c, rc, 1, pbi, ccx * 4, ccy * 4, bd, dst, f.work.pitch[c]);
        }
    }
}

// shared memory of one CTA without the residual area
static size_t recon_fixed_smem(int log2_ctb, int nplanes, int hs, int vs, int cfi)
{
    const int cs = 1 << log2_ctb;
    size_t words = 0;
    for (int c = 0; c < nplanes; c++) {
        const int h = c ? hs : 0, v = c ? vs : 0;
        words += (size_t)(((cs >> h) >> 1) + 2) * (cs >> v);
    }
    words += 18 + MAX_CTB_TUS / 2 + 4 + 4;
    return words * 4 + (cfi == 3 ? RC_WARPS * sizeof(WarpScratch) + 16 : 0);
}

// Units (128 B) of the residual area: per-warp scratch + a full CTB of coefficients.
static int rc_area_units(int log2_ctb, int nplanes, int hs, int vs)
{
    const int need = (1 << (2 * log2_ctb)) + (nplanes == 3 ? 2 * ((1 << (2 * log2_ctb)) >> (hs + vs)) : 0);
    return (RC_WARPS * (int)sizeof(WarpScratch) + 2 * need + POOL_UNIT - 1) / POOL_UNIT;
}

size_t recon_smem_bytes(const FrameDev &f, int area_units)
{
    return (size_t)area_units * POOL_UNIT + recon_fixed_smem(f.log2_ctb, f.nplanes, f.hs, f.vs, f.cfi);
}

int recon_sparse_ctb_coefs(int log2_ctb, int nplanes, int hs, int vs) { return coef_area(rc_area_units(log2_ctb, nplanes, hs, vs)); }

// Function attributes are per DEVICE: ohcu_create calls this after cudaSetDevice, once per context
// (a process-wide flag left the second GPU of a process without the > 48 KB opt-in).
int recon_set_attrs()
{
    if (cudaFuncSetAttribute(residual_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024) != cudaSuccess ||
        cudaFuncSetAttribute(residual_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024) != cudaSuccess)
        return -1;
    return 0;
}

int launch_recon(const FrameDev &f, int flags, cudaStream_t st)
{
    const int nctb = f.ctb_w * f.ctb_h;
    int n = 0;
    if ((flags & 1) && f.n_pu > 0) {
        if (f.bd > 8) mc_kernel<true><<<nctb, RC_THREADS, 0, st>>>(f);
        else          mc_kernel<false><<<nctb, RC_THREADS, 0, st>>>(f);
        n++;
    }
    if ((flags & 2) && f.n_tu > 0) {
        const int hu = rc_area_units(f.log2_ctb, f.nplanes, f.hs, f.vs);
        const size_t smem = recon_smem_bytes(f, hu);
        if (f.bd > 8) residual_kernel<true><<<nctb, RC_THREADS, smem, st>>>(f, hu);
        else          residual_kernel<false><<<nctb, RC_THREADS, smem, st>>>(f, hu);
        n++;
    }
    return n;
}

// ================================================================= per-entry test shims
// One block through the SAME device code as the frame kernels (mc_filter + epi_pair,
// tb_residual, idct_mma, tile_add_pair), behind entry points with the argument meaning of
// the reference's table (ohcu.h: ohcu_dsp_*).  Test infrastructure for tests/test_gpu_leaves.py.
struct DspMc {
    const uint8_t *win;       // device copy of the source window: rows h + NT - 1, first needed sample at column 0
    int pitch;                // bytes, multiple of 16
    int w, h, mx, my, luma, family, bd;        // family: 0 put, 1 uni, 2 uni_w, 3 bi, 4 bi_w
    Epi e;
    const int16_t *src2;      // bi: first list's intermediates, stride 64 (MAX_PB_SIZE, hevcdec.h:51)
    int16_t *out16;           // put: stride 64
    uint8_t *dst;             // w x h pixels, pitch dpitch bytes
    int dpitch;
};

template <bool HIGH, int NT, int OW, int OH>
__device__ __forceinline__ void dsp_mc_cell(const DspMc &a, int cx, int cy)
{
    uint32_t res[OH][OW / 2];
    const int sh1 = a.bd - 8;
    if (NT == 8)
        mc_filter<HIGH, NT, OW, OH>(a.win, a.pitch, cx * OW, cy * OH, c_qpel4[a.mx][0], c_qpel4[a.mx][1], c_qpel4[a.my][0],
                                    c_qpel4[a.my][1], sh1, res);
    else
        mc_filter<HIGH, NT, OW, OH>(a.win, a.pitch, cx * OW, cy * OH, c_epel4[a.mx][0], 0, c_epel4[a.my][0], 0, sh1, res);
#pragma unroll
    for (int r = 0; r < OH; r++)
#pragma unroll
        for (int k = 0; k < OW / 2; k++) {
            const int x = cx * OW + 2 * k, y = cy * OH + r;
            if (x >= a.w || y >= a.h) continue;
            if (a.family == 0) {
                *(uint32_t *)&a.out16[y * 64 + x] = res[r][k];
                continue;
            }
            const uint32_t prev = a.family >= 3 ? *(const uint32_t *)&a.src2[y * 64 + x] : 0u;
            uint32_t o;
            epi_pair(a.e, res[r][k], prev, &o);
            if (HIGH) *(uint32_t *)(a.dst + (size_t)y * a.dpitch + 2 * x) = o;
            else *(uint16_t *)(a.dst + (size_t)y * a.dpitch + x) = (uint16_t)((o & 0xffu) | ((o >> 8) & 0xff00u));
        }
}

template <bool HIGH>
__global__ void dsp_mc_kernel(const DspMc a)
{
    const int ow = a.luma ? 4 : 2, oh = a.luma ? 4 : 2;
    const int ncx = (a.w + ow - 1) / ow, ncy = (a.h + oh - 1) / oh;
    for (int i = threadIdx.x; i < ncx * ncy; i += blockDim.x) {
        const int cy = i / ncx, cx = i - cy * ncx;
        if (a.luma) dsp_mc_cell<HIGH, 8, 4, 4>(a, cx, cy);
        else dsp_mc_cell<HIGH, 4, 2, 2>(a, cx, cy);
    }
}

// scratch shared by the shim entry points (tests are single-threaded)
static uint8_t *dsp_scratch(size_t bytes)
{
    static uint8_t *buf = nullptr;
    static size_t cap = 0;
    if (bytes > cap) {
        if (buf) cudaFree(buf);
        cap = bytes < (1u << 20) ? (1u << 20) : bytes;
        if (cudaMalloc((void **)&buf, cap) != cudaSuccess) { buf = nullptr; cap = 0; }
    }
    return buf;
}
static int dsp_tables()
{
    static bool done = false;
    if (!done) {
        if (upload_mc_taps() || upload_dct_table()) return -1;
        done = true;
    }
    return 0;
}

int dsp_mc(int bd, int luma, int family, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src, ptrdiff_t srcstride,
           const int16_t *src2, int16_t *out16, int h, int denom, int wx0, int wx1, int ox0, int ox1, int mx, int my, int w)
{
    if (dsp_tables()) return -1;
    const int P = bd > 8 ? 2 : 1, NT = luma ? 8 : 4, BEFORE = luma ? 3 : 1;
    if (w < 2 || w > 64 || h < 2 || h > 64 || (w & 1) || (h & 1) || mx < 0 || my < 0 || mx > (luma ? 3 : 7) || my > (luma ? 3 : 7)) return -1;
    const int rows = h + NT - 1, pitch = (((w + NT + 8) * P) + 15) & ~15;
    const size_t win_bytes = (size_t)rows * pitch, dpitch = (size_t)((w * P + 15) & ~15);
    uint8_t *base = dsp_scratch(win_bytes + 64 * 64 * 2 * 2 + dpitch * h + 256);
    if (!base) return -1;
    uint8_t *d_win = base, *d_src2 = base + ((win_bytes + 255) & ~(size_t)255), *d_out16 = d_src2 + 64 * 64 * 2, *d_dst = d_out16 + 64 * 64 * 2;
    cudaMemset(d_win, 0, win_bytes);
    // the window as the reference reads it: BEFORE rows / columns ahead of src, NT - 1 - BEFORE behind
    const uint8_t *s0 = src - (ptrdiff_t)BEFORE * srcstride - BEFORE * P;
    if (cudaMemcpy2D(d_win, pitch, s0, srcstride, (size_t)(w + NT - 1) * P, rows, cudaMemcpyHostToDevice) != cudaSuccess) return -1;
    if (family >= 3 && cudaMemcpy(d_src2, src2, 64 * 64 * 2, cudaMemcpyHostToDevice) != cudaSuccess) return -1;
    DspMc a{};
    a.win = d_win; a.pitch = pitch; a.w = w; a.h = h; a.mx = mx; a.my = my; a.luma = luma; a.family = family; a.bd = bd;
    a.src2 = (const int16_t *)d_src2; a.out16 = (int16_t *)d_out16; a.dst = d_dst; a.dpitch = (int)dpitch;
    Epi &e = a.e;                                  // the table of `Epi` above
    e.maxv = (1 << bd) - 1; e.wB = 0; e.post = 0;
    const int sc = 1 << (bd - 8);
    if (family == 1) { e.wA = 1; e.sh = 14 - bd; e.add = bd < 14 ? 1 << (13 - bd) : 0; }
    else if (family == 2) { e.wA = wx0; e.sh = denom + 14 - bd; e.add = bd < 14 ? 1 << (e.sh - 1) : 0; e.post = ox0 * sc; }
    else if (family == 3) { e.wA = 1; e.wB = 1; e.sh = 15 - bd; e.add = bd < 14 ? 1 << (14 - bd) : 0; }
    else if (family == 4) {
        const int log2Wd = denom + 14 - bd;
        e.wA = wx1; e.wB = wx0; e.sh = log2Wd + 1; e.add = (ox0 * sc + ox1 * sc + 1) * (1 << log2Wd);
    } else { e.wA = 1; e.sh = 0; e.add = 0; }
    if (bd > 8) dsp_mc_kernel<true><<<1, 256>>>(a);
    else dsp_mc_kernel<false><<<1, 256>>>(a);
    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
    if (family == 0) {
        // rows of 64 int16; only the block is defined
        return cudaMemcpy2D(out16, 64 * 2, d_out16, 64 * 2, (size_t)w * 2, h, cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -1;
    }
    return cudaMemcpy2D(dst, dststride, d_dst, dpitch, (size_t)w * P, h, cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -1;
}

// residual entries: idct[], idct_dc[], idct_4x4_luma, transform_skip, transform_rdpcm (in place on coeffs)
template <int LOG2>
__global__ void dsp_residual_kernel(int16_t *coef, ohcu_tu tu, int bd, int use_mma)
{
    constexpr int N = 1 << LOG2, S = Stride<N>::S;
    __shared__ __align__(16) int16_t src[N * N];
    __shared__ __align__(16) int16_t t1[N * (N + 8)];
    __shared__ __align__(16) int16_t blk[N * S];
    __shared__ uint32_t outw[N * N / 2];
    const int lane = threadIdx.x;
    if constexpr (LOG2 >= 4) {
        if (use_mma) {
            for (int i = lane; i < N * N; i += 32) src[i] = coef[i];
            __syncwarp();
            idct_mma<LOG2>(src, t1, outw, N / 2, tu.col_limit, bd, lane, true);
            __syncwarp();
            for (int i = lane; i < N * N / 2; i += 32) ((uint32_t *)coef)[i] = outw[i];
            return;
        }
    }
    if (lane < N) {
        const unsigned mask = N == 32 ? 0xffffffffu : ((1u << N) - 1u);
        tb_residual<LOG2>(bd, tu.mode, tu.flags, tu.col_limit, coef, blk, lane, mask);
        for (int k = 0; k < N; k++) coef[lane * N + k] = blk[lane * S + k];
    }
}

int dsp_residual(int bd, int mode, int flags, int16_t *coeffs, int log2, int col_limit, int use_mma)
{
    if (dsp_tables() || log2 < 2 || log2 > 5) return -1;
    const size_t bytes = (size_t)2 << (2 * log2);
    int16_t *d = (int16_t *)dsp_scratch(bytes);
    if (!d || cudaMemcpy(d, coeffs, bytes, cudaMemcpyHostToDevice) != cudaSuccess) return -1;
    ohcu_tu tu{};
    tu.log2_size = (uint8_t)log2; tu.mode = (uint8_t)mode; tu.flags = (uint8_t)flags; tu.col_limit = (int8_t)col_limit;
    switch (log2) {
    case 2: dsp_residual_kernel<2><<<1, 32>>>(d, tu, bd, 0); break;
    case 3: dsp_residual_kernel<3><<<1, 32>>>(d, tu, bd, 0); break;
    case 4: dsp_residual_kernel<4><<<1, 32>>>(d, tu, bd, use_mma); break;
    default: dsp_residual_kernel<5><<<1, 32>>>(d, tu, bd, use_mma); break;
    }
    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
    return cudaMemcpy(coeffs, d, bytes, cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -1;
}

// transform_add[] (template :45-84): dst block += coeffs, clipped
template <bool HIGH>
__global__ void dsp_transform_add_kernel(uint8_t *dst, int dpitch, const int16_t *coef, int n, int bd)
{
    const int maxv = (1 << bd) - 1;
    for (int i = threadIdx.x; i < n * n / 2; i += blockDim.x) {
        const int y = i / (n / 2), x = 2 * (i - y * (n / 2));
        uint8_t *p = dst + (size_t)y * dpitch + x * (HIGH ? 2 : 1);
        uint32_t w = HIGH ? *(uint32_t *)p : ((uint32_t)p[0] | ((uint32_t)p[1] << 16));
        tile_add_pair(&w, coef[y * n + x], coef[y * n + x + 1], maxv);
        if (HIGH) *(uint32_t *)p = w;
        else { p[0] = (uint8_t)w; p[1] = (uint8_t)(w >> 16); }
    }
}

int dsp_transform_add(int bd, uint8_t *dst, const int16_t *coeffs, ptrdiff_t stride, int n)
{
    const int P = bd > 8 ? 2 : 1;
    if (n != 4 && n != 8 && n != 16 && n != 32) return -1;
    const size_t dpitch = (size_t)((n * P + 15) & ~15);
    uint8_t *base = dsp_scratch(dpitch * n + (size_t)n * n * 2 + 256);
    if (!base) return -1;
    int16_t *d_c = (int16_t *)(base + ((dpitch * n + 255) & ~(size_t)255));
    if (cudaMemcpy2D(base, dpitch, dst, stride, (size_t)n * P, n, cudaMemcpyHostToDevice) != cudaSuccess) return -1;
    if (cudaMemcpy(d_c, coeffs, (size_t)n * n * 2, cudaMemcpyHostToDevice) != cudaSuccess) return -1;
    if (bd > 8) dsp_transform_add_kernel<true><<<1, 128>>>(base, (int)dpitch, d_c, n, bd);
    else dsp_transform_add_kernel<false><<<1, 128>>>(base, (int)dpitch, d_c, n, bd);
    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
    return cudaMemcpy2D(dst, stride, base, dpitch, (size_t)n * P, n, cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -1;
}

}  // namespace ohcu

// ================================================================= ohcu_dsp_idct_noalloc
// The idct[] table entry once more (ohcu_dsp_idct above), for callers that must not allocate: the device
// staging of the blocks is a file-scope array instead of dsp_scratch()'s cudaMalloc.  No counterpart in the
// reference (include/ohcu.h).
//   * g_noalloc_scratch lives in every GPU that loads the library until the process exits (64 KB).
//   * In HOST code its name is a host-side stand-in, not a device address: the address a kernel may be given
//     comes from cudaGetSymbolAddress, looked up once per device with that device current.
//   * Every call shares the array and finds the previous call's contents in it.  Nothing has to start at zero:
//     a call copies its blocks in before the kernel reads them and reads back only what it wrote.  Calls that
//     overlap on the GPU (two threads, two streams) race on it - this entry point is serialised by a mutex and
//     runs synchronously on the default stream.
static __device__ __align__(256) unsigned char g_noalloc_scratch[OHCU_NOALLOC_MAX_BYTES];

namespace {
std::mutex g_noalloc_mu;
constexpr int NOALLOC_MAX_DEVICES = 64;
int16_t *g_noalloc_ptr[NOALLOC_MAX_DEVICES];        // device address of g_noalloc_scratch, per device (nullptr = not set up)
}

extern "C" int ohcu_dsp_idct_noalloc(int bd, int16_t *coeffs, int log2_size, int col_limit, int nblocks)
{
    using namespace ohcu;
    if (!coeffs || log2_size < 2 || log2_size > 5 || nblocks < 1)
        return set_error(OHCU_EINVAL, "ohcu_dsp_idct_noalloc: bad argument (log2_size %d, nblocks %d)", log2_size, nblocks);
    if (bd != 8 && bd != 9 && bd != 10 && bd != 12 && bd != 14)
        return set_error(OHCU_EINVAL, "ohcu_dsp_idct_noalloc: unsupported bit depth %d", bd);
    const size_t block = (size_t)2 << (2 * log2_size), bytes = block * (size_t)nblocks;
    if (bytes > (size_t)OHCU_NOALLOC_MAX_BYTES)
        return set_error(OHCU_EINVAL, "ohcu_dsp_idct_noalloc: %d blocks of %dx%d need %zu bytes, the static scratch holds %d "
                         "(OHCU_NOALLOC_MAX_BYTES); split the call or use ohcu_dsp_idct", nblocks, 1 << log2_size, 1 << log2_size,
                         bytes, OHCU_NOALLOC_MAX_BYTES);
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return set_error(OHCU_ENODEV, "ohcu_dsp_idct_noalloc: no CUDA device");
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return set_error(OHCU_EINVAL, "ohcu_dsp_idct_noalloc: device %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        // set-up, once per device, with that device current: the transform constants (per-device symbols) and the
        // scratch array's DEVICE address
        void *ptr = nullptr;
        if (upload_dct_table() || upload_mc_taps()) return set_error(OHCU_ECUDA, "ohcu_dsp_idct_noalloc: constant upload failed");
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return set_error(OHCU_ECUDA, "ohcu_dsp_idct_noalloc: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
        g_noalloc_ptr[dev] = (int16_t *)ptr;
    }
    int16_t *d = g_noalloc_ptr[dev];
    cudaError_t e = cudaMemcpy(d, coeffs, bytes, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return set_error(OHCU_ECUDA, "ohcu_dsp_idct_noalloc: upload -> %s", cudaGetErrorString(e));
    ohcu_tu tu{};
    tu.log2_size = (uint8_t)log2_size; tu.mode = OHCU_TU_IDCT; tu.col_limit = (int8_t)col_limit;
    const size_t n16 = block / 2;
    for (int b = 0; b < nblocks; b++) {
        switch (log2_size) {
        case 2: dsp_residual_kernel<2><<<1, 32>>>(d + b * n16, tu, bd, 0); break;
        case 3: dsp_residual_kernel<3><<<1, 32>>>(d + b * n16, tu, bd, 0); break;
        case 4: dsp_residual_kernel<4><<<1, 32>>>(d + b * n16, tu, bd, 1); break;
        default: dsp_residual_kernel<5><<<1, 32>>>(d + b * n16, tu, bd, 1); break;
        }
    }
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(coeffs, d, bytes, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return set_error(OHCU_ECUDA, "ohcu_dsp_idct_noalloc: %s", cudaGetErrorString(e));
    return 0;
}
