// residual.cu — batched inverse transforms + transform_add for every coded TB of a picture.
//
// Replaces the per-TB calls of ff_hevc_hls_transform / _c (hevc_cabac.c:1695-1906):
// transform_rdpcm / transform_skip / idct_4x4_luma / idct_dc / idct (col_limit zero-skip)
// / cross-component prediction, followed by transform_add (hevcdsp_template.c:45-296).
//
// TBs are binned by size on the host (append_tu); one launch per size class.  A group of
// N lanes owns one NxN block held in shared memory: lane g transforms column g in the
// first pass and row g in the second, with the partial butterflies run on CUDA cores out
// of registers; int32 accumulation and the int16 clip after each pass as in the reference.
#include "ohcu_internal.h"

namespace ohcu {

// transform[32][32] (hevcdsp.c:32-97) as 32-bit constants so that unrolled butterflies
// take them straight from the constant bank.  Filled at load time by the host (api.cu)
// from the cosine index rule (see oracle/hevc_oracle.c or_dct_coef).
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

// N-point inverse DCT of in[0..N) (stride-1 registers) into out[0..N): even/odd
// decomposition, exact integer sums (TR_4..TR_32, template :180-239).
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

// one TB: coefficients -> residual, in place in blk[N][S]
template <int BD, int LOG2>
__device__ void tb_residual(const FrameDev &f, const ohcu_tu &tu, int16_t *blk, int g, unsigned mask)
{
    constexpr int N = 1 << LOG2, S = Stride<N>::S;
    // load N*N int16 (pairs)
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
#pragma unroll
            for (int j = 0; j < N; j++) in[j] = idct_reads<N>(j, limit) ? blk[g * S + j] : 0;
            idct1d<N>(in, out);
#pragma unroll
            for (int k = 0; k < N; k++)
                blk[g * S + k] = (int16_t)clip_s16((out[k] + (1 << (19 - BD))) >> (20 - BD));
        }
    } else if (mode == OHCU_TU_DC) {                           // IDCT_DC, template :273-286
        const int v = (((blk[0] + 1) >> 1) + (1 << (13 - BD))) >> (14 - BD);
        __syncwarp(mask);
#pragma unroll
        for (int k = 0; k < N; k++) blk[g * S + k] = (int16_t)v;
    } else if (mode == OHCU_TU_DST4) {                         // transform_4x4_luma, template :140-177
        if constexpr (N == 4) {
#pragma unroll
            for (int pass = 0; pass < 2; pass++) {
                const int st = pass ? 1 : S, base = pass ? g * S : g;
                const int shift = pass ? 20 - BD : 7, add = 1 << (shift - 1);
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
            constexpr int shift = 15 - BD - LOG2;              // transform_skip, template :109-133
#pragma unroll
            for (int k = 0; k < N; k++) {
                const int v = blk[g * S + k];
                blk[g * S + k] = shift > 0 ? (int16_t)((v + (1 << (shift > 0 ? shift - 1 : 0))) >> (shift > 0 ? shift : 0))
                                           : (int16_t)(v * (1 << (shift > 0 ? 0 : -shift)));
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

template <int BD, int LOG2>
__global__ void __launch_bounds__(128) residual_kernel(const __grid_constant__ FrameDev f)
{
    typedef typename Px<BD>::T pixel;
    constexpr int N = 1 << LOG2, S = Stride<N>::S, GROUPS = 128 / N;
    __shared__ __align__(16) int16_t sblk[GROUPS][2][N * S];
    const int grp = threadIdx.x >> LOG2, g = threadIdx.x & (N - 1);
    const int ti = blockIdx.x * GROUPS + grp;
    if (ti >= f.n_bin[LOG2 - 2]) return;
    const unsigned mask = N == 32 ? 0xffffffffu : (((1u << N) - 1u) << ((threadIdx.x & 31) & ~(N - 1)));
    const ohcu_tu tu = f.tus[f.tu_bin[LOG2 - 2][ti]];
    int16_t *blk = sblk[grp][0];

    tb_residual<BD, LOG2>(f, tu, blk, g, mask);
    if (tu.flags & OHCU_TUF_CROSS) {                           // hevc_cabac.c:1810-1817
        int16_t *by = sblk[grp][1];
        const ohcu_tu ty = f.tus[tu.cross_tu];
        if (ty.log2_size == LOG2) {
            tb_residual<BD, LOG2>(f, ty, by, g, mask);
            const int scale = tu.res_scale_val;
#pragma unroll
            for (int k = 0; k < N; k++)
                blk[g * S + k] = (int16_t)(blk[g * S + k] + ((scale * by[g * S + k]) >> 3));
            __syncwarp(mask);
        }
    }
    // transform_add (transquant_bypass, template :45-60): one row per step, lanes = columns
    const int c = tu.c_idx;
    uint8_t *plane = f.work.p[c];
    const int pitch = f.work.pitch[c];
#pragma unroll 4
    for (int r = 0; r < N; r++) {
        pixel *d = (pixel *)(plane + (size_t)(tu.y0 + r) * pitch) + tu.x0 + g;
        *d = (pixel)clip_px<BD>((int)*d + blk[r * S + g]);
    }
}

template <int BD>
static int launch_residual_bd(const FrameDev &f, cudaStream_t st)
{
    int n = 0;
    if (f.n_bin[0] > 0) { residual_kernel<BD, 2><<<(f.n_bin[0] + 31) / 32, 128, 0, st>>>(f); n++; }
    if (f.n_bin[1] > 0) { residual_kernel<BD, 3><<<(f.n_bin[1] + 15) / 16, 128, 0, st>>>(f); n++; }
    if (f.n_bin[2] > 0) { residual_kernel<BD, 4><<<(f.n_bin[2] + 7) / 8, 128, 0, st>>>(f); n++; }
    if (f.n_bin[3] > 0) { residual_kernel<BD, 5><<<(f.n_bin[3] + 3) / 4, 128, 0, st>>>(f); n++; }
    return n;
}

int launch_residual(const FrameDev &f, cudaStream_t st)
{
    if (f.n_tu <= 0) return 0;
    int n = 0;
    OHCU_DISPATCH_BD(f.bd, (n = launch_residual_bd<BD>(f, st)));
    return n;
}

}  // namespace ohcu
