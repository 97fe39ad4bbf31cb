// intra.cu — intra prediction of one transform block on the device (first step).
//
// Replaces, per block, HEVCPredContext.intra_pred[log2 - 2](s, x0, y0, c_idx) and the three predictors it dispatches
// to (hevcpred.h:30-41; hevcpred_template.c: reference samples and substitution :30-272, smoothing :274-326,
// pred_planar :359-372, pred_dc :388-417, pred_angular :419-505; call sites hevcdec.c:1335-1526).
// constrained_intra_pred_flag == 0 (the CIP substitution, :112-150 / :172-243, stays with the caller for now).
//
// intra_pred_tb is ONE WARP's work: the 4N + 1 reference samples are gathered into shared memory in parallel,
// the substitution rules run as a handful of warp-uniform steps, smoothing and the predictor are one sample per
// lane and step.  A block depends on the reconstructed samples left of and above it, so the frame-level kernel will
// walk the CTBs as a wavefront with the residual add in between; this file holds the block
// function that kernel will call and the per-entry shim that tests it 1:1 against the reference table.
#include "ohcu_internal.h"

namespace ohcu {

enum { CAND_BL = 1, CAND_L = 2, CAND_UL = 4, CAND_U = 8, CAND_UR = 16 };

struct IntraScratch {                 // per warp
    int16_t raw[2][66];               // [0] left, [1] top; index i + 1 holds sample i, index 0 the corner
    int16_t flt[2][66];               // smoothed copies
    int16_t ref[100];                 // angular: ref[k + 32], k = -32 .. 2N
};

__constant__ int8_t c_intra_angle[33] = {32, 26, 21, 17, 13, 9, 5, 2, 0, -2, -5, -9, -13, -17, -21, -26, -32,
                                         -26, -21, -17, -13, -9, -5, -2, 0, 2, 5, 9, 13, 17, 21, 26, 32};
__constant__ int16_t c_intra_inv_angle[15] = {-4096, -1638, -910, -630, -482, -390, -315, -256, -315, -390, -482, -630, -910, -1638, -4096};

// plane: sample (0, 0) of the plane, pitch in samples; (x, y) the block; cand: CAND_* final availability;
// filter_ok: !intra_smoothing_disabled && (c_idx == 0 || chroma_format_idc == 3); strong: sps strong smoothing flag;
// res: the block's residual (first sample, pitch in int16) added to the prediction with the clip of transform_add
// (hevcdsp_template.c:45-84), or nullptr.  Neighbours are read past the L1 (another SM may have written them).
template <typename pixel>
__device__ void intra_pred_tb(pixel *plane, int pitch, int pw, int ph, int x, int y, int log2_size, int c_idx, int mode, int cand,
                              int filter_ok, int strong, int bd, IntraScratch &s, int lane, const int16_t *res = nullptr,
                              int res_pitch = 0)
{
    const int n = 1 << log2_size, n2 = 2 * n;
    pixel *blk = plane + (size_t)y * pitch + x;
    const int maxv = (1 << bd) - 1;
    auto put = [&](int xx, int yy, int v) {
        if (res) v = min(max(v + (int)__ldcg(res + (size_t)yy * res_pitch + xx), 0), maxv);
        blk[(size_t)yy * pitch + xx] = (pixel)v;
    };
    const int bl_size = max(0, min(y + n2, ph) - (y + n)), tr_size = max(0, min(x + n2, pw) - (x + n));
    int16_t *left = s.raw[0] + 1, *top = s.raw[1] + 1;
    bool bl = cand & CAND_BL, lf = cand & CAND_L, ul = cand & CAND_UL, up = cand & CAND_U, ur = cand & CAND_UR;

    // ---- what exists (:152-171); runs that leave the picture repeat their last sample
    for (int i = lane; i < n2; i += 32) {
        int t = 0, l = 0;
        if (i < n) {
            if (up) t = __ldcg(blk - pitch + i);
            if (lf) l = __ldcg(blk + (ptrdiff_t)i * pitch - 1);
        } else {
            if (ur) t = __ldcg(blk - pitch + n + min(i - n, tr_size - 1));
            if (bl) l = __ldcg(blk + (ptrdiff_t)(n + min(i - n, bl_size - 1)) * pitch - 1);
        }
        top[i] = (int16_t)t;
        left[i] = (int16_t)l;
    }
    int corner = ul ? (int)__ldcg(blk - pitch - 1) : 0;
    __syncwarp();
    // ---- substitution (:245-272), warp-uniform decisions
    if (!bl) {
        if (lf) {
            const int v = left[n - 1];
            for (int i = n + lane; i < n2; i += 32) left[i] = (int16_t)v;
        } else {
            int v;
            const bool any = ul || up || ur;
            if (ul) v = corner;
            else if (up) { v = top[0]; ul = true; }
            else if (ur) { v = top[n]; up = ul = true; for (int i = lane; i < n; i += 32) top[i] = (int16_t)v; }
            else { v = 1 << (bd - 1); for (int i = lane; i < n2; i += 32) top[i] = (int16_t)v; }
            corner = v;
            for (int i = lane; i < n2; i += 32) left[i] = (int16_t)v;
            if (any) lf = true;
        }
        __syncwarp();
    }
    if (!lf) {
        const int v = left[n];
        for (int i = lane; i < n; i += 32) left[i] = (int16_t)v;
        __syncwarp();
    }
    if (!ul) corner = left[0];
    if (!up) for (int i = lane; i < n; i += 32) top[i] = (int16_t)corner;
    if (!ur) {
        __syncwarp();
        const int v = top[n - 1];
        for (int i = n + lane; i < n2; i += 32) top[i] = (int16_t)v;
    }
    if (lane == 0) left[-1] = top[-1] = (int16_t)corner;
    __syncwarp();

    // ---- smoothing (:274-326)
    if (filter_ok && mode != 1 && n != 4) {
        const int dist = min(abs(mode - 26), abs(mode - 10));
        const int thresh = log2_size == 3 ? 7 : (log2_size == 4 ? 1 : 0);
        if (dist > thresh) {
            int16_t *fl = s.flt[0] + 1, *ft = s.flt[1] + 1;
            const bool bilinear = strong && c_idx == 0 && log2_size == 5 && abs(top[-1] + top[63] - 2 * top[31]) < (1 << (bd - 5)) &&
                                  abs(left[-1] + left[63] - 2 * left[31]) < (1 << (bd - 5));
            for (int i = lane - 1; i < n2; i += 32) {
                int a, b;
                if (bilinear) {
                    a = (i < 0 || i == 63) ? left[i] : ((63 - i) * left[-1] + (i + 1) * left[63] + 32) >> 6;
                    b = (i < 0 || i == 63) ? top[i] : ((63 - i) * top[-1] + (i + 1) * top[63] + 32) >> 6;
                } else if (i < 0) {
                    a = b = (left[0] + 2 * left[-1] + top[0] + 2) >> 2;
                } else if (i == n2 - 1) {
                    a = left[i]; b = top[i];
                } else {
                    a = (left[i + 1] + 2 * left[i] + left[i - 1] + 2) >> 2;
                    b = (top[i + 1] + 2 * top[i] + top[i - 1] + 2) >> 2;
                }
                fl[i] = (int16_t)a;
                ft[i] = (int16_t)b;
            }
            __syncwarp();
            left = fl; top = ft;
        }
    }

    // ---- predictors: lanes run along rows of the block (coalesced stores)
    const bool edge = c_idx == 0 && n < 32;
    if (mode == 0) {                                       // planar, :359-372
        for (int i = lane; i < n * n; i += 32) {
            const int xx = i & (n - 1), yy = i >> log2_size;
            put(xx, yy, ((n - 1 - xx) * left[yy] + (xx + 1) * top[n] + (n - 1 - yy) * top[xx] + (yy + 1) * left[n] + n) >> (log2_size + 1));
        }
    } else if (mode == 1) {                                // DC, :388-417
        int sum = 0;
        for (int i = lane; i < n; i += 32) sum += left[i] + top[i];
#pragma unroll
        for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const int dc = (sum + n) >> (log2_size + 1);
        for (int i = lane; i < n * n; i += 32) {
            const int xx = i & (n - 1), yy = i >> log2_size;
            int v = dc;
            if (edge) {
                if (!(xx | yy)) v = (left[0] + 2 * dc + top[0] + 2) >> 2;
                else if (!yy) v = (top[xx] + 3 * dc + 2) >> 2;
                else if (!xx) v = (left[yy] + 3 * dc + 2) >> 2;
            }
            put(xx, yy, v);
        }
    } else {                                               // angular, :419-505
        const int angle = c_intra_angle[mode - 2];
        const bool vertical = mode >= 18;
        const int16_t *mn = vertical ? top : left, *sd = vertical ? left : top;
        int16_t *ref = s.ref + 32;
        const int last = (n * angle) >> 5;
        for (int k = lane; k <= n2; k += 32) ref[k] = mn[k - 1];
        if (angle < 0 && last < -1) {
            const int inv = c_intra_inv_angle[mode - 11];
            for (int k = last + lane; k <= -1; k += 32) ref[k] = sd[-1 + ((k * inv + 128) >> 8)];
        }
        __syncwarp();
        for (int i = lane; i < n * n; i += 32) {
            const int xx = i & (n - 1), yy = i >> log2_size;
            const int a = vertical ? yy : xx, b = vertical ? xx : yy;
            const int idx = ((a + 1) * angle) >> 5, fact = ((a + 1) * angle) & 31;
            int v = fact ? ((32 - fact) * ref[b + idx + 1] + fact * ref[b + idx + 2] + 16) >> 5 : ref[b + idx + 1];
            if (edge) {
                if (mode == 26 && xx == 0) v = min(max(top[0] + ((left[yy] - left[-1]) >> 1), 0), maxv);
                if (mode == 10 && yy == 0) v = min(max(left[0] + ((top[xx] - top[-1]) >> 1), 0), maxv);
            }
            put(xx, yy, v);
        }
    }
    __syncwarp();
}


// ================================================================= frame level: the CTB wavefront
// hls_transform_unit predicts a block right before it adds the block's residual (hevcdec.c:1335-1526), in decoding
// order; a block reads only samples of blocks decoded earlier: its own CTB, and the CTBs left, up-left, up and
// up-right of it (the z-scan availability test, hevcpred_template.c:103-107).  Everything that is not intra has been
// reconstructed by mc_kernel / residual_kernel / pcm_kernel before this kernel starts, and residual_kernel has left
// the intra blocks' residuals in the job's residual planes.  What remains is the dependency between intra blocks:
//   * one CTA per CTB row walks its CTBs left to right; before a CTB with intra blocks it waits until the row above
//     has finished the CTB up-right of it (progress counter per row: CTBs finished) - the wavefront the reference's
//     WPP threads run (hls_decode_entry_wpp, hevcdec.c:2966-3051);
//   * inside a CTB the three planes are independent (no cross-component prediction between predicted samples):
//     warp c takes plane c's blocks in list order.
// A CTB without intra blocks costs a counter store.
constexpr int INTRA_THREADS = 96;
constexpr int INTRA_MAX_IU = 768;          // per CTB: 256 4x4 blocks in each of three planes

__device__ __forceinline__ int ld_acquire(const int *p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int *p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <typename pixel>
__global__ void __launch_bounds__(INTRA_THREADS) intra_kernel(const __grid_constant__ FrameDev f)
{
    __shared__ IntraScratch scratch[3];
    __shared__ ohcu_iu s_iu[INTRA_MAX_IU];
    extern __shared__ uint2 s_index[];                     // the row's ohcu_iu_index entries
    if (f.err && *f.err) return;                           // malformed lists: every CTA leaves, nobody waits
    const int row = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < f.ctb_w; i += INTRA_THREADS) {
        const ohcu_iu_index ix = f.iu_index[row * f.ctb_w + i];
        s_index[i] = make_uint2(ix.first_iu, ix.n_iu);
    }
    __syncthreads();
    int *const progress = f.intra_progress;
    const int filter_luma = !f.intra_smoothing_disabled, filter_chroma = filter_luma && f.cfi == 3;
    for (int cx = 0; cx < f.ctb_w; cx++) {
        const uint2 ix = s_index[cx];
        if (ix.y) {
            for (uint32_t i = tid; i < ix.y; i += INTRA_THREADS) s_iu[i] = f.ius[ix.x + i];
            if (row > 0 && tid == 0) {
                const int need = min(cx + 2, f.ctb_w);
                while (ld_acquire(progress + row - 1) < need) __nanosleep(40);
                __threadfence();
            }
            __syncthreads();
            if (warp < f.nplanes) {
                pixel *plane = (pixel *)f.work.p[warp];
                const int pitch = f.work.pitch[warp] / (int)sizeof(pixel), rp = f.resid_pitch[warp] >> 1;
                for (uint32_t i = 0; i < ix.y; i++) {
                    const ohcu_iu u = s_iu[i];
                    if (u.c_idx != warp) continue;
                    const int16_t *res = (u.cand & OHCU_IU_CBF)
                                             ? (const int16_t *)((const uint8_t *)f.resid[warp] + (size_t)u.y0 * f.resid_pitch[warp]) + u.x0
                                             : nullptr;
                    intra_pred_tb<pixel>(plane, pitch, f.pw[warp], f.ph[warp], u.x0, u.y0, u.log2_size, warp, u.mode, u.cand & 31,
                                         warp ? filter_chroma : filter_luma, f.strong_intra_smoothing, f.bd, scratch[warp], lane, res, rp);
                }
            }
            __syncthreads();                               // every plane of the CTB is written
        }
        if (tid == 0) {
            __threadfence();
            st_release(progress + row, cx + 1);
        }
    }
}

int launch_intra(const FrameDev &f, cudaStream_t st)
{
    if (f.n_iu <= 0) return 0;
    if (!f.ius || !f.iu_index || !f.intra_progress || !f.resid[0]) return -1;
    const size_t smem = sizeof(uint2) * f.ctb_w;
    if (f.bd > 8) intra_kernel<uint16_t><<<f.ctb_h, INTRA_THREADS, smem, st>>>(f);
    else          intra_kernel<uint8_t><<<f.ctb_h, INTRA_THREADS, smem, st>>>(f);
    return 1;
}

// ================================================================= per-entry test shim
template <typename pixel>
__global__ void dsp_intra_kernel(pixel *plane, int pitch, int pw, int ph, int x, int y, int log2_size, int c_idx, int mode, int cand,
                                 int filter_ok, int strong, int bd)
{
    __shared__ IntraScratch s;
    intra_pred_tb<pixel>(plane, pitch, pw, ph, x, y, log2_size, c_idx, mode, cand, filter_ok, strong, bd, s, threadIdx.x);
}

int dsp_intra_pred(int bd, uint8_t *plane, ptrdiff_t stride, int pw, int ph, int x, int y, int log2_size, int c_idx, int mode, int cand,
                   int filter_ok, int strong)
{
    const int P = bd > 8 ? 2 : 1, n = 1 << log2_size;
    if (log2_size < 2 || log2_size > 5 || mode < 0 || mode > 34 || c_idx < 0 || c_idx > 2 || x < 0 || y < 0 || x + n > pw || y + n > ph ||
        stride < (ptrdiff_t)pw * P || (stride % P) != 0)
        return -1;
    // neighbours named available must lie inside the plane
    if (((cand & (CAND_L | CAND_BL | CAND_UL)) && x < 1) || ((cand & (CAND_U | CAND_UR | CAND_UL)) && y < 1)) return -1;
    uint8_t *d = nullptr;
    const size_t bytes = (size_t)stride * ph;
    if (cudaMalloc((void **)&d, bytes) != cudaSuccess) return -1;
    int rc = -1;
    if (cudaMemcpy(d, plane, bytes, cudaMemcpyHostToDevice) == cudaSuccess) {
        if (bd > 8) dsp_intra_kernel<uint16_t><<<1, 32>>>((uint16_t *)d, (int)(stride / 2), pw, ph, x, y, log2_size, c_idx, mode, cand, filter_ok, strong, bd);
        else dsp_intra_kernel<uint8_t><<<1, 32>>>(d, (int)stride, pw, ph, x, y, log2_size, c_idx, mode, cand, filter_ok, strong, bd);
        if (cudaDeviceSynchronize() == cudaSuccess && cudaMemcpy(plane, d, bytes, cudaMemcpyDeviceToHost) == cudaSuccess) rc = 0;
    }
    cudaFree(d);
    return rc;
}

}  // namespace ohcu
