// intra.cu — intra prediction on the device: the block function and the frame-level wavefront.
//
// Replaces, per block, HEVCPredContext.intra_pred[log2 - 2](s, x0, y0, c_idx) and the three predictors it dispatches
// to (hevcpred.h:30-41; hevcpred_template.c: reference samples and substitution :30-272, smoothing :274-326,
// pred_planar :359-372, pred_dc :388-417, pred_angular :419-505; call sites hevcdec.c:1335-1526).
// constrained_intra_pred_flag == 0 (the CIP substitution, :112-150 / :172-243, stays with the caller for now).
//
// intra_pred_tb is ONE WARP's work: the 4N + 1 reference samples are gathered into shared memory in parallel,
// the substitution rules run as a handful of warp-uniform steps, smoothing and the predictor are one sample per
// lane and step.  A block depends on the reconstructed samples left of and above it, so the frame-level kernel
// (intra_kernel below) walks the CTBs as a wavefront, with the CTB's samples and residuals staged in shared memory;
// the per-entry shim tests the block function 1:1 against the reference table.
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

// Where a block's samples live.  ld / st take coordinates relative to the block's first sample.
template <typename pixel> struct PlaneSurf {           // a plane in global memory (per-entry shim)
    pixel *blk; int pitch;                             // first sample of the block, pitch in samples
    __device__ __forceinline__ int ld(int x, int y) const { return (int)__ldcg(blk + (ptrdiff_t)y * pitch + x); }
    __device__ __forceinline__ void st(int x, int y, int v) const { blk[(ptrdiff_t)y * pitch + x] = (pixel)v; }
};
// The CTB's samples in shared memory (intra_kernel): row -1 (corner, top, top-right of the CTB) in `top`, the CTB's own
// rows with the column left of it in `body`; sample x of a row sits at index x + TILE_X0.  Stores go to the tile and
// through to the picture.
constexpr int TILE_X0 = 8;
template <typename pixel> struct TileSurf {
    uint16_t *top, *body; int bpitch;
    int bx, by;                                        // the block inside the CTB
    pixel *blk; int pitch;                             // the block in the picture
    __device__ __forceinline__ int ld(int x, int y) const
    {
        const int X = bx + x + TILE_X0, Y = by + y;
        return Y < 0 ? top[X] : body[Y * bpitch + X];
    }
    __device__ __forceinline__ void st(int x, int y, int v) const
    {
        body[(by + y) * bpitch + bx + x + TILE_X0] = (uint16_t)v;
        blk[(ptrdiff_t)y * pitch + x] = (pixel)v;
    }
};

// sf: the block's surface; tr_size / bl_size: samples of the top-right / bottom-left run inside the picture (:109-112);
// cand: CAND_* final availability; filter_ok: !intra_smoothing_disabled && (c_idx == 0 || chroma_format_idc == 3);
// strong: sps strong smoothing flag; res: the block's residual (first sample, pitch in int16) added to the prediction with
// the clip of transform_add (hevcdsp_template.c:45-84), or nullptr.
template <typename Surf>
__device__ void intra_pred_tb(const Surf &sf, int tr_size, int bl_size, int log2_size, int c_idx, int mode, int cand,
                              int filter_ok, int strong, int bd, IntraScratch &s, int lane, const int16_t *res = nullptr,
                              int res_pitch = 0)
{
    const int n = 1 << log2_size, n2 = 2 * n;
    const int maxv = (1 << bd) - 1;
    auto put = [&](int xx, int yy, int v) {
        if (res) v = min(max(v + (int)res[yy * res_pitch + xx], 0), maxv);
        sf.st(xx, yy, v);
    };
    int16_t *left = s.raw[0] + 1, *top = s.raw[1] + 1;
    bool bl = cand & CAND_BL, lf = cand & CAND_L, ul = cand & CAND_UL, up = cand & CAND_U, ur = cand & CAND_UR;

    // ---- what exists (:152-171); runs that leave the picture repeat their last sample
    for (int i = lane; i < n2; i += 32) {
        int t = 0, l = 0;
        if (i < n) {
            if (up) t = sf.ld(i, -1);
            if (lf) l = sf.ld(-1, i);
        } else {
            if (ur) t = sf.ld(n + min(i - n, tr_size - 1), -1);
            if (bl) l = sf.ld(-1, n + min(i - n, bl_size - 1));
        }
        top[i] = (int16_t)t;
        left[i] = (int16_t)l;
    }
    int corner = ul ? sf.ld(-1, -1) : 0;
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
//   * the CTB's samples (with the row above and the column left of it) and its residuals are staged in shared memory by
//     the whole CTA with 16-byte loads, so that the chain of blocks inside the CTB - each one reads what the one before
//     wrote - never waits for global memory; predicted samples are written to the tile and through to the picture;
//   * inside a CTB the three planes are independent: warp c takes plane c's blocks in list order.
// A CTB without intra blocks costs a counter store.
constexpr int INTRA_THREADS = 256;
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

// eight samples of a picture row (16-byte / 8-byte load past the L1) widened into the tile
template <typename pixel> __device__ __forceinline__ void load8(uint16_t *dst, const pixel *src)
{
    if (sizeof(pixel) == 2) {
        *(uint4 *)dst = __ldcg((const uint4 *)src);
    } else {
        const uint2 v = __ldcg((const uint2 *)src);
        uint4 o;
        o.x = __byte_perm(v.x, 0, 0x4140); o.y = __byte_perm(v.x, 0, 0x4342);
        o.z = __byte_perm(v.y, 0, 0x4140); o.w = __byte_perm(v.y, 0, 0x4342);
        *(uint4 *)dst = o;
    }
}

struct IntraTile {                         // one plane's staging (pointers into dynamic shared memory)
    uint16_t *top, *body;
    int16_t *res;
    int sw, sh, bpitch;                    // CTB size in samples of the plane, tile pitch
};

template <typename pixel>
__global__ void __launch_bounds__(INTRA_THREADS) intra_kernel(const __grid_constant__ FrameDev f)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ IntraScratch scratch[3];
    __shared__ ohcu_iu s_iu[INTRA_MAX_IU];
    if (f.err && *f.err) return;                           // malformed lists: every CTA leaves, nobody waits
    const int row = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = 1 << f.log2_ctb;
    // carve: per plane top (TILE_X0 + 2 sw), body (sh rows of sw + TILE_X0), residual (sh x sw); then the row's index
    IntraTile tile[3];
    size_t off = 0;
    for (int c = 0; c < 3; c++) {
        IntraTile &t = tile[c];
        t.sw = S >> (c ? f.hs : 0); t.sh = S >> (c ? f.vs : 0); t.bpitch = t.sw + TILE_X0;
        if (c >= f.nplanes) { t.top = t.body = nullptr; t.res = nullptr; continue; }
        t.top = (uint16_t *)(smem + off);  off += (size_t)(TILE_X0 + 2 * t.sw) * 2;
        t.body = (uint16_t *)(smem + off); off += (size_t)t.sh * t.bpitch * 2;
        t.res = (int16_t *)(smem + off);   off += (size_t)t.sh * t.sw * 2;
    }
    uint2 *s_index = (uint2 *)(smem + off);
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
            // ---- what does not depend on the row above: records, residuals, the CTB's own rows, the column left of it
            int cbf = 0;
            for (uint32_t i = tid; i < ix.y; i += INTRA_THREADS) {
                const ohcu_iu u = f.ius[ix.x + i];
                s_iu[i] = u;
                cbf |= u.cand & OHCU_IU_CBF;
            }
            if (row > 0 && tid == 0) {                     // meanwhile: the row above must have finished the CTB up-right
                const int need = min(cx + 2, f.ctb_w);
                while (ld_acquire(progress + row - 1) < need) __nanosleep(20);
                __threadfence();
            }
            cbf = __syncthreads_or(cbf);
            for (int c = 0; c < f.nplanes; c++) {
                const IntraTile &t = tile[c];
                const pixel *pl = (const pixel *)f.work.p[c];
                const int pitch = f.work.pitch[c] / (int)sizeof(pixel), pw = f.pw[c], ph = f.ph[c];
                const int ox = cx * t.sw, oy = row * t.sh, rows = min(t.sh, ph - oy), cpr = t.sw >> 3;   // chunks per row
                for (int i = tid; i < rows * cpr; i += INTRA_THREADS) {
                    const int y = i / cpr, k = i - y * cpr, x = ox + 8 * k;
                    if (x < pw) load8<pixel>(t.body + y * t.bpitch + TILE_X0 + 8 * k, pl + (size_t)(oy + y) * pitch + x);
                    if (cbf && x < pw)
                        *(uint4 *)(t.res + y * t.sw + 8 * k) =
                            __ldcg((const uint4 *)((const uint8_t *)f.resid[c] + (size_t)(oy + y) * f.resid_pitch[c]) + ((ox >> 3) + k));
                }
                if (ox > 0)
                    for (int y = tid; y < rows; y += INTRA_THREADS)
                        t.body[y * t.bpitch + TILE_X0 - 1] = __ldcg(pl + (size_t)(oy + y) * pitch + ox - 1);
                // ---- the row above the CTB: corner, top, top-right (the wait above covers it)
                if (oy > 0) {
                    for (int k = tid; k < 2 * cpr; k += INTRA_THREADS)
                        if (ox + 8 * k < pw) load8<pixel>(t.top + TILE_X0 + 8 * k, pl + (size_t)(oy - 1) * pitch + ox + 8 * k);
                    if (ox > 0 && tid == INTRA_THREADS - 1) t.top[TILE_X0 - 1] = __ldcg(pl + (size_t)(oy - 1) * pitch + ox - 1);
                }
            }
            __syncthreads();
            if (warp < f.nplanes) {
                const IntraTile &t = tile[warp];
                pixel *plane = (pixel *)f.work.p[warp];
                const int pitch = f.work.pitch[warp] / (int)sizeof(pixel), pw = f.pw[warp], ph = f.ph[warp];
                const int ox = cx * t.sw, oy = row * t.sh;
                TileSurf<pixel> sf;
                sf.top = t.top; sf.body = t.body; sf.bpitch = t.bpitch; sf.pitch = pitch;
                for (uint32_t i = 0; i < ix.y; i++) {
                    const ohcu_iu u = s_iu[i];
                    if (u.c_idx != warp) continue;
                    const int n = 1 << u.log2_size;
                    sf.bx = u.x0 - ox; sf.by = u.y0 - oy;
                    sf.blk = plane + (size_t)u.y0 * pitch + u.x0;
                    const int16_t *res = (u.cand & OHCU_IU_CBF) ? t.res + sf.by * t.sw + sf.bx : nullptr;
                    intra_pred_tb(sf, max(0, min(u.x0 + 2 * n, pw) - (u.x0 + n)), max(0, min(u.y0 + 2 * n, ph) - (u.y0 + n)), u.log2_size,
                                  warp, u.mode, u.cand & 31, warp ? filter_chroma : filter_luma, f.strong_intra_smoothing, f.bd,
                                  scratch[warp], lane, res, t.sw);
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

static size_t intra_smem_bytes(const FrameDev &f)
{
    const int S = 1 << f.log2_ctb;
    size_t b = 0;
    for (int c = 0; c < f.nplanes; c++) {
        const int sw = S >> (c ? f.hs : 0), sh = S >> (c ? f.vs : 0);
        b += (size_t)(TILE_X0 + 2 * sw) * 2 + (size_t)sh * (sw + TILE_X0) * 2 + (size_t)sh * sw * 2;
    }
    return b + sizeof(uint2) * f.ctb_w;
}

int intra_set_attrs()
{
    // three 64x64 planes of samples and residuals, the row index of a 16384-wide picture with 16x16 CTBs
    const int bytes = 80 * 1024;
    if (cudaFuncSetAttribute(intra_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess ||
        cudaFuncSetAttribute(intra_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
        return -1;
    return 0;
}

int launch_intra(const FrameDev &f, cudaStream_t st)
{
    if (f.n_iu <= 0) return 0;
    if (!f.ius || !f.iu_index || !f.intra_progress || !f.resid[0]) return -1;
    const size_t smem = intra_smem_bytes(f);
    if (smem > 80 * 1024) return -1;
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
    const int n = 1 << log2_size;
    PlaneSurf<pixel> sf;
    sf.blk = plane + (size_t)y * pitch + x; sf.pitch = pitch;
    intra_pred_tb(sf, max(0, min(x + 2 * n, pw) - (x + n)), max(0, min(y + 2 * n, ph) - (y + n)), log2_size, c_idx, mode, cand, filter_ok,
                  strong, bd, s, threadIdx.x);
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
