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
#include <algorithm>

#include "ohcu_internal.h"

namespace ohcu {

// -DINTRA_PROFILE: cycle counters (tools/intra_phases.py): [0..5] the phases of a CTB in intra_kernel (thread 0),
// [8..13] the steps of a 4x4 / 8x8 block in intra_pred_small (lane 0 of the luma warp), [7] CTBs, [14] blocks
#ifdef INTRA_PROFILE
__device__ unsigned long long g_intra_prof[16];
#define IB_START() long long ib_t = clock64()
#define IB_MARK(k) do { if (lane == 0 && c_idx == 0) { const long long t_ = clock64(); atomicAdd(&g_intra_prof[k], (unsigned long long)(t_ - ib_t)); ib_t = t_; } } while (0)
#else
#define IB_START() do { } while (0)
#define IB_MARK(k) do { } while (0)
#endif

enum { CAND_BL = 1, CAND_L = 2, CAND_UL = 4, CAND_U = 8, CAND_UR = 16 };      // bits 0..4: the order the reference walks the groups in

struct IntraScratch {                 // per warp
    int16_t raw[2][66];               // [0] left, [1] top; index i + 1 holds sample i, index 0 the corner
    int16_t flt[2][66];               // smoothed copies
};

__constant__ int8_t c_intra_angle[33] = {32, 26, 21, 17, 13, 9, 5, 2, 0, -2, -5, -9, -13, -17, -21, -26, -32,
                                         -26, -21, -17, -13, -9, -5, -2, 0, 2, 5, 9, 13, 17, 21, 26, 32};
__constant__ int16_t c_intra_inv_angle[15] = {-4096, -1638, -910, -630, -482, -390, -315, -256, -315, -390, -482, -630, -910, -1638, -4096};

// An int16 array in shared memory addressed with ld.shared / st.shared (see TileSurf); index -1 is valid where the
// array has an element in front of it.
struct SArr {
    uint32_t a;                                            // shared address of element 0
    __device__ __forceinline__ int operator[](int i) const
    {
        int v;
        asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(a + 2u * (uint32_t)i) : "memory");
        return v;
    }
    __device__ __forceinline__ void set(int i, int v) const { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a + 2u * (uint32_t)i), "r"(v) : "memory"); }
};

// Where a block's samples live.  ld / st take coordinates relative to the block's first sample.
template <typename pixel> struct PlaneSurf {           // a plane in global memory (per-entry shim)
    pixel *blk; int pitch;                             // first sample of the block, pitch in samples
    __device__ __forceinline__ int ld(int x, int y) const { return (int)__ldcg(blk + (ptrdiff_t)y * pitch + x); }
    __device__ __forceinline__ void st(int x, int y, int v) const { blk[(ptrdiff_t)y * pitch + x] = (pixel)v; }
    __device__ __forceinline__ int res(int, int) const { return 0; }           // (the shim predicts only)
};
// The CTB's samples in shared memory (intra_kernel): row -1 (corner, top, top-right of the CTB) in `top`, the CTB's own
// rows with the column left of it in `body`; sample x of a row sits at index x + TILE_X0.  Stores go to the tile and
// through to the picture.
constexpr int TILE_X0 = 8;
template <typename pixel> struct TileSurf {
    // 32-bit shared-memory addresses and ld.shared / st.shared: the structs that carry the tile travel through shared
    // memory and function calls, where the compiler only knows generic pointers - a generic load that resolves to shared
    // memory takes the L1 path's latency, and the chain of blocks is made of these loads
    uint32_t top_s; int body_off, bpitch;              // `top` (shared address), where `body` starts behind it (samples)
    uint32_t res_s; int res_pitch;                     // the block's residual (shared address of its first value), pitch in int16
    int bx, by;                                        // the block inside the CTB
    pixel *blk; int pitch;                             // the block in the picture
    __device__ __forceinline__ int ld(int x, int y) const
    {
        // one load whichever array the sample is in (`body` follows `top` in the tile: an offset, not a branch)
        const int X = bx + x + TILE_X0, Y = by + y;
        uint32_t v;
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(top_s + 2u * (uint32_t)(X + (Y < 0 ? 0 : body_off + Y * bpitch))) : "memory");
        return (int)v;
    }
    __device__ __forceinline__ void st(int x, int y, int v) const
    {
        asm volatile("st.shared.u16 [%0], %1;" ::"r"(top_s + 2u * (uint32_t)(body_off + (by + y) * bpitch + bx + x + TILE_X0)), "r"(v) : "memory");
        blk[(ptrdiff_t)y * pitch + x] = (pixel)v;
    }
    __device__ __forceinline__ int res(int x, int y) const
    {
        int v;
        asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(res_s + 2u * (uint32_t)(y * res_pitch + x)) : "memory");
        return v;
    }
};

// ---- reference samples with the substitution of unavailable ones folded into the fetch (:152-272).
// The reference walks from the bottom-left end up the left column, over the corner and along the top row: a group
// that is missing (BL, L, UL, U, UR = bits 0..4 of cand) repeats the last sample of the nearest available group before
// it, and the groups before the first available one repeat that group's first sample; nothing available: mid-grey.
// Every reference sample therefore IS one sample of the picture (or the constant).  Which one, per group, is a function of
// the five flags alone: c_ref_src[cand] holds six bits per group - bit 0 the group's own samples, bit 1 mid-grey, else the
// coordinates of the one sample it repeats as x = (bit 2 ? 0 : -1) + (bit 3 ? N : 0), y = (bit 4 ? 0 : -1) + (bit 5 ? N : 0):
//   last sample of the nearest available group before it   BL (-1, N)   L (-1, 0)   UL (-1, -1)   U (N - 1, -1)
//   first sample of the first available group              L (-1, N - 1)   UL (-1, -1)   U (0, -1)   UR (N, -1)
__constant__ uint32_t c_ref_src[32] = {
    0x02082082u, 0x30c30c01u, 0x10410060u, 0x10410041u, 0x00001000u, 0x00001c01u, 0x00001060u, 0x00001041u,
    0x08044104u, 0x08070c01u, 0x08050060u, 0x08050041u, 0x08041000u, 0x08041c01u, 0x08041060u, 0x08041041u,
    0x0130c30cu, 0x01c30c01u, 0x01410060u, 0x01410041u, 0x01001000u, 0x01001c01u, 0x01001060u, 0x01001041u,
    0x01044104u, 0x01070c01u, 0x01050060u, 0x01050041u, 0x01041000u, 0x01041c01u, 0x01041060u, 0x01041041u};
// side 0: left column (i = row), side 1: top row (i = column); i = -1: the corner; runs that leave the picture repeat
// their last sample (:109-112, :158, :168).  src = c_ref_src[cand & 31]
template <int n, typename Surf>
__device__ __forceinline__ int ref_fetch(const Surf &sf, uint32_t src, int side, int i, int tr_size, int bl_size, int bd)
{
    const int g = i < 0 ? 2 : 2 + (side ? 1 : -1) * (i < n ? 1 : 2);       // BL 0, L 1, UL 2, U 3, UR 4
    const uint32_t w = src >> (6 * g);
    int X = (int)((w >> 2) & 1u) - 1 + ((w & 8u) ? n : 0), Y = (int)((w >> 4) & 1u) - 1 + ((w & 32u) ? n : 0);
    if (w & 1u) {
        const int j = i < n ? i : n + min(i - n, (side ? tr_size : bl_size) - 1);
        X = side ? j : -1; Y = side ? -1 : j;
    }
    int v = 1 << (bd - 1);
    if (!(w & 2u)) v = sf.ld(X, Y);
    return v;
}

// ---- 4x4 and 8x8 blocks: the 4N reference samples live in ONE register across the warp (lanes 0..15 the left column,
// lanes 16..31 the top row, the corner in every lane), smoothing and the predictors read them with shuffles: no scratch
// array, no warp barrier, a third of the dependent instructions of the general path.  These are most of a picture's blocks.
template <int log2_size, typename Surf>
__device__ __forceinline__ void intra_pred_small(const Surf &sf, uint32_t src, int tr_size, int bl_size, int c_idx, int mode, int filter_ok,
                                              int bd, int lane, bool has_res)
{
    constexpr int n = 1 << log2_size, n2 = 2 * n;
    constexpr unsigned FULL = 0xffffffffu;
    static_assert(n2 <= 16, "one register holds both arrays");
    const int maxv = (1 << bd) - 1;
    IB_START();
    IB_MARK(8);
    // the block's residual first: these loads depend on nothing and are back before the prediction needs them
    int rv[(n * n + 31) / 32];
#pragma unroll
    for (int s = 0; s < (n * n + 31) / 32; s++) {
        const int idx = lane + 32 * s;
        rv[s] = (has_res && idx < n * n) ? sf.res(idx & (n - 1), idx >> log2_size) : 0;
    }
    const int side = lane >> 4, i = lane & 15;
    int r = i < n2 ? ref_fetch<n>(sf, src, side, i, tr_size, bl_size, bd) : 0;
    int corner = ref_fetch<n>(sf, src, 0, -1, tr_size, bl_size, bd);
    IB_MARK(9);                                            // fetched
    // smoothing (:274-326): 8x8 only (4x4 is never smoothed, the strong form is for 32x32)
    if (n == 8 && filter_ok && mode != 1 && min(abs(mode - 26), abs(mode - 10)) > 7) {
        const int up = __shfl_down_sync(FULL, r, 1), dn = __shfl_up_sync(FULL, r, 1);
        const int l0 = __shfl_sync(FULL, r, 0), t0 = __shfl_sync(FULL, r, 16);
        if (i != n2 - 1) r = (up + 2 * r + (i == 0 ? corner : dn) + 2) >> 2;
        corner = (l0 + 2 * corner + t0 + 2) >> 2;
    }
    IB_MARK(10);                                           // smoothed
    auto L = [&](int j) { return __shfl_sync(FULL, r, j & 15); };             // left[j], j >= 0
    auto T = [&](int j) { return __shfl_sync(FULL, r, 16 + (j & 15)); };      // top[j], j >= 0
    const bool edge = c_idx == 0;
#pragma unroll
    for (int s = 0; s < (n * n + 31) / 32; s++) {
        const int idx = lane + 32 * s, xx = idx & (n - 1), yy = (idx >> log2_size) & (n - 1);
        int v;
        if (mode == 0) {                                   // planar, :359-372
            const int ly = L(yy), tn = T(n), tx = T(xx), ln = L(n);
            v = ((n - 1 - xx) * ly + (xx + 1) * tn + (n - 1 - yy) * tx + (yy + 1) * ln + n) >> (log2_size + 1);
        } else if (mode == 1) {                            // DC, :388-417
            int sum = i < n ? r : 0;
#pragma unroll
            for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(FULL, sum, o);
            const int dc = (sum + n) >> (log2_size + 1);
            const int ly = L(yy), tx = T(xx), l0 = L(0), t0 = T(0);
            v = dc;
            if (edge) {
                if (!(xx | yy)) v = (l0 + 2 * dc + t0 + 2) >> 2;
                else if (!yy) v = (tx + 3 * dc + 2) >> 2;
                else if (!xx) v = (ly + 3 * dc + 2) >> 2;
            }
        } else {                                           // angular, :419-505
            const int angle = c_intra_angle[mode - 2];
            const bool vertical = mode >= 18;
            const int mbase = vertical ? 16 : 0, sbase = vertical ? 0 : 16;
            const int inv = angle < 0 ? c_intra_inv_angle[mode - 11] : 0;
            const int a = vertical ? yy : xx, b = vertical ? xx : yy;
            const int pos = (a + 1) * angle, k = b + (pos >> 5) + 1, fact = pos & 31;
            // ref[k] = main[k - 1] (k >= 0) or the side array through the inverse angle (:440-447); index -1 = the corner
            auto ref = [&](int kk) {
                const int j = kk >= 0 ? kk - 1 : -1 + ((kk * inv + 128) >> 8);
                const int got = __shfl_sync(FULL, r, ((kk >= 0 ? mbase : sbase) + j) & 31);
                return j < 0 ? corner : got;
            };
            const int r0 = ref(k), r1 = ref(k + 1);
            v = fact ? ((32 - fact) * r0 + fact * r1 + 16) >> 5 : r0;
            if (edge && (mode == 26 || mode == 10)) {      // (uniform)
                const int ly = L(yy), tx = T(xx), l0 = L(0), t0 = T(0);
                if (mode == 26 && xx == 0) v = min(max(t0 + ((ly - corner) >> 1), 0), maxv);
                if (mode == 10 && yy == 0) v = min(max(l0 + ((tx - corner) >> 1), 0), maxv);
            }
        }
        IB_MARK(11);                                       // predicted
        if (idx < n * n) {
            if (has_res) v = min(max(v + rv[s], 0), maxv);
            sf.st(xx, yy, v);
        }
        IB_MARK(12);                                       // residual added, stored
    }
    __syncwarp();
    IB_MARK(13);
#ifdef INTRA_PROFILE
    if (lane == 0 && c_idx == 0) atomicAdd(&g_intra_prof[14], 1ull);
#endif
}

// sf: the block's surface; tr_size / bl_size: samples of the top-right / bottom-left run inside the picture (:109-112);
// cand: CAND_* final availability; filter_ok: !intra_smoothing_disabled && (c_idx == 0 || chroma_format_idc == 3);
// strong: sps strong smoothing flag; has_res: the surface's residual (Surf::res) is added to the prediction with the clip of
// transform_add (hevcdsp_template.c:45-84).
// (one instance per block size: trip counts and index arithmetic fold at compile time)
template <int log2_size, typename Surf>
__device__ __forceinline__ void intra_pred_size(const Surf &sf, uint32_t src, int tr_size, int bl_size, int c_idx, int mode, int filter_ok,
                                             int strong, int bd, IntraScratch &s, int lane, bool has_res)
{
    constexpr int n = 1 << log2_size, n2 = 2 * n;
    const int maxv = (1 << bd) - 1;
    auto put = [&](int xx, int yy, int v) {
        if (has_res) v = min(max(v + sf.res(xx, yy), 0), maxv);
        sf.st(xx, yy, v);
    };
    // the scratch arrays through their shared addresses: element i of an array sits at index i + 1, the corner at 0
    const SArr raw[2] = {{(uint32_t)__cvta_generic_to_shared(s.raw[0]) + 2u}, {(uint32_t)__cvta_generic_to_shared(s.raw[1]) + 2u}};
    const SArr flt[2] = {{(uint32_t)__cvta_generic_to_shared(s.flt[0]) + 2u}, {(uint32_t)__cvta_generic_to_shared(s.flt[1]) + 2u}};
    SArr left = raw[0], top = raw[1];

    auto fetch = [&](int side, int i) -> int { return ref_fetch<n>(sf, src, side, i, tr_size, bl_size, bd); };
    for (int t = lane; t < 2 * n2; t += 32) {
        const int side = t >= n2, i = t - (side ? n2 : 0);
        (side ? raw[1] : raw[0]).set(i, fetch(side, i));
    }
    const int corner = fetch(0, -1);
    if (lane == 0) { left.set(-1, corner); top.set(-1, corner); }
    __syncwarp();

    // ---- smoothing (:274-326)
    if (filter_ok && mode != 1 && n != 4) {
        const int dist = min(abs(mode - 26), abs(mode - 10));
        const int thresh = log2_size == 3 ? 7 : (log2_size == 4 ? 1 : 0);
        if (dist > thresh) {
            const bool bilinear = strong && c_idx == 0 && log2_size == 5 && abs(corner + top[63] - 2 * top[31]) < (1 << (bd - 5)) &&
                                  abs(corner + left[63] - 2 * left[31]) < (1 << (bd - 5));
            for (int t = lane; t < 2 * n2; t += 32) {
                const int side = t >= n2, i = t - (side ? n2 : 0);
                const SArr r = side ? raw[1] : raw[0];
                int v;
                if (i == n2 - 1) v = r[i];
                else if (bilinear) v = ((63 - i) * corner + (i + 1) * r[63] + 32) >> 6;
                else v = (r[i + 1] + 2 * r[i] + r[i - 1] + 2) >> 2;
                (side ? flt[1] : flt[0]).set(i, v);
            }
            if (lane == 0) {
                const int fc = bilinear ? corner : (left[0] + 2 * corner + top[0] + 2) >> 2;
                flt[0].set(-1, fc); flt[1].set(-1, fc);
            }
            __syncwarp();
            left = flt[0]; top = flt[1];
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
        for (int t = lane; t < n2; t += 32) sum += t < n ? left[t] : top[t - n];
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
        const SArr mn = vertical ? top : left, sd = vertical ? left : top;
        const int inv = angle < 0 ? c_intra_inv_angle[mode - 11] : 0;
        const bool edge26 = edge && mode == 26, edge10 = edge && mode == 10;
#pragma unroll
        for (int i = lane; i < n * n; i += 32) {
            const int xx = i & (n - 1), yy = i >> log2_size;
            const int a = vertical ? yy : xx, b = vertical ? xx : yy;
            const int pos = (a + 1) * angle, k = b + (pos >> 5) + 1, fact = pos & 31;
            // ref[k] = main[k - 1] for k >= 0, and for the negative angles that reach below 0 the side array projected
            // through the inverse angle (:440-447): taken straight from the two arrays
            int v = k >= 0 ? mn[k - 1] : sd[-1 + ((k * inv + 128) >> 8)];
            if (fact) v = ((32 - fact) * v + fact * (k + 1 >= 0 ? mn[k] : sd[-1 + (((k + 1) * inv + 128) >> 8)]) + 16) >> 5;
            if (edge26 && xx == 0) v = min(max(top[0] + ((left[yy] - left[-1]) >> 1), 0), maxv);
            if (edge10 && yy == 0) v = min(max(left[0] + ((top[xx] - top[-1]) >> 1), 0), maxv);
            put(xx, yy, v);
        }
    }
    __syncwarp();
}

// src: c_ref_src[cand & 31]
template <typename Surf>
__device__ __forceinline__ void intra_pred_tb(const Surf &sf, uint32_t src, int tr_size, int bl_size, int log2_size, int c_idx, int mode,
                                              int filter_ok, int strong, int bd, IntraScratch &s, int lane, bool has_res = false)
{
    switch (log2_size) {
    case 2: intra_pred_small<2>(sf, src, tr_size, bl_size, c_idx, mode, filter_ok, bd, lane, has_res); break;
    case 3: intra_pred_small<3>(sf, src, tr_size, bl_size, c_idx, mode, filter_ok, bd, lane, has_res); break;
    case 4: intra_pred_size<4>(sf, src, tr_size, bl_size, c_idx, mode, filter_ok, strong, bd, s, lane, has_res); break;
    default: intra_pred_size<5>(sf, src, tr_size, bl_size, c_idx, mode, filter_ok, strong, bd, s, lane, has_res); break;
    }
}

// ================================================================= frame level: CTBs in dependency order
// hls_transform_unit predicts a block right before it adds the block's residual (hevcdec.c:1335-1526), in decoding
// order; a block reads only samples of blocks decoded earlier: its own CTB, and the CTBs left, up-left, up and
// up-right of it (the z-scan availability test, hevcpred_template.c:103-107).  Everything that is not intra has been
// reconstructed by mc_kernel / residual_kernel / pcm_kernel before this kernel starts, and residual_kernel has left
// the intra blocks' residuals in the job's residual planes.  What remains is the dependency between intra blocks:
//   * a CTA takes one CTB at a time, CTBs handed out along the anti-diagonals x + 2 y by a ticket (everything a CTB can
//     read lies on an earlier diagonal: a CTA only ever waits for CTBs taken before its own, which are running or done);
//     CTBs without intra blocks are passed over.  It stages the CTB's samples
//     and residuals in shared memory with 16-byte loads - all in flight together, and ahead of any wait - and look at
//     their own records to see WHICH neighbour CTBs they read (a block on the CTB's left column with its left
//     neighbour available, ...): only those, and only if they hold intra blocks themselves, are waited for (one flag
//     per CTB, release / acquire).  Intra CTBs that do not touch each other run side by side; an all-intra picture
//     degenerates to the wavefront the reference's WPP threads run (hls_decode_entry_wpp, hevcdec.c:2966-3051);
//   * after the wait: the column left of the CTB and the row above it (with the corner and the top-right run), one
//     more round trip; then the chain of blocks inside the CTB - each one reads what the one before wrote - runs
//     out of shared memory, predicted samples written to the tile and through to the picture;
//   * inside a CTB the three planes are independent: warp c takes plane c's blocks in list order.
constexpr int INTRA_THREADS = 256;
constexpr int INTRA_MAX_IU = 768;          // per CTB: 256 4x4 blocks in each of three planes
enum { NEED_L = 1, NEED_UL = 2, NEED_U = 4, NEED_UR = 8 };

__device__ __forceinline__ int ld_acquire(const int *p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// The waiting thread sits in the CTA's LAST warp: warps 0..2 predict, and a warp's scheduler is its number modulo 4, so
// the poll loops of the CTAs that share an SM with a working one never compete with its predicting warps for issue
// slots (measured: with thread 0 polling every 20 ns, 89% of the kernel's executed instructions were poll loops).
constexpr int INTRA_SPIN_TID = 224;
__device__ __forceinline__ void spin_until_set(const int *flag)
{
    unsigned ns = 100;
    while (!ld_acquire(flag)) {
        __nanosleep(ns);
        if (ns < 800) ns += ns;
    }
}
__device__ __forceinline__ void st_release(int *p, int v)
{
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// eight samples of a picture row: a 16-byte / 8-byte load past the L1 into registers, widened into the tile later -
// a thread issues ALL its loads of a CTB before the first store, so that they are in flight together
template <typename pixel> __device__ __forceinline__ uint4 ld8(const pixel *src)
{
    if (sizeof(pixel) == 2) return __ldcg((const uint4 *)src);
    const uint2 v = __ldcg((const uint2 *)src);
    return make_uint4(v.x, v.y, 0, 0);
}
template <typename pixel> __device__ __forceinline__ void st8(uint16_t *dst, uint4 v)
{
    if (sizeof(pixel) == 1)
        v = make_uint4(__byte_perm(v.x, 0, 0x4140), __byte_perm(v.x, 0, 0x4342), __byte_perm(v.y, 0, 0x4140), __byte_perm(v.y, 0, 0x4342));
    *(uint4 *)dst = v;
}

struct IntraTile {                         // one plane's staging (pointers into dynamic shared memory)
    uint16_t *top, *body;
    int16_t *res;
    int sw, sh, bpitch, cshift;            // CTB size in samples of the plane, tile pitch, log2(chunks of 8 per row)
};

// -DINTRA_PROFILE: thread 0 of every CTA adds the cycles between phase marks to g_intra_prof (tools/intra_phases.py)
#ifdef INTRA_PROFILE
#define IP_MARK(k) do { if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_intra_prof[k], (unsigned long long)(t_ - ip_t)); ip_t = t_; } } while (0)
extern "C" int ohcu_intra_profile(unsigned long long *out, int reset)
{
    if (cudaMemcpyFromSymbol(out, g_intra_prof, sizeof(g_intra_prof)) != cudaSuccess) return -1;
    if (reset) { unsigned long long z[16] = {}; cudaMemcpyToSymbol(g_intra_prof, z, sizeof(z)); }
    return 0;
}
#else
#define IP_MARK(k) do { } while (0)
#endif

template <typename pixel>
__global__ void __launch_bounds__(INTRA_THREADS) intra_kernel(const __grid_constant__ FrameDev f)
{
#ifdef INTRA_PROFILE
    long long ip_t = clock64();
#endif
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ IntraScratch scratch[3];
    __shared__ __align__(16) uint4 s_blk[INTRA_MAX_IU];      // per block, prepared in parallel: see the record loop below
    __shared__ IntraTile tile[3];                          // (indexed by run-time plane numbers: not registers)
    __shared__ int rows[3], oy[3];
    __shared__ int s_ctb, s_need;
    if (f.err && *f.err) return;                           // malformed lists: every CTA leaves, nobody waits
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    int *const ticket = f.intra_progress, *const done = f.intra_progress + 1;
    const int nctb = f.ctb_w * f.ctb_h;
  for (;;) {                                               // persistent: at most one CTA per SM, the rest of the GPU stays free for the other frames in flight
    __syncthreads();                                       // the previous CTB is finished with the shared state
    if (tid == 0) {
        int c;
        do {                                               // (nobody waits for a CTB without intra blocks)
            c = atomicAdd(ticket, 1);
            c = c < nctb ? f.intra_order[c] : nctb;
        } while (c < nctb && !f.iu_index[c].n_iu);
        s_ctb = c; s_need = 0;
    }
    __syncthreads();
    const int ctb = s_ctb, cx = ctb % f.ctb_w, row = ctb / f.ctb_w;
    if (ctb >= nctb) return;
    IP_MARK(0);                                            // ticket
    const ohcu_iu_index ixr = f.iu_index[ctb];
    const uint2 ix = make_uint2(ixr.first_iu, ixr.n_iu);
    // which neighbour CTBs hold intra blocks (thread 0 waits for those this CTB reads)
    int nb_intra = 0;
    if (tid == INTRA_SPIN_TID) {
        if (cx > 0) nb_intra |= f.iu_index[ctb - 1].n_iu ? NEED_L : 0;
        if (row > 0) {
            if (cx > 0) nb_intra |= f.iu_index[ctb - f.ctb_w - 1].n_iu ? NEED_UL : 0;
            nb_intra |= f.iu_index[ctb - f.ctb_w].n_iu ? NEED_U : 0;
            if (cx + 1 < f.ctb_w) nb_intra |= f.iu_index[ctb - f.ctb_w + 1].n_iu ? NEED_UR : 0;
        }
    }
    const int S = 1 << f.log2_ctb;
    // carve: per plane top (TILE_X0 + 2 sw), body (sh rows of sw + TILE_X0), residual (sh x sw)
    size_t off = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        IntraTile t;
        t.sw = S >> (c ? f.hs : 0); t.sh = S >> (c ? f.vs : 0); t.bpitch = t.sw + TILE_X0; t.cshift = 31 - __clz(t.sw) - 3;
        t.top = t.body = nullptr; t.res = nullptr;
        if (c < f.nplanes) {
            t.top = (uint16_t *)(smem + off);  off += (size_t)(TILE_X0 + 2 * t.sw) * 2;
            t.body = (uint16_t *)(smem + off); off += (size_t)t.sh * t.bpitch * 2;
            t.res = (int16_t *)(smem + off);   off += (size_t)t.sh * t.sw * 2;
        } else {
            t.sh = 0;
        }
        if (tid == 0) {
            tile[c] = t;
            oy[c] = row * t.sh;
            rows[c] = c < f.nplanes ? min(t.sh, f.ph[c] - row * t.sh) : 0;
        }
    }
    __syncthreads();
    const int filter_luma = !f.intra_smoothing_disabled, filter_chroma = filter_luma && f.cfi == 3;
    const int nb0 = rows[0] << tile[0].cshift, nb1 = rows[1] << tile[1].cshift, nb2 = rows[2] << tile[2].cshift;
    constexpr int KB = 6;                                  // body chunks per thread: 3 planes x 64 rows x 8 chunks / 256

    // ---- loads that depend on no other intra CTB, all issued before the wait: records, the CTB's own rows, their residuals
    uint2 rec[3];
#pragma unroll
    for (int q = 0; q < 3; q++) {
        const uint32_t i = tid + q * INTRA_THREADS;
        rec[q] = i < ix.y ? __ldg((const uint2 *)(f.ius + ix.x + i)) : make_uint2(0, 0);
    }
    uint4 va[KB], vb[KB];
#pragma unroll
    for (int q = 0; q < KB; q++) {
        int i = tid + q * INTRA_THREADS;
        const int c = i < nb0 ? 0 : (i < nb0 + nb1 ? 1 : 2);
        i -= c == 0 ? 0 : (c == 1 ? nb0 : nb0 + nb1);
        va[q] = vb[q] = make_uint4(0, 0, 0, 0);
        if (i < (c == 2 ? nb2 : 1 << 30)) {
            const IntraTile &t = tile[c];
            const int y = i >> t.cshift, k = i & ((1 << t.cshift) - 1), x = cx * t.sw + 8 * k;
            if (x < f.pw[c]) {
                va[q] = ld8<pixel>((const pixel *)(f.work.p[c] + (size_t)(oy[c] + y) * f.work.pitch[c]) + x);
                vb[q] = __ldcg((const uint4 *)((const uint8_t *)f.resid[c] + (size_t)(oy[c] + y) * f.resid_pitch[c]) + (x >> 3));
            }
        }
    }
    // ---- the records: into shared memory, and which neighbour CTBs their blocks read
    int need = 0;
#pragma unroll
    for (int q = 0; q < 3; q++) {
        const uint32_t i = tid + q * INTRA_THREADS;
        if (i < ix.y) {
            union { uint2 w; ohcu_iu u; } cv;
            cv.w = rec[q];
            const ohcu_iu u = cv.u;
            const IntraTile &t = tile[u.c_idx];
            const int n = 1 << u.log2_size, bx = u.x0 & (t.sw - 1), by = u.y0 & (t.sh - 1);
            // everything about the block that does not depend on samples, so that the chain of blocks starts from one load:
            //   x  where its reference samples come from (c_ref_src)        y  offset of its first sample in the plane
            //   z  bx | by << 8 | top-right run << 16 | bottom-left run << 24  (samples inside the picture, :109-112)
            //   w  mode | log2 size << 8 | coded << 12 | plane << 16
            {
                const int c = u.c_idx, pw = f.pw[c], ph = f.ph[c];
                const int tr = max(0, min(u.x0 + 2 * n, pw) - (u.x0 + n)), bl = max(0, min(u.y0 + 2 * n, ph) - (u.y0 + n));
                s_blk[i] = make_uint4(c_ref_src[u.cand & 31], (uint32_t)u.y0 * (uint32_t)(f.work.pitch[c] / (int)sizeof(pixel)) + u.x0,
                                      (uint32_t)bx | (uint32_t)by << 8 | (uint32_t)tr << 16 | (uint32_t)bl << 24,
                                      (uint32_t)u.mode | (uint32_t)u.log2_size << 8 | ((u.cand & OHCU_IU_CBF) ? 1u << 12 : 0u) | (uint32_t)c << 16);
            }
            if (bx == 0) {
                if (u.cand & (CAND_L | CAND_BL)) need |= NEED_L;
                if (u.cand & CAND_UL) need |= by ? NEED_L : NEED_UL;
            }
            if (by == 0) {
                if ((u.cand & CAND_U) || (bx && (u.cand & CAND_UL))) need |= NEED_U;
                if (u.cand & CAND_UR) need |= (bx + n < t.sw ? NEED_U : 0) | (bx + 2 * n > t.sw ? NEED_UR : 0);
            }
        }
    }
    need = __reduce_or_sync(0xffffffffu, need);
    if (lane == 0 && need) atomicOr(&s_need, need);
#pragma unroll
    for (int q = 0; q < KB; q++) {
        int i = tid + q * INTRA_THREADS;
        const int c = i < nb0 ? 0 : (i < nb0 + nb1 ? 1 : 2);
        i -= c == 0 ? 0 : (c == 1 ? nb0 : nb0 + nb1);
        if (i < (c == 2 ? nb2 : 1 << 30)) {
            const IntraTile &t = tile[c];
            const int y = i >> t.cshift, k = i & ((1 << t.cshift) - 1);
            st8<pixel>(t.body + y * t.bpitch + TILE_X0 + 8 * k, va[q]);
            *(uint4 *)(t.res + y * t.sw + 8 * k) = vb[q];
        }
    }
    __syncthreads();
    IP_MARK(1);                                            // records, own rows, residuals
    // ---- wait for the neighbour CTBs this one reads intra samples of
    if (tid == INTRA_SPIN_TID) {
        const int w = s_need & nb_intra;
        if (w) {
            if (w & NEED_L) spin_until_set(done + ctb - 1);
            if (w & NEED_UL) spin_until_set(done + ctb - f.ctb_w - 1);
            if (w & NEED_U) spin_until_set(done + ctb - f.ctb_w);
            if (w & NEED_UR) spin_until_set(done + ctb - f.ctb_w + 1);
            __threadfence();
        }
    }
    __syncthreads();
    IP_MARK(2);                                            // wait
    // ---- the column left of the CTB, the row above it: corner, top, top-right
    {
        int i = tid;
        const int c = i < rows[0] ? 0 : (i < rows[0] + rows[1] ? 1 : 2);
        i -= c == 0 ? 0 : (c == 1 ? rows[0] : rows[0] + rows[1]);
        if (cx > 0 && i < rows[c])
            tile[c].body[i * tile[c].bpitch + TILE_X0 - 1] =
                (uint16_t)__ldcg((const pixel *)(f.work.p[c] + (size_t)(oy[c] + i) * f.work.pitch[c]) + cx * tile[c].sw - 1);
    }
    if (row > 0 && warp >= 5) {                            // warps 5, 6, 7: one plane each
        const int c = warp - 5;
        if (c < f.nplanes) {
            const IntraTile &t = tile[c];
            const pixel *above = (const pixel *)(f.work.p[c] + (size_t)(oy[c] - 1) * f.work.pitch[c]) + cx * t.sw;
            const int nk = 2 << t.cshift;                  // <= 16 chunks
            uint4 v = make_uint4(0, 0, 0, 0);
            int corner = 0;
            if (lane < nk && cx * t.sw + 8 * lane < f.pw[c]) v = ld8<pixel>(above + 8 * lane);
            if (lane == 31 && cx > 0) corner = __ldcg(above - 1);
            if (lane < nk) st8<pixel>(t.top + TILE_X0 + 8 * lane, v);
            if (lane == 31) t.top[TILE_X0 - 1] = (uint16_t)corner;
        }
    }
    __syncthreads();
    IP_MARK(3);                                            // left column, top row
    if (warp < f.nplanes) {
        const IntraTile &t = tile[warp];
        pixel *plane = (pixel *)f.work.p[warp];
        const int pitch = f.work.pitch[warp] / (int)sizeof(pixel);
        TileSurf<pixel> sf;
        sf.top_s = (uint32_t)__cvta_generic_to_shared(t.top); sf.body_off = (int)(t.body - t.top); sf.bpitch = t.bpitch; sf.pitch = pitch;
        sf.res_pitch = t.sw;
        const uint32_t res0 = (uint32_t)__cvta_generic_to_shared(t.res);
        const int filt = warp ? filter_chroma : filter_luma;
        for (uint32_t i = 0; i < ix.y; i++) {
            const uint4 b = s_blk[i];
            if ((int)(b.w >> 16) != warp) continue;
            sf.bx = b.z & 255; sf.by = (b.z >> 8) & 255;
            sf.blk = plane + b.y;
            sf.res_s = res0 + 2u * (uint32_t)(sf.by * t.sw + sf.bx);
            intra_pred_tb(sf, b.x, (int)((b.z >> 16) & 255), (int)(b.z >> 24), (int)((b.w >> 8) & 15), warp, (int)(b.w & 255), filt,
                          f.strong_intra_smoothing, f.bd, scratch[warp], lane, (b.w >> 12) & 1);
        }
    }
    __syncthreads();                                       // every plane of the CTB is written
    IP_MARK(4);                                            // the blocks (thread 0: plane 0's, then the wait for the others)
    if (tid == 0) {
        __threadfence();
        st_release(done + ctb, 1);
    }
    IP_MARK(5);                                            // fence + release
#ifdef INTRA_PROFILE
    if (tid == 0) atomicAdd(&g_intra_prof[7], 1ull);
#endif
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
    return b;
}

int intra_set_attrs()
{
    // three 64x64 planes of samples and residuals
    const int bytes = 64 * 1024;
    if (cudaFuncSetAttribute(intra_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess ||
        cudaFuncSetAttribute(intra_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
        return -1;
    return 0;
}

int launch_intra(const FrameDev &f, cudaStream_t st)
{
    if (f.n_iu <= 0) return 0;
    if (!f.ius || !f.iu_index || !f.intra_progress || !f.intra_order || !f.resid[0]) return -1;
    const size_t smem = intra_smem_bytes(f);
    if (smem > 64 * 1024) return -1;
    // the wavefront of an all-intra picture is at most min(ctb_h, ctb_w / 2) + 1 CTBs wide; pictures with scattered intra
    // CTBs use every CTA.  One CTA per SM at most: spinning CTAs must not take the registers of the other frames' kernels
    // (with one CTA per CTB, 444 resident CTAs of an I picture held 94 % of the register files: 8 frames in flight
    // with one I picture among them ran at 564 frame/s)
    static int sm_count[64];                               // per device, looked up once
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
    if (!sm_count[dev] && cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
    const int sms = sm_count[dev];
    const int grid = std::min(f.ctb_w * f.ctb_h, sms);
    if (f.bd > 8) intra_kernel<uint16_t><<<grid, INTRA_THREADS, smem, st>>>(f);
    else          intra_kernel<uint8_t><<<grid, INTRA_THREADS, smem, st>>>(f);
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
    intra_pred_tb(sf, c_ref_src[cand & 31], max(0, min(x + 2 * n, pw) - (x + n)), max(0, min(y + 2 * n, ph) - (y + n)), log2_size, c_idx, mode,
                  filter_ok, strong, bd, s, threadIdx.x);
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
