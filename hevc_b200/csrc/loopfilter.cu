// loopfilter.cu — deblocking + SAO of one picture in ONE pass: out = SAO(deblock(recon)).
//
// Replaces, per picture, ff_hevc_hls_filter / ff_hevc_hls_filters (hevc_filter.c:1211-1245):
//   deblocking_filter_CTB (hevc_filter.c:525-762) and the table entries
//     hevc_{h,v}_loop_filter_{luma,chroma}(_c)             hevcdsp_template.c:3377-3553
//   sao_filter_CTB (hevc_filter.c:264-502), restore_tqb_pixels (:229-260) and
//     sao_band_filter / sao_edge_filter / sao_edge_restore   hevcdsp_template.c:310-531
//
// One CTA owns a 64x64 luma region (and its chroma).  The region plus a halo is fetched
// from the reconstructed picture into shared memory by TMA (one 2-D box per plane, zero
// fill outside the picture), then
//   1. every 4-line segment of every VERTICAL edge in the tile is filtered in place,
//   2. every 4-line segment of every HORIZONTAL edge, on the result (the standard's order:
//      all vertical edges of the picture before the horizontal ones; edges sit on an 8x8
//      grid and touch at most 3 samples per side, so segments of one direction are
//      independent and a 4-sample halo reproduces the picture-wide result),
//   3. SAO reads the deblocked tile (+1 sample of halo) and writes the DPB slot with
//      full-width stores.
// The deblocked picture never goes to HBM: one read and one write per sample for both
// filters.  tc / beta / pcm flags are derived on the device from the bs planes, qp_y_tab
// and the per-CTB offsets exactly as the reference driver does, including its choice of
// which CTB's offsets apply (hevc_filter.c:666, 699, 712, 721).
#include "ohcu_internal.h"

namespace ohcu {

#ifndef LF_NTHREADS
#define LF_NTHREADS 256
#endif
constexpr int LF_THREADS = LF_NTHREADS;
constexpr int LF_REGION = 64;                 // luma samples per side of a CTA's region

__constant__ uint8_t c_tctable[54] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3,
    3, 3, 3, 4, 4, 4, 5, 5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 22, 24};
__constant__ uint8_t c_betatable[52] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18,
    20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64};
__constant__ uint8_t c_qpc420[14] = {29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37};

// ---------------------------------------------------------------- TMA / mbarrier (as recon.cu)
__device__ __forceinline__ uint32_t lf_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void lf_mbar_wait(uint64_t *b, uint32_t parity)
{
    uint32_t ok = 0;
    long long t0 = 0;
    for (int spin = 0; !ok; spin++) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(lf_smem_u32(b)), "r"(parity), "r"(20000) : "memory");
        if (!ok && (spin & 63) == 63) {
            if (!t0) t0 = clock64();
            else if (clock64() - t0 > 4000000000ll) __trap();
        }
    }
}

// ---------------------------------------------------------------- side information
__device__ __forceinline__ int qpy_at(const FrameDev &f, int x, int y)
{
    return f.qp_y[(x >> f.log2_min_cb) + (y >> f.log2_min_cb) * f.min_cb_w];
}
__device__ __forceinline__ int pcm_at(const FrameDev &f, int x, int y)     // get_pcm, hevc_filter.c:504-518
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
__device__ __forceinline__ int chroma_tc(const FrameDev &f, int qp_y, int c, int tc_offset)   // hevc_filter.c:65-92
{
    const int qi = clip3i(qp_y + (c == 1 ? f.cb_qp_offset : f.cr_qp_offset), 0, 57);
    int qp;
    if (f.cfi == 1) qp = qi < 30 ? qi : (qi > 43 ? qi - 6 : c_qpc420[qi - 30]);
    else qp = min(qi, 51);
    return c_tctable[clip3i(qp + 2 + tc_offset, 0, 53)];
}

struct SegParam { int beta, tc, no_p, no_q; bool on; };

// parameters of the 4-row segment starting at plane row ys of the VERTICAL edge at plane column ex
__device__ __forceinline__ SegParam vseg_param(const FrameDev &f, int c, int ex, int ys)
{
    SegParam p = {0, 0, 0, 0, false};
    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
    const int lx = ex << hs, ysl = ys << vs;
    const int bs = f.vbs[(lx + ysl * f.bs_width) >> 2];
    const int cb = ctb_at(f, lx, ysl);
    if (c == 0) {
        if (!bs) return p;
        const int yu = ysl & ~7;                                   // 8-row unit, qp at its first row (:571)
        const int qp = (qpy_at(f, lx - 1, yu) + qpy_at(f, lx, yu) + 1) >> 1;
        p.beta = c_betatable[clip3i(qp + f.db[cb].beta_offset, 0, 51)];
        const int tco = f.db[cb].tc_offset;
        p.tc = c_tctable[clip3i(qp + 2 * (bs - 1) + (tco >> 1 << 1), 0, 53)];
    } else {
        if (bs != 2) return p;
        const int qp = (qpy_at(f, lx - 1, ysl) + qpy_at(f, lx, ysl) + 1) >> 1;   // per segment (:614-615)
        p.tc = chroma_tc(f, qp, c, f.db[cb].tc_offset);
    }
    if (f.pcm_or_bypass) { p.no_p = pcm_at(f, lx - 1, ysl); p.no_q = pcm_at(f, lx, ysl); }
    p.on = true;
    return p;
}

// parameters of the 4-column segment starting at plane column xs of the HORIZONTAL edge at plane row ey
__device__ __forceinline__ SegParam hseg_param(const FrameDev &f, int c, int xs, int ey)
{
    SegParam p = {0, 0, 0, 0, false};
    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
    const int xsl = xs << hs, ly = ey << vs;
    const int bs = f.hbs[(xsl + ly * f.bs_width) >> 2];
    const int unit = 8 << hs;                        // edge unit length in luma columns
    const int xu = xsl & ~(unit - 1);
    // CTB that runs this unit in the reference: the one owning column xu + unit, the last CTB
    // column taking the remainder (hevc_filter.c:661-667, 715-718)
    const int xproc = (xu + unit < f.width) ? xu + unit : xu;
    const int cb_proc = ctb_at(f, xproc, ly), cb_here = ctb_at(f, xu, ly);
    if (c == 0) {
        if (!bs) return p;
        const int qp = (qpy_at(f, xu, ly - 1) + qpy_at(f, xu, ly) + 1) >> 1;
        p.beta = c_betatable[clip3i(qp + f.db[cb_here].beta_offset, 0, 51)];
        const int tco = f.db[cb_proc].tc_offset;
        p.tc = c_tctable[clip3i(qp + 2 * (bs - 1) + (tco >> 1 << 1), 0, 53)];
    } else {
        if (bs != 2) return p;
        const int qp = (qpy_at(f, xsl, ly - 1) + qpy_at(f, xsl, ly) + 1) >> 1;
        const bool first = xsl == xu;                // first segment of the unit
        p.tc = chroma_tc(f, qp, c, f.db[first ? cb_here : cb_proc].tc_offset);
    }
    if (f.pcm_or_bypass) { p.no_p = pcm_at(f, xsl, ly - 1); p.no_q = pcm_at(f, xsl, ly); }
    p.on = true;
    return p;
}

// one 4-line luma segment: l[line][0..7] = P3 P2 P1 P0 Q0 Q1 Q2 Q3 (template :3390-3470)
__device__ __forceinline__ void luma_segment(int (&l)[4][8], int bd, int beta, int tc, int no_p, int no_q)
{
    const int maxv = (1 << bd) - 1;
    beta <<= bd - 8;
    tc <<= bd - 8;
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
                if (!no_p) l[d][3] = clip3i(p0 + delta0, 0, maxv);
                if (!no_q) l[d][4] = clip3i(q0 - delta0, 0, maxv);
                if (!no_p && nd_p2)
                    l[d][2] = clip3i(p1 + clip3i((((p2 + p0 + 1) >> 1) - p1 + delta0) >> 1, -tc_2, tc_2), 0, maxv);
                if (!no_q && nd_q2)
                    l[d][5] = clip3i(q1 + clip3i((((q2 + q0 + 1) >> 1) - q1 - delta0) >> 1, -tc_2, tc_2), 0, maxv);
            }
        }
    }
}

// one 4-line chroma segment: l[line][2..5] = P1 P0 Q0 Q1 (template :3473-3505)
__device__ __forceinline__ void chroma_segment(int (&l)[4][8], int bd, int tc, int no_p, int no_q)
{
    const int maxv = (1 << bd) - 1;
    tc <<= bd - 8;
    if (tc <= 0) return;
#pragma unroll
    for (int d = 0; d < 4; d++) {
        const int p1 = l[d][2], p0 = l[d][3], q0 = l[d][4], q1 = l[d][5];
        const int delta0 = clip3i((((q0 - p0) * 4) + p1 - q1 + 4) >> 3, -tc, tc);
        if (!no_p) l[d][3] = clip3i(p0 + delta0, 0, maxv);
        if (!no_q) l[d][4] = clip3i(q0 - delta0, 0, maxv);
    }
}

// four consecutive samples, aligned to 4 * sizeof(pixel)
__device__ __forceinline__ void ld4(const uint16_t *p, int *v)
{
    const uint2 w = *(const uint2 *)p;
    v[0] = w.x & 0xffff; v[1] = w.x >> 16; v[2] = w.y & 0xffff; v[3] = w.y >> 16;
}
__device__ __forceinline__ void ld4(const uint8_t *p, int *v)
{
    const uint32_t w = *(const uint32_t *)p;
    v[0] = w & 0xff; v[1] = (w >> 8) & 0xff; v[2] = (w >> 16) & 0xff; v[3] = w >> 24;
}
__device__ __forceinline__ void st4(uint16_t *p, const int *v)
{
    *(uint2 *)p = make_uint2((uint32_t)v[0] | ((uint32_t)v[1] << 16), (uint32_t)v[2] | ((uint32_t)v[3] << 16));
}
__device__ __forceinline__ void st4(uint8_t *p, const int *v)
{
    *(uint32_t *)p = (uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16) | ((uint32_t)v[3] << 24);
}

// eight consecutive samples, aligned to 8 * sizeof(pixel)
__device__ __forceinline__ void ld8(const uint16_t *p, int *v)
{
    const uint4 w = *(const uint4 *)p;
    v[0] = w.x & 0xffff; v[1] = w.x >> 16; v[2] = w.y & 0xffff; v[3] = w.y >> 16;
    v[4] = w.z & 0xffff; v[5] = w.z >> 16; v[6] = w.w & 0xffff; v[7] = w.w >> 16;
}
__device__ __forceinline__ void ld8(const uint8_t *p, int *v)
{
    const uint2 w = *(const uint2 *)p;
    v[0] = w.x & 0xff; v[1] = (w.x >> 8) & 0xff; v[2] = (w.x >> 16) & 0xff; v[3] = w.x >> 24;
    v[4] = w.y & 0xff; v[5] = (w.y >> 8) & 0xff; v[6] = (w.y >> 16) & 0xff; v[7] = w.y >> 24;
}
__device__ __forceinline__ void st8(uint16_t *p, const int *v)
{
    *(uint4 *)p = make_uint4((uint32_t)v[0] | ((uint32_t)v[1] << 16), (uint32_t)v[2] | ((uint32_t)v[3] << 16),
                             (uint32_t)v[4] | ((uint32_t)v[5] << 16), (uint32_t)v[6] | ((uint32_t)v[7] << 16));
}
__device__ __forceinline__ void st8(uint8_t *p, const int *v)
{
    *(uint2 *)p = make_uint2((uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16) | ((uint32_t)v[3] << 24),
                             (uint32_t)v[4] | ((uint32_t)v[5] << 8) | ((uint32_t)v[6] << 16) | ((uint32_t)v[7] << 24));
}

// ---------------------------------------------------------------- SAO of eight samples
// Offsets of a CTB packed as signed bytes and picked with one PRMT (selector nibble bit 3
// replicates the sign of the chosen byte, so bytes 1..3 of the result sign-extend byte 0):
//   edge  byte[2 + sign(v-a) + sign(v-b)] = offset_val[edge_idx[..]],  edge_idx = {1,2,0,3,4}
//   band  byte[k] = offset_val[k + 1] for k < 4, byte 4 = 0
__device__ __forceinline__ int packed_offset(uint32_t lo, uint32_t hi, int idx)
{
    // prmt.b32 directly: __byte_perm() drops selector bit 3
    int r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"((uint32_t)idx * 0x1111u + 0x8880u));
    return r;
}

// CTB away from every border rule: neighbours exist, nothing is restored.
// rowa / rowb point at sample x0 of the rows holding neighbour a / b; DXA = column step to a.
template <typename pixel, int DXA>
__device__ __forceinline__ void sao_edge_fast(const pixel *rowa, const pixel *rowb, const int (&v)[8], uint32_t lo,
                                              uint32_t hi, int maxv, int (&o)[8])
{
    int A[10], B[10];                                  // A[k] = rowa[k - 1]
    ld8(rowa, A + 1);
    ld8(rowb, B + 1);
    if (DXA < 0) { A[0] = rowa[-1]; B[9] = rowb[8]; }
    if (DXA > 0) { A[9] = rowa[8]; B[0] = rowb[-1]; }
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int a = A[i + 1 + DXA], b = B[i + 1 - DXA];
        const int sgn = clip3i(v[i] - a, -1, 1) + clip3i(v[i] - b, -1, 1);
        o[i] = clip3i(v[i] + packed_offset(lo, hi, sgn + 2), 0, maxv);
    }
}

// sao_band_filter_0 (template :310-335) on eight samples; lo = offsets of bands 0..3 packed, hi = 0
__device__ __forceinline__ void sao_band_fast(const int (&v)[8], uint32_t lo, uint32_t hi, int pos, int sh, int maxv, int (&o)[8])
{
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int kb = ((v[i] >> sh) - pos) & 31;
        o[i] = clip3i(v[i] + packed_offset(lo, hi, min(kb, 4)), 0, maxv);
    }
}

// per (plane, CTB of the region): what the SAO step needs, built once per CTA
struct SaoRec {
    uint32_t lo, hi;          // packed offsets (fast paths)
    uint32_t misc;            // type | eo << 2 | fast << 4 | band_position << 8
    uint32_t pad;
};

// Any CTB: picture borders, slice / tile restore flags, pcm / bypass restore, wide offsets.
// Follows sao_edge_filter + sao_edge_restore_{0,1} (template :342-529), sao_band_filter_0
// (:310-335) and restore_tqb_pixels (hevc_filter.c:229-260) sample by sample.
// srow = tile row of plane row y, indexed by plane column; bw = tile pitch in samples.
template <typename pixel>
__device__ __noinline__ void sao_slow(const FrameDev &f, int c, int x0, int y, const pixel *srow, int bw, pixel *drow)
{
    const int bd = f.bd, maxv = (1 << bd) - 1;
    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
    const int pw = f.pw[c], ph = f.ph[c];
    const int lcw = f.log2_ctb - hs, lch = f.log2_ctb - vs;
    const int cw = 1 << lcw, chh = 1 << lch;
    const int ctbx = x0 >> lcw, ctby = y >> lch;
    const int rs = ctby * f.ctb_w + ctbx;
    const ohcu_sao sp = f.sao[rs];
    const int type = sp.type_idx[c];
    int off[5];
#pragma unroll
    for (int i = 0; i < 5; i++) off[i] = sp.offset_val[c][i];
    const int cx0 = ctbx * cw, cy0 = ctby * chh;          // CTB origin in the plane
    const int w = min(cw, pw - cx0), h = min(chh, ph - cy0);
    const int ly = y - cy0;
    const bool tqb = f.pcm_or_bypass && f.is_pcm;
    // the min-PU range is taken from the LUMA origin of the CTB plus the PLANE's width/height, as in the reference
    const int lx_ctb = ctbx << f.log2_ctb, ly_ctb = ctby << f.log2_ctb;
    const int pu_xmax = (lx_ctb + w) >> f.log2_min_pu, pu_ymax = (ly_ctb + h) >> f.log2_min_pu;
    const int eo = sp.eo_class[c];
    const int dxa = (eo == 0 || eo == 2) ? -1 : (eo == 3 ? 1 : 0);
    const int dya = eo == 0 ? 0 : -1;
    const int fl = f.sao_ctb_flags[rs];
    const bool b0 = ctbx == 0, b1 = ctby == 0, b2 = ctbx == f.ctb_w - 1, b3 = ctby == f.ctb_h - 1;
    const bool notv = eo != OHCU_SAO_EO_VERT, noth = eo != OHCU_SAO_EO_HORIZ;
    const int wr = w - ((notv && b2) ? 1 : 0);             // width / height after the border
    const int hr = h - ((noth && b3) ? 1 : 0);             // columns / rows were peeled off
    const int init_x = (notv && b0) ? 1 : 0;
    const bool restore = fl & SAOF_RESTORE;
    const bool ve0 = fl & SAOF_VERT0, ve1 = fl & SAOF_VERT1;
    const bool he0 = fl & SAOF_HORIZ0, he1 = fl & SAOF_HORIZ1;
    const bool d0 = fl & SAOF_DIAG0, d1 = fl & SAOF_DIAG1, d2 = fl & SAOF_DIAG2, d3 = fl & SAOF_DIAG3;
    const int keep_ul = !d0 && eo == 2 && !b0 && !b1, keep_ur = !d1 && eo == 3 && !b1 && !b2;
    const int keep_lr = !d2 && eo == 2 && !b2 && !b3, keep_ll = !d3 && eo == 3 && !b0 && !b3;
    const int ya = y + dya, yb = y - dya;
    const bool ya_ok = ya >= 0 && ya < ph, yb_ok = yb >= 0 && yb < ph;
    const pixel *rowa = srow + dya * bw, *rowb = srow - dya * bw;
#pragma unroll 1
    for (int i = 0; i < 8; i++) {
        const int x = x0 + i, lx = x - cx0, v = srow[x];
        int q;
        if (type == OHCU_SAO_BAND) {
            const int k = ((v >> (bd - 5)) - sp.band_position[c]) & 31;
            q = clip3i(v + (k < 4 ? off[k + 1] : 0), 0, maxv);
        } else {
            const int xa = x + dxa, xb = x - dxa;
            // out-of-picture neighbours only occur where the border rules below override the result
            const int a = (ya_ok && xa >= 0 && xa < pw) ? rowa[xa] : v;
            const int b = (yb_ok && xb >= 0 && xb < pw) ? rowb[xb] : v;
            const int sgn = ((v > a) - (v < a)) + ((v > b) - (v < b));   // -2..2
            // edge_idx[] = {1, 2, 0, 3, 4}
            const int cat = sgn == 0 ? 0 : (sgn == -2 ? 1 : (sgn == -1 ? 2 : (sgn == 1 ? 3 : 4)));
            q = clip3i(v + off[cat], 0, maxv);
            const int keep = clip3i(v + off[0], 0, maxv);
            // picture borders (restore_0 part)
            if (notv && ((b0 && lx == 0) || (b2 && lx == w - 1))) q = keep;
            if (noth && ((b1 && ly == 0) || (b3 && ly == h - 1)) && lx >= init_x && lx < wr) q = keep;
            if (restore) {                                     // slice / tile boundaries (restore_1)
                if (ve0 && notv && lx == 0 && ly >= keep_ul && ly < hr - keep_ll) q = v;
                if (ve1 && notv && lx == wr - 1 && ly >= keep_ur && ly < hr - keep_lr) q = v;
                if (he0 && noth && ly == 0 && lx >= init_x + keep_ul && lx < wr - keep_ur) q = v;
                if (he1 && noth && ly == hr - 1 && lx >= init_x + keep_ll && lx < wr - keep_lr) q = v;
                if (d0 && eo == 2 && lx == 0 && ly == 0) q = v;
                if (d1 && eo == 3 && lx == wr - 1 && ly == 0) q = v;
                if (d2 && eo == 2 && lx == wr - 1 && ly == hr - 1) q = v;
                if (d3 && eo == 3 && lx == 0 && ly == hr - 1) q = v;
            }
        }
        if (tqb) {
            const int pux = (x << hs) >> f.log2_min_pu, puy = (y << vs) >> f.log2_min_pu;
            if (pux < pu_xmax && puy < pu_ymax && f.is_pcm[puy * f.min_pu_w + pux]) q = v;
        }
        drow[x] = (pixel)q;
    }
}

constexpr int LF_MAX_SEGS = 512;               // per direction: 3 x 162 for 4:4:4

// ---------------------------------------------------------------- one edge pass, a thread per LINE
// The four lines of a segment sit on four neighbouring lanes.  The decisions of the reference
// (template :3390-3420) read lines 0 and 3 only: every lane computes dp, dq and the strong-filter
// test of its own line, the quad's lanes 0 and 3 hand theirs out by shuffle / ballot, and each
// lane then filters its line.  All 256 threads work on 4 n lines instead of n threads on n
// segments (n is 130-190 per pass of a 64x64 region: half the CTA used to wait at the barrier).
template <typename pixel, bool HORIZ>
__device__ __forceinline__ void edge_pass(uint8_t *smem, uint32_t toff0, uint32_t toff1, uint32_t toff2, int bw0, int bwc,
                                          const uint2 *list, int nl, int nch, int bd, int tid)
{
    const int n = nl + nch, n4 = 4 * n;
    const int q0 = tid & 28;                                   // first lane of this quad within the warp
    const int maxv = (1 << bd) - 1;
    for (int base = 0; base < n4; base += LF_THREADS) {
        const int k4 = base + tid, k = k4 >> 2, d = k4 & 3;
        const bool act = k < n;
        const uint2 r = act ? list[k < nl ? k : LF_MAX_SEGS - 1 - (k - nl)] : make_uint2(0u, 0u);
        const int c = (r.x >> 16) & 3;
        const int bw = c == 0 ? bw0 : bwc;
        pixel *p = (pixel *)(smem + (c == 0 ? toff0 : (c == 1 ? toff1 : toff2))) + (r.x & 0xffff);
        int l[8];
        if (act) {
            if (!HORIZ) { ld4(p + d * bw, l); ld4(p + d * bw + 4, l + 4); }
            else {
#pragma unroll
                for (int q = 0; q < 8; q++) l[q] = p[q * bw + d];
            }
        } else {
#pragma unroll
            for (int q = 0; q < 8; q++) l[q] = 0;
        }
        const int beta = (int)(r.y & 0xffff) << (bd - 8), tc = (int)(r.y >> 16) << (bd - 8);
        const int dp = abs(l[1] - 2 * l[2] + l[3]), dq = abs(l[6] - 2 * l[5] + l[4]);
        const int ds = dp + dq;
        const bool sline = abs(l[0] - l[3]) + abs(l[7] - l[4]) < (beta >> 3) && abs(l[3] - l[4]) < ((tc * 5 + 1) >> 1) &&
                           (ds << 1) < (beta >> 2);
        const int ds0 = __shfl_sync(0xffffffffu, ds, q0), ds3 = __shfl_sync(0xffffffffu, ds, q0 + 3);
        const int dp0 = __shfl_sync(0xffffffffu, dp, q0), dp3 = __shfl_sync(0xffffffffu, dp, q0 + 3);
        const unsigned sb = __ballot_sync(0xffffffffu, sline);
        if (!act) continue;
        const int no_p = (r.x >> 18) & 1, no_q = (r.x >> 19) & 1;
        if (c == 0) {
            if (ds0 + ds3 >= beta) continue;
            const int p3 = l[0], p2 = l[1], p1 = l[2], p0 = l[3], q0_ = l[4], q1 = l[5], q2 = l[6], q3 = l[7];
            if (((sb >> q0) & 1u) && ((sb >> (q0 + 3)) & 1u)) {
                const int tc2 = tc << 1;
                if (!no_p) {
                    l[3] = p0 + clip3i(((p2 + 2 * p1 + 2 * p0 + 2 * q0_ + q1 + 4) >> 3) - p0, -tc2, tc2);
                    l[2] = p1 + clip3i(((p2 + p1 + p0 + q0_ + 2) >> 2) - p1, -tc2, tc2);
                    l[1] = p2 + clip3i(((2 * p3 + 3 * p2 + p1 + p0 + q0_ + 4) >> 3) - p2, -tc2, tc2);
                }
                if (!no_q) {
                    l[4] = q0_ + clip3i(((p1 + 2 * p0 + 2 * q0_ + 2 * q1 + q2 + 4) >> 3) - q0_, -tc2, tc2);
                    l[5] = q1 + clip3i(((p0 + q0_ + q1 + q2 + 2) >> 2) - q1, -tc2, tc2);
                    l[6] = q2 + clip3i(((2 * q3 + 3 * q2 + q1 + q0_ + p0 + 4) >> 3) - q2, -tc2, tc2);
                }
            } else {
                const int side = (beta + (beta >> 1)) >> 3, tc_2 = tc >> 1;
                const bool nd_p2 = dp0 + dp3 < side, nd_q2 = (ds0 - dp0) + (ds3 - dp3) < side;
                int delta0 = (9 * (q0_ - p0) - 3 * (q1 - p1) + 8) >> 4;
                if (abs(delta0) >= 10 * tc) continue;
                delta0 = clip3i(delta0, -tc, tc);
                if (!no_p) l[3] = clip3i(p0 + delta0, 0, maxv);
                if (!no_q) l[4] = clip3i(q0_ - delta0, 0, maxv);
                if (!no_p && nd_p2) l[2] = clip3i(p1 + clip3i((((p2 + p0 + 1) >> 1) - p1 + delta0) >> 1, -tc_2, tc_2), 0, maxv);
                if (!no_q && nd_q2) l[5] = clip3i(q1 + clip3i((((q2 + q0_ + 1) >> 1) - q1 - delta0) >> 1, -tc_2, tc_2), 0, maxv);
            }
        } else {
            const int p1 = l[2], p0 = l[3], q0_ = l[4], q1 = l[5];
            const int delta0 = clip3i((((q0_ - p0) * 4) + p1 - q1 + 4) >> 3, -tc, tc);
            if (!no_p) l[3] = clip3i(p0 + delta0, 0, maxv);
            if (!no_q) l[4] = clip3i(q0_ - delta0, 0, maxv);
        }
        if (!HORIZ) { st4(p + d * bw, l); st4(p + d * bw + 4, l + 4); }
        else {
#pragma unroll
            for (int q = 1; q < 7; q++) p[q * bw + d] = (pixel)l[q];
        }
    }
}

// ---------------------------------------------------------------- the kernel
// Edge segments that actually filter, found while the tile is in flight.
//   w0  tile offset of the segment's first sample (P3 of line 0) | plane << 16 | no_p << 18 | no_q << 19
//   w1  beta | tc << 16   (table values, scaled to the bit depth by the filter)

struct LfGeom {
    int TW, TH, BW, BH, px0, py0, lw;          // lw = log2(TW)
};

// flags: bit 0 deblock, bit 1 SAO.  Source: the job's reconstructed picture through `maps`
// (= f.work); destination f.out (the DPB slot).
#ifndef LF_MINB
#define LF_MINB 4
#endif
// -DLF_PROFILE: thread LF_PROF_TID of every CTA adds the cycles between phase marks to g_lf_prof (tools/lf_phases.py)
#ifdef LF_PROFILE
#ifndef LF_PROF_TID
#define LF_PROF_TID 0
#endif
__device__ unsigned long long g_lf_prof[16];
#define LF_MARK(i) do { if (threadIdx.x == LF_PROF_TID) { const long long t_ = clock64(); atomicAdd(&g_lf_prof[i], (unsigned long long)(t_ - lf_t)); lf_t = t_; } } while (0)
#else
#define LF_MARK(i) do { } while (0)
#endif
template <typename pixel>
__global__ void __launch_bounds__(LF_THREADS, LF_MINB) loopfilter_kernel(const __grid_constant__ FrameDev f,
                                                                 const __grid_constant__ LfMaps maps, int flags)
{
    extern __shared__ __align__(128) uint8_t lf_smem[];
#ifdef LF_PROFILE
    long long lf_t = clock64();
#endif
    constexpr int AX = 16 / (int)sizeof(pixel);           // samples per 16 bytes: TMA boxes start on them
    if (f.err && *f.err) return;                          // malformed lists (prepare.cu)
    const int tid = threadIdx.x;
    const int bd = f.bd;
    const int np = f.nplanes;
    LfGeom g[3];
    uint32_t toff[3], total = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
        g[c].lw = 6 - hs;
        g[c].TW = LF_REGION >> hs; g[c].TH = LF_REGION >> vs;
        g[c].BW = g[c].TW + 2 * AX; g[c].BH = g[c].TH + 8;
        g[c].px0 = (blockIdx.x * LF_REGION) >> hs;
        g[c].py0 = (blockIdx.y * LF_REGION) >> vs;
        toff[c] = total;
        if (c < np) total += (uint32_t)(g[c].BW * g[c].BH * sizeof(pixel));
    }
    uint64_t *mbar = (uint64_t *)(lf_smem + total);
    int *cnt = (int *)(mbar + 1);                              // [dir][luma, chroma]
    SaoRec *srec = (SaoRec *)(lf_smem + total + 32);           // [plane][CTB of the region], <= 3 x 16
    uint2 *seg = (uint2 *)(srec + 48);                         // [dir][LF_MAX_SEGS]
#define TILE(c) ((pixel *)(lf_smem + toff[c]))

    // ---- fetch: one box per plane, origin (px0 - AX, py0 - 4); outside the picture reads as zero
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(lf_smem_u32(mbar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(lf_smem_u32(mbar)), "r"(total) : "memory");
#pragma unroll
        for (int c = 0; c < 3; c++)
            if (c < np)
                asm volatile(
                    "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                    ::"r"(lf_smem_u32(TILE(c))), "l"(&maps.m[c]), "r"(g[c].px0 - AX), "r"(g[c].py0 - 4), "r"(lf_smem_u32(mbar)) : "memory");
    }
    if (tid < 4) cnt[tid] = 0;
    __syncthreads();
    LF_MARK(0);

    // ---- while the tile is in flight: side information -> segment lists and SAO records
    const bool do_db = (flags & 1) && f.vbs;
    if (do_db) {
        // item space: [V: plane 0, 1, 2][H: plane 0, 1, 2]
        int nV[3], nH[3], nvt = 0, nht = 0;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            nV[c] = c < np ? (g[c].TW / 8 + 1) * (g[c].BH / 4) : 0;
            nH[c] = c < np ? (g[c].TH / 8 + 1) * ((g[c].TW + 8) / 4) : 0;
            nvt += nV[c]; nht += nH[c];
        }
        for (int it = tid; it < nvt + nht; it += LF_THREADS) {
            const int dir = it >= nvt;
            int k = dir ? it - nvt : it, c = 0;
#pragma unroll
            for (int cc = 0; cc < 2; cc++) {
                const int n = dir ? nH[cc] : nV[cc];
                if (c == cc && k >= n) { k -= n; c = cc + 1; }
            }
            // planes 1 and 2 share their geometry; the two grid widths are 64 and 32 samples, so the
            // divisions are by constants (9, 5 edges per row of items; 18, 10 segments per edge row)
            const int TW = c == 0 ? g[0].TW : g[1].TW, BW = c == 0 ? g[0].BW : g[1].BW;
            const int px0 = c == 0 ? g[0].px0 : g[1].px0, py0 = c == 0 ? g[0].py0 : g[1].py0;
            SegParam p;
            int base;
            if (!dir) {
                const int s = TW == 64 ? k / 9 : k / 5, i = k - s * (TW / 8 + 1);
                const int ex = px0 + 8 * i, ys = py0 - 4 + 4 * s;
                if (ex <= 0 || ex >= f.pw[c] || ys < 0 || ys >= f.ph[c]) continue;
                p = vseg_param(f, c, ex, ys);
                base = (4 * s) * BW + 8 * i + AX - 4;
            } else {
                const int j = TW == 64 ? k / 18 : k / 10, s = k - j * ((TW + 8) / 4);
                const int ey = py0 + 8 * j, xs = px0 - 4 + 4 * s;
                if (ey <= 0 || ey >= f.ph[c] || xs < 0 || xs >= f.pw[c]) continue;
                p = hseg_param(f, c, xs, ey);
                base = (8 * j) * BW + 4 * s + AX - 4;             // row py0 - 4 is tile row 0
            }
            if (!p.on || (c && p.tc == 0)) continue;
            const int slot = atomicAdd(&cnt[dir * 2 + (c ? 1 : 0)], 1);
            // luma segments from the front, chroma from the back: warps stay on one filter
            seg[dir * LF_MAX_SEGS + (c ? LF_MAX_SEGS - 1 - slot : slot)] =
                make_uint2((uint32_t)base | (uint32_t)c << 16 | (p.no_p ? 1u << 18 : 0u) | (p.no_q ? 1u << 19 : 0u),
                           (uint32_t)p.beta | (uint32_t)p.tc << 16);
        }
    }
    LF_MARK(9);
    const int lctb = min(f.log2_ctb, 6);                       // CTBs of the region: nreg x nreg
    const int nreg = LF_REGION >> lctb;
    // the SAO records go to the LAST threads: the first ones carry one segment item more (524 items on
    // 256 threads at 4:2:0) and were the CTA's critical path with the records on top
    const int stid = LF_THREADS - 1 - tid;
    if (stid < np * nreg * nreg) {
        const int c = stid / (nreg * nreg), k = stid - c * nreg * nreg;
        const int ctbx = ((blockIdx.x * LF_REGION) >> f.log2_ctb) + (k % nreg);
        const int ctby = ((blockIdx.y * LF_REGION) >> f.log2_ctb) + (k / nreg);
        SaoRec r = {0, 0, OHCU_SAO_NOT_APPLIED, 0};
        if ((flags & 2) && f.sao && ctbx < f.ctb_w && ctby < f.ctb_h) {
            const int rs = ctby * f.ctb_w + ctbx;
            const ohcu_sao sp = f.sao[rs];
            const int type = sp.type_idx[c];
            if (type == OHCU_SAO_BAND || type == OHCU_SAO_EDGE) {
                int off[5];
                bool fast = !(f.pcm_or_bypass && f.is_pcm);
#pragma unroll
                for (int i = 0; i < 5; i++) {
                    off[i] = sp.offset_val[c][i];
                    if (off[i] < -128 || off[i] > 127) fast = false;
                }
                const int eo = sp.eo_class[c];
                if (type == OHCU_SAO_EDGE) {
                    const bool notv = eo != OHCU_SAO_EO_VERT, noth = eo != OHCU_SAO_EO_HORIZ;
                    const bool bx = ctbx == 0 || ctbx == f.ctb_w - 1, by = ctby == 0 || ctby == f.ctb_h - 1;
                    const int fl = f.sao_ctb_flags[rs];
                    if ((notv && bx) || (noth && by) || ((fl & SAOF_RESTORE) && (fl & ~SAOF_RESTORE))) fast = false;
                    r.lo = (off[1] & 0xff) | (off[2] & 0xff) << 8 | (off[0] & 0xff) << 16 | (uint32_t)(off[3] & 0xff) << 24;
                    r.hi = off[4] & 0xff;
                } else {
                    r.lo = (off[1] & 0xff) | (off[2] & 0xff) << 8 | (off[3] & 0xff) << 16 | (uint32_t)(off[4] & 0xff) << 24;
                    r.hi = 0;
                }
                r.misc = (uint32_t)type | (uint32_t)eo << 2 | (fast ? 1u << 4 : 0u) | (uint32_t)sp.band_position[c] << 8;
            }
        }
        srec[stid] = r;
    }
    LF_MARK(1);
    __syncthreads();
    LF_MARK(2);
    lf_mbar_wait(mbar, 0);
    LF_MARK(3);

    if (do_db) {
        // ---- 1. vertical edges: segment = 4 rows x (P3..Q3); lines are rows
        edge_pass<pixel, false>(lf_smem, toff[0], toff[1], toff[2], g[0].BW, g[1].BW, seg, cnt[0], cnt[1], bd, tid);
        LF_MARK(4);
        __syncthreads();
        LF_MARK(5);
        // ---- 2. horizontal edges: segment = (P3..Q3) rows x 4 columns; lines are columns
        edge_pass<pixel, true>(lf_smem, toff[0], toff[1], toff[2], g[0].BW, g[1].BW, seg + LF_MAX_SEGS, cnt[2], cnt[3], bd, tid);
        LF_MARK(6);
        __syncthreads();
        LF_MARK(7);
    }

    // ---- 3. SAO (or plain copy) of the region into the DPB slot, eight samples per thread step
    const int maxv = (1 << bd) - 1;
    int nU[3], nut = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) { nU[c] = c < np ? (g[c].TW >> 3) * g[c].TH : 0; nut += nU[c]; }
    for (int it = tid; it < nut; it += LF_THREADS) {
        int k = it, c = 0;
        if (k >= nU[0]) { k -= nU[0]; c = 1; if (k >= nU[1]) { k -= nU[1]; c = 2; } }
        const LfGeom G = c == 0 ? g[0] : (c == 1 ? g[1] : g[2]);
        const int r = k >> (G.lw - 3), u = k & ((1 << (G.lw - 3)) - 1);
        const int x0 = G.px0 + 8 * u, y = G.py0 + r;
        if (x0 >= f.pw[c] || y >= f.ph[c]) continue;
        const int bw = G.BW;
        const pixel *trow = (const pixel *)(lf_smem + (c == 0 ? toff[0] : (c == 1 ? toff[1] : toff[2]))) + (r + 4) * bw + AX;
        const pixel *sp8 = trow + 8 * u;                               // deblocked samples (x0.., y)
        pixel *drow = (pixel *)(f.out.p[c] + (size_t)y * f.out.pitch[c]);
        // CTB of the unit inside the region
        const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
        const int ci = (r >> (lctb - vs)) * nreg + ((8 * u) >> (lctb - hs));
        const SaoRec rec = srec[c * nreg * nreg + ci];
        const int type = rec.misc & 3;
        int v[8], o[8];
        if (type == OHCU_SAO_NOT_APPLIED) {
            ld8(sp8, o);
        } else if (!(rec.misc & 16)) {
            sao_slow<pixel>(f, c, x0, y, trow - G.px0, bw, drow);
            continue;
        } else if (type == OHCU_SAO_BAND) {                            // sao_band_filter_0, template :310-335
            ld8(sp8, v);
            sao_band_fast(v, rec.lo, rec.hi, (rec.misc >> 8) & 31, bd - 5, maxv, o);
        } else {                                                       // sao_edge_filter, template :342-378
            ld8(sp8, v);
            const int eo = (rec.misc >> 2) & 3;
            if (eo == 0) sao_edge_fast<pixel, -1>(sp8, sp8, v, rec.lo, rec.hi, maxv, o);
            else if (eo == 1) sao_edge_fast<pixel, 0>(sp8 - bw, sp8 + bw, v, rec.lo, rec.hi, maxv, o);
            else if (eo == 2) sao_edge_fast<pixel, -1>(sp8 - bw, sp8 + bw, v, rec.lo, rec.hi, maxv, o);
            else sao_edge_fast<pixel, 1>(sp8 - bw, sp8 + bw, v, rec.lo, rec.hi, maxv, o);
        }
        st8(drow + x0, o);
    }
    LF_MARK(8);
#undef TILE
}

#ifdef LF_PROFILE
extern "C" int ohcu_lf_profile(unsigned long long *out, int reset)
{
    if (cudaMemcpyFromSymbol(out, g_lf_prof, sizeof(g_lf_prof)) != cudaSuccess) return -1;
    if (reset) { unsigned long long z[16] = {}; cudaMemcpyToSymbol(g_lf_prof, z, sizeof(z)); }
    return 0;
}
#endif

size_t loopfilter_smem_bytes(const FrameDev &f)
{
    const int P = f.bd > 8 ? 2 : 1, AX = 16 / P;
    size_t total = 0;
    for (int c = 0; c < f.nplanes; c++) {
        const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
        total += (size_t)((LF_REGION >> hs) + 2 * AX) * ((LF_REGION >> vs) + 8) * P;
    }
    return total + 32 + 48 * sizeof(SaoRec) + 2 * LF_MAX_SEGS * sizeof(uint2);
}

// box of plane c that loopfilter_kernel fetches (api.cu encodes one tensor map per job and plane)
void loopfilter_box(const FrameDev &f, int c, int *bw, int *bh)
{
    const int P = f.bd > 8 ? 2 : 1, AX = 16 / P;
    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
    *bw = (LF_REGION >> hs) + 2 * AX;
    *bh = (LF_REGION >> vs) + 8;
}

// per device, from ohcu_create (see recon_set_attrs)
int loopfilter_set_attrs()
{
    if (cudaFuncSetAttribute(loopfilter_kernel<uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024) != cudaSuccess ||
        cudaFuncSetAttribute(loopfilter_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024) != cudaSuccess)
        return -1;
    return 0;
}

int launch_loopfilter(const FrameDev &f, const LfMaps &maps, int flags, cudaStream_t st)
{
    if (!(flags & 3)) return 0;
    const size_t smem = loopfilter_smem_bytes(f);
    dim3 grid((f.width + LF_REGION - 1) / LF_REGION, (f.height + LF_REGION - 1) / LF_REGION);
    if (f.bd > 8) loopfilter_kernel<uint16_t><<<grid, LF_THREADS, smem, st>>>(f, maps, flags);
    else          loopfilter_kernel<uint8_t><<<grid, LF_THREADS, smem, st>>>(f, maps, flags);
    return 1;
}

// ================================================================= per-entry test shims
// hevc_{h,v}_loop_filter_{luma,chroma} (template :3377-3553): 8 lines = two 4-line segments
// through luma_segment / chroma_segment; pix = first Q sample.
template <typename pixel>
__global__ void dsp_deblock_kernel(pixel *pix, int stride, int luma, int vertical, int bd, int beta, int tc0, int tc1, int nop0,
                                   int nop1, int noq0, int noq1)
{
    const int s = threadIdx.x;                    // segment
    if (s >= 2) return;
    const int xs = vertical ? 1 : stride, ys = vertical ? stride : 1;      // along the line / to the next line
    pixel *base = pix + 4 * s * ys - 4 * xs;
    int l[4][8];
    for (int d = 0; d < 4; d++)
        for (int k = 0; k < 8; k++) l[d][k] = base[d * ys + k * xs];
    const int tc = s ? tc1 : tc0, no_p = s ? nop1 : nop0, no_q = s ? noq1 : noq0;
    if (luma) luma_segment(l, bd, beta, tc, no_p, no_q);
    else chroma_segment(l, bd, tc, no_p, no_q);
    for (int d = 0; d < 4; d++)
        for (int k = 0; k < 8; k++) base[d * ys + k * xs] = (pixel)l[d][k];
}

int dsp_loop_filter(int bd, int luma, int vertical, uint8_t *pix, ptrdiff_t stride, int beta, const int *tc, const uint8_t *no_p,
                    const uint8_t *no_q)
{
    const int P = bd > 8 ? 2 : 1;
    // 8 samples across the edge x 8 lines along it, copied as a 8 x 8 patch whose centre is pix
    uint8_t *d = nullptr;
    if (cudaMalloc((void **)&d, 16 * 16 * 2) != cudaSuccess) return -1;
    uint8_t *h0 = vertical ? pix - 4 * P : pix - 4 * stride;
    const int rows = 8, wbytes = 8 * P;
    int rc = -1;
    if (cudaMemcpy2D(d, 32, h0, stride, wbytes, rows, cudaMemcpyHostToDevice) == cudaSuccess) {
        // device patch: pitch 32 bytes; pix sits at column 4 (vertical) or row 4 (horizontal)
        uint8_t *dp = vertical ? d + 4 * P : d + 4 * 32;
        if (bd > 8) dsp_deblock_kernel<uint16_t><<<1, 32>>>((uint16_t *)dp, 16, luma, vertical, bd, beta, tc[0], tc[1], no_p[0], no_p[1], no_q[0], no_q[1]);
        else dsp_deblock_kernel<uint8_t><<<1, 32>>>(dp, 32, luma, vertical, bd, beta, tc[0], tc[1], no_p[0], no_p[1], no_q[0], no_q[1]);
        if (cudaDeviceSynchronize() == cudaSuccess &&
            cudaMemcpy2D(h0, stride, d, 32, wbytes, rows, cudaMemcpyDeviceToHost) == cudaSuccess) rc = 0;
    }
    cudaFree(d);
    return rc;
}

// sao_band_filter / sao_edge_filter (template :310-378) on a w x h block through the 8-sample
// fast paths of the frame kernel.  src carries a 1-sample halo (edge); tile pitch TP samples.
template <typename pixel>
__global__ void dsp_sao_kernel(const pixel *tile, int tp, pixel *dst, int dp, int w, int h, int bd, int edge, int eo, int pos,
                               uint32_t lo, uint32_t hi)
{
    const int maxv = (1 << bd) - 1, units = (w + 7) >> 3;
    for (int it = threadIdx.x; it < units * h; it += blockDim.x) {
        const int y = it / units, u = it - y * units;
        const pixel *sp8 = tile + (y + 1) * tp + 8 + 8 * u;       // block origin at tile (1, 8): aligned, halo around
        int v[8], o[8];
        ld8(sp8, v);
        if (!edge) sao_band_fast(v, lo, hi, pos, bd - 5, maxv, o);
        else if (eo == 0) sao_edge_fast<pixel, -1>(sp8, sp8, v, lo, hi, maxv, o);
        else if (eo == 1) sao_edge_fast<pixel, 0>(sp8 - tp, sp8 + tp, v, lo, hi, maxv, o);
        else if (eo == 2) sao_edge_fast<pixel, -1>(sp8 - tp, sp8 + tp, v, lo, hi, maxv, o);
        else sao_edge_fast<pixel, 1>(sp8 - tp, sp8 + tp, v, lo, hi, maxv, o);
        for (int i = 0; i < 8 && 8 * u + i < w; i++) dst[y * dp + 8 * u + i] = (pixel)o[i];
    }
}

int dsp_sao(int bd, int edge, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst, ptrdiff_t stride_src, const int16_t *offset_val,
            int eo_or_band, int w, int h)
{
    const int P = bd > 8 ? 2 : 1;
    if (w < 1 || w > 64 || h < 1 || h > 64) return -1;
    for (int i = 0; i < 5; i++) if (offset_val[i] < -128 || offset_val[i] > 127) return -1;
    uint32_t lo, hi;
    const int16_t *ov = offset_val;
    if (edge) { lo = (ov[1] & 0xff) | (ov[2] & 0xff) << 8 | (ov[0] & 0xff) << 16 | (uint32_t)(ov[3] & 0xff) << 24; hi = ov[4] & 0xff; }
    else { lo = (ov[1] & 0xff) | (ov[2] & 0xff) << 8 | (ov[3] & 0xff) << 16 | (uint32_t)(ov[4] & 0xff) << 24; hi = 0; }
    const int tp = 96;                                   // samples per tile row: 8 + 64 + slack, rows 16-byte aligned
    const size_t tbytes = (size_t)tp * P * (h + 2), dbytes = (size_t)64 * P * h;
    uint8_t *d = nullptr;
    if (cudaMalloc((void **)&d, tbytes + dbytes + 256) != cudaSuccess) return -1;
    uint8_t *d_dst = d + ((tbytes + 255) & ~(size_t)255);
    cudaMemset(d, 0, tbytes);
    int rc = -1;
    // edge: rows -1..h and columns -1..w of src exist (the caller's halo); band: the block only
    const int halo = edge ? 1 : 0;
    const uint8_t *s0 = src - halo * stride_src - halo * P;
    uint8_t *t0 = d + ((size_t)(1 - halo) * tp + (8 - halo)) * P;
    if (cudaMemcpy2D(t0, (size_t)tp * P, s0, stride_src, (size_t)(w + 2 * halo) * P, h + 2 * halo, cudaMemcpyHostToDevice) == cudaSuccess) {
        if (bd > 8) dsp_sao_kernel<uint16_t><<<1, 256>>>((const uint16_t *)d, tp, (uint16_t *)d_dst, 64, w, h, bd, edge, eo_or_band, eo_or_band, lo, hi);
        else dsp_sao_kernel<uint8_t><<<1, 256>>>(d, tp, d_dst, 64, w, h, bd, edge, eo_or_band, eo_or_band, lo, hi);
        if (cudaDeviceSynchronize() == cudaSuccess &&
            cudaMemcpy2D(dst, stride_dst, d_dst, (size_t)64 * P, (size_t)w * P, h, cudaMemcpyDeviceToHost) == cudaSuccess) rc = 0;
    }
    cudaFree(d);
    return rc;
}

}  // namespace ohcu
