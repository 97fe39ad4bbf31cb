// bs.cu — boundary strengths of one picture, derived on the device from its work lists
// (SURVEY section 8f rank 2).
//
// Replaces, per picture:
//   ff_hevc_deblocking_boundary_strengths   hevc_filter.c:989-1125  (called per transform-tree
//                                           leaf at hevcdec.c:1677, :1707, :2603, :2687)
//   boundary_strength                       hevc_filter.c:764-880   (the TEST_MV_POC variant)
//
// The reference visits every edge unit (4 samples on the 8x8 grid) exactly once, from the leaf on
// its Q side: on the leaf's own left / top boundary the unit gets 2 if either side is intra, else
// 1 if the transform block on either side has cbf_luma, else the motion rule; inside the leaf
// only the motion rule.  "A transform-block boundary with cbf_luma on a side" is the same as
// "on the perimeter of a coded luma TB", so the coded TBs of the TU list are enough and the
// transform tree itself need not cross the bus.  4x4 blocks no PU covers count as intra (2), except inside one
// intra leaf - the luma ohcu_iu blocks are the leaves of the intra CUs - which is no boundary.
//
// One CTA per CTB.  The PUs and coded luma TBs of the CTB, of its left and of its upper neighbour
// enter a (cells + 1)^2 map of 4x4 cells in shared memory (PU index, TB index); then one thread
// per edge unit (16 x 8 vertical and 8 x 16 horizontal units of a 64x64 CTB).
// The planes are the ones loopfilter_kernel reads (index (x + y * bs_width) >> 2, hevcdec.c:127).
#include "ohcu_internal.h"

namespace ohcu {

#ifndef BS_NTHREADS
#define BS_NTHREADS 128
#endif
constexpr int BS_THREADS = BS_NTHREADS;      // 128: all CTBs of a 4K picture are resident at once (16 CTAs per SM)
constexpr int BS_CELLS = 17;                 // 16 cells of a 64x64 CTB + the neighbour's column / row

struct BsCell { int pu, tb, iu; };

__device__ __forceinline__ bool bs_far(const int16_t a[2], const int16_t b[2])
{
    return abs((int)a[0] - (int)b[0]) >= 4 || abs((int)a[1] - (int)b[1]) >= 4;
}

// boundary_strength(s, curr, neigh), hevc_filter.c:765-880; memcmp-equal MvFields give 0 through
// the same comparisons.  Pictures are compared by DPB slot (the reference: by POC).
__device__ int bs_motion(const FrameDev &f, const ohcu_pu &c, const ohcu_pu &n)
{
    const ohcu_slice_tab &sc = f.slices[c.slice], &sn = f.slices[n.slice];
    const int pc0 = sc.ref_slot[0][c.ref_idx[0] & (OHCU_MAX_REFS - 1)], pc1 = sc.ref_slot[1][c.ref_idx[1] & (OHCU_MAX_REFS - 1)];
    const int pn0 = sn.ref_slot[0][n.ref_idx[0] & (OHCU_MAX_REFS - 1)], pn1 = sn.ref_slot[1][n.ref_idx[1] & (OHCU_MAX_REFS - 1)];
    if (c.pred_flag == OHCU_PF_BI && n.pred_flag == OHCU_PF_BI) {
        if (pc0 == pn0 && pc0 == pc1 && pn0 == pn1)
            return (bs_far(n.mv[0], c.mv[0]) || bs_far(n.mv[1], c.mv[1])) && (bs_far(n.mv[1], c.mv[0]) || bs_far(n.mv[0], c.mv[1]));
        if (pn0 == pc0 && pn1 == pc1) return bs_far(n.mv[0], c.mv[0]) || bs_far(n.mv[1], c.mv[1]);
        if (pn1 == pc0 && pn0 == pc1) return bs_far(n.mv[1], c.mv[0]) || bs_far(n.mv[0], c.mv[1]);
        return 1;
    }
    if (c.pred_flag != OHCU_PF_BI && n.pred_flag != OHCU_PF_BI) {
        const bool c0 = c.pred_flag & 1, n0 = n.pred_flag & 1;
        const int rc = c0 ? pc0 : pc1, rn = n0 ? pn0 : pn1;
        if (rc != rn) return 1;
        return bs_far(c0 ? c.mv[0] : c.mv[1], n0 ? n.mv[0] : n.mv[1]);
    }
    return 1;
}

__global__ void __launch_bounds__(BS_THREADS) bs_kernel(const __grid_constant__ FrameDev f, uint8_t *vbs, uint8_t *hbs)
{
    __shared__ BsCell cell[BS_CELLS * BS_CELLS];          // [1 + y][1 + x], row / column 0 = the neighbours' last cells
    if (f.err && *f.err) return;                          // malformed lists (prepare.cu)
    const int tid = threadIdx.x;
    const int ctb = blockIdx.x, ctbx = ctb % f.ctb_w, ctby = ctb / f.ctb_w;
    const int cs = 1 << f.log2_ctb, n4 = cs >> 2;
    const int gx0 = (ctbx << f.log2_ctb) >> 2, gy0 = (ctby << f.log2_ctb) >> 2;     // cell coordinates of the CTB's origin
    for (int i = tid; i < BS_CELLS * BS_CELLS; i += BS_THREADS) cell[i] = BsCell{-1, -1, -1};
    __syncthreads();

    // ---- the cell map: this CTB, its left and its upper neighbour (clipped to the window)
    const int nb[3] = {ctb, ctbx > 0 ? ctb - 1 : -1, ctby > 0 ? ctb - f.ctb_w : -1};
    auto paint = [&](int x4, int y4, int w4, int h4, int idx, int what) {      // what: 0 PU, 1 coded luma TB, 2 intra leaf
        // rectangle in picture cells -> window cells [-1, n4)
        const int xa = max(x4 - gx0, -1), xb = min(x4 + w4 - gx0, n4), ya = max(y4 - gy0, -1), yb = min(y4 + h4 - gy0, n4);
        for (int y = ya; y < yb; y++)
            for (int x = xa; x < xb; x++) {
                BsCell &c = cell[(y + 1) * BS_CELLS + x + 1];
                if (what == 1) c.tb = idx; else if (what == 2) c.iu = idx; else c.pu = idx;
            }
    };
    int2 pr[3], tr[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {                          // the six ranges first: independent loads
        pr[k] = nb[k] >= 0 ? f.ctb_pu_range[nb[k]] : make_int2(0, 0);
        tr[k] = nb[k] >= 0 && f.n_tu > 0 ? f.ctb_tu_range[nb[k]] : make_int2(0, 0);
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const int np = min(pr[k].y & 0xffffff, 128);
        for (int i = tid; i < np; i += BS_THREADS) {
            const ohcu_pu &pu = f.pus[pr[k].x + i];
            paint(pu.x0 >> 2, pu.y0 >> 2, pu.w >> 2, pu.h >> 2, pr[k].x + i, 0);
        }
        for (int i = tid; i < tr[k].y; i += BS_THREADS) {
            const ohcu_tu &tu = f.tus[tr[k].x + i];
            if (tu.c_idx == 0) paint(tu.x0 >> 2, tu.y0 >> 2, 1 << (tu.log2_size - 2), 1 << (tu.log2_size - 2), tr[k].x + i, 1);
        }
        if (f.n_iu > 0 && nb[k] >= 0) {                    // the luma intra blocks are the leaves of the intra CUs
            const ohcu_iu_index ii = f.iu_index[nb[k]];
            for (uint32_t i = tid; i < min(ii.n_iu, 768u); i += BS_THREADS) {
                const ohcu_iu &u = f.ius[ii.first_iu + i];
                if (u.c_idx == 0) paint(u.x0 >> 2, u.y0 >> 2, 1 << (u.log2_size - 2), 1 << (u.log2_size - 2), (int)(ii.first_iu + i), 2);
            }
        }
    }
    __syncthreads();

    // ---- one thread per edge unit: 2 * n4 * (n4 / 2) of them
    const int per_dir = n4 * (n4 >> 1);
    const int flags = f.bs_ctb_flags[ctb];
    for (int it = tid; it < 2 * per_dir; it += BS_THREADS) {
        const int dir = it >= per_dir, k = dir ? it - per_dir : it;
        // vertical edges: e = edge column (8-sample steps), u = unit along it (4-sample steps); horizontal: transposed
        const int e = k / n4, u = k - e * n4;
        const int cx = dir ? u : 2 * e, cy = dir ? 2 * e : u;                  // Q cell inside the CTB
        const int x = (gx0 + cx) << 2, y = (gy0 + cy) << 2;
        if (x >= f.width || y >= f.height || (dir ? y : x) == 0) continue;    // picture edges are never filtered
        int bs = 0;
        const bool on_ctb_edge = e == 0;
        if (!(flags & BSF_DISABLED) && !(on_ctb_edge && (flags & (dir ? BSF_NO_UP : BSF_NO_LEFT)))) {
            const BsCell q = cell[(cy + 1) * BS_CELLS + cx + 1];
            const BsCell p = dir ? cell[cy * BS_CELLS + cx + 1] : cell[(cy + 1) * BS_CELLS + cx];
            if (q.pu < 0 || p.pu < 0) bs = (q.iu >= 0 && q.iu == p.iu) ? 0 : 2;      // inside one intra leaf: no edge
            else if (q.tb != p.tb && (q.tb >= 0 || p.tb >= 0)) bs = 1;
            else if (q.pu != p.pu) bs = bs_motion(f, f.pus[q.pu], f.pus[p.pu]);
        }
        (dir ? hbs : vbs)[(x + y * f.bs_width) >> 2] = (uint8_t)bs;
    }
}

int launch_bs_derive(const FrameDev &f, cudaStream_t st)
{
    if (!f.bs_ctb_flags || !f.vbs) return 0;
    bs_kernel<<<f.ctb_w * f.ctb_h, BS_THREADS, 0, st>>>(f, const_cast<uint8_t *>(f.vbs), const_cast<uint8_t *>(f.hbs));
    return 1;
}

}  // namespace ohcu
