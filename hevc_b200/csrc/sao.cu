// sao.cu — picture-level sample adaptive offset, out of place: out = SAO(deblocked).
//
// Replaces sao_filter_CTB (hevc_filter.c:264-502), restore_tqb_pixels (:229-260) and the
// table entries sao_band_filter / sao_edge_filter / sao_edge_restore[0,1]
// (hevcdsp_template.c:310-531).  The reference filters CTB by CTB in place and keeps the
// pre-SAO border samples in line buffers; reading the whole deblocked picture and writing
// a different one gives every sample the same inputs without the buffers.  Samples of
// CTBs whose SAO is off are copied.  Each thread produces 4 consecutive samples.
#include "ohcu_internal.h"

namespace ohcu {

template <int BD>
__global__ void __launch_bounds__(256) sao_kernel(const __grid_constant__ FrameDev f, int c)
{
    typedef typename Px<BD>::T pixel;
    const int pw = f.pw[c], ph = f.ph[c];
    const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
    const int x0 = (blockIdx.x * 32 + (threadIdx.x & 31)) * 4;
    const int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x0 >= pw || y >= ph) return;

    const uint8_t *src = f.work.p[c];
    const int spitch = f.work.pitch[c];
    uint8_t *dst = f.out.p[c];
    const int dpitch = f.out.pitch[c];
    const pixel *srow = (const pixel *)(src + (size_t)y * spitch);
    pixel *drow = (pixel *)(dst + (size_t)y * dpitch);

    // CTB of these samples (CTB plane width is a multiple of 8, so 4 samples share it)
    const int cw = (1 << f.log2_ctb) >> hs, chh = (1 << f.log2_ctb) >> vs;
    const int ctbx = x0 / cw, ctby = y / chh;
    const int rs = ctby * f.ctb_w + ctbx;
    const int type = f.sao ? f.sao[rs].type_idx[c] : OHCU_SAO_NOT_APPLIED;
    const int nv = min(4, pw - x0);

    if (type != OHCU_SAO_BAND && type != OHCU_SAO_EDGE) {
        for (int i = 0; i < nv; i++) drow[x0 + i] = srow[x0 + i];
        return;
    }
    const ohcu_sao sp = f.sao[rs];
    int off[5];
#pragma unroll
    for (int i = 0; i < 5; i++) off[i] = sp.offset_val[c][i];
    const int cx0 = ctbx * cw, cy0 = ctby * chh;          // CTB origin in the plane
    const int w = min(cw, pw - cx0), h = min(chh, ph - cy0);
    const int ly = y - cy0;

    // restore_tqb_pixels (:229-260): the min-PU range is taken from the LUMA origin of the
    // CTB plus the PLANE's width/height, as in the reference
    const bool tqb = f.pcm_or_bypass && f.is_pcm;
    const int lx_ctb = ctbx << f.log2_ctb, ly_ctb = ctby << f.log2_ctb;   // luma origin
    const int pu_xmax = (lx_ctb + w) >> f.log2_min_pu, pu_ymax = (ly_ctb + h) >> f.log2_min_pu;

    if (type == OHCU_SAO_BAND) {                           // sao_band_filter_0, template :310-335
        const int pos = sp.band_position[c];
        for (int i = 0; i < nv; i++) {
            const int x = x0 + i;
            const int v = srow[x];
            const int k = ((v >> (BD - 5)) - pos) & 31;
            int o = clip_px<BD>(v + (k < 4 ? off[k + 1] : 0));
            if (tqb) {
                const int pux = (x << hs) >> f.log2_min_pu, puy = (y << vs) >> f.log2_min_pu;
                if (pux < pu_xmax && puy < pu_ymax && f.is_pcm[puy * f.min_pu_w + pux]) o = v;
            }
            drow[x] = (pixel)o;
        }
        return;
    }

    // ---- edge offset: sao_edge_filter (:342-378) then sao_edge_restore_{0,1} (:380-529)
    const int eo = sp.eo_class[c];
    const int dxa = (eo == 0 || eo == 2) ? -1 : (eo == 3 ? 1 : 0);
    const int dya = eo == 0 ? 0 : -1;
    const int flags = f.sao_ctb_flags[rs];
    const bool b0 = ctbx == 0, b1 = ctby == 0, b2 = ctbx == f.ctb_w - 1, b3 = ctby == f.ctb_h - 1;
    const bool notv = eo != OHCU_SAO_EO_VERT, noth = eo != OHCU_SAO_EO_HORIZ;
    const int wr = w - ((notv && b2) ? 1 : 0);             // width / height after the border
    const int hr = h - ((noth && b3) ? 1 : 0);             // columns / rows were peeled off
    const int init_x = (notv && b0) ? 1 : 0;
    const bool restore = flags & SAOF_RESTORE;
    const bool ve0 = flags & SAOF_VERT0, ve1 = flags & SAOF_VERT1;
    const bool he0 = flags & SAOF_HORIZ0, he1 = flags & SAOF_HORIZ1;
    const bool d0 = flags & SAOF_DIAG0, d1 = flags & SAOF_DIAG1, d2 = flags & SAOF_DIAG2, d3 = flags & SAOF_DIAG3;
    const int keep_ul = !d0 && eo == 2 && !b0 && !b1, keep_ur = !d1 && eo == 3 && !b1 && !b2;
    const int keep_lr = !d2 && eo == 2 && !b2 && !b3, keep_ll = !d3 && eo == 3 && !b0 && !b3;

    const int ya = y + dya, yb = y - dya;
    const bool ya_ok = ya >= 0 && ya < ph, yb_ok = yb >= 0 && yb < ph;
    const pixel *rowa = (const pixel *)(src + (size_t)(ya_ok ? ya : y) * spitch);
    const pixel *rowb = (const pixel *)(src + (size_t)(yb_ok ? yb : y) * spitch);

    for (int i = 0; i < nv; i++) {
        const int x = x0 + i, lx = x - cx0;
        const int v = srow[x];
        const int xa = x + dxa, xb = x - dxa;
        // out-of-picture neighbours only occur where the border rules below override the result
        const int a = (ya_ok && xa >= 0 && xa < pw) ? rowa[xa] : v;
        const int b = (yb_ok && xb >= 0 && xb < pw) ? rowb[xb] : v;
        const int sgn = ((v > a) - (v < a)) + ((v > b) - (v < b));   // -2..2
        // edge_idx[] = {1, 2, 0, 3, 4}
        const int cat = sgn == 0 ? 0 : (sgn == -2 ? 1 : (sgn == -1 ? 2 : (sgn == 1 ? 3 : 4)));
        int o = clip_px<BD>(v + off[cat]);
        const int keep = clip_px<BD>(v + off[0]);
        // picture borders (restore_0 part)
        if (notv && ((b0 && lx == 0) || (b2 && lx == w - 1))) o = keep;
        if (noth && ((b1 && ly == 0) || (b3 && ly == h - 1)) && lx >= init_x && lx < wr) o = keep;
        if (restore) {                                     // slice / tile boundaries (restore_1)
            if (ve0 && notv && lx == 0 && ly >= keep_ul && ly < hr - keep_ll) o = v;
            if (ve1 && notv && lx == wr - 1 && ly >= keep_ur && ly < hr - keep_lr) o = v;
            if (he0 && noth && ly == 0 && lx >= init_x + keep_ul && lx < wr - keep_ur) o = v;
            if (he1 && noth && ly == hr - 1 && lx >= init_x + keep_ll && lx < wr - keep_lr) o = v;
            if (d0 && eo == 2 && lx == 0 && ly == 0) o = v;
            if (d1 && eo == 3 && lx == wr - 1 && ly == 0) o = v;
            if (d2 && eo == 2 && lx == wr - 1 && ly == hr - 1) o = v;
            if (d3 && eo == 3 && lx == 0 && ly == hr - 1) o = v;
        }
        if (tqb) {
            const int pux = (x << hs) >> f.log2_min_pu, puy = (y << vs) >> f.log2_min_pu;
            if (pux < pu_xmax && puy < pu_ymax && f.is_pcm[puy * f.min_pu_w + pux]) o = v;
        }
        drow[x] = (pixel)o;
    }
}

template <int BD>
__global__ void copy_picture_kernel(uint8_t *dst, int dpitch, const uint8_t *src, int spitch, int wbytes, int h)
{
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    const int y = blockIdx.y;
    if (x >= wbytes || y >= h) return;
    if (x + 16 <= wbytes)
        *(uint4 *)(dst + (size_t)y * dpitch + x) = *(const uint4 *)(src + (size_t)y * spitch + x);
    else
        for (int i = x; i < wbytes; i++) dst[(size_t)y * dpitch + i] = src[(size_t)y * spitch + i];
}

// Replicated halo of a DPB picture (PAD_X / PAD_Y, ohcu_internal.h): one CTA per padded row.
// Rows inside the picture only get their left/right margins; rows above/below copy the
// first/last picture row with its margins.  Equivalent of ff_emulated_edge_mc
// (videodsp_template.c:22-100) applied once to the whole picture.
template <typename pixel>
__global__ void __launch_bounds__(256) pad_kernel(uint8_t *p0, uint8_t *p1, uint8_t *p2, int pitch0, int pitch1, int pitch2,
                                                   int pw0, int ph0, int pwc, int phc)
{
    int row = blockIdx.x;
    const int rows0 = ph0 + 2 * PAD_Y, rowsc = phc + 2 * PAD_Y;
    uint8_t *plane; int pitch, pw, ph;
    if (row < rows0) { plane = p0; pitch = pitch0; pw = pw0; ph = ph0; }
    else if (row < rows0 + rowsc) { row -= rows0; plane = p1; pitch = pitch1; pw = pwc; ph = phc; }
    else { row -= rows0 + rowsc; plane = p2; pitch = pitch2; pw = pwc; ph = phc; }
    const int y = row - PAD_Y;
    const int ys = min(max(y, 0), ph - 1);
    const pixel *src = (const pixel *)(plane + (ptrdiff_t)ys * pitch);
    pixel *dst = (pixel *)(plane + (ptrdiff_t)y * pitch);
    if (y >= 0 && y < ph) {
        const pixel l = src[0], r = src[pw - 1];
        for (int t = threadIdx.x; t < 2 * PAD_X; t += blockDim.x) {
            if (t < PAD_X) dst[t - PAD_X] = l;
            else dst[pw + t - PAD_X] = r;
        }
    } else {
        for (int x = -PAD_X + (int)threadIdx.x; x < pw + PAD_X; x += blockDim.x) dst[x] = src[min(max(x, 0), pw - 1)];
    }
}

int launch_pad_picture(const FrameDev &f, const PlaneSet &pic, cudaStream_t st)
{
    const int rows = f.ph[0] + 2 * PAD_Y + (f.nplanes == 3 ? 2 * (f.ph[1] + 2 * PAD_Y) : 0);
    if (f.bd > 8)
        pad_kernel<uint16_t><<<rows, 256, 0, st>>>(pic.p[0], pic.p[1], pic.p[2], pic.pitch[0], pic.pitch[1], pic.pitch[2],
                                                   f.pw[0], f.ph[0], f.pw[1], f.ph[1]);
    else
        pad_kernel<uint8_t><<<rows, 256, 0, st>>>(pic.p[0], pic.p[1], pic.p[2], pic.pitch[0], pic.pitch[1], pic.pitch[2],
                                                  f.pw[0], f.ph[0], f.pw[1], f.ph[1]);
    return 1;
}

int launch_copy_picture(const FrameDev &f, const PlaneSet &dst, const PlaneSet &src, cudaStream_t st)
{
    const int P = f.bd > 8 ? 2 : 1;
    for (int c = 0; c < f.nplanes; c++) {
        const int wbytes = f.pw[c] * P;
        dim3 grid((wbytes / 16 + 128) / 128, f.ph[c]);
        copy_picture_kernel<8><<<grid, 128, 0, st>>>(dst.p[c], dst.pitch[c], src.p[c], src.pitch[c], wbytes, f.ph[c]);
    }
    return f.nplanes;
}

int launch_sao(const FrameDev &f, cudaStream_t st)
{
    int n = 0;
    for (int c = 0; c < f.nplanes; c++) {
        dim3 grid((f.pw[c] + 127) / 128, (f.ph[c] + 7) / 8);
        OHCU_DISPATCH_BD(f.bd, (sao_kernel<BD><<<grid, 256, 0, st>>>(f, c)));
        n++;
    }
    return n;
}

__global__ void l2_flush_kernel(uint4 *buf, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) { uint4 v = buf[i]; v.x += 1; buf[i] = v; }
}

int launch_l2_flush(void *buf, size_t bytes, cudaStream_t st)
{
    l2_flush_kernel<<<148 * 8, 256, 0, st>>>((uint4 *)buf, bytes / 16);
    return 1;
}

}  // namespace ohcu
