// prepare.cu — the device half of the in-place (zero-copy) host path, and the PCM sample placement.
//
// With ohcu_frame_arenas_get / ohcu_frame_commit_lists (include/ohcu.h) the decoder's parser threads write
// PU / TU records, coefficient pairs and the per-CTB / per-edge tables straight into the job's pinned
// staging; nothing is validated or copied on the host.  prepare_kernel does, on the job's stream and
// ahead of the frame's kernels, what ohcu_append_pu / ohcu_append_tu_* / ohcu_set_sao /
// ohcu_set_deblock_derived / ohcu_frame_stage do on the submitting thread in the copying path:
//   * every record of every CTB checked (bounds, 4-sample alignment, CTB membership, block sizes, modes,
//     reference slots, coefficient areas, quantisers, cross-component references; intra blocks: size, mode,
//     CTB membership, available neighbours inside the plane),
//   * per-CTB ranges for mc_kernel / residual_kernel / bs_kernel, with the "PUs cover the CTB" flag,
//   * the SAO restore flags of sao_filter_CTB (hevc_filter.c:277-325) and the CTB-edge flags of
//     lc->boundary_flags (hevcdec.c:2822-2837) for the bs derivation.
// A malformed list sets the job's error word (first error wins); every kernel of the frame returns at
// once when it is set, and the host reads it back with the job's results.
//
// pcm_kernel places PCM samples (hls_pcm_sample -> put_pcm, hevcdec.c:1686-1721, template :30-43).
#include "ohcu_internal.h"

namespace ohcu {

constexpr int PREP_THREADS = 128;

__device__ __forceinline__ void prep_fail(int *err, int code, uint32_t rec)
{
    atomicCAS(err, 0, code | (int)(min(rec, 0xffffffu) << 8));
}

// per-CTB flags from the slice / tile tables (one thread per CTB)
__device__ __forceinline__ void prep_ctb_flags(const FrameDev &f, const PrepDev &p, int rs, int ctbx, int ctby)
{
    const int cw = f.ctb_w, chh = f.ctb_h;
    auto slice_at = [&](int o) { return p.tsa ? p.tsa[o] : 0; };
    auto tile_at = [&](int o) { return p.tile ? p.tile[o] : 0; };
    const bool lfase = p.fse ? p.fse[rs] != 0 : true;
    const bool no_tile = p.no_tile_filter != 0;
    if (p.has_sao) {                                       // sao_filter_CTB, hevc_filter.c:277-325
        unsigned fl = 0;
        if (no_tile || !lfase) {
            fl |= SAOF_RESTORE;
            auto dslice = [&](int dx, int dy) { return !lfase && slice_at(rs) != slice_at(rs + dy * cw + dx); };
            auto dtile = [&](int dx, int dy) { return no_tile && tile_at(rs) != tile_at(rs + dy * cw + dx); };
            const bool e0 = ctbx == 0, e1 = ctby == 0, e2 = ctbx == cw - 1, e3 = ctby == chh - 1;
            bool lt = false, rt = false, ut = false, bt = false;
            if (!e0) { lt = dtile(-1, 0); if (dslice(-1, 0) || lt) fl |= SAOF_VERT0; }
            if (!e2) { rt = dtile(1, 0);  if (dslice(1, 0) || rt)  fl |= SAOF_VERT1; }
            if (!e1) { ut = dtile(0, -1); if (dslice(0, -1) || ut) fl |= SAOF_HORIZ0; }
            if (!e3) { bt = dtile(0, 1);  if (dslice(0, 1) || bt)  fl |= SAOF_HORIZ1; }
            if (!e0 && !e1 && (dslice(-1, -1) || lt || ut)) fl |= SAOF_DIAG0;
            if (!e1 && !e2 && (dslice(1, -1) || rt || ut))  fl |= SAOF_DIAG1;
            if (!e2 && !e3 && (dslice(1, 1) || rt || bt))   fl |= SAOF_DIAG2;
            if (!e0 && !e3 && (dslice(-1, 1) || lt || bt))  fl |= SAOF_DIAG3;
        }
        p.sao_flags[rs] = (uint16_t)fl;
    }
    if (p.derive) {                                        // lc->boundary_flags, hevcdec.c:2822-2837
        auto differs = [&](int other) {
            return (p.tsa && p.tsa[rs] != p.tsa[other] && !lfase) || (no_tile && p.tile && p.tile[rs] != p.tile[other]);
        };
        unsigned fl = 0;
        if (ctbx > 0 && differs(rs - 1)) fl |= BSF_NO_LEFT;
        if (ctby > 0 && differs(rs - cw)) fl |= BSF_NO_UP;
        if (p.dbk_disabled && p.dbk_disabled[rs]) fl |= BSF_DISABLED;
        p.bs_flags[rs] = (uint8_t)fl;
    }
}

__global__ void __launch_bounds__(PREP_THREADS) prepare_kernel(const __grid_constant__ FrameDev f, const __grid_constant__ PrepDev p)
{
    __shared__ int s_area, s_pairs;
    const int tid = threadIdx.x, ctb = blockIdx.x;
    const int ctbx = ctb % f.ctb_w, ctby = ctb / f.ctb_w, l2 = f.log2_ctb, cs = 1 << l2;
    const ohcu_ctb_index ix = p.idx[ctb];
    if (tid == 0) { s_area = 0; s_pairs = 0; }
    __syncthreads();
    const bool idx_ok = ix.n_pu <= 128u && (uint64_t)ix.first_pu + ix.n_pu <= p.n_pu && ix.n_tu <= 768u &&
                        (uint64_t)ix.first_tu + ix.n_tu <= p.n_tu &&
                        (p.coef_form < 2 || (uint64_t)ix.first_coef + ix.n_coef <= p.n_coef);
    if (!idx_ok) {
        if (tid == 0) {
            prep_fail(p.err, PE_INDEX, (uint32_t)ctb);
            p.pu_range[ctb] = make_int2(0, 0); p.tu_range[ctb] = make_int2(0, 0); p.coef_range[ctb] = make_int2(0, 0);
        }
        return;
    }
    // ---- PUs: what ohcu_append_pu checks, plus the reference slots
    int area = 0;
    for (uint32_t i = tid; i < ix.n_pu; i += PREP_THREADS) {
        const uint32_t r = ix.first_pu + i;
        const ohcu_pu pu = f.pus[r];
        const int x1 = pu.x0 + pu.w - 1, y1 = pu.y0 + pu.h - 1;
        bool ok = pu.pred_flag >= 1 && pu.pred_flag <= 3 && pu.w >= 4 && pu.h >= 4 && pu.w <= 64 && pu.h <= 64 &&
                  !((pu.x0 | pu.y0 | pu.w | pu.h) & 3) && x1 < f.width && y1 < f.height &&
                  (pu.x0 >> l2) == ctbx && (x1 >> l2) == ctbx && (pu.y0 >> l2) == ctby && (y1 >> l2) == ctby &&
                  (int)pu.slice < p.n_slices && pu.ref_idx[0] < OHCU_MAX_REFS && pu.ref_idx[1] < OHCU_MAX_REFS;
        if (ok) {
            const ohcu_slice_tab &sl = f.slices[pu.slice];
#pragma unroll
            for (int l = 0; l < 2; l++)
                if (pu.pred_flag & (1 << l)) {
                    const int slot = sl.ref_slot[l][pu.ref_idx[l]];
                    ok = ok && slot < p.dpb_slots && slot != p.out_slot;      // a picture never predicts from itself
                }
        }
        if (!ok) prep_fail(p.err, PE_PU, r);
        area += pu.w * pu.h;
    }
    if (area) atomicAdd(&s_area, area);
    // ---- coded TBs: what add_tus / append_pairs check
    int pairs = 0;
    const int max_c = f.nplanes == 3 ? 2 : 0;
    for (uint32_t i = tid; i < ix.n_tu; i += PREP_THREADS) {
        const uint32_t r = ix.first_tu + i;
        const ohcu_tu tu = f.tus[r];
        const unsigned l = (unsigned)tu.log2_size - 2u;
        bool ok = l <= 3u && tu.c_idx <= max_c && tu.mode <= OHCU_TU_BYPASS && !(tu.mode == OHCU_TU_DST4 && l != 0) &&
                  (!(tu.flags & OHCU_TUF_INTRA) || f.resid[0]);
        if (ok) {
            const int n = 1 << tu.log2_size, c = tu.c_idx;
            const int hs = c ? f.hs : 0, vs = c ? f.vs : 0;
            const int x1 = tu.x0 + n - 1, y1 = tu.y0 + n - 1;
            ok = !((tu.x0 | tu.y0) & 3) && x1 < f.pw[c] && y1 < f.ph[c] &&
                 ((tu.x0 << hs) >> l2) == ctbx && ((x1 << hs) >> l2) == ctbx &&
                 ((tu.y0 << vs) >> l2) == ctby && ((y1 << vs) >> l2) == ctby;
            if (p.coef_form >= 2)
                ok = ok && !(tu.coeff_off & 7) && tu.coeff_off + (uint32_t)(n * n) <= (uint32_t)p.area && tu.pad <= n * n;
            else
                ok = ok && !(tu.coeff_off & 1) && (uint64_t)tu.coeff_off + (uint64_t)(n * n) <= p.n_coef;
            if (ok && p.coef_form == 3) {
                const ohcu_tu_quant q = f.tu_quant[r];
                ok = q.qp < 76 && (q.matrix_id <= 5 || q.matrix_id == OHCU_QM_FLAT) && (q.matrix_id == OHCU_QM_FLAT || p.has_sl);
            }
            if (ok && (tu.flags & OHCU_TUF_CROSS))
                ok = tu.cross_tu < p.n_tu && f.tus[tu.cross_tu].log2_size == tu.log2_size && f.tus[tu.cross_tu].c_idx == 0;
        }
        if (!ok) prep_fail(p.err, PE_TU, r);
        pairs += tu.pad;
    }
    if (pairs) atomicAdd(&s_pairs, pairs);
    // ---- intra blocks: what ohcu_append_iu checks
    if (f.n_iu) {
        const ohcu_iu_index ii = f.iu_index[ctb];
        if (ii.n_iu > 768u || (uint64_t)ii.first_iu + ii.n_iu > (uint64_t)f.n_iu) {
            if (tid == 0) prep_fail(p.err, PE_IU, 0xffffffu);
        } else {
            for (uint32_t i = tid; i < ii.n_iu; i += PREP_THREADS) {
                const uint32_t r = ii.first_iu + i;
                const ohcu_iu u = f.ius[r];
                bool ok = iu_ok(u, f.nplanes, f.pw, f.ph, f.hs, f.vs, l2);
                if (ok) {
                    const int hs = u.c_idx ? f.hs : 0, vs = u.c_idx ? f.vs : 0;
                    ok = ((u.x0 << hs) >> l2) == ctbx && ((u.y0 << vs) >> l2) == ctby;
                }
                if (!ok) prep_fail(p.err, PE_IU, r);
            }
        }
    }
    __syncthreads();
    if (tid != 0) return;
    const int valid = min(cs, f.width - (ctbx << l2)) * min(cs, f.height - (ctby << l2));
    if (s_area > valid) prep_fail(p.err, PE_OVERLAP, (uint32_t)ctb);
    const bool covered = s_area == valid;
    if (p.all_inter && !covered) prep_fail(p.err, PE_NOT_COVERED, (uint32_t)ctb);
    if (p.coef_form >= 2 && (uint32_t)s_pairs != ix.n_coef) prep_fail(p.err, PE_PAIRS, (uint32_t)ctb);
    p.pu_range[ctb] = make_int2((int)ix.first_pu, (int)ix.n_pu | (covered ? CTB_COVERED : 0));
    p.tu_range[ctb] = make_int2((int)ix.first_tu, (int)ix.n_tu);
    p.coef_range[ctb] = make_int2((int)ix.first_coef, (int)ix.n_coef);
    prep_ctb_flags(f, p, ctb, ctbx, ctby);
}

int launch_prepare(const FrameDev &f, const PrepDev &p, cudaStream_t st)
{
    prepare_kernel<<<f.ctb_w * f.ctb_h, PREP_THREADS, 0, st>>>(f, p);
    return 1;
}

const char *prepare_error_text(int code)
{
    switch (code & 0xff) {
    case PE_INDEX: return "CTB index entry outside the lists (ohcu_ctb_index of CTB";
    case PE_PU: return "malformed PU record (size, alignment, outside its CTB or the picture, slice or reference slot; record";
    case PE_TU: return "malformed TU record (size, mode, outside its CTB or plane, coefficient area, quantiser or cross reference; record";
    case PE_OVERLAP: return "PUs overlap or exceed the CTB (CTB";
    case PE_NOT_COVERED: return "all_inter was promised but the PUs do not cover the CTB (CTB";
    case PE_PAIRS: return "the TU records' pair counts do not add up to the CTB's n_coef (CTB";
    case PE_IU: return "malformed intra block record (size, mode, outside its CTB or plane, a neighbour named available outside the plane; record";
    default: return "unknown error (";
    }
}

// ---------------------------------------------------------------- PCM samples
// One CTA per record: the block's samples, row-major in the job's sample arena, into plane c_idx of the
// picture being reconstructed.  Runs after residual_kernel and before loopfilter_kernel.
template <typename pixel>
__global__ void __launch_bounds__(128) pcm_kernel(const __grid_constant__ FrameDev f, const PcmRec *recs, const uint8_t *samples)
{
    if (f.err && *f.err) return;
    const PcmRec r = recs[blockIdx.x];
    const int n = 1 << r.log2_size;
    const pixel *src = (const pixel *)(samples + r.off);
    pixel *dst = (pixel *)(f.work.p[r.c_idx] + (size_t)r.y0 * f.work.pitch[r.c_idx]) + r.x0;
    const int pitch = f.work.pitch[r.c_idx] / (int)sizeof(pixel);
    for (int i = threadIdx.x; i < n * n; i += blockDim.x) {
        const int y = i >> r.log2_size, x = i & (n - 1);
        dst[y * pitch + x] = src[i];
    }
}

int launch_pcm(const FrameDev &f, const PcmRec *recs, int n, const uint8_t *samples, cudaStream_t st)
{
    if (n <= 0) return 0;
    if (f.bd > 8) pcm_kernel<uint16_t><<<n, 128, 0, st>>>(f, recs, samples);
    else          pcm_kernel<uint8_t><<<n, 128, 0, st>>>(f, recs, samples);
    return 1;
}

}  // namespace ohcu
