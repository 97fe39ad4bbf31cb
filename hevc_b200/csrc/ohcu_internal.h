// ohcu_internal.h — device-side parameter blocks and kernel launchers shared by the
// sm_100a kernels and the C-ABI host layer (api.cu).  Not part of the public ABI.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ohcu.h"

namespace ohcu {

struct PlaneSet {
    uint8_t *p[3];         // sample (0, 0) of each plane; DPB planes carry a replicated halo around it
    int pitch[3];          // bytes
};

// ---- halo-padded DPB planes ------------------------------------------------------------
// Every DPB plane is allocated with PAD_X samples left/right and PAD_Y rows above/below
// that replicate the border sample (the reference's ff_emulated_edge_mc, done once per
// picture instead of once per block).  Motion compensation reads its windows straight from
// these planes; a window is clamped so that it starts at most one window size outside the
// picture, and the 8-byte loads of the filter may begin up to 3 samples left of it.
constexpr int PAD_X = 96;                  // samples; >= 71 needed + alignment slack of the loads
constexpr int PAD_Y = 80;                  // rows;    >= 71 needed

// Everything a launch needs, passed by value as one kernel parameter.
struct FrameDev {
    // sequence
    int width, height, bd, cfi, hs, vs;
    int log2_ctb, ctb_w, ctb_h;
    int log2_min_cb, min_cb_w;
    int log2_min_pu, min_pu_w, min_pu_h;
    int bs_width;
    int pcm_or_bypass, cb_qp_offset, cr_qp_offset;
    int nplanes;
    int pw[3], ph[3];      // plane dimensions in samples
    // pictures
    PlaneSet work;         // picture being reconstructed (pre-SAO)
    PlaneSet out;          // SAO destination (the DPB slot)
    PlaneSet dpb[OHCU_MAX_DPB];
    // MC
    const ohcu_pu *pus;
    int n_pu;
    const int2 *ctb_pu_range;     // per CTB (raster): first PU, number of PUs | CTB_COVERED
    const ohcu_slice_tab *slices;
    // residual
    const ohcu_tu *tus;
    int n_tu;
    const int16_t *coeffs;
    const int2 *ctb_tu_range;     // per CTB (raster): first TU, number of TUs
    const int2 *ctb_coef_range;   // sparse coefficients: per CTB first (pos, level) pair in `coeffs`, number of pairs; nullptr = dense blocks
    const ohcu_tu_quant *tu_quant; // levels form: dequantisation parameters per TU (the pairs carry un-scaled levels); else nullptr
    const uint8_t *scaling;        // ScalingList: sl[4][6][64] then sl_dc[2][6]; nullptr = none set
    // deblock
    const uint8_t *vbs, *hbs;
    const int8_t *qp_y;
    const ohcu_dbparams *db;
    const uint8_t *is_pcm;
    const uint8_t *bs_ctb_flags;     // bs derived on the device: per CTB, BSF_* below; nullptr = the planes came from the host
    // SAO
    const ohcu_sao *sao;
    const uint16_t *sao_ctb_flags;   // per CTB, see SAOF_* below
    // intra blocks (intra.cu): records in decoding order, where each CTB's sit, the residual planes residual_kernel
    // leaves the OHCU_TUF_INTRA blocks in (int16, pitch in bytes), a ticket followed by one done flag per CTB
    const ohcu_iu *ius;
    const ohcu_iu_index *iu_index;
    int n_iu;
    int16_t *resid[3];
    int resid_pitch[3];
    int *intra_progress;
    const int *intra_order;          // CTBs in wavefront order (x + 2 y ascending): the order the tickets hand them out in
    int intra_smoothing_disabled, strong_intra_smoothing;
    // the job's error word (prepare.cu): non-zero = the lists are malformed, every kernel of the frame returns at once
    const int *err;
};

// ---- in-place lists: what prepare_kernel reads and writes besides FrameDev (prepare.cu)
struct PrepDev {
    const ohcu_ctb_index *idx;       // per CTB: where its records sit (written by the parser)
    int2 *pu_range, *tu_range, *coef_range;      // out: FrameDev::ctb_pu_range / ctb_tu_range / ctb_coef_range
    uint16_t *sao_flags;             // out: FrameDev::sao_ctb_flags
    uint8_t *bs_flags;               // out: FrameDev::bs_ctb_flags
    const uint8_t *fse;              // slice_loop_filter_across_slices_enabled_flag per CTB (nullptr = 1)
    const int32_t *tsa, *tile;       // tab_slice_address / tile id per CTB (nullptr = one slice / tile)
    const uint8_t *dbk_disabled;     // slice_deblocking_filter_disabled_flag per CTB (nullptr = 0)
    int *err;
    uint32_t n_pu, n_tu;
    uint64_t n_coef;
    int n_slices, coef_form, dpb_slots, out_slot, all_inter, no_tile_filter, has_sao, derive, area, has_sl;
};
enum { PE_INDEX = 1, PE_PU, PE_TU, PE_OVERLAP, PE_NOT_COVERED, PE_PAIRS, PE_IU };
int launch_prepare(const FrameDev &f, const PrepDev &p, cudaStream_t st);
const char *prepare_error_text(int code);

// PCM samples of one block (ohcu_append_pcm): plane coordinates, byte offset of its samples in the job's arena
struct PcmRec { uint16_t x0, y0; uint8_t log2_size, c_idx; uint16_t pad; uint32_t off; };
int launch_pcm(const FrameDev &f, const PcmRec *recs, int n, const uint8_t *samples, cudaStream_t st);

// set in ctb_pu_range[].y when the CTB's PUs cover every sample of it
constexpr int CTB_COVERED = 1 << 30;

// per-CTB SAO flags precomputed on the host from slice addresses / tile ids
// (sao_filter_CTB, hevc_filter.c:277-325)
enum {
    SAOF_RESTORE = 1 << 0,    // restore = no_tile_filter || !lfase
    SAOF_VERT0 = 1 << 1, SAOF_VERT1 = 1 << 2,
    SAOF_HORIZ0 = 1 << 3, SAOF_HORIZ1 = 1 << 4,
    SAOF_DIAG0 = 1 << 5, SAOF_DIAG1 = 1 << 6, SAOF_DIAG2 = 1 << 7, SAOF_DIAG3 = 1 << 8,
};

// per-CTB flags of the bs derivation (ohcu_set_deblock_derived), from the slice / tile tables
enum { BSF_NO_LEFT = 1, BSF_NO_UP = 2, BSF_DISABLED = 4 };

// kernel launchers (each returns the number of kernels it enqueued, < 0 on error)
int launch_recon(const FrameDev &f, int flags, cudaStream_t st);   // flags: 1 MC, 2 residual
int recon_sparse_ctb_coefs(int log2_ctb, int nplanes, int hs, int vs);   // int16 a CTB's rebuilt coefficient area may hold (sparse form)
int recon_set_attrs();                     // per-device function attributes (dynamic shared memory opt-in)
int loopfilter_set_attrs();
int intra_set_attrs();
int upload_mc_taps();
int upload_dct_table();
struct LfMaps { CUtensorMap m[3]; };        // the job's reconstructed picture, one box per plane
int launch_loopfilter(const FrameDev &f, const LfMaps &maps, int flags, cudaStream_t st);   // flags: 1 deblock, 2 SAO
int launch_bs_derive(const FrameDev &f, cudaStream_t st);      // vbs / hbs from the PU and TU lists (bs.cu)
int launch_intra(const FrameDev &f, cudaStream_t st);          // the intra blocks, CTB wavefront (intra.cu)
// an intra block record is well formed for this sequence (shared by ohcu_append_iu and prepare_kernel): inside its
// plane and one CTB, and no neighbour named available that the decoding order cannot have produced yet - outside the
// plane, below the CTB, or right of the CTB on any row but the one above it (intra_kernel stages exactly that much)
__host__ __device__ inline bool iu_ok(const ohcu_iu &u, int nplanes, const int *pw, const int *ph, int hs, int vs, int log2_ctb)
{
    if (u.c_idx >= nplanes || u.log2_size < 2 || u.log2_size > 5 || u.mode > 34 || (u.cand & 0xc0)) return false;
    const int n = 1 << u.log2_size;
    if (((u.x0 | u.y0) & 3) || u.x0 + n > pw[u.c_idx] || u.y0 + n > ph[u.c_idx]) return false;
    const int sw = (1 << log2_ctb) >> (u.c_idx ? hs : 0), sh = (1 << log2_ctb) >> (u.c_idx ? vs : 0);
    const int bx = u.x0 & (sw - 1), by = u.y0 & (sh - 1);
    if (bx + n > sw || by + n > sh) return false;
    if ((u.cand & (OHCU_CAND_LEFT | OHCU_CAND_BOTTOM_LEFT | OHCU_CAND_UP_LEFT)) && u.x0 == 0) return false;
    if ((u.cand & (OHCU_CAND_UP | OHCU_CAND_UP_RIGHT | OHCU_CAND_UP_LEFT)) && u.y0 == 0) return false;
    if ((u.cand & OHCU_CAND_BOTTOM_LEFT) && (u.y0 + n >= ph[u.c_idx] || by + n >= sh)) return false;
    if ((u.cand & OHCU_CAND_UP_RIGHT) && (u.x0 + n >= pw[u.c_idx] || (bx + n >= sw && by > 0))) return false;
    return true;
}
// per-entry test shims (recon.cu, loopfilter.cu)
int dsp_mc(int bd, int luma, int family, uint8_t *dst, ptrdiff_t dststride, const uint8_t *src, ptrdiff_t srcstride,
           const int16_t *src2, int16_t *out16, int h, int denom, int wx0, int wx1, int ox0, int ox1, int mx, int my, int w);
int dsp_residual(int bd, int mode, int flags, int16_t *coeffs, int log2, int col_limit, int use_mma);
int dsp_transform_add(int bd, uint8_t *dst, const int16_t *coeffs, ptrdiff_t stride, int n);
int dsp_loop_filter(int bd, int luma, int vertical, uint8_t *pix, ptrdiff_t stride, int beta, const int *tc, const uint8_t *no_p,
                    const uint8_t *no_q);
int dsp_sao(int bd, int edge, uint8_t *dst, const uint8_t *src, ptrdiff_t stride_dst, ptrdiff_t stride_src, const int16_t *offset_val,
            int eo_or_band, int w, int h);
int dsp_intra_pred(int bd, uint8_t *plane, ptrdiff_t stride, int pw, int ph, int x, int y, int log2_size, int c_idx, int mode, int cand,
                   int filter_ok, int strong);                    // intra.cu
void loopfilter_box(const FrameDev &f, int c, int *bw, int *bh);
int launch_pad_picture(const FrameDev &f, const PlaneSet &pic, cudaStream_t st);   // fill the replicated halo
int launch_copy_picture(const FrameDev &f, const PlaneSet &dst, const PlaneSet &src, cudaStream_t st);
// the window of every plane packed without row gaps (ohcu_slot_output_cropped)
int launch_crop_pack(const PlaneSet &pic, int nplanes, int P, const int left[3], const int top[3], const int w[3],
                     const int h[3], uint8_t *const dst[3], cudaStream_t st);
int launch_md5(const PlaneSet &pic, int nplanes, int P, const int pw[3], const int ph[3], uint8_t *out48, cudaStream_t st);
int launch_l2_flush(void *buf, size_t bytes, cudaStream_t st);
int set_error(int code, const char *fmt, ...);     // text for ohcu_last_error() of the calling thread; returns code

// ---- small device helpers -------------------------------------------------------
template <int BD> struct Px { typedef uint16_t T; };
template <> struct Px<8> { typedef uint8_t T; };

__device__ __forceinline__ int clip3i(int v, int lo, int hi) { return min(max(v, lo), hi); }
template <int BD> __device__ __forceinline__ int clip_px(int v) { return min(max(v, 0), (1 << BD) - 1); }
__device__ __forceinline__ int clip_s16(int v) { return min(max(v, -32768), 32767); }

}  // namespace ohcu

// dispatch a template<int BD> launcher on the run-time bit depth
#define OHCU_DISPATCH_BD(bd, CALL)                 \
    switch (bd) {                                  \
    case 8:  { constexpr int BD = 8;  CALL; } break; \
    case 9:  { constexpr int BD = 9;  CALL; } break; \
    case 10: { constexpr int BD = 10; CALL; } break; \
    case 12: { constexpr int BD = 12; CALL; } break; \
    default: return -1;                            \
    }
