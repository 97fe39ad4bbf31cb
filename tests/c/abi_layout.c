/* abi_layout.c — prints sizeof and every field offset of the structs of include/ohcu.h as JSON, compiled by
 * tests/test_abi.py with the plain C compiler and compared with the ctypes mirror (hevc_b200/abi.py): layout drift
 * between the header and its Python description fails the CPU test suite.  Also proves the header is valid C99
 * on its own (no C++ / CUDA types). */
#include <stddef.h>
#include <stdio.h>

#include "ohcu.h"

#define S(T) printf("%s\"%s\": {\"sizeof\": %zu, \"fields\": {", first_s++ ? ",\n" : "", #T, sizeof(T)); first_f = 0;
#define F(T, f) printf("%s\"%s\": %zu", first_f++ ? ", " : "", #f, offsetof(T, f));
#define E() printf("}}");

int main(void)
{
    int first_s = 0, first_f = 0;
    printf("{");
    S(ohcu_seq_params) F(ohcu_seq_params, width) F(ohcu_seq_params, height) F(ohcu_seq_params, bit_depth)
        F(ohcu_seq_params, chroma_format_idc) F(ohcu_seq_params, log2_ctb_size) F(ohcu_seq_params, log2_min_cb_size)
        F(ohcu_seq_params, log2_min_pu_size) F(ohcu_seq_params, dpb_slots) F(ohcu_seq_params, jobs)
        F(ohcu_seq_params, sao_enabled) F(ohcu_seq_params, pcm_or_bypass) F(ohcu_seq_params, cb_qp_offset)
        F(ohcu_seq_params, cr_qp_offset) F(ohcu_seq_params, no_tile_filter) F(ohcu_seq_params, max_pus)
        F(ohcu_seq_params, max_tus) F(ohcu_seq_params, max_coeffs) F(ohcu_seq_params, intra) F(ohcu_seq_params, max_ius)
        F(ohcu_seq_params, intra_smoothing_disabled) F(ohcu_seq_params, strong_intra_smoothing) E()
    S(ohcu_slice_tab) F(ohcu_slice_tab, ref_slot) F(ohcu_slice_tab, weight_flag) F(ohcu_slice_tab, luma_log2_weight_denom)
        F(ohcu_slice_tab, chroma_log2_weight_denom) F(ohcu_slice_tab, pad) F(ohcu_slice_tab, luma_weight)
        F(ohcu_slice_tab, luma_offset) F(ohcu_slice_tab, chroma_weight) F(ohcu_slice_tab, chroma_offset) E()
    S(ohcu_pu) F(ohcu_pu, x0) F(ohcu_pu, y0) F(ohcu_pu, w) F(ohcu_pu, h) F(ohcu_pu, pred_flag) F(ohcu_pu, slice)
        F(ohcu_pu, mv) F(ohcu_pu, ref_idx) F(ohcu_pu, pad) E()
    S(ohcu_tu) F(ohcu_tu, x0) F(ohcu_tu, y0) F(ohcu_tu, c_idx) F(ohcu_tu, log2_size) F(ohcu_tu, mode) F(ohcu_tu, flags)
        F(ohcu_tu, col_limit) F(ohcu_tu, res_scale_val) F(ohcu_tu, pad) F(ohcu_tu, coeff_off) F(ohcu_tu, cross_tu) E()
    S(ohcu_iu) F(ohcu_iu, x0) F(ohcu_iu, y0) F(ohcu_iu, c_idx) F(ohcu_iu, log2_size) F(ohcu_iu, mode) F(ohcu_iu, cand) E()
    S(ohcu_iu_index) F(ohcu_iu_index, first_iu) F(ohcu_iu_index, n_iu) E()
    S(ohcu_dbparams) F(ohcu_dbparams, beta_offset) F(ohcu_dbparams, tc_offset) E()
    S(ohcu_sao) F(ohcu_sao, type_idx) F(ohcu_sao, eo_class) F(ohcu_sao, band_position) F(ohcu_sao, pad)
        F(ohcu_sao, offset_val) E()
    S(ohcu_deblock_in) F(ohcu_deblock_in, vertical_bs) F(ohcu_deblock_in, horizontal_bs) F(ohcu_deblock_in, qp_y_tab)
        F(ohcu_deblock_in, deblock) F(ohcu_deblock_in, is_pcm) E()
    S(ohcu_deblock_derive_in) F(ohcu_deblock_derive_in, qp_y_tab) F(ohcu_deblock_derive_in, deblock)
        F(ohcu_deblock_derive_in, is_pcm) F(ohcu_deblock_derive_in, filter_slice_edges)
        F(ohcu_deblock_derive_in, tab_slice_address) F(ohcu_deblock_derive_in, tile_id)
        F(ohcu_deblock_derive_in, deblock_disabled) E()
    S(ohcu_sao_in) F(ohcu_sao_in, sao) F(ohcu_sao_in, filter_slice_edges) F(ohcu_sao_in, tab_slice_address)
        F(ohcu_sao_in, tile_id) E()
    S(ohcu_crop) F(ohcu_crop, left) F(ohcu_crop, top) F(ohcu_crop, width) F(ohcu_crop, height) E()
    S(ohcu_coef) F(ohcu_coef, pos) F(ohcu_coef, level) E()
    S(ohcu_tu_quant) F(ohcu_tu_quant, qp) F(ohcu_tu_quant, matrix_id) E()
    S(ohcu_ctb_index) F(ohcu_ctb_index, first_pu) F(ohcu_ctb_index, n_pu) F(ohcu_ctb_index, first_tu)
        F(ohcu_ctb_index, n_tu) F(ohcu_ctb_index, first_coef) F(ohcu_ctb_index, n_coef) E()
    S(ohcu_frame_arenas) F(ohcu_frame_arenas, pus) F(ohcu_frame_arenas, max_pus) F(ohcu_frame_arenas, max_tus)
        F(ohcu_frame_arenas, tus) F(ohcu_frame_arenas, quant) F(ohcu_frame_arenas, coefs) F(ohcu_frame_arenas, coeffs)
        F(ohcu_frame_arenas, max_coefs) F(ohcu_frame_arenas, ctbs) F(ohcu_frame_arenas, slices)
        F(ohcu_frame_arenas, vertical_bs) F(ohcu_frame_arenas, horizontal_bs) F(ohcu_frame_arenas, qp_y_tab)
        F(ohcu_frame_arenas, deblock) F(ohcu_frame_arenas, is_pcm) F(ohcu_frame_arenas, sao)
        F(ohcu_frame_arenas, filter_slice_edges) F(ohcu_frame_arenas, tab_slice_address) F(ohcu_frame_arenas, tile_id)
        F(ohcu_frame_arenas, deblock_disabled) F(ohcu_frame_arenas, scaling_list)
        F(ohcu_frame_arenas, ius) F(ohcu_frame_arenas, iu_ctbs) F(ohcu_frame_arenas, max_ius) F(ohcu_frame_arenas, reserved) E()
    S(ohcu_frame_counts) F(ohcu_frame_counts, n_pu) F(ohcu_frame_counts, n_tu) F(ohcu_frame_counts, n_coef)
        F(ohcu_frame_counts, n_slices) F(ohcu_frame_counts, coef_form) F(ohcu_frame_counts, deblock)
        F(ohcu_frame_counts, sao) F(ohcu_frame_counts, pcm) F(ohcu_frame_counts, scaling_list)
        F(ohcu_frame_counts, all_inter) F(ohcu_frame_counts, n_iu) E()
    S(ohcu_pic_params) F(ohcu_pic_params, pcm_or_bypass) F(ohcu_pic_params, cb_qp_offset)
        F(ohcu_pic_params, cr_qp_offset) F(ohcu_pic_params, no_tile_filter) E()
    printf("}\n");
    return 0;
}
